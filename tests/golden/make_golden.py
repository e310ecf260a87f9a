"""Regenerates tests/golden/cases.txt + expected.txt by running the UNMODIFIED reference
(oracle/_ref/ref_harness, built from /root/reference by oracle/Makefile) on seeded cases.
Run here (the container that has /root/reference); the outputs are committed."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tests import refcases as rc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
orc, synth = rc.orc, rc.synth


def main():
    rng = np.random.default_rng(20261104)
    path = os.path.join(HERE, "cases.txt")
    with open(path, "w") as f:
        # SeqAn's own golden vectors (tests/align/test_alignment_algorithms_global.h:2084-2151)
        f.write("ALIGNX 2 -1 -1 -3 ATGT ATAGAT\n")
        f.write("ALIGNX 2 -1 -1 -3 AAAAAGGGGTTTT AAAGTT\n")
        f.write("ALIGNX 2 -1 -1 -3 AAAAAATTTTTGGG TTTTTTTTGGGGGGGG\n")
        for i in range(150):
            L = int(rng.integers(20, 160)); m = int(rng.integers(1, 60))
            ref = rng.integers(0, 4, L).astype(np.uint8)
            st = int(rng.integers(0, max(1, L - m)))
            rd = synth.mutate(rng, ref[st:st + m], 0.15)
            if len(rd) == 0:
                rd = np.array([0], np.uint8)
            if rng.random() < 0.5:
                rd = synth.revcomp(rd)
            if rng.random() < 0.2:
                rd = rng.integers(0, 5, m).astype(np.uint8)
            if rng.random() < 0.2:
                ref[rng.random(L) < 0.1] = 4
            q = orc.norm_quals(rd, rng.integers(0, 42, len(rd)))
            f.write("ALIGN %s %s %s\n" % (rc.s(ref), rc.s(rd), rc.qs(q)))
        for sd in range(8):
            v = rc.small_variant(sd, n_rate=0.01 if sd == 3 else 0, dummy_last=(sd == 2), err=0.04 + 0.02 * (sd % 4))
            rc.write_round(f, v, sd % 2)
        for sd in range(4):
            rc.write_polish(f, rc.small_variant(100 + sd, err=0.08))
        for i in range(60):
            n = int(rng.integers(5, 300))
            a = rng.integers(0, 4, n).astype(np.uint8); b = a.copy(); t = i % 6
            if t == 1 and n > 40:
                b = np.concatenate([a[:n // 3], a[n // 3 + int(rng.integers(1, n // 3)):]])
            if t == 2:
                b = np.concatenate([a[:n // 2], rng.integers(0, 4, int(rng.integers(1, 60))).astype(np.uint8), a[n // 2:]])
            if t == 3:
                b = synth.mutate(rng, a, 0.1)
            if t == 4:
                b = rng.integers(0, 4, int(rng.integers(1, 100))).astype(np.uint8)
            if t == 5:
                b = synth.mutate(rng, np.concatenate([a[:n // 3], a[n // 2:]]), 0.03); a[rng.random(n) < 0.02] = 4
            if len(b) == 0:
                b = np.array([1], np.uint8)
            f.write("FINAL %s %s\n" % (rc.s(a), rc.s(b)))
    out = rc.run_harness(path)
    with open(os.path.join(HERE, "expected.txt"), "w") as f:
        f.write(out)
    print("wrote", path, len(out.splitlines()), "results")


if __name__ == "__main__":
    main()
