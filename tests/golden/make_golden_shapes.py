"""Regenerates tests/golden/shapes_cases.txt + shapes_expected.txt: the UNMODIFIED reference (oracle/_ref/ref_harness) at the
REAL shapes of BASELINE.json's configs -- 150 bp reads against 1.8 kb windows, the 152 / 153 bp boundary between the two row
classes of the score kernel, 250 / 256 bp reads (the device limit), strand ties and palindromes, reads of Ns, whole variants of
configs[1] / configs[2] shape (30x, -k 400, DEL and INS) through polish_to_perfection, and final alignments of 1.8 kb against
2.8 kb with a 1 kb deletion / insertion. cases.txt holds miniatures so that the reference answers in milliseconds; these take
it about a minute. Run here (the container that has /root/reference); the outputs are committed."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tests import refcases as rc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
orc, synth = rc.orc, rc.synth


def main():
    rng = np.random.default_rng(20261105)
    path = os.path.join(HERE, "shapes_cases.txt")
    with open(path, "w") as f:
        # map_single_read at real shapes (mapping.h:84-137)
        for i, (m, L) in enumerate([(150, 1800)] * 8 + [(152, 1800), (153, 1800), (152, 900), (153, 2000)] + [(250, 2000)] * 3
                                   + [(256, 2000)] * 3 + [(150, 13000)] * 2):
            ref = rng.integers(0, 4, L).astype(np.uint8)
            st = int(rng.integers(0, L - m))
            rd = synth.mutate(rng, ref[st:st + m], 0.01 if i % 3 else 0.08)[:256]
            if i % 2:
                rd = synth.revcomp(rd)
            if i % 7 == 3:                                   # hangs off the left / right window edge
                rd = np.concatenate([rng.integers(0, 4, 40).astype(np.uint8), ref[:m - 40]]) if i % 2 else \
                    np.concatenate([ref[L - (m - 40):], rng.integers(0, 4, 40).astype(np.uint8)])
            q = orc.norm_quals(rd, rng.integers(2, 42, len(rd)))
            f.write("ALIGN %s %s %s\n" % (rc.s(ref), rc.s(rd), rc.qs(q)))
        # strand ties: a reverse-palindromic read scores the same on both strands (mapping.h:121 picks forward on a tie)
        for L in (600, 1800):
            ref = rng.integers(0, 4, L).astype(np.uint8)
            half = rng.integers(0, 4, 75).astype(np.uint8)
            pal = np.concatenate([half, synth.revcomp(half)])
            ref[200:350] = pal
            f.write("ALIGN %s %s %s\n" % (rc.s(ref), rc.s(pal), rc.qs(orc.norm_quals(pal, np.full(150, 30)))))
            at = np.tile(np.array([0, 3], np.uint8), 75)                               # (AT)n: its own reverse complement
            ref[400:560] = np.tile(np.array([0, 3], np.uint8), 80)
            f.write("ALIGN %s %s %s\n" % (rc.s(ref), rc.s(at), rc.qs(orc.norm_quals(at, np.full(150, 30)))))
            ns = np.full(150, 4, np.uint8)                                             # a read of Ns (N == N matches)
            ref[50:60] = 4
            f.write("ALIGN %s %s %s\n" % (rc.s(ref), rc.s(ns), rc.qs(orc.norm_quals(ns, np.zeros(150)))))
            unrelated = rng.integers(0, 4, 150).astype(np.uint8)
            f.write("ALIGN %s %s %s\n" % (rc.s(ref), rc.s(unrelated), rc.qs(orc.norm_quals(unrelated, np.full(150, 20)))))
        # one mapping + pileup round and whole polish_to_perfection runs at configs[1] / configs[2] shape
        vd = synth.make_variant(7001, "DEL", svlen=300, k=400, cov=30, cons_err=0.05)
        vi = synth.make_variant(7002, "INS", svlen=500, k=400, cov=30, cons_err=0.05)
        rc.write_round(f, vd, 0)
        rc.write_round(f, vi, 1)
        rc.write_polish(f, vd)
        rc.write_polish(f, vi)
        # final alignment (sviper.h:457-466) at real size: polished sequence against the reference part
        for kind in ("DEL", "INS", "DEL"):
            a = rng.integers(0, 4, 2800).astype(np.uint8)
            b = np.concatenate([a[:900], a[1900:]]) if kind == "DEL" else a
            if kind == "INS":
                a = np.concatenate([b[:900], b[1900:]])
            b = synth.mutate(rng, b, 0.003)
            f.write("FINAL %s %s\n" % (rc.s(a), rc.s(b)))
    out = rc.run_harness(path)
    with open(os.path.join(HERE, "shapes_expected.txt"), "w") as f:
        f.write(out)
    print("wrote", path, len(out.splitlines()), "results")


if __name__ == "__main__":
    main()
