"""Regenerates tests/golden/msa_cases.txt + msa_expected.txt by running the UNMODIFIED reference (oracle/_ref/ref_harness) on
seeded cases of SURVEY 8f row f2 (build_consensus): the two pairwise DP fills of the T-Coffee library generation (SWFILL: the
legacy clumping Smith-Waterman, GFILL: the Gotoh trace matrix + the fragments of the library call) and whole build_consensus
calls (MSA). Run here (the container that has /root/reference); the outputs are committed."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tests import refcases as rc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
synth = rc.synth


def crop_set(rng, n_reads, length, err):
    truth = rng.integers(0, 4, length).astype(np.uint8)
    out = []
    for _ in range(n_reads):
        r = synth.mutate(rng, truth, err, frac_sub=0.3, frac_ins=0.35)
        out.append(r if len(r) else np.array([0], np.uint8))
    return out


def main():
    rng = np.random.default_rng(20261105)
    path = os.path.join(HERE, "msa_cases.txt")
    with open(path, "w") as f:
        for i in range(70):
            n = int(rng.integers(1, 330)); a = rng.integers(0, 4, n).astype(np.uint8)
            t = i % 5
            b = synth.mutate(rng, a, 0.1)
            if t == 1:
                b = rng.integers(0, 4, int(rng.integers(1, 200))).astype(np.uint8)
            if t == 2:
                b = np.concatenate([rng.integers(0, 4, 30).astype(np.uint8), synth.mutate(rng, a[n // 4:], 0.15)])
            if t == 3 and rng.random() < 0.5:
                a = np.concatenate([a, a[:n // 2]])                      # repeats: several good local alignments
            if t == 4:
                b[rng.random(len(b)) < 0.05] = 4
            if len(b) == 0:
                b = np.array([2], np.uint8)
            f.write("GFILL %s %s\n" % (rc.s(a), rc.s(b)))
            # forbidden cells: none, a diagonal stretch (an earlier local alignment), or scattered
            cells = []
            if i % 3 == 1:
                c0, r0 = int(rng.integers(1, len(a) + 1)), int(rng.integers(1, len(b) + 1))
                k = 0
                while c0 + k <= len(a) and r0 + k <= len(b) and k < 120:
                    cells.append((c0 + k, r0 + k)); k += 1
            if i % 3 == 2:
                for _ in range(int(rng.integers(1, 40))):
                    cells.append((int(rng.integers(1, len(a) + 1)), int(rng.integers(1, len(b) + 1))))
            f.write("SWFILL %s %s %d %s\n" % (rc.s(a), rc.s(b), len(cells), " ".join("%d %d" % c for c in cells)))
        shapes = [(2, 150), (3, 200), (5, 250), (4, 320), (5, 400), (1, 120), (5, 180)]
        for i in range(110):
            n_reads, length = shapes[i % len(shapes)]
            if i >= 104:
                n_reads, length = 5, 900 + 150 * (i - 104)                 # k = 400 crops: 800 bp and more
            err = [0.1, 0.1, 0.05, 0.15, 0.02][i % 5]
            reads = crop_set(rng, n_reads, length, err)
            if i % 11 == 3:
                reads[0] = reads[0][: len(reads[0]) // 2]                  # a crop that covers only half the region
            if i % 13 == 5:
                reads[-1][rng.random(len(reads[-1])) < 0.03] = 4           # N bases
            f.write("MSA %d\n" % len(reads))
            for r in reads:
                f.write("%s %d\n" % (rc.s(r), int(rng.integers(20, 61))))
    out = rc.run_harness(path)
    with open(os.path.join(HERE, "msa_expected.txt"), "w") as f:
        f.write(out)
    print("wrote", path, len(out.splitlines()), "results")


if __name__ == "__main__":
    main()
