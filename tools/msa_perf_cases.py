#!/usr/bin/env python
"""Writes a case file of build_consensus calls on 5 long-read crops of a given length at 10 % error (the long-read model of
SURVEY 8d) for timing sviper_b200/host/msa_harness against SVIPER_B200_MSA=host.  usage: msa_perf_cases.py out.txt len n_cases"""
import sys

import numpy as np

sys.path.insert(0, ".")
from tests import refcases as rc  # noqa: E402


def main():
    out, length, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    rng = np.random.default_rng(length)
    with open(out, "w") as f:
        for _ in range(n):
            truth = rng.integers(0, 4, length).astype(np.uint8)
            f.write("MSA 5\n")
            for _ in range(5):
                f.write("%s %d\n" % (rc.s(rc.synth.mutate(rng, truth, 0.10, frac_sub=0.3, frac_ins=0.35)), int(rng.integers(20, 61))))


if __name__ == "__main__":
    main()
