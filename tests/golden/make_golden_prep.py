"""Regenerates tests/golden/prep_cases.txt + prep_expected.txt: the CIGAR walks of the prepare side of polish_variant
 run by the UNMODIFIED reference (oracle/_ref/ref_harness SUPPORT / REGION / CUTCOV):
  record_supports_variant      include/sviper/variant.h:239-275   (long-read selection, sviper.h:236-249)
  get_read_region_boundaries   include/sviper/basics.h:189-241    (long-read crop, sviper.h:318-340)
  cut_down_high_coverage       include/sviper/basics.h:403-442    (short-read thinning, sviper.h:405)
Run here (the container that has /root/reference); the outputs are committed."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tests import refcases as rc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def cig(el):
    return "".join("%d%s" % (c, o) for o, c in el)


def long_read_cigar(rng, begin, end, sv=None, clip=True, ops="MID"):
    """a noisy long-read alignment from `begin` to about `end`, with an optional (pos, op, len) SV element"""
    el = []
    if clip and rng.random() < 0.5:
        el.append(("S" if rng.random() < 0.7 else "H", int(rng.integers(1, 3000))))
    ref = begin
    placed = sv is None
    while ref < end:
        if not placed and ref >= sv[0]:
            el.append((sv[1], sv[2]))
            if sv[1] == "D":
                ref += sv[2]
            placed = True
            continue
        stop = end if placed else min(end, max(sv[0], ref + 1))
        run = int(min(stop - ref, rng.integers(1, 60)))
        el.append(("M=X"[int(rng.integers(0, 3))] if "=" in ops else "M", run)); ref += run
        if ref < stop and rng.random() < 0.6:
            op = "ID"[int(rng.integers(0, 2))]
            c = int(rng.integers(1, 6))
            if rng.random() < 0.02:
                c = int(rng.integers(20, 400))
            el.append((op, c))
            if op == "D":
                ref += c
    if clip and rng.random() < 0.5:
        el.append(("S" if rng.random() < 0.7 else "H", int(rng.integers(1, 3000))))
    return el


def support_case(rng, i):
    typ = "DEL" if rng.random() < 0.5 else "INS"
    svlen = int(np.exp(rng.uniform(np.log(30), np.log(3000))))
    pos = int(rng.integers(20000, 500000))
    mode = i % 8
    true_len = svlen
    if mode == 1:
        true_len = max(1, int(svlen * rng.uniform(0.3, 0.5)))            # around the lower size bound (0.4)
    elif mode == 2:
        true_len = int(svlen * rng.uniform(1.5, 1.7))                    # around the upper size bound (1.6)
    shift = int(rng.integers(-20, 21))
    if mode == 3:
        shift = int(svlen * rng.uniform(-0.7, -0.3))                     # around the start of the region of interest
    elif mode == 4:
        shift = int(svlen * rng.uniform(1.2, 1.8))                       # around its end
    begin = pos - int(rng.integers(200, 8000))
    if mode == 5:
        begin = pos + int(svlen * rng.uniform(-0.6, 2.0))                # record starts inside / behind the region
    end = max(begin + 50, pos + svlen + int(rng.integers(-svlen, 8000)))
    sv = None
    if mode != 6:
        op = "D" if typ == "DEL" else "I"
        if mode == 7:
            op = "I" if typ == "DEL" else "D"                            # element of the other type
        sv = (max(begin + 1, pos + shift), op, max(1, true_len))
    el = long_read_cigar(rng, begin, end, sv, ops="M=X" if rng.random() < 0.3 else "MID")
    return "SUPPORT %s %d %d %d %s\n" % (typ, pos, svlen, begin, cig(el))


def region_case(rng, i):
    begin = int(rng.integers(1000, 400000))
    span = int(rng.integers(300, 12000))
    el = long_read_cigar(rng, begin, begin + span, None, ops="M=X" if rng.random() < 0.3 else "MID")
    mode = i % 6
    rb = begin + int(rng.integers(0, span // 2))
    re = rb + int(rng.integers(1, span // 2))
    if mode == 1:
        rb = begin - int(rng.integers(0, 500))                           # region starts before the record
        re = begin + 2 + int(rng.integers(0, span // 2))                 # (a region that ENDS before it is undefined
                                                                         #  behaviour in the reference: cigar[-1])
    elif mode == 2:
        re = begin + span + int(rng.integers(0, 500))                    # region ends behind the record
    elif mode == 3:
        rb, re = re, rb                                                  # start >= end: error path, (0, 0)
    elif mode == 4:
        re = rb                                                          # empty region: error path
    return "REGION %d %s %d %d\n" % (begin, cig(el), rb, re)


def cutcov_case(rng, i):
    n = int(rng.integers(1, 400)) if i % 10 else 0
    cov = int(rng.integers(1, 60))
    start = int(rng.integers(1000, 300000))
    width = int(rng.integers(200, 3000))
    lines = []
    for j in range(n):
        b = start + int(rng.integers(0, width))
        if i % 7 == 3 and j % 5 == 0:
            b = start + int(rng.integers(0, 4)) * 50                     # many records on the same position
        L = 150 if rng.random() < 0.8 else int(rng.integers(100, 251))
        r = rng.random()
        if r < 0.6:
            el = [("M", L)]
        elif r < 0.8:
            s = int(rng.integers(1, L // 2))
            el = [("S", s), ("M", L - s)] if rng.random() < 0.5 else [("M", L - s), ("S", s)]
        else:
            a = int(rng.integers(10, L - 20)); c = int(rng.integers(1, 8))
            el = [("M", a), ("D" if rng.random() < 0.5 else "I", c), ("M", L - a - c if L - a - c > 0 else 5)]
        lines.append("r%d_%d %d %d %s" % (i, j, b, int(rng.integers(0, 61)) if rng.random() < 0.7 else 60, cig(el)))
    return "CUTCOV %d %d\n%s" % (cov, n, "".join(x + "\n" for x in lines))


def main():
    rng = np.random.default_rng(20261105)
    cases = [support_case(rng, i) for i in range(240)] + [region_case(rng, i) for i in range(180)] + \
            [cutcov_case(rng, i) for i in range(60)]
    path = os.path.join(HERE, "prep_cases.txt")
    with open(path, "w") as f:
        f.write("".join(cases))
    out = rc.run_harness(path)
    with open(os.path.join(HERE, "prep_expected.txt"), "w") as f:
        f.write(out)
    lines = out.strip().split("\n")
    print(len(lines), "lines;", sum(x == "SUPPORT supports=1" for x in lines), "supporting;",
          sum(x == "REGION begin=0 end=0" for x in lines), "region error paths")


if __name__ == "__main__":
    main()
