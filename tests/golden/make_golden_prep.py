"""Regenerates tests/golden/prep_cases.txt + prep_expected.txt: the CIGAR walks of the prepare side of polish_variant
 run by the UNMODIFIED reference (oracle/_ref/ref_harness SUPPORT / REGION / CUTCOV):
  record_supports_variant      include/sviper/variant.h:239-275   (long-read selection, sviper.h:236-249)
  get_read_region_boundaries   include/sviper/basics.h:189-241    (long-read crop, sviper.h:318-340)
  cut_down_high_coverage       include/sviper/basics.h:403-442    (short-read thinning, sviper.h:405)
  merge_alignments (--merge)   include/sviper/merge_split_alignments.h:378-561 (split long reads, sviper.h:262-266)
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


if __name__ == "__main__" and "--merge" not in sys.argv:
    main()


# ---- merge_alignments (merge_split_alignments.h:378-561): split long reads ------------------------------------------
def body(rng, n_read, noisy=True):
    """CIGAR body that consumes exactly n_read read bases, starts and ends with M; returns (elements, ref length)"""
    el, left, ref = [], n_read, 0
    while left > 0:
        run = int(min(left, rng.integers(20, 200)))
        el.append(("M", run)); left -= run; ref += run
        if left > 30 and noisy and rng.random() < 0.5:
            c = int(rng.integers(1, 6))
            if rng.random() < 0.5:
                el.append(("I", c)); left -= c
            else:
                el.append(("D", c)); ref += c
    if el[-1][0] != "M":
        el.append(("M", 1))          # never reached (left > 30 guard), kept for safety
    return el, ref


def rec_line(name, flag, rid, begin, mq, el, read, clip_l, clip_r):
    """clip_l / clip_r: (kind, length) with kind in 'S', 'H' or None; read: full read string"""
    n = len(read)
    a = clip_l[1] if clip_l[0] else 0
    b = clip_r[1] if clip_r[0] else 0
    full = ([(clip_l[0], a)] if clip_l[0] and a else []) + el + ([(clip_r[0], b)] if clip_r[0] and b else [])
    seq = read[(a if clip_l[0] == "H" else 0):(n - b if clip_r[0] == "H" else n)]
    return "%s %d %d %d %d %s %s\n" % (name, flag, rid, begin, mq, cig(full), seq)


def merge_case(rng, i):
    lines = []
    n_groups = int(rng.integers(1, 4))
    for gi in range(n_groups):
        name = "read%03d_%d" % (i, gi)
        N = int(rng.integers(400, 1600))
        read = "".join("ACGT"[int(x)] for x in rng.integers(0, 4, N))
        strand = 16 if rng.random() < 0.5 else 0
        kind = (i + gi) % 7
        if kind == 6:                                                    # a read with one alignment only
            el, _ = body(rng, N)
            lines.append(rec_line(name, strand, 0, int(rng.integers(1000, 90000)), int(rng.integers(0, 61)), el, read, (None, 0), (None, 0)))
            continue
        a = int(rng.integers(0, 60)) if rng.random() < 0.4 else 0        # outer clips shared by both records
        b = int(rng.integers(0, 60)) if rng.random() < 0.4 else 0
        x = int(rng.integers(a + 150, N - b - 150))                      # end of the left part in the read
        ov = int(rng.integers(0, 41)) if kind in (0, 1, 2) else -int(rng.integers(0, 21))
        xr = x - ov                                                      # start of the right part in the read
        el_l, ref_l = body(rng, x - a, noisy=rng.random() < 0.8)
        el_r, ref_r = body(rng, N - b - xr, noisy=rng.random() < 0.8)
        B = int(rng.integers(1000, 90000))
        if kind in (0, 3):
            gap = int(rng.integers(30, 3000))                            # deletion between the parts
        elif kind in (1, 4):
            gap = -int(rng.integers(1, min(120, ref_l - 20)))            # parts overlap on the reference: insertion
        else:
            gap = 0
        Br = B + ref_l + gap
        ck = lambda: "H" if rng.random() < 0.5 else "S"
        # unaligned read bases between the parts: with hard clips the reference copies prefix(seq, n) beyond the other
        # record's sequence (a heap over-read, garbage in its output; the C ABI returns SVP_ERR_ARG) -- soft clips there
        ik = (lambda: "S") if ov < 0 else ck
        left_primary = rng.random() < 0.5
        mq_p, mq_s = int(rng.integers(0, 61)), int(rng.integers(0, 61))
        k_ol, k_or = ck(), ck()
        recs = [("L", rec_line(name, strand | (0 if left_primary else 2048), 0, B, mq_p if left_primary else mq_s, el_l, read,
                               (k_ol if a else None, a), (ik(), N - x))),
                ("R", rec_line(name, strand | (2048 if left_primary else 0), 0, Br, mq_s if left_primary else mq_p, el_r, read,
                               (ik(), xr), (k_or if b else None, b)))]
        extra = rng.random()
        if extra < 0.15:                                                 # a secondary alignment: never merged
            el, _ = body(rng, N)
            recs.append(("X", rec_line(name, strand | 256, 0, int(rng.integers(1000, 90000)), int(rng.integers(0, 61)), el, read, (None, 0), (None, 0))))
        elif extra < 0.3:                                                # supplementary on another contig / strand
            el, _ = body(rng, N - 300)
            other = rng.random() < 0.5
            recs.append(("X", rec_line(name, (strand ^ (0 if other else 16)) | 2048, 1 if other else 0, int(rng.integers(1000, 90000)),
                                       int(rng.integers(0, 61)), el, read, ("H", 300), (None, 0))))
        elif extra < 0.4:                                                # a supplementary that is completely covered
            el, _ = body(rng, 100, noisy=False)
            recs.append(("X", rec_line(name, strand | 2048, 0, B + 20, int(rng.integers(0, 61)), el, read, ("H", a + 20), ("H", N - a - 120))))
        order = rng.permutation(len(recs))
        lines.extend(recs[j][1] for j in order)
    return "MERGE %d\n%s" % (len(lines), "".join(lines))


def main_merge():
    rng = np.random.default_rng(20261106)
    cases = [merge_case(rng, i) for i in range(110)]
    path = os.path.join(HERE, "merge_cases.txt")
    with open(path, "w") as f:
        f.write("".join(cases))
    out = rc.run_harness(path).strip("\n").split("\n")
    # A "completely covered" supplementary in FRONT of the primary is not recognised as such by the reference and its hard
    # clips then index beyond the other record's sequence: the reference's output holds heap garbage there (non-ACGT
    # characters). Such cases are dropped from the fixture; tests/test_prep_golden.py checks that the C ABI answers
    # SVP_ERR_ARG for a hand-made one.
    import re
    keep = [i for i, line in enumerate(out) if not any(re.search(r"[^ACGT]", t.split(":")[-1]) for t in line.split()[2:])]
    with open(path, "w") as f:
        f.write("".join(cases[i] for i in keep))
    with open(os.path.join(HERE, "merge_expected.txt"), "w") as f:
        f.write("".join(out[i] + "\n" for i in keep))
    print(len(keep), "merge lines kept,", len(out) - len(keep), "with undefined behaviour in the reference")


if __name__ == "__main__" and "--merge" in sys.argv:
    main_merge()
