"""Regenerates tests/golden/eval_cases.txt + eval_expected.txt: the tail of polish_variant
(assign_quality + refine_variant, sviper.h:468-489) run by the UNMODIFIED reference
(oracle/_ref/ref_harness EVAL, built from /root/reference by oracle/Makefile) on seeded final records.
Run here (the container that has /root/reference); the outputs are committed."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tests import refcases as rc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def info_field(rng, typ, pos, end, svlen):
    parts = ["SVTYPE=" + typ]
    r = rng.random()
    if r < 0.35:
        parts.append("SVLEN=-%d" % svlen)
    elif r < 0.7:
        parts.append("SVLEN=%d" % svlen)
    elif r < 0.75:
        parts.append("SVLEN=NA")
    if rng.random() < 0.7:
        parts.append("END=%d" % end)
    if rng.random() < 0.3:
        parts.append("CIPOS=-10,10;CIEND=-%d,%d" % (rng.integers(1, 50), rng.integers(1, 50)))
    if rng.random() < 0.25:
        parts.append("SEQ=" + "".join("ACGTacgtN"[int(x)] for x in rng.integers(0, 9, int(rng.integers(0, 12)))))
    if rng.random() < 0.1:
        parts.append("SVLEN=%d" % (svlen + 1))      # a second occurrence: regex_replace rewrites every match
    if rng.random() < 0.2:
        parts.append("RE=%d;IMPRECISE" % rng.integers(1, 40))
    order = rng.permutation(len(parts))
    return ";".join(parts[i] for i in order)


def one_case(rng, i):
    typ = "DEL" if rng.random() < 0.5 else "INS"
    k = 400 if rng.random() < 0.7 else 1000
    svlen = int(np.exp(rng.uniform(np.log(30), np.log(1500))))
    pos = int(rng.integers(5000, 200000))
    end = pos + svlen if typ == "DEL" else pos
    begin = max(0, pos - k - 500)
    # CIGAR of the polished sequence vs the reference part: flank, small noise, the SV element (or not), flank
    mode = i % 8
    true_len = svlen
    if mode == 1:
        true_len = max(1, int(svlen * rng.uniform(0.3, 0.45)))        # just around the lower size bound
    elif mode == 2:
        true_len = int(svlen * rng.uniform(1.55, 1.7))                # around the upper size bound
    elif mode == 3:
        true_len = max(1, int(svlen * rng.uniform(0.9, 1.1)))
    shift = int(rng.integers(-10, 11))
    if mode == 4:
        shift = int(svlen * rng.uniform(-0.7, 0.7)) + int(rng.integers(-3, 4))   # around the position bounds
    if mode == 5:
        shift = int(svlen * rng.uniform(1.3, 1.8))                    # past the region of interest
    sv_at = max(begin + 10, pos + shift)
    el = []
    ref = begin

    def noisy_match(upto):
        nonlocal ref
        while ref < upto:
            run = int(min(upto - ref, rng.integers(20, 400)))
            el.append(("M", run)); ref += run
            if ref < upto and rng.random() < 0.5:
                op = "ID"[int(rng.integers(0, 2))]
                c = int(rng.integers(1, 4))
                if rng.random() < 0.05:
                    c = int(rng.integers(svlen // 2 + 1, svlen + 2))     # a decoy of the other (or same) type
                el.append((op, c))
                if op == "D":
                    ref += c
    if mode == 6 and rng.random() < 0.5:
        el.append(("S", int(rng.integers(1, 30))))
    noisy_match(sv_at)
    if mode != 7:                                                       # mode 7: polished away
        el.append(("D" if typ == "DEL" else "I", true_len))
        if typ == "DEL":
            ref += true_len
    noisy_match(ref + k + 500 - int(rng.integers(0, 20)))
    if mode == 6:
        el = [(("=" if rng.random() < 0.5 else "X") if o == "M" and rng.random() < 0.3 else o, c) for o, c in el]
        if rng.random() < 0.5:
            el.append(("H", 5))
    # merge equal neighbours (getCigarString never emits them) and build a consistent sequence
    cig = []
    for o, c in el:
        if cig and cig[-1][0] == o:
            cig[-1] = (o, cig[-1][1] + c)
        else:
            cig.append((o, c))
    n_seq = sum(c for o, c in cig if o in "MIS=X")
    seq = "".join("ACGTN"[int(x)] for x in rng.choice(5, n_seq, p=[0.2475, 0.2475, 0.2475, 0.2475, 0.01]))
    if typ == "DEL" and i % 16:
        seq = "*"                                   # refine_variant reads the sequence for insertions only
    return "EVAL %s %d %d %d %d %d %s %s %s\n" % (typ, pos, end, svlen, k, begin, info_field(rng, typ, pos, end, svlen),
                                                 "".join("%d%s" % (c, o) for o, c in cig), seq)


def main():
    rng = np.random.default_rng(20261105)
    path = os.path.join(HERE, "eval_cases.txt")
    with open(path, "w") as f:
        # hand-made edges: single element, everything before the region of interest, immediate stop, CIEND quirk
        f.write("EVAL DEL 1000 1100 100 400 500 SVTYPE=DEL 1000M *\n")
        f.write("EVAL DEL 1000 1100 100 400 0 SVTYPE=DEL;END=1100;SVLEN=-100 100M *\n")
        f.write("EVAL DEL 1000 1100 100 400 1000 END=1100;CIEND=-5,5;SVLEN=100 100D900M *\n")
        f.write("EVAL DEL 1000 1100 100 400 960 SVLEN=-100;SVLEN=-7 10M40D900M *\n")
        f.write("EVAL DEL 1000 1100 100 400 960 SVLEN=-100 10M41D900M *\n")
        f.write("EVAL DEL 1000 1100 100 400 960 SVLEN=-100 10M160D900M *\n")
        f.write("EVAL DEL 1000 1100 100 400 960 SVLEN=-100 10M161D900M *\n")
        f.write("EVAL INS 1000 1000 6 400 990 SVTYPE=INS;SVLEN=6 12M5I300M ACGTACGTACGTTTTTTACGT" + "A" * 296 + "\n")
        f.write("EVAL INS 1000 1000 6 400 990 SVTYPE=INS;SEQ=;SVLEN=6;SEQ=acgtNNN 12M5I300M ACGTACGTACGTTNTTTACGT" + "C" * 296 + "\n")
        f.write("EVAL INS 1000 1000 100 400 500 SVTYPE=INS 449M100I *\n".replace("*", "G" * 549))
        for i in range(320):
            f.write(one_case(rng, i))
    out = rc.run_harness(path)
    with open(os.path.join(HERE, "eval_expected.txt"), "w") as f:
        f.write(out)
    lines = out.strip().split("\n")
    print("wrote", path, len(lines), "results;", sum(" has=1 " in x for x in lines), "refined,",
          sum("FAIL5" in x for x in lines), "FAIL5")


if __name__ == "__main__":
    main()
