#!/usr/bin/env python
"""Randomised parity sweep on a GPU box: small variants of widely varying shape (read lengths 20..256, coverage,
flank, error rates, N bases, insert windows, garbage / same-strand / mate-less pairs), polished in batches through
the C ABI and compared with the oracle's literal restatement of polish_to_perfection + the convex final
alignment. Bit-exact or it prints the offending seed and exits 1.
usage: tools/fuzz_parity.py [seconds] [seed0] [report.json]
The report (written also when a mismatch stops the campaign) holds the seed range -- random_variant(seed) regenerates any
case -- the batch sizes, per-shape counts of what was covered, and the offending seed if there was one."""
import collections
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from tests import refcases as rc  # noqa: E402

orc, synth = rc.orc, rc.synth


SHAPES = collections.Counter()


def random_variant(seed):
    rng = np.random.default_rng(seed)
    read_len = int(rng.choice([20, 36, 50, 75, 100, 150, 151, 153, 200, 256]))
    flank = int(rng.integers(5, 60))
    L_core = max(int(rng.integers(30, 400)), 2 * read_len + 140 - 2 * flank)   # room for one fragment
    n_pairs = int(rng.integers(4, 60))
    err = float(rng.choice([0.0, 0.01, 0.03, 0.06, 0.12]))
    ins_mean = float(rng.choice([30.0, 60.0, 120.0, 280.054]))
    ins_sd = float(rng.choice([5.0, 25.0, 145.162]))
    v = rc.small_variant(seed, L_core=L_core, flank=flank, read_len=read_len, n_pairs=n_pairs, err=err,
                         ins_mean=ins_mean, ins_sd=ins_sd, n_rate=float(rng.choice([0.0, 0.0, 0.01])),
                         dummy_last=bool(rng.random() < 0.3))
    # unrelated pairs, ragged lengths
    for _ in range(int(rng.integers(0, 4))):
        la, lb = int(rng.integers(1, read_len + 1)), int(rng.integers(1, read_len + 1))
        v["reads1"].append(rng.integers(0, 4, la).astype(np.uint8)); v["quals1"].append(rng.integers(0, 42, la).astype(np.uint8))
        v["reads2"].append(rng.integers(0, 4, lb).astype(np.uint8)); v["quals2"].append(rng.integers(0, 42, lb).astype(np.uint8))
    # a reference part that differs from the truth by an indel, for the final alignment
    t = v["truth"]
    cut = int(rng.integers(flank, len(t) - flank))
    if rng.random() < 0.5:
        v["ref_part"] = np.concatenate([t[:cut], rng.integers(0, 4, int(rng.integers(1, 80))).astype(np.uint8), t[cut:]])
    else:
        v["ref_part"] = np.concatenate([t[:cut], t[min(len(t) - flank, cut + int(rng.integers(1, 40))):]])
    SHAPES["read_len=%d" % read_len] += 1
    SHAPES["cons_err=%g" % err] += 1
    SHAPES["insert=%g+-%g" % (ins_mean, ins_sd)] += 1
    SHAPES["with_N" if any((r == 4).any() for r in v["reads1"]) else "no_N"] += 1
    SHAPES["dummy_mate" if any(len(r) == 0 for r in v["reads2"]) else "all_mated"] += 1
    SHAPES["ref_part_ins" if len(v["ref_part"]) > len(t) else "ref_part_del"] += 1
    SHAPES["window<=200" if len(v["consensus"]) <= 200 else ("window<=400" if len(v["consensus"]) <= 400 else "window>400")] += 1
    return v


OUTCOMES = collections.Counter()


def check(ctx, vs, seeds):
    got = ctx.polish_batch(vs, with_final=True)
    for v, g, sd in zip(vs, got, seeds):
        r1p, r2p = rc.packed(v)
        cfg = rc.oracle_cfg(v, r1p, r2p)
        seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
        ok = (np.array_equal(seq, g["polished"]) and np.array_equal(cov, g["cov"]) and
              (cfg.rounds, cfg.subst, cfg.ins, cfg.del_) == (g["rounds"], g["subst"], g["ins"], g["dele"]) and
              cfg.mappQ_mean == g["mappQ_mean"])
        if ok:
            sc, ops = orc.align_convex(v["ref_part"], seq)
            ok = sc == g["final_score"] and orc.cigar(ops) == g.get("cigar", [])
        OUTCOMES["rounds=%d" % g["rounds"]] += 1
        OUTCOMES["realigned_rounds=%d" % g["alignment_rounds"]] += 1
        OUTCOMES["cigar_elements<=3" if g["n_cigar"] <= 3 else "cigar_elements>3"] += 1
        OUTCOMES["edited" if (g["subst"] + g["ins"] + g["dele"]) else "unchanged"] += 1
        if not ok:
            print("MISMATCH at seed", sd)
            return sd
    return None


def main():
    from sviper_b200 import api
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    report = sys.argv[3] if len(sys.argv) > 3 else None
    ctx = api.Context(0)
    t0, n, seed0, batches, bad = time.time(), 0, seed, collections.Counter(), None
    while time.time() - t0 < budget and bad is None:
        k = int(np.random.default_rng(seed).integers(1, 40))          # batch sizes on both sides of the lane split
        seeds = list(range(seed, seed + k))
        vs = [random_variant(s) for s in seeds]
        bad = check(ctx, vs, seeds)
        batches["batch<8" if k < 8 else ("batch<16" if k < 16 else "batch>=16")] += 1
        seed += k; n += k
    if report:
        with open(report, "w") as f:
            json.dump({"tool": "tools/fuzz_parity.py", "seconds": time.time() - t0, "variants": n, "first_seed": seed0,
                       "last_seed": seed - 1, "regenerate": "tools/fuzz_parity.py: random_variant(seed)",
                       "mismatch_at_seed": bad, "batches": dict(sorted(batches.items())),
                       "shapes": dict(sorted(SHAPES.items())), "outcomes": dict(sorted(OUTCOMES.items())),
                       "compared": "polished sequence, coverage profile, rounds / subst / ins / del counters, mappQ_mean, "
                                   "final fp64 score (==) and CIGAR against the oracle, bit for bit"}, f, indent=1)
    if bad is not None:
        sys.exit(1)
    print("fuzz_parity: %d variants, all bit-exact (%.0f s)" % (n, time.time() - t0))


if __name__ == "__main__":
    main()
