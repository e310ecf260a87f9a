"""Case files for oracle/_ref/ref_harness (the unmodified reference) and parsers for its output.
Shared by tests/golden/make_golden.py (writes the committed fixtures) and the parity tests."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402  (tests may import the oracle)
from sviper_b200 import synth  # noqa: E402

HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def s(a):
    return orc.dec(a) if len(a) else "*"


def qs(q):
    return "".join(chr(33 + int(x)) for x in q) if len(q) else "*"


def small_variant(seed, L_core=120, flank=30, read_len=40, n_pairs=30, err=0.06, ins_mean=60.0, ins_sd=25.0,
                  n_rate=0.0, dummy_last=False):
    """miniature of synth.make_variant so that the reference answers in milliseconds"""
    rng = np.random.default_rng(seed)
    margin = 60
    alt = rng.integers(0, 4, margin + flank + L_core + flank + margin).astype(np.uint8)
    truth = alt[margin:len(alt) - margin]
    mid = synth.mutate(rng, truth[flank:len(truth) - flank], err)
    cons = np.concatenate([truth[:flank], mid, truth[len(truth) - flank:]])
    r1, q1, r2, q2 = [], [], [], []
    while len(r1) < n_pairs:
        fl = int(np.clip(rng.normal(2 * read_len + ins_mean * 0.6, 10), 2 * read_len + 2, 2 * read_len + 120))
        st = int(rng.integers(0, len(alt) - fl))
        a = alt[st:st + read_len].copy(); b = synth.revcomp(alt[st + fl - read_len:st + fl])
        for r in (a, b):
            m = rng.random(read_len) < 0.01
            r[m] = (r[m] + 1) % 4
            if n_rate:
                r[rng.random(read_len) < n_rate] = 4
        qa = rng.integers(2, 41, read_len).astype(np.uint8); qb = rng.integers(2, 41, read_len).astype(np.uint8)
        qa[a == 4] = 0; qb[b == 4] = 0
        if rng.random() < 0.5:
            a, b, qa, qb = b, a, qb, qa
        if rng.random() < 0.1:      # same-strand pair: never proper
            b = synth.revcomp(b); qb = qb[::-1].copy()
        r1.append(a); q1.append(qa); r2.append(b); q2.append(qb)
    if dummy_last:
        r2[-1] = np.zeros(0, np.uint8); q2[-1] = np.zeros(0, np.uint8)
    return dict(flank=flank, consensus=cons, truth=truth, reads1=r1, quals1=q1, reads2=r2, quals2=q2,
                ins_mean=ins_mean, ins_sd=ins_sd)


def write_pairs(f, v):
    for a, qa, b, qb in zip(v["reads1"], v["quals1"], v["reads2"], v["quals2"]):
        f.write("%s %s %s %s\n" % (s(a), qs(qa), s(b), qs(qb)))


def write_round(f, v, fix):
    f.write("ROUND %d %d %r %r %d %s\n" % (fix, v["flank"], v["ins_mean"], v["ins_sd"], len(v["reads1"]),
                                           s(v["consensus"])))
    write_pairs(f, v)


def write_polish(f, v):
    f.write("POLISH %d %r %r %d %s\n" % (v["flank"], v["ins_mean"], v["ins_sd"], len(v["reads1"]), s(v["consensus"])))
    write_pairs(f, v)


def run_harness(case_path):
    return subprocess.check_output([HARNESS, case_path], text=True)


def parse_line(line):
    parts = line.split()
    d = {"cmd": parts[0]}
    for kv in parts[1:]:
        k, _, v = kv.partition("=")
        d[k] = v
    return d


def hexlist(v):
    return np.array([float.fromhex(x) for x in v.split(",")], np.float64) if v else np.zeros(0)


def intlist(v):
    return np.array([int(x) for x in v.split(",")], np.int64) if v else np.zeros(0, np.int64)


# ---- the same quantities through the oracle -------------------------------------------------

def oracle_cfg(v, r1p, r2p):
    bq = orc.baseq_mean(r1p[1], r1p[2], r2p[1], r2p[2])
    return orc.make_config(v["flank"], v["ins_mean"], v["ins_sd"], bq)


def packed(v):
    return orc.pack_reads(v["reads1"], v["quals1"]), orc.pack_reads(v["reads2"], v["quals2"])


def check_round(v, fix, exp):
    r1p, r2p = packed(v)
    cfg = oracle_cfg(v, r1p, r2p)
    cfg.fix_indels = fix
    out = orc.polish_round(cfg, *r1p, *r2p, v["consensus"])
    assert float.fromhex(exp["baseQ_mean"]) == cfg.baseQ_mean
    assert float.fromhex(exp["mappQ_mean"]) == cfg.mappQ_mean
    assert "".join(str(int(x)) for x in out["proper"]) == exp["proper"]
    assert "".join("-" if g else "x" for g in out["isgap"]) == "".join("-" if c == "-" else "x" for c in exp["gaps"])
    assert np.array_equal(out["cov"], intlist(exp["cov"]))
    prof = hexlist(exp["profile"]).reshape(-1, 6)
    assert prof.shape == out["profile"].shape
    assert np.array_equal(prof.view(np.uint64), out["profile"].view(np.uint64))   # bit-exact doubles
    assert s(out["out"]) == exp["seq"]
    assert (cfg.subst, cfg.ins, cfg.del_) == (int(exp["subst"]), int(exp["ins"]), int(exp["del"]))


def check_polish(v, exp):
    r1p, r2p = packed(v)
    cfg = oracle_cfg(v, r1p, r2p)
    seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
    assert s(seq) == exp["seq"]
    assert (cfg.rounds, cfg.subst, cfg.ins, cfg.del_) == tuple(int(exp[k]) for k in ("rounds", "subst", "ins", "del"))
    assert float.fromhex(exp["mappQ_mean"]) == cfg.mappQ_mean
    assert np.array_equal(cov, intlist(exp["cov"]))


# ---- reading a case file back (so committed fixtures are self-contained) --------------------

def unq(q):
    return np.array([ord(c) - 33 for c in q], np.uint8) if q != "*" else np.zeros(0, np.uint8)


def uns(x):
    return orc.enc(x) if x != "*" else np.zeros(0, np.uint8)


def read_cases(path):
    cases = []
    with open(path) as f:
        lines = iter(f.read().split("\n"))
    for line in lines:
        p = line.split()
        if not p:
            continue
        if p[0] == "ALIGNX":
            cases.append(dict(cmd="ALIGNX", sc=tuple(int(x) for x in p[1:5]), ref=uns(p[5]), read=uns(p[6])))
        elif p[0] == "ALIGN":
            cases.append(dict(cmd="ALIGN", ref=uns(p[1]), read=uns(p[2]), qual=unq(p[3])))
        elif p[0] == "FINAL":
            cases.append(dict(cmd="FINAL", a=uns(p[1]), b=uns(p[2])))
        elif p[0] in ("ROUND", "POLISH"):
            k = 1
            fix = 0
            if p[0] == "ROUND":
                fix = int(p[1]); k = 2
            v = dict(flank=int(p[k]), ins_mean=float(p[k + 1]), ins_sd=float(p[k + 2]), consensus=uns(p[k + 4]),
                     reads1=[], quals1=[], reads2=[], quals2=[])
            for _ in range(int(p[k + 3])):
                a, qa, b, qb = next(lines).split()
                v["reads1"].append(uns(a)); v["quals1"].append(unq(qa))
                v["reads2"].append(uns(b)); v["quals2"].append(unq(qb))
            cases.append(dict(cmd=p[0], fix=fix, v=v))
    return cases


def check_case(c, e):
    """compare the oracle with one parsed line of reference output"""
    assert c["cmd"] == e["cmd"]
    if c["cmd"] == "ALIGNX":
        sc, ops, _ = orc.align_semiglobal(c["ref"], c["read"], *c["sc"])
        rr, qq = orc.rows_from_ops(ops, c["ref"], c["read"])
        assert (sc, rr, qq) == (int(e["score"]), e["ref"], e["read"])
    elif c["cmd"] == "ALIGN":
        r = orc.map_read(c["ref"], c["read"], c["qual"])
        rd = synth.revcomp(c["read"]) if r["is_rc"] else c["read"]
        rr, qq = orc.rows_from_ops(r["ops"], c["ref"], rd)
        best = r["score_rc"] if r["is_rc"] else r["score"]
        assert (r["is_rc"], max(0, best), r["lead"], r["end"], rr, qq) == \
            (int(e["rc"]), int(e["mapQ"]), int(e["lead"]), int(e["end"]), e["ref"], e["read"])
    elif c["cmd"] == "FINAL":
        sc, ops = orc.align_convex(c["a"], c["b"])
        assert sc == float.fromhex(e["score"])
        assert orc.cigar_str(ops) == e["cigar"] and len(orc.cigar(ops)) == int(e["ncigar"])
    elif c["cmd"] == "ROUND":
        check_round(c["v"], c["fix"], e)
    else:
        check_polish(c["v"], e)


def oracle_full(v):
    """polish_to_perfection + final alignment of one variant through the oracle (ctypes releases the GIL: callers run
    several of these on a thread pool). Returns what the parity tests compare."""
    r1p, r2p = packed(v)
    cfg = oracle_cfg(v, r1p, r2p)
    seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
    sc, ops = orc.align_convex(v["ref_part"], seq)
    return dict(polished=seq, cov=cov, counters=(cfg.rounds, cfg.subst, cfg.ins, cfg.del_), final_score=sc, cigar=orc.cigar(ops))


def assert_matches_oracle(vs, got, workers=None):
    """every variant of a GPU result list against the oracle, bit for bit; the oracle runs on all host cores"""
    import concurrent.futures as cf
    import numpy as np
    with cf.ThreadPoolExecutor(max_workers=workers or min(len(vs), os.cpu_count() or 1)) as ex:
        want = list(ex.map(oracle_full, vs))
    for i, (w, g) in enumerate(zip(want, got)):
        assert np.array_equal(w["polished"], g["polished"]), "variant %d: polished sequence" % i
        assert np.array_equal(w["cov"], g["cov"]), "variant %d: coverage profile" % i
        assert w["counters"] == (g["rounds"], g["subst"], g["ins"], g["dele"]), "variant %d: counters" % i
        assert w["final_score"] == g["final_score"] and w["cigar"] == g["cigar"], "variant %d: final alignment" % i
    return want
