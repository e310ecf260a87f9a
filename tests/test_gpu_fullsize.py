"""GPU tests at BASELINE.json's full shapes. The oracle is too slow to re-run whole batches at these sizes, so
besides a few direct oracle comparisons the CUDA path is checked through size-independent properties:
batch-composition independence, idempotence of polishing, CIGAR/sequence length accounting, determinism."""
import numpy as np
import pytest

from tests import refcases as rc

pytestmark = pytest.mark.gpu
orc, synth = rc.orc, rc.synth


@pytest.fixture(scope="module")
def ctx():
    from sviper_b200 import api
    c = api.Context(0)
    yield c
    c.close()


def same(a, b):
    return (np.array_equal(a["polished"], b["polished"]) and np.array_equal(a["cov"], b["cov"]) and
            (a["rounds"], a["subst"], a["ins"], a["dele"]) == (b["rounds"], b["subst"], b["ins"], b["dele"]) and
            a["mappQ_mean"] == b["mappQ_mean"] and a.get("cigar") == b.get("cigar") and
            a["final_score"] == b["final_score"])


@pytest.fixture(scope="module")
def config1_batch():
    """configs[1] shape: DELs 50..1000 bp, 30x 2x150 reads, -k 400, 5% noisy consensus (bench.py's workload)"""
    import bench
    return bench.make_batch(24, seed=77)


def test_config1_oracle_on_full_size_variants(ctx, config1_batch):
    vs = config1_batch[:3]
    got = ctx.polish_batch(vs, with_final=True)
    for v, g in zip(vs, got):
        r1p, r2p = rc.packed(v)
        cfg = rc.oracle_cfg(v, r1p, r2p)
        seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
        assert np.array_equal(seq, g["polished"]) and np.array_equal(cov, g["cov"])
        assert (cfg.rounds, cfg.subst, cfg.ins, cfg.del_) == (g["rounds"], g["subst"], g["ins"], g["dele"])
        sc, ops = orc.align_convex(v["ref_part"], seq)
        assert sc == g["final_score"] and orc.cigar(ops) == g["cigar"]


def test_config1_properties(ctx, config1_batch):
    vs = config1_batch
    got = ctx.polish_batch(vs, with_final=True)
    # determinism: the same batch again (different arena contents, same results)
    again = ctx.polish_batch(vs, with_final=True)
    assert all(same(a, b) for a, b in zip(got, again))
    # batch-composition independence: a variant alone, and in a shuffled sub-batch, gives identical bits
    for i in (0, 7, 23):
        assert same(ctx.polish_batch([vs[i]], with_final=True)[0], got[i])
    order = [5, 3, 11, 2]
    sub = ctx.polish_batch([vs[i] for i in order], with_final=True)
    assert all(same(s, got[i]) for s, i in zip(sub, order))
    n_truth = 0
    for v, g in zip(vs, got):
        # CIGAR accounting (getCigarString, bam_io/cigar.h:354-420): M+D spans the reference part, M+I the polished seq
        ref_span = sum(n for o, n in g["cigar"] if o in "MD")
        seq_span = sum(n for o, n in g["cigar"] if o in "MI")
        assert ref_span == len(v["ref_part"]) and seq_span == len(g["polished"])
        assert g["n_cigar"] == len(g["cigar"]) and all(n > 0 for _, n in g["cigar"])
        assert all(a[0] != b[0] for a, b in zip(g["cigar"], g["cigar"][1:]))            # run-length encoded
        # the flanks are copied verbatim (polishing.h:247,287)
        f = v["flank"]
        assert np.array_equal(g["polished"][:f], v["consensus"][:f]) and np.array_equal(g["polished"][-f:], v["consensus"][-f:])
        assert 2 <= g["rounds"] <= 30 and len(g["cov"]) >= len(g["polished"]) - 1
        # the first indel round reuses the mapping of the unchanged consensus unless phase 1 hit its round limit
        assert g["alignment_rounds"] == g["rounds"] - (1 if g["rounds"] < 21 or g["alignment_rounds"] < g["rounds"] else 0)
        assert g["rounds"] - 1 <= g["alignment_rounds"] <= g["rounds"]
        n_truth += int(np.array_equal(g["polished"], v["truth"]))
    assert n_truth >= 0.7 * len(vs), "polishing should recover most of the synthetic truths (%d/%d)" % (n_truth, len(vs))
    # idempotence: polishing an already polished window changes nothing (one round per phase, polishing.h:327-343)
    vs2 = [dict(v, consensus=g["polished"]) for v, g in zip(vs, got)]
    got2 = ctx.polish_batch(vs2, with_final=True)
    for g, g2 in zip(got, got2):
        if g["rounds"] < 30:
            assert np.array_equal(g2["polished"], g["polished"])
            assert (g2["rounds"], g2["subst"], g2["ins"], g2["dele"]) == (2, 0, 0, 0)
            assert g2["cigar"] == g["cigar"] and g2["final_score"] == g["final_score"]


def test_config4_large_insertion_window(ctx):
    """configs[3] shape: kb-scale insertion with -k 1000 (window ~5 kb, final DP ~5k x 3k, multi-band rows)"""
    v = synth.make_variant(41, "INS", svlen=2000, k=1000, cov=5, cons_err=0.02)
    g = ctx.polish_batch([v], with_final=True)[0]
    r1p, r2p = rc.packed(v)
    cfg = rc.oracle_cfg(v, r1p, r2p)
    seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
    assert np.array_equal(seq, g["polished"]) and np.array_equal(cov, g["cov"])
    assert (cfg.rounds, cfg.subst, cfg.ins, cfg.del_) == (g["rounds"], g["subst"], g["ins"], g["dele"])
    sc, ops = orc.align_convex(v["ref_part"], seq)
    assert sc == g["final_score"] and orc.cigar(ops) == g["cigar"]
    assert any(o == "I" and n > 1500 for o, n in g["cigar"])


def test_config4_real_shape_against_the_oracle(ctx):
    """configs[3] at its real shape: insertions of 5, 7.5 and 10 kb with -k 1000 at 30x -- windows of 8-13 kb,
    700-1,200 read pairs each, final alignment ~3 k x 13 k over several row bands -- complete polish_to_perfection + final
    alignment + CIGAR, bit for bit against the oracle (which spreads the pairs of one window over the host cores)."""
    vs = [synth.make_variant(4100 + i, "INS", svlen=sv, k=1000, cov=30, cons_err=0.05) for i, sv in enumerate((5000, 7500, 10000))]
    assert [len(v["consensus"]) for v in vs] >= [8000, 10000, 12500]
    got = ctx.polish_batch(vs, with_final=True)
    rc.assert_matches_oracle(vs, got, workers=1)            # one window at a time: the oracle parallelises inside
    for v, g in zip(vs, got):
        assert max(n for o, n in g["cigar"] if o == "I") >= 0.95 * v["svlen"]


def test_final_alignment_with_the_table_outside_shared_memory(monkeypatch):
    """k5_final_kernel<false>: the gap-step table read through the read-only cache instead of shared memory. Chosen by the
    library for sequences beyond ~24,000 bases (table > 190 KB); SVP_K5_GLOBAL_TABLE=1 forces it, so that the path is
    checked at sizes the oracle answers in seconds, and one genuine case (25,000 x 600) runs it unforced."""
    from sviper_b200 import api
    rng = np.random.default_rng(99)
    pairs = []
    for i in range(4):
        a = rng.integers(0, 4, int(rng.integers(1800, 4200))).astype(np.uint8)
        cut = len(a) // 3
        b = synth.mutate(rng, np.concatenate([a[:cut], a[cut + int(rng.integers(20, 700)):]]), 0.003 * i)
        pairs.append((a, b) if i % 2 == 0 else (b, a))
    monkeypatch.setenv("SVP_K5_GLOBAL_TABLE", "1")
    c = api.Context(0)
    monkeypatch.delenv("SVP_K5_GLOBAL_TABLE")
    try:
        got = c.final_align_batch(pairs)
        for (a, b), g in zip(pairs, got):
            sc, ops = orc.align_convex(a, b)
            assert sc == g["score"] and orc.cigar(ops) == g["cigar"]
    finally:
        c.close()
    long_ref = rng.integers(0, 4, 25000).astype(np.uint8)
    short = synth.mutate(rng, long_ref[12000:12600], 0.01)
    c = api.Context(0)
    try:
        g = c.final_align_batch([(long_ref, short)], cigar_cap=8192)[0]
        sc, ops = orc.align_convex(long_ref, short)
        assert sc == g["score"] and orc.cigar(ops) == g["cigar"]
    finally:
        c.close()


def test_mixed_read_lengths_and_ns_in_one_batch(ctx):
    """ragged batch: 150 bp and 250 bp libraries, N bases, a dummy mate, windows of different size"""
    vs = [synth.make_variant(51, "DEL", svlen=80, k=200, cov=8, cons_err=0.03, n_rate=0.004),
          synth.make_variant(52, "INS", svlen=60, k=150, cov=8, read_len=250, cons_err=0.03, frag_mean=560.0),
          synth.make_variant(53, "DEL", svlen=400, k=300, cov=6, cons_err=0.05)]
    vs[0]["reads2"][-1] = np.zeros(0, np.uint8); vs[0]["quals2"][-1] = np.zeros(0, np.uint8)
    vs[1]["ins_mean"], vs[1]["ins_sd"] = 300.0, 150.0
    # pairs that cannot be proper (the traced pass skips them, plan_kernel): unrelated sequence, mates on the same
    # strand, reverse-forward order, and an insert far outside mean +- 3 sd
    rng = np.random.default_rng(5)
    v2, q = vs[2], np.full(150, 30, np.uint8)
    cons = v2["consensus"]
    for _ in range(6):
        v2["reads1"].append(rng.integers(0, 4, 150).astype(np.uint8)); v2["quals1"].append(q)
        v2["reads2"].append(rng.integers(0, 4, 150).astype(np.uint8)); v2["quals2"].append(q)
    a, b = cons[520:670].copy(), cons[900:1050].copy()
    extra = [(a, b), (synth.revcomp(a), synth.revcomp(b)), (synth.revcomp(a), b), (a, synth.revcomp(cons[700:850])),
             (cons[510:660].copy(), synth.revcomp(cons[len(cons) - 660:len(cons) - 510]))]
    for x, y in extra:
        v2["reads1"].append(np.ascontiguousarray(x)); v2["quals1"].append(q)
        v2["reads2"].append(np.ascontiguousarray(y)); v2["quals2"].append(q)
    v2["ins_mean"], v2["ins_sd"] = 120.0, 30.0
    got = ctx.polish_batch(vs, with_final=True)
    for v, g in zip(vs, got):
        r1p, r2p = rc.packed(v)
        cfg = rc.oracle_cfg(v, r1p, r2p)
        seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
        assert np.array_equal(seq, g["polished"]) and np.array_equal(cov, g["cov"])
        assert (cfg.rounds, cfg.subst, cfg.ins, cfg.del_) == (g["rounds"], g["subst"], g["ins"], g["dele"])
        sc, ops = orc.align_convex(v["ref_part"], seq)
        assert sc == g["final_score"] and orc.cigar(ops) == g["cigar"]


def test_lanes_do_not_change_results(ctx, config1_batch):
    """svp_set_lanes: the batch split into 1, 3 or 6 concurrent parts gives the same bits, also on a rerun"""
    vs = config1_batch
    try:
        ctx.set_lanes(1)
        base = ctx.polish_batch(vs, with_final=True)
        for n in (3, 6, 4):
            ctx.set_lanes(n)
            got = ctx.polish_batch(vs, with_final=True)
            assert all(same(a, b) for a, b in zip(base, got)), "lanes=%d" % n
            ctx.polish_rerun()
            n_launch, _ = ctx.stats(reset=1)
            assert n_launch > 0
    finally:
        ctx.set_lanes(4)


def test_config3_mixed_del_ins_batch(ctx):
    """configs[2] shape (a shard of it): mixed DEL / INS of 50..1000 bp, -k 400, 30x reads; every variant against the
    oracle, plus the properties that do not need it"""
    rng = np.random.default_rng(2024)
    vs = []
    for i in range(16):
        kind = "DEL" if i % 2 == 0 else "INS"
        svlen = int(np.exp(rng.uniform(np.log(50), np.log(1000))))
        vs.append(synth.make_variant(3000 + i, kind, svlen=svlen, k=400, cov=30, cons_err=0.05))
    got = ctx.polish_batch(vs, with_final=True)
    rc.assert_matches_oracle(vs, got)                                      # all 16: 8 DEL + 8 INS, oracle on all host cores
    n_truth = 0
    for v, g in zip(vs, got):
        ref_span = sum(n for o, n in g["cigar"] if o in "MD")
        seq_span = sum(n for o, n in g["cigar"] if o in "MI")
        assert ref_span == len(v["ref_part"]) and seq_span == len(g["polished"])
        # the structural variant itself shows up as the longest gap of the right kind and about the right length
        want = "D" if v["kind"] == "DEL" else "I"
        longest = max((n for o, n in g["cigar"] if o == want), default=0)
        assert abs(longest - v["svlen"]) <= 0.1 * v["svlen"] + 5, (v["kind"], v["svlen"], longest)
        n_truth += int(np.array_equal(g["polished"], v["truth"]))
    assert n_truth >= 0.7 * len(vs)
    # one variant at a time == the batch (different lane split, different arenas)
    for i in (0, 5):
        assert same(ctx.polish_batch([vs[i]], with_final=True)[0], got[i])


def test_vcf_fields_end_to_end(ctx):
    """POS / END / SVLEN / SEQ / QUAL / FILTER from the device CIGAR: candidate records with
    jittered POS and SVLEN go through polish + final alignment on the GPU and the host tail
    (svp_assign_quality / svp_refine_variant / svp_refine_info). Three variants are compared field by field with the
    oracle run from the same inputs; all of them must recover the simulated variant; a candidate of the wrong size is
    'polished away' (FAIL5)."""
    from sviper_b200 import api
    rng = np.random.default_rng(4242)
    begin = 100000                                     # final_record.beginPos (sviper.h:461) of every window here
    vs, cands = [], []
    for i in range(12):
        kind = "DEL" if i % 2 == 0 else "INS"
        svlen = int(np.exp(rng.uniform(np.log(50), np.log(1000))))
        v = synth.make_variant(5000 + i, kind, svlen=svlen, k=400, cov=30, cons_err=0.05)
        true_pos = begin + v["flank"] + v["k"]
        pos = true_pos + int(rng.integers(-10, 11))
        cl = int(round(svlen * rng.uniform(0.9, 1.1)))
        end = pos + cl if kind == "DEL" else pos
        info = "SVTYPE=%s;END=%d;SVLEN=%s%d;RE=12" % (kind, end, "-" if kind == "DEL" else "", cl)
        vs.append(v)
        cands.append(dict(sv_type=api.SV_DEL if kind == "DEL" else api.SV_INS, ref_pos=pos, ref_pos_end=end, sv_length=cl,
                          info=info, filter="PASS"))
    got = ctx.polish_batch(vs, with_final=True)
    for i, (v, c, g) in enumerate(zip(vs, cands, got)):
        out = api.evaluate_final_mapping(c, g["cigar"], g["polished"], begin, v["k"])
        assert out["filter"] == "PASS", (i, v["kind"], v["svlen"], g["cigar"])
        assert out["sv_length"] == v["svlen"]
        # random flanks can share a base or two with the variant's ends: the breakpoint may sit a few bases off
        assert abs(out["ref_pos"] - (begin + v["flank"] + v["k"])) <= 8
        assert out["ref_pos_end"] == (out["ref_pos"] + out["sv_length"] if v["kind"] == "DEL" else out["ref_pos"])
        assert "SVLEN=%s%d" % ("-" if v["kind"] == "DEL" else "", v["svlen"]) in out["info"]
        assert "END=%d" % out["ref_pos_end"] in out["info"]
        if v["kind"] == "INS":
            seq = out["info"].split("SEQ=")[1].split(";")[0]
            assert len(seq) == v["svlen"]
        assert out["quality"] == api.assign_quality(len(g["cigar"]), v["k"]) and 0.0 <= out["quality"] <= 100.0
        if i in (0, 1, 5):                              # DEL, INS, INS through the oracle from the same inputs
            r1p, r2p = rc.packed(v)
            cfg = rc.oracle_cfg(v, r1p, r2p)
            seq_o, _ = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
            _, ops = orc.align_convex(v["ref_part"], seq_o)
            o = orc.evaluate_final_mapping(c["sv_type"], c["ref_pos"], c["ref_pos_end"], c["sv_length"], c["info"],
                                           orc.cigar(ops), orc.dec(seq_o), begin, v["k"])
            assert (o["pos"], o["end"], o["svlen"], o["filter"], o["info"], o["qual"]) == \
                (out["ref_pos"], out["ref_pos_end"], out["sv_length"], out["filter"], out["info"], out["quality"])
        # a candidate three times the size no longer matches the polished sequence: FAIL5, fields untouched
        wrong = dict(c, sv_length=3 * c["sv_length"] + 50)
        bad = api.evaluate_final_mapping(wrong, g["cigar"], g["polished"], begin, v["k"])
        assert bad["filter"] == "FAIL5" and bad["info"] == c["info"] and bad["ref_pos"] == c["ref_pos"]


def test_mixed_strand_rounds_are_exact_and_save_cells(ctx, config1_batch):
    """Rounds >= 2 run the two known strands of a pair as ONE packed task where the strand certificate holds
    (k1_score.cuh, K1Mix): same results as with SVP_NO_MIXED-style regular rounds (here: against the oracle-checked
    results of the other tests through batch equality), and the score kernel computes clearly fewer cells than the
    reference's 2 x |read| x |window| per read and alignment round."""
    ctx.k1_cells(reset=1)
    res = ctx.polish_batch(config1_batch, with_final=True)
    cells = ctx.k1_cells(reset=1)
    algorithmic = 0
    for v, r in zip(config1_batch, res):
        nreads = sum(1 for x in v["reads1"] if len(x)) + sum(1 for x in v["reads2"] if len(x))
        algorithmic += 2 * 150 * len(v["consensus"]) * nreads * r["alignment_rounds"]
    assert 0.5 * algorithmic < cells < 0.9 * algorithmic
    # every variant again as a batch of its own (other batch composition, other lists): identical results
    for v, r in list(zip(config1_batch, res))[:6]:
        one = ctx.polish_batch([v], with_final=True)[0]
        assert same(one, r)
    # A/B: a context whose rounds are all regular (both strands of every read, SVP_NO_MIXED=1)
    import os
    from sviper_b200 import api
    os.environ["SVP_NO_MIXED"] = "1"
    try:
        plain = api.Context(0)
    finally:
        del os.environ["SVP_NO_MIXED"]
    plain.polish_batch(config1_batch, with_final=True)      # sizes the per-round arenas (a round that does not fit is repeated)
    plain.k1_cells(reset=1)
    res0 = plain.polish_batch(config1_batch, with_final=True)
    cells0 = plain.k1_cells(reset=1)
    plain.close()
    assert all(same(a, b) for a, b in zip(res, res0))
    assert abs(cells0 / algorithmic - 1.0) < 0.02
