"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the committed
reference fixtures. Bit-exact bar: scores, strand, alignment ops, fp64 profiles, consensus, CIGAR."""
import os

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


@pytest.fixture(scope="module")
def ctx_full():
    """same kernels, but the traced pass covers the whole window (no sub-window restriction)"""
    from sviper_b200 import api
    os.environ["SVP_FULL_WINDOW"] = "1"
    c = api.Context(0)
    del os.environ["SVP_FULL_WINDOW"]
    yield c
    c.close()


def check_align(ctx, reads, wins, tr, tw, tag=""):
    got = ctx.align_batch(reads, wins, tr, tw, traceback=True, ops_cap=1024)
    bad = []
    for t in range(len(tr)):
        rd, win = reads[tr[t]], wins[tw[t]]
        o = orc.map_read(win, rd, np.zeros(len(rd), np.uint8))
        want = (o["score"], o["score_rc"], o["is_rc"], o["lead"], o["end"], o["ops"][o["lead"]:o["end"]])
        have = (int(got["score"][t]), int(got["score_rc"][t]), int(got["is_rc"][t]), int(got["lead"][t]),
                int(got["endpos"][t]), got["ops"][t])
        if want != have:
            bad.append((t, len(rd), len(win), want, have))
    assert not bad, "%s: %d/%d tasks differ, first: %r" % (tag, len(bad), len(tr), bad[0])


def random_tasks(seed, n_win, reads_per_win, read_lens, win_lo, win_hi, err=0.08, n_rate=0.0, junk=0.1):
    rng = np.random.default_rng(seed)
    wins, reads, tr, tw = [], [], [], []
    for w in range(n_win):
        L = int(rng.integers(win_lo, win_hi))
        win = rng.integers(0, 4, L).astype(np.uint8)
        if n_rate:
            win[rng.random(L) < n_rate] = 4
        wins.append(win)
        for _ in range(reads_per_win):
            R = int(read_lens[rng.integers(0, len(read_lens))])
            if rng.random() < junk or L < R + 12:
                rd = rng.integers(0, 4, R).astype(np.uint8)
            else:
                st = int(rng.integers(-20, L - R + 20))
                src = win[max(0, st):max(0, st) + R + 10]
                rd = synth.mutate(rng, src, err * rng.random() * 2)[:R]
                if len(rd) < R:
                    rd = np.concatenate([rd, rng.integers(0, 4, R - len(rd)).astype(np.uint8)])
                if st < 0 or st + R > L:   # hang off the window edge
                    rd = np.concatenate([rng.integers(0, 4, 8).astype(np.uint8), rd])[:R]
            if n_rate:
                rd[rng.random(len(rd)) < n_rate] = 4
            if rng.random() < 0.5:
                rd = synth.revcomp(rd)
            reads.append(np.ascontiguousarray(rd)); tr.append(len(reads) - 1); tw.append(w)
    return reads, wins, np.array(tr, np.int32), np.array(tw, np.int32)


def test_align_golden_fixture(ctx):
    cases = [c for c in rc.read_cases(os.path.join(os.path.dirname(__file__), "golden", "cases.txt")) if c["cmd"] == "ALIGN"]
    reads = [c["read"] for c in cases]; wins = [c["ref"] for c in cases]
    idx = np.arange(len(cases), dtype=np.int32)
    check_align(ctx, reads, wins, idx, idx, "golden ALIGN")


@pytest.mark.parametrize("seed,lens", [(1, [150]), (2, [151, 152, 149, 100, 37, 1, 2]), (3, [153, 160, 200, 256]),
                                       (4, [150, 250])])
def test_align_random(ctx, seed, lens):
    reads, wins, tr, tw = random_tasks(seed, 6, 40, lens, 300, 2200, n_rate=0.01 if seed == 2 else 0.0)
    check_align(ctx, reads, wins, tr, tw, "random seed %d" % seed)


def test_align_short_windows_and_edges(ctx):
    reads, wins, tr, tw = random_tasks(11, 12, 12, [150, 30, 5], 1, 160, junk=0.3)
    check_align(ctx, reads, wins, tr, tw, "short windows")


def test_align_full_window_equals_subwindow(ctx, ctx_full):
    reads, wins, tr, tw = random_tasks(21, 4, 30, [150, 151], 900, 1900, junk=0.25)
    check_align(ctx_full, reads, wins, tr, tw, "full window")
    a = ctx.align_batch(reads, wins, tr, tw, traceback=True, ops_cap=1024)
    b = ctx_full.align_batch(reads, wins, tr, tw, traceback=True, ops_cap=1024)
    assert a["ops"] == b["ops"] and all(np.array_equal(a[k], b[k]) for k in ("score", "score_rc", "lead", "endpos"))


def check_round(ctx, v, fix):
    r1p, r2p = rc.packed(v)
    cfg = rc.oracle_cfg(v, r1p, r2p)
    cfg.fix_indels = fix
    want = orc.polish_round(cfg, *r1p, *r2p, v["consensus"])
    got = ctx.debug_round(v, fix)
    assert got["baseQ_mean"] == cfg.baseQ_mean
    assert np.array_equal(got["proper"], want["proper"])
    assert got["mappQ_mean"] == cfg.mappQ_mean
    assert np.array_equal(got["isgap"], want["isgap"])
    assert np.array_equal(got["cov"], want["cov"])
    assert np.array_equal(got["profile"].view(np.uint64), want["profile"].view(np.uint64))
    assert np.array_equal(got["out"], want["out"])
    assert (got["subst"], got["ins"], got["dele"]) == (cfg.subst, cfg.ins, cfg.del_)


@pytest.mark.parametrize("seed", range(6))
def test_round_small(ctx, seed):
    v = rc.small_variant(seed, n_rate=0.01 if seed == 3 else 0, dummy_last=(seed == 2), err=0.04 + 0.02 * (seed % 4))
    check_round(ctx, v, seed % 2)


def test_round_golden_fixture(ctx):
    cases = [c for c in rc.read_cases(os.path.join(os.path.dirname(__file__), "golden", "cases.txt")) if c["cmd"] == "ROUND"]
    for c in cases:
        check_round(ctx, c["v"], c["fix"])


def test_round_realistic(ctx):
    v = synth.make_variant(5, "DEL", svlen=200, k=400, cov=20, cons_err=0.05)
    check_round(ctx, v, 0)
    check_round(ctx, v, 1)


def check_polish(ctx, vs, with_final=True):
    got = ctx.polish_batch(vs, with_final=with_final)
    for v, g in zip(vs, got):
        r1p, r2p = rc.packed(v)
        cfg = rc.oracle_cfg(v, r1p, r2p)
        seq, cov = orc.polish_to_perfection(cfg, *r1p, *r2p, v["consensus"])
        assert np.array_equal(seq, g["polished"])
        assert (cfg.rounds, cfg.subst, cfg.ins, cfg.del_) == (g["rounds"], g["subst"], g["ins"], g["dele"])
        assert cfg.mappQ_mean == g["mappQ_mean"]
        assert np.array_equal(cov, g["cov"])
        if with_final and v.get("ref_part") is not None:
            sc, ops = orc.align_convex(v["ref_part"], seq)
            assert sc == g["final_score"] and orc.cigar(ops) == g["cigar"]
            assert orc.lib().orc_assign_quality(g["n_cigar"], 400) == __import__("sviper_b200").api.assign_quality(g["n_cigar"], 400)


def test_polish_small_batch(ctx):
    vs = [rc.small_variant(100 + s, err=0.08, dummy_last=(s == 1)) for s in range(6)]
    for v in vs:
        v["ref_part"] = v["truth"]
    check_polish(ctx, vs)


def test_polish_strand_certificate_corner_cases(ctx):
    """Inputs aimed at the strand certificate of the mixed rounds (k1_score.cuh, K1Mix): reverse-complement palindromes
    and near-palindromes (both strands tie or differ by one mismatch: mapping.h:121 must keep choosing the same one),
    reads unrelated to the window (both strands score alike and low), low-complexity reads, and a consensus so noisy
    that the windows change a lot between rounds. Everything against the oracle's literal both-strands restatement."""
    rng = np.random.default_rng(4711)
    vs = []
    for s_ in range(8):
        v = rc.small_variant(900 + s_, L_core=260, flank=40, read_len=60, n_pairs=40, err=[0.05, 0.12, 0.2][s_ % 3])
        R = 60
        for i in range(0, len(v["reads1"]), 5):
            half = rng.integers(0, 4, R // 2).astype(np.uint8)
            pal = np.concatenate([half, synth.revcomp(half)])                   # read == revcomp(read)
            if i % 10 == 0:
                pal[int(rng.integers(0, R))] = (pal[int(rng.integers(0, R))] + 1) % 4   # near-palindrome
            v["reads1"][i] = pal
        for i in range(1, len(v["reads2"]), 7):
            if len(v["reads2"][i]):
                v["reads2"][i] = rng.integers(0, 4, R).astype(np.uint8)         # unrelated mate
        for i in range(2, len(v["reads1"]), 9):
            v["reads1"][i] = np.tile(np.array([0, 3], np.uint8), R // 2)        # ATAT...: its own reverse complement, shifted
        # palindromic stretch inside the window itself: reads on it align equally well on both strands
        t = v["consensus"]
        mid = len(t) // 2
        half = t[mid - 30:mid].copy()
        v["consensus"] = np.concatenate([t[:mid], synth.revcomp(half), t[mid + 30:]])
        v["truth"] = v["consensus"].copy()
        v["ref_part"] = v["truth"]
        vs.append(v)
    check_polish(ctx, vs)


def test_polish_mateless_reads(ctx):
    """batches of mate-less reads (dummy mates, basics.h:317,398): no pair can be mixed, every round stays regular -- a few
    of them (wide shape in rounds >= 2) and more than 3,000 (throughput shape)"""
    def strip_mates(v):
        v["reads2"] = [np.zeros(0, np.uint8) for _ in v["reads2"]]
        v["quals2"] = [np.zeros(0, np.uint8) for _ in v["quals2"]]
        v["ref_part"] = v["truth"]
        return v
    check_polish(ctx, [strip_mates(rc.small_variant(1200 + s_, n_pairs=30, err=0.08)) for s_ in range(3)])
    many = [strip_mates(rc.small_variant(1300 + s_, n_pairs=150, err=0.08)) for s_ in range(22)]
    assert sum(len(v["reads1"]) for v in many) > 3000
    check_polish(ctx, many)


def test_polish_golden_fixture(ctx):
    cases = rc.read_cases(os.path.join(os.path.dirname(__file__), "golden", "cases.txt"))
    with open(os.path.join(os.path.dirname(__file__), "golden", "expected.txt")) as f:
        exp = [rc.parse_line(x) for x in f.read().strip().split("\n")]
    vs = [(c["v"], e) for c, e in zip(cases, exp) if c["cmd"] == "POLISH"]
    got = ctx.polish_batch([v for v, _ in vs], with_final=False)
    for (v, e), g in zip(vs, got):
        assert rc.s(g["polished"]) == e["seq"]
        assert (g["rounds"], g["subst"], g["ins"], g["dele"]) == tuple(int(e[k]) for k in ("rounds", "subst", "ins", "del"))
        assert g["mappQ_mean"] == float.fromhex(e["mappQ_mean"])
        assert np.array_equal(g["cov"], rc.intlist(e["cov"]))


def test_polish_realistic_del_ins(ctx):
    vs = [synth.make_variant(31, "DEL", svlen=300, k=400, cov=12, cons_err=0.04),
          synth.make_variant(32, "INS", svlen=120, k=400, cov=12, cons_err=0.04)]
    check_polish(ctx, vs)
    # the polished windows should be (near) the truth: sanity that the workload is meaningful
    got = ctx.polish_batch(vs, with_final=True)
    assert [c[0] for c in got[0]["cigar"]].count("D") >= 1 and [c[0] for c in got[1]["cigar"]].count("I") >= 1


def test_final_golden_and_random(ctx):
    cases = rc.read_cases(os.path.join(os.path.dirname(__file__), "golden", "cases.txt"))
    with open(os.path.join(os.path.dirname(__file__), "golden", "expected.txt")) as f:
        exp = [rc.parse_line(x) for x in f.read().strip().split("\n")]
    fin = [(c, e) for c, e in zip(cases, exp) if c["cmd"] == "FINAL"]
    got = ctx.final_align_batch([(c["a"], c["b"]) for c, _ in fin])
    for (c, e), g in zip(fin, got):
        assert g["score"] == float.fromhex(e["score"])
        assert "".join("%d%s" % (n, o) for o, n in g["cigar"]) == e["cigar"] and g["n_cigar"] == int(e["ncigar"])
    rng = np.random.default_rng(3)
    pairs = []
    for i in range(6):
        a = rng.integers(0, 4, int(rng.integers(1500, 5000))).astype(np.uint8)
        cut = len(a) // 3
        b = synth.mutate(rng, np.concatenate([a[:cut], a[cut + int(rng.integers(20, 700)):]]), 0.002 * i)
        pairs.append((a, b) if i % 2 == 0 else (b, a))
    got = ctx.final_align_batch(pairs)
    for (a, b), g in zip(pairs, got):
        sc, ops = orc.align_convex(a, b)
        assert sc == g["score"] and orc.cigar(ops) == g["cigar"]


def test_async_scheduler(ctx):
    """svp_submit_polish / svp_flush / svp_wait_any (asynchronous submissions with an explicit wait) return the
    same results as one svp_polish_batch, whatever device batches the lanes form: several submissions in flight,
    small batch targets, waiting threads"""
    import threading
    vs = [rc.small_variant(300 + s) for s in range(23)]
    for v in vs:
        v["ref_part"] = v["truth"]
    want = ctx.polish_batch(vs, with_final=True)
    assert ctx.wait_any() is None                          # nothing outstanding
    # (a) everything queued, one flush
    pb = ctx.prepare(vs, True)
    ctx.submit_prepared(pb, tag_base=1000)
    assert sorted(ctx.wait_n(len(vs))) == list(range(1000, 1000 + len(vs))) and ctx.wait_any() is None
    for g, w in zip(ctx.collect(pb), want):
        assert np.array_equal(g["polished"], w["polished"]) and g["cigar"] == w["cigar"] and g["rounds"] == w["rounds"]
        assert np.array_equal(g["cov"], w["cov"]) and g["final_score"] == w["final_score"]
    # (b) a small batch target: the lanes start while the submitter is still queueing; a second thread waits
    ctx.engine_config(batch_target=3)
    before = ctx.engine_stats()
    pb = ctx.prepare(vs, True)
    tags = []
    th = threading.Thread(target=lambda: tags.extend(ctx.wait_n(len(vs))))
    ctx.submit_prepared(pb, tag_base=0, flush=False)
    th.start()                                             # the waiter flushes what is left
    th.join(timeout=120)
    assert not th.is_alive() and sorted(tags) == list(range(len(vs)))
    st = ctx.engine_stats()
    assert st["variants"] - before["variants"] == len(vs) and st["batches"] - before["batches"] >= 2
    for g, w in zip(ctx.collect(pb), want):
        assert np.array_equal(g["polished"], w["polished"]) and g["cigar"] == w["cigar"]
    ctx.engine_config(batch_target=256)


def test_round_loop_variants_agree(ctx, monkeypatch):
    """The round loop sizes the per-round arenas from an
    estimate. Neither may change a result: (a) SVP_SPECULATIVE_ROUNDS=1 queues round r + 1 before round r's outcome is known; (b) SVP_SMALL_ARENAS=1 starts from
    arenas that are far too small, so the first round of every batch overflows, leaves the state untouched and is repeated
    after the arenas have grown (RoundCtl::ovf); (c) 64- and 128-thread blocks of the score kernel."""
    from sviper_b200 import api
    vs = [rc.small_variant(900 + s, n_pairs=30 + 7 * s) for s in range(9)]
    for v in vs:
        v["ref_part"] = v["truth"]
    want = ctx.polish_batch(vs, with_final=True)

    def same(got):
        for g, w in zip(got, want):
            assert np.array_equal(g["polished"], w["polished"]) and g["cigar"] == w["cigar"] and g["rounds"] == w["rounds"]
            assert np.array_equal(g["cov"], w["cov"]) and (g["subst"], g["ins"], g["dele"]) == (w["subst"], w["ins"], w["dele"])

    monkeypatch.setenv("SVP_SPECULATIVE_ROUNDS", "1")
    c = api.Context(0)
    same(c.polish_batch(vs, with_final=True))
    monkeypatch.setenv("SVP_SMALL_ARENAS", "1")
    same(c.polish_batch(vs, with_final=True))
    monkeypatch.delenv("SVP_SMALL_ARENAS")
    c.close()
    monkeypatch.delenv("SVP_SPECULATIVE_ROUNDS")
    for lanes in (1, 2):
        c = api.Context(0)
        c.set_lanes(lanes)
        monkeypatch.setenv("SVP_SMALL_ARENAS", "1")
        same(c.polish_batch(vs, with_final=True))
        r = c.align_batch([v["reads1"][0] for v in vs], [v["consensus"] for v in vs], np.arange(9), np.arange(9), ops_cap=700)
        monkeypatch.delenv("SVP_SMALL_ARENAS")
        r2 = ctx.align_batch([v["reads1"][0] for v in vs], [v["consensus"] for v in vs], np.arange(9), np.arange(9), ops_cap=700)
        assert r["ops"] == r2["ops"] and np.array_equal(r["lead"], r2["lead"])
        c.close()
    for kt in ("64", "128"):
        monkeypatch.setenv("SVP_K1_THREADS", kt)
        c = api.Context(0)
        same(c.polish_batch(vs, with_final=True))
        c.close()
    monkeypatch.delenv("SVP_K1_THREADS")
