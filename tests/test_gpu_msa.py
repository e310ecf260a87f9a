"""the pairwise DP fills of build_consensus on the device (svp_msa_*; k6_msa.cuh) against the oracle
(oracle/oracle_msa.c, pinned to the reference by tests/test_msa_golden.py) and against the committed reference fixtures:
scores, end cells and COMPLETE trace matrices, byte for byte. Plus the whole build_consensus through the host side that a
SViper build uses (sviper_b200/host/msa_harness: device fills + SeqAn's unchanged traceback / graph / progressive alignment)
against the consensus strings the unmodified reference printed."""
import os
import subprocess

import numpy as np
import pytest

from tests import refcases as rc
from tests.test_msa_golden import load

pytestmark = pytest.mark.gpu
orc, synth = rc.orc, rc.synth
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MSA_HARNESS = os.path.join(ROOT, "sviper_b200", "host", "msa_harness")
G = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def ctx():
    from sviper_b200 import api
    c = api.Context(0)
    yield c
    c.close()


def check_global(got, h, v):
    trace, res = got
    want, sc, (S, H, V) = orc.msa_global_fill(h, v)
    assert (res["score"], res["S"], res["H"], res["V"]) == (sc, S, H, V)
    assert np.array_equal(trace, want)


def check_local(got, h, v, cells):
    trace, res = got
    want, sc, row, col, d = orc.msa_sw_fill(h, v, cells)
    assert (res["score"], res["row"], res["col"]) == (sc, row, col)
    if sc > 0:
        assert res["dir"] == d
    assert np.array_equal(trace, want)


def test_fills_match_the_reference_fixtures(ctx):
    cases, exp = load()
    gf = [(c, e) for c, e in zip(cases, exp) if c[0] == "GFILL"]
    sw = [(c, e) for c, e in zip(cases, exp) if c[0] == "SWFILL"]
    ctx.msa_begin([(c[1], c[2]) for c, _ in gf])
    for (c, e), (trace, res) in zip(gf, ctx.msa_global_fill(list(range(len(gf))))):
        assert res["score"] == int(e["score"]) and (res["S"], res["H"], res["V"]) == (int(e["S"]), int(e["H"]), int(e["V"]))
        assert len(trace) == int(e["tracelen"]) and orc.fnv1a(trace) == int(e["hash"])
    ctx.msa_begin([(c[1], c[2]) for c, _ in sw])
    for (c, e), (trace, res) in zip(sw, ctx.msa_local_fill(list(range(len(sw))), [c[3] for c, _ in sw])):
        assert (res["score"], res["row"], res["col"]) == (int(e["score"]), int(e["row"]), int(e["col"]))
        assert len(trace) == int(e["tracelen"]) and orc.fnv1a(trace) == int(e["hash"])


def test_fills_match_the_oracle_on_random_and_edge_shapes(ctx):
    rng = np.random.default_rng(606)
    pairs = [(np.array([1], np.uint8), np.array([1], np.uint8)), (np.array([1], np.uint8), np.array([2, 1, 1], np.uint8)),
             (rng.integers(0, 4, 9).astype(np.uint8), rng.integers(0, 4, 8).astype(np.uint8))]
    for n in (7, 8, 255, 256, 257, 2047, 2048, 2049, 2600):            # thread, block and band boundaries (2,048 rows per band)
        a = rng.integers(0, 4, n).astype(np.uint8)
        pairs.append((synth.mutate(rng, a, 0.1), a))                   # rows = n
        pairs.append((a[: max(1, n // 3)], synth.mutate(rng, a, 0.12)))
    a = rng.integers(0, 5, 700).astype(np.uint8)                       # N bases: N == N scores as a match
    pairs.append((a, synth.mutate(rng, a, 0.05)))
    rep = rng.integers(0, 4, 150).astype(np.uint8)
    pairs.append((np.concatenate([rep, rep, rep]), np.concatenate([rep, rng.integers(0, 4, 40).astype(np.uint8), rep])))
    ctx.msa_begin(pairs)
    idx = list(range(len(pairs)))
    for got, (h, v) in zip(ctx.msa_global_fill(idx), pairs):
        check_global(got, h, v)
    # local: four rounds as the library does them; the cells of a diagonal run through the best end cell are forbidden next
    cells = [[] for _ in pairs]
    new = [[] for _ in pairs]
    for rnd in range(4):
        got = ctx.msa_local_fill(idx, new)
        new = []
        for k, (g, (h, v)) in enumerate(zip(got, pairs)):
            check_local(g, h, v, cells[k])
            r, c, add = g[1]["row"], g[1]["col"], []
            while r >= 1 and c >= 1 and len(add) < 60 + 7 * rnd:
                add.append((c, r)); r -= 1; c -= 1
            cells[k] = cells[k] + add
            new.append(add)
    # a subset of the pairs, in another order, after the bitmaps have been used (with the cells of the last round still to send)
    sub = [5, 2, len(pairs) - 1]
    for g, k in zip(ctx.msa_local_fill(sub, [new[k] for k in sub]), sub):
        check_local(g, pairs[k][0], pairs[k][1], cells[k])


def test_fill_arguments_are_checked(ctx):
    from sviper_b200 import api
    import ctypes as C
    a = np.array([0, 1, 2], np.uint8)
    bad = (api.MsaPair * 1)(api.MsaPair(api._p(a), 3, api._p(a), 0))
    assert ctx.L.svp_msa_begin(ctx.h, 1, bad, 5, -3, -1, -3) == -2
    ctx.msa_begin([(a, a)])
    tp = (api.u8p * 1)(api._p(np.zeros(16, np.uint8)))
    res = (api.MsaFillResult * 1)()
    assert ctx.L.svp_msa_local_fill(ctx.h, 1, api._p(np.array([3], np.int32), api.i32p), None, None, tp, res) == -2   # no such pair
    with pytest.raises(RuntimeError):
        ctx.msa_local_fill([0], [[(9, 1)]])
    trace, res = ctx.msa_local_fill([0])[0]
    assert res["score"] == 15 and (res["row"], res["col"]) == (3, 3)


def test_build_consensus_through_the_host_side_matches_the_reference():
    """every MSA fixture (110 sets of 1-5 long-read crops, 120-1,650 bp, 2-15 % error, N bases, half-length crops) through
    sviper_b200/host/msa_harness: the consensus string must equal what the unmodified reference's build_consensus returned"""
    assert os.path.exists(MSA_HARNESS), "build() must have produced sviper_b200/host/msa_harness"
    out = subprocess.check_output([MSA_HARNESS, os.path.join(G, "msa_cases.txt")], text=True).strip().split("\n")
    exp = [x for x in open(os.path.join(G, "msa_expected.txt")).read().strip().split("\n") if x.startswith("MSA ")]
    assert len(out) == len(exp) == 110
    assert out == exp
