"""Error behaviour of the C ABI on the GPU (include/sviper_b200.h: SVP_ERR_*). The reference's functions cannot
fail; here a bad argument or an exceeded capacity is reported, never papered over, and the context stays usable."""
import ctypes as C

import numpy as np
import pytest

from tests import refcases as rc

pytestmark = pytest.mark.gpu
synth = rc.synth
SVP_OK, SVP_ERR_ARG, SVP_ERR_CAPACITY, SVP_ERR_UNSUPPORTED = 0, -2, -3, -4


@pytest.fixture(scope="module")
def ctx():
    from sviper_b200 import api
    c = api.Context(0)
    yield c
    c.close()


def small(seed, **kw):
    return synth.make_variant(seed, "DEL", svlen=60, k=150, cov=6, cons_err=0.03, **kw)


def call(ctx, pb):
    return ctx.L.svp_polish_batch(ctx.h, pb["n"], pb["win"], pb["res"])


def test_bad_arguments_are_refused(ctx):
    vs = [small(1)]
    pb = ctx.prepare(vs, True)
    assert ctx.L.svp_polish_batch(None, 1, pb["win"], pb["res"]) == SVP_ERR_ARG
    assert ctx.L.svp_polish_batch(ctx.h, -1, pb["win"], pb["res"]) == SVP_ERR_ARG
    assert ctx.L.svp_polish_batch(ctx.h, 1, None, pb["res"]) == SVP_ERR_ARG
    assert ctx.L.svp_polish_batch(ctx.h, 0, pb["win"], pb["res"]) == SVP_OK          # empty batch: nothing to do
    assert ctx.L.svp_set_lanes(ctx.h, 0) == SVP_ERR_ARG and ctx.L.svp_set_lanes(ctx.h, 1000) == SVP_ERR_ARG
    # per-variant limits are per-variant results: the call succeeds, the variant carries the status
    # window shorter than its two flanks (sviper.h:363 guarantees consensus >= 2 * ref_flank_length)
    bad = dict(small(2)); bad["flank"] = len(bad["consensus"])
    pb2 = ctx.prepare([bad], True)
    assert call(ctx, pb2) == SVP_OK and pb2["res"][0].status == SVP_ERR_ARG and b"flank" in ctx.L.svp_last_error(ctx.h)
    # empty first read of a pair (the dummy is always the MATE, basics.h:317,398)
    bad = small(3); bad["reads1"][0] = np.zeros(0, np.uint8); bad["quals1"][0] = np.zeros(0, np.uint8)
    pb2 = ctx.prepare([bad], True)
    assert call(ctx, pb2) == SVP_OK and pb2["res"][0].status == SVP_ERR_ARG
    # a read longer than SVP_MAX_READ_LEN
    bad = small(4); bad["reads1"][0] = np.zeros(700, np.uint8); bad["quals1"][0] = np.full(700, 30, np.uint8)
    pb2 = ctx.prepare([bad], True)
    assert call(ctx, pb2) == SVP_OK and pb2["res"][0].status == SVP_ERR_UNSUPPORTED
    # the context is still good
    assert call(ctx, pb) == SVP_OK and pb["res"][0].status == SVP_OK


def test_caller_buffer_too_small_is_a_per_variant_status(ctx):
    vs = [small(11), small(12), small(13)]
    want = ctx.polish_batch(vs, True)
    pb = ctx.prepare(vs, True)
    pb["res"][1].polished_cap = 10                       # too small for the polished sequence
    assert call(ctx, pb) == SVP_OK
    assert [pb["res"][i].status for i in range(3)] == [SVP_OK, SVP_ERR_CAPACITY, SVP_OK]
    got = [pb["res"][i] for i in (0, 2)]
    for r, w in zip(got, (want[0], want[2])):
        assert r.rounds == w["rounds"] and r.polished_len == len(w["polished"]) and r.n_cigar == w["n_cigar"]
    # the coverage buffer is optional and may be short: it is filled up to its capacity
    pb = ctx.prepare(vs, True)
    pb["res"][0].cov_cap = 7
    assert call(ctx, pb) == SVP_OK and pb["res"][0].status == SVP_OK and pb["res"][0].cov_len == len(want[0]["cov"])
    assert np.array_equal(np.ctypeslib.as_array(pb["res"][0].cov_profile, (7,)), want[0]["cov"][:7])


def test_one_bad_variant_does_not_fail_its_batch(ctx):
    """an out-of-limit variant must not take the co-batched variants down (the reference polishes the others)"""
    vs = [small(21 + i) for i in range(9)]
    want = ctx.polish_batch(vs, True)
    bad = small(40); bad["reads1"][2] = np.zeros(700, np.uint8); bad["quals1"][2] = np.full(700, 30, np.uint8)
    short = dict(small(41)); short["flank"] = len(short["consensus"])
    mixed = vs[:3] + [bad] + vs[3:7] + [short] + vs[7:]
    pb = ctx.prepare(mixed, True)
    assert call(ctx, pb) == SVP_OK
    st = [pb["res"][i].status for i in range(len(mixed))]
    assert st == [0, 0, 0, SVP_ERR_UNSUPPORTED, 0, 0, 0, 0, SVP_ERR_ARG, 0, 0]
    got = [g for g in ctx.collect(pb, allow_failed=True) if "status" not in g]
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert np.array_equal(g["polished"], w["polished"]) and g["cigar"] == w["cigar"] and g["rounds"] == w["rounds"]
    # the same through the asynchronous scheduler
    pb = ctx.prepare(mixed, True)
    ctx.submit_prepared(pb, tag_base=500)
    tags = ctx.wait_n(len(mixed))
    assert sorted(tags) == list(range(500, 500 + len(mixed))) and ctx.wait_any() is None
    assert [pb["res"][i].status for i in range(len(mixed))] == st
    got = [g for g in ctx.collect(pb, allow_failed=True) if "status" not in g]
    for g, w in zip(got, want):
        assert np.array_equal(g["polished"], w["polished"]) and g["cigar"] == w["cigar"]


def test_rerun_without_a_resident_batch(ctx):
    from sviper_b200 import api
    c = api.Context(0)
    try:
        assert c.L.svp_polish_rerun(c.h) == SVP_ERR_ARG
    finally:
        c.close()
