"""svp_final_align_static (include/sviper_b200.h): the final alignment for callers that must not allocate -- scratch in a
file-scope device array, inputs / outputs as DEVICE pointers of the caller. No counterpart in the reference.
Checked against the existing entry point svp_final_align_batch (which the other GPU tests
pin to the oracle and the reference fixtures) with the bar of those tests: identical fp64 score bits and identical CIGAR.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SVP_OK, SVP_ERR_ARG, SVP_ERR_CAPACITY = 0, -2, -3
MAX_LEN = 4096                                             # SVP_STATIC_MAX_LEN


class StaticResult(C.Structure):
    _fields_ = [("n_cigar", C.c_int32), ("stored", C.c_int32), ("score", C.c_double)]


def make_pair(rng, n, del_len, err):
    """a reference part of n bases and a 'polished' sequence: the reference with a deletion and a few point differences"""
    a = rng.integers(0, 4, n).astype(np.uint8)
    cut = n // 3
    b = np.concatenate([a[:cut], a[cut + del_len:]]).copy()
    m = rng.random(len(b)) < err
    b[m] = (b[m] + 1 + rng.integers(0, 3, int(m.sum()))) % 4
    return a, b


def call_static(ctx, torch, a, b, cap=4096):
    ta = torch.from_numpy(a).cuda(); tb = torch.from_numpy(b).cuda()
    op = torch.full((cap,), 0x7F, dtype=torch.uint8, device="cuda")       # poisoned: the call must write what it reports
    cnt = torch.full((cap,), -1, dtype=torch.int32, device="cuda")
    res = torch.full((16,), 0xEE, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    rc = ctx.L.svp_final_align_static(ctx.h, ta.data_ptr(), len(a), tb.data_ptr(), len(b), op.data_ptr(), cnt.data_ptr(), cap,
                                      res.data_ptr())
    torch.cuda.synchronize()
    if rc != SVP_OK:
        return rc, None
    r = StaticResult.from_buffer_copy(res.cpu().numpy().tobytes())
    k = min(r.n_cigar, r.stored)
    ops = op[:k].cpu().numpy().tobytes().decode()
    cnts = cnt[:k].cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    return rc, dict(score=r.score, n_cigar=r.n_cigar, stored=r.stored, cigar=[(o, int(c)) for o, c in zip(ops, cnts)])


def test_static_final_alignment_matches_the_batch_entry_point():
    import torch
    from sviper_b200 import api
    torch.cuda.set_device(0)
    ctx = api.Context(0, torch.cuda.current_stream().cuda_stream)    # everything is queued on torch's current stream
    try:
        rng = np.random.default_rng(20261105)
        # two calls in a row on the same context; the second is SMALLER, so anything the first left in the shared scratch
        # (trace words, band rows, end cell) would show; then a two-band case (> 2,048 rows) and the largest size
        pairs = [make_pair(rng, 2800, 700, 0.004), make_pair(rng, 1300, 90, 0.01), make_pair(rng, 4000, 300, 0.002),
                 make_pair(rng, MAX_LEN, 10, 0.0)]
        pairs.append((pairs[1][1], pairs[1][0]))                                    # roles swapped: an insertion
        want = ctx.final_align_batch(pairs, cigar_cap=8192)
        for rep in range(2):
            for (a, b), w in zip(pairs, want):
                rc, got = call_static(ctx, torch, a, b)
                assert rc == SVP_OK
                assert np.float64(got["score"]).view(np.uint64) == np.float64(w["score"]).view(np.uint64)
                assert got["n_cigar"] == w["n_cigar"] == got["stored"] and got["cigar"] == w["cigar"]
        assert [o for o, _ in want[0]["cigar"]].count("D") >= 1 and [o for o, _ in want[4]["cigar"]].count("I") >= 1
        # a CIGAR buffer that is too small: the count is still right, only `stored` elements are written
        rc, got = call_static(ctx, torch, *pairs[0], cap=2)
        assert rc == SVP_OK and got["n_cigar"] == want[0]["n_cigar"] > 2 and got["stored"] == 2
        # sizes beyond the static scratch are refused with a clear error, nothing is launched
        big = rng.integers(0, 4, MAX_LEN + 1).astype(np.uint8)
        rc, _ = call_static(ctx, torch, big, pairs[1][1])
        assert rc == SVP_ERR_CAPACITY and b"SVP_STATIC_MAX_LEN" in ctx.L.svp_last_error(ctx.h)
        rc, _ = call_static(ctx, torch, pairs[1][0], big)
        assert rc == SVP_ERR_CAPACITY
        assert ctx.L.svp_final_align_static(ctx.h, None, 5, None, 5, None, None, 1, None) == SVP_ERR_ARG
        # and the context still works afterwards, through both entry points
        rc, got = call_static(ctx, torch, *pairs[1])
        assert rc == SVP_OK and got["cigar"] == want[1]["cigar"]
        again = ctx.final_align_batch(pairs[:2], cigar_cap=8192)
        assert [x["cigar"] for x in again] == [x["cigar"] for x in want[:2]]
    finally:
        ctx.close()


def test_static_entry_point_does_not_allocate_device_memory():
    """free device memory is the same before and after a series of calls (after the one-time set-up of the first call)"""
    import torch
    from sviper_b200 import api
    torch.cuda.set_device(0)
    ctx = api.Context(0, torch.cuda.current_stream().cuda_stream)
    try:
        rng = np.random.default_rng(7)
        a, b = make_pair(rng, 3000, 500, 0.003)
        ta = torch.from_numpy(a).cuda(); tb = torch.from_numpy(b).cuda()
        op = torch.zeros(4096, dtype=torch.uint8, device="cuda"); cnt = torch.zeros(4096, dtype=torch.int32, device="cuda")
        res = torch.zeros(16, dtype=torch.uint8, device="cuda")
        call = lambda: ctx.L.svp_final_align_static(ctx.h, ta.data_ptr(), len(a), tb.data_ptr(), len(b), op.data_ptr(),
                                                    cnt.data_ptr(), 4096, res.data_ptr())
        assert call() == SVP_OK
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        for _ in range(5):
            assert call() == SVP_OK
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 == free0
    finally:
        ctx.close()
