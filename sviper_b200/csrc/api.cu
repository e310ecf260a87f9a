// C ABI of sviper_b200 (see include/sviper_b200.h). Host orchestration only: every base that is aligned,
// piled up, voted on or traced back is processed by the sm_100a kernels in this directory.
// There is deliberately no CPU implementation of any of it here (no fallback).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <mutex>
#include <thread>
#include <cfloat>
#include <condition_variable>
#include <deque>
#include <cub/device/device_scan.cuh>
#include <nvtx3/nvToolsExt.h>
#include "ctx.cuh"
#include "k1_score.cuh"

static svp_ctx *make_ctx(int cuda_device, void *cuda_stream, bool is_lane);

#include "api_align.inc"
#include "api_polish.inc"
#include "api_final.inc"
#include "api_msa.inc"
#include "api_eval.inc"
#include "api_prepare.inc"
#include "api_merge.inc"

extern "C" {

const char *svp_version(void) { return "sviper_b200 0.1 (sm_100a)"; }

}  // extern "C"

// Dynamic shared memory opt-in of the kernels that need more than 48 KB: once per device (the attribute is
// per function and device; setting it per call would race between the lanes of a split batch).
static int init_kernel_attrs(int device)
{
    static std::mutex mu;
    static uint64_t done = 0;
    std::lock_guard<std::mutex> lk(mu);
    if (device < 64 && (done >> device & 1)) return SVP_OK;
    int optin = 0;
    if (cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device) != cudaSuccess) return SVP_ERR_CUDA;
    auto set_max = [&](const void *fn, int want) {
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, fn) != cudaSuccess) return false;
        int room = optin - (int)fa.sharedSizeBytes;
        return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, want < 0 ? room : std::min(want, room)) ==
               cudaSuccess;
    };
    bool ok = set_max((const void *)svp::k1_score_kernel<svp::K1_G, 19, false, 128>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 32, false, 128>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_GW, 5, false, 128>, svp::K1Shape<svp::K1_GW, 5>::SMEM_BYTES) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_GW, 8, false, 128>, svp::K1Shape<svp::K1_GW, 8>::SMEM_BYTES) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 19, true, 128>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 32, true, 128>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_GW, 5, true, 128>, svp::K1Shape<svp::K1_GW, 5>::SMEM_BYTES) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_GW, 8, true, 128>, svp::K1Shape<svp::K1_GW, 8>::SMEM_BYTES) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 19, false, 64>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 32, false, 64>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 19, true, 64>, -1) &&
              set_max((const void *)svp::k1_score_kernel<svp::K1_G, 32, true, 64>, -1) &&
              set_max((const void *)svp::k3_pileup_vote_kernel, -1) &&
              set_max((const void *)svp::k5_final_kernel<true>, -1);
    if (!ok) return SVP_ERR_CUDA;
    if (device < 64) done |= (uint64_t)1 << device;
    return SVP_OK;
}

// is_lane: a sub-context of a split polish batch (own stream and arenas, no lanes of its own)
static svp_ctx *make_ctx(int cuda_device, void *cuda_stream, bool is_lane)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || cuda_device < 0 || cuda_device >= ndev)
    {
        fprintf(stderr, "[sviper_b200] no usable CUDA device %d (found %d) -- this library has no CPU path\n",
                cuda_device, ndev);
        return nullptr;
    }
    if (cudaSetDevice(cuda_device) != cudaSuccess) return nullptr;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cuda_device) != cudaSuccess) return nullptr;
    if (prop.major < 10)
    {
        fprintf(stderr, "[sviper_b200] device %d is sm_%d%d; this build targets sm_100a only\n", cuda_device,
                prop.major, prop.minor);
        return nullptr;
    }
    if (init_kernel_attrs(cuda_device) != SVP_OK)
    {
        fprintf(stderr, "[sviper_b200] cannot opt in to large shared memory on device %d\n", cuda_device);
        return nullptr;
    }
    svp_ctx *ctx = new svp_ctx();
    ctx->device = cuda_device;
    ctx->is_lane = is_lane;
    if (cuda_stream) ctx->stream = (cudaStream_t)cuda_stream;
    else
    {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return nullptr; }
        ctx->own_stream = true;
    }
    cudaEventCreate(&ctx->ev0);
    cudaEventCreate(&ctx->ev1);
    cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ctx->ev_call, cudaEventDisableTiming);
    for (int c = 0; c < svp::TC_N; ++c)
    {
        if (cudaStreamCreateWithFlags(&ctx->side[c], cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return nullptr; }
        cudaEventCreateWithFlags(&ctx->ev_side[c], cudaEventDisableTiming);
    }
    if (cudaHostAlloc((void **)&ctx->h_ctl, 2 * sizeof(svp::RoundCtl), cudaHostAllocDefault) != cudaSuccess) { delete ctx; return nullptr; }
    memset(ctx->h_ctl, 0, 2 * sizeof(svp::RoundCtl));
    // how host threads wait for the device (svp_wait_event, ctx.cuh)
    if (const char *sm = getenv("SVP_SYNC"))
        ctx->sync_mode = !strcmp(sm, "spin") ? SVP_SYNC_SPIN : !strcmp(sm, "block") ? SVP_SYNC_BLOCK : SVP_SYNC_YIELD;
    const unsigned evf = cudaEventDisableTiming | (ctx->sync_mode == SVP_SYNC_BLOCK ? cudaEventBlockingSync : 0);
    cudaEventCreateWithFlags(&ctx->ev_sync, evf);
    for (int i = 0; i < 2; ++i) cudaEventCreateWithFlags(&ctx->ev_ctl[i], evf);
    if (const char *sr = getenv("SVP_SPECULATIVE_ROUNDS")) ctx->sync_rounds = atoi(sr) != 1;
    if (const char *kt = getenv("SVP_K1_THREADS")) ctx->k1_threads = atoi(kt);
    if (const char *np_ = getenv("SVP_NO_PREDICT")) ctx->predict_certificates = atoi(np_) != 1;
    if (const char *mr = getenv("SVP_MIX_REGULAR")) ctx->mix_reg_wide = std::string(mr) == "wide";
    if (const char *kg = getenv("SVP_K5_GLOBAL_TABLE")) ctx->k5_global_table = atoi(kg) == 1;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) ctx->arena_budget = (size_t)(free_b * 0.6);
    const char *fw = getenv("SVP_FULL_WINDOW");
    if (fw && fw[0] == '1') ctx->full_window = 1;
    const char *nm = getenv("SVP_NO_MIXED");
    if (nm && atoi(nm) == 1) ctx->mixed_rounds = false;
    const char *ln = getenv("SVP_LANES");
    if (ln && atoi(ln) >= 1) ctx->n_lanes = std::min(atoi(ln), SVP_MAX_LANES);
    if (is_lane) ctx->n_lanes = 1;
    return ctx;
}

extern "C" {

svp_ctx *svp_create(int cuda_device, void *cuda_stream) { return make_ctx(cuda_device, cuda_stream, false); }

int svp_set_lanes(svp_ctx *ctx, int32_t n_lanes)
{
    if (!ctx || n_lanes < 1 || n_lanes > SVP_MAX_LANES) return SVP_ERR_ARG;
    ctx->n_lanes = n_lanes;
    return SVP_OK;
}

void svp_destroy(svp_ctx *ctx)
{
    if (!ctx) return;
    engine_stop(ctx);
    if (cudaSetDevice(ctx->device) != cudaErrorCudartUnloading) msa_release(ctx);
    if (cudaSetDevice(ctx->device) == cudaErrorCudartUnloading)
    {
        // process exit: the CUDA runtime is already gone (a context held by a static object); nothing left to free
        for (svp_ctx *l : ctx->lanes) delete l;
        delete ctx->resident;
        delete ctx;
        return;
    }
    cudaStreamSynchronize(ctx->stream);
    for (svp_ctx *l : ctx->lanes) svp_destroy(l);
    ctx->lanes.clear();
    DevBuf *bufs[] = {&ctx->reads.bytes, &ctx->reads.w2, &ctx->reads.wn, &ctx->reads.off32, &ctx->reads.len,
                      &ctx->reads.unit_seq, &ctx->reads.byte_off, &ctx->wins.bytes, &ctx->wins.w2, &ctx->wins.wn,
                      &ctx->wins.off32, &ctx->wins.len, &ctx->wins.unit_seq, &ctx->wins.byte_off, &ctx->read_quals,
                      &ctx->qual_off, &ctx->task_read, &ctx->task_win, &ctx->task_mate, &ctx->win_ins, &ctx->task_idx19, &ctx->task_idx32,
                      &ctx->score_out, &ctx->plans, &ctx->units, &ctx->E, &ctx->tunits, &ctx->TE, &ctx->cls_count, &ctx->cls_list, &ctx->scan_tmp, &ctx->trace, &ctx->ops,
                      &ctx->alns, &ctx->rec_code, &ctx->rec_qual, &ctx->rec_inslen, &ctx->rec_insq, &ctx->descs,
                      &ctx->states, &ctx->pair_task, &ctx->cons0, &ctx->cons1, &ctx->inslen, &ctx->voff, &ctx->profile,
                      &ctx->cov, &ctx->isgap, &ctx->decision, &ctx->proper, &ctx->fjobs, &ctx->fref, &ctx->fseq,
                      &ctx->ftrace, &ctx->fbndS, &ctx->fbndV, &ctx->fbndVL, &ctx->fends, &ctx->fcig_op, &ctx->fcig_cnt,
                      &ctx->fcig_out, &ctx->gapstep, &ctx->edge};
    for (DevBuf *b : bufs) b->release();
    ctx->cons_init.release(); ctx->states_init.release();
    ctx->win_done.release(); ctx->ctl.release(); ctx->gather_desc.release(); ctx->gather_out.release();
    ctx->fail19.release(); ctx->fail32.release(); ctx->sstate.release(); ctx->reg19.release(); ctx->reg32.release();
    ctx->mix19.release(); ctx->mix32.release(); ctx->mixcnt.release(); ctx->edit.release(); ctx->emax.release();
    for (int i = 0; i < 2; ++i) if (ctx->ev_ctl[i]) cudaEventDestroy(ctx->ev_ctl[i]);
    if (ctx->ev_sync) cudaEventDestroy(ctx->ev_sync);
    ctx->stage.release();
    if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
    delete ctx->resident;
    for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaEventDestroy(ctx->ev_fork);
    cudaEventDestroy(ctx->ev_call);
    for (int c = 0; c < svp::TC_N; ++c) { cudaEventDestroy(ctx->ev_side[c]); cudaStreamDestroy(ctx->side[c]); }
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *svp_last_error(const svp_ctx *ctx) { return ctx ? ctx->err.c_str() : "no context"; }

int svp_align_batch(svp_ctx *ctx, const svp_align_in *in, svp_align_out *out, int with_traceback)
{
    if (!ctx || !in || !out) return SVP_ERR_ARG;
    int rc = begin_call(ctx);
    if (rc) return rc;
    std::vector<int64_t> roff(in->n_reads + 1), woff(in->n_windows + 1);
    std::vector<int32_t> rlen(in->n_reads);
    for (int i = 0; i <= in->n_reads; ++i) roff[i] = in->read_off[i];
    for (int i = 0; i <= in->n_windows; ++i) woff[i] = in->win_off[i];
    for (int i = 0; i < in->n_reads; ++i) rlen[i] = in->read_off[i + 1] - in->read_off[i];
    for (int t = 0; t < in->n_tasks; ++t)
        if (in->task_read[t] < 0 || in->task_read[t] >= in->n_reads || in->task_win[t] < 0 ||
            in->task_win[t] >= in->n_windows || in->win_off[in->task_win[t] + 1] - in->win_off[in->task_win[t]] < 1)
            return svp_fail_msg(ctx, SVP_ERR_ARG, "task refers to a missing read/window");
    rc = upload_seqs(ctx, ctx->reads, in->n_reads, in->read_bases, roff.data(), nullptr, true);
    if (rc) return rc;
    rc = upload_seqs(ctx, ctx->wins, in->n_windows, in->win_bases, woff.data(), nullptr, true);
    if (rc) return rc;
    rc = upload_tasks(ctx, in->n_tasks, in->task_read, in->task_win, rlen.data());
    if (rc) return rc;
    ctx->win_bytes = ctx->wins.bytes.as<uint8_t>();
    rc = run_score_pass(ctx, nullptr);
    if (rc) return rc;
    const int n = in->n_tasks;
    if (n == 0) return SVP_OK;
    STAGE(sc, svp::ScoreOut, n);
    CK(cudaMemcpyAsync(sc, ctx->score_out.p, (size_t)n * sizeof(svp::ScoreOut), cudaMemcpyDeviceToHost, ctx->stream));
    svp::AlnOut *al = nullptr;
    uint8_t *ops = nullptr;
    int64_t *E = nullptr;
    if (with_traceback)
    {
        rc = run_trace_pass_checked(ctx, nullptr, false);
        if (rc) return rc;
        al = ctx->stage.get<svp::AlnOut>(n);
        if (!al) return svp_fail_msg(ctx, SVP_ERR_CUDA, "cannot allocate pinned staging memory");
        CK(cudaMemcpyAsync(al, ctx->alns.p, (size_t)n * sizeof(svp::AlnOut), cudaMemcpyDeviceToHost, ctx->stream));
        if (out->ops)
        {
            size_t nops = (size_t)ctx->E_total + (size_t)n * SVP_MAX_READ_LEN;
            E = ctx->stage.get<int64_t>(n + 1);
            ops = ctx->stage.get<uint8_t>(nops);
            if (!E || !ops) return svp_fail_msg(ctx, SVP_ERR_CUDA, "cannot allocate pinned staging memory");
            CK(cudaMemcpyAsync(E, ctx->E.p, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaMemcpyAsync(ops, ctx->ops.p, nops, cudaMemcpyDeviceToHost, ctx->stream));
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    for (int t = 0; t < n; ++t)
    {
        if (out->score) out->score[t] = sc[t].score_f;
        if (out->score_rc) out->score_rc[t] = sc[t].score_r;
        if (out->is_rc) out->is_rc[t] = !(sc[t].score_f > sc[t].score_r);
        if (!with_traceback) continue;
        if (al[t].err) return svp_fail_msg(ctx, SVP_ERR_CUDA, "traceback invariant violated (err " + std::to_string(al[t].err) + ")");
        if (out->lead) out->lead[t] = al[t].lead;
        if (out->endpos) out->endpos[t] = al[t].endpos;
        if (out->ops)
        {
            // view columns [lead, endpos): leading 'I's, then the path (stored reversed) without its trailing 'D's
            const uint8_t *p = ops + E[t] + (int64_t)t * SVP_MAX_READ_LEN;
            int len = al[t].endpos - al[t].lead, k = 0;
            char *dst = out->ops + (size_t)t * out->ops_cap;
            for (int i = 0; i < al[t].lead_i && k < len; ++i, ++k) if (k < out->ops_cap) dst[k] = 'I';
            for (int i = al[t].npath - 1; i >= 0 && k < len; --i, ++k) if (k < out->ops_cap) dst[k] = (char)p[i];
            if (out->ops_len) out->ops_len[t] = len;
        }
    }
    return SVP_OK;
}

int svp_align_rerun(svp_ctx *ctx, int with_traceback, float *kernel_ms)
{
    if (!ctx || ctx->ntasks == 0) return SVP_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    bool t = ctx->timing; ctx->timing = false;
    int rc = run_score_pass(ctx, nullptr);
    ctx->timing = t;
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    if (with_traceback) { rc = run_trace_pass_checked(ctx, nullptr, false); if (rc) return rc; }
    CK(cudaStreamSynchronize(ctx->stream));
    if (kernel_ms) CK(cudaEventElapsedTime(kernel_ms, ctx->ev0, ctx->ev1));
    return SVP_OK;
}

int svp_probe_dpx(svp_ctx *ctx, double *viaddmax_lane_ops, double *vimax3_lane_ops, double *mix_lane_ops)
{
    if (!ctx) return SVP_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    DevBuf o; CK(o.ensure((size_t)blocks * threads * 4));
    double *outs[3] = {viaddmax_lane_ops, vimax3_lane_ops, mix_lane_ops};
    const double ops_per_iter[3] = {8.0, 8.0, 8.0 * 5.0};
    for (int mode = 0; mode < 3; ++mode)
    {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep)
        {
            CK(cudaEventRecord(ctx->ev0, ctx->stream));
            if (mode == 0) svp::dpx_probe_kernel<0><<<blocks, threads, 0, ctx->stream>>>(o.as<uint32_t>(), iters);
            else if (mode == 1) svp::dpx_probe_kernel<1><<<blocks, threads, 0, ctx->stream>>>(o.as<uint32_t>(), iters);
            else svp::dpx_probe_kernel<2><<<blocks, threads, 0, ctx->stream>>>(o.as<uint32_t>(), iters);
            CK(cudaGetLastError());
            CK(cudaEventRecord(ctx->ev1, ctx->stream));
            CK(cudaEventSynchronize(ctx->ev1));
            float ms = 0; CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
            if (rep > 0 && ms < best) best = ms;
        }
        if (outs[mode]) *outs[mode] = (double)blocks * threads * iters * ops_per_iter[mode] / (best * 1e-3);
    }
    o.release();
    return SVP_OK;
}

int svp_probe_fp64(svp_ctx *ctx, double *dadd_lane_ops)
{
    if (!ctx || !dadd_lane_ops) return SVP_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    const int blocks = 148 * 8, threads = 256, iters = 2048;
    DevBuf o; CK(o.ensure((size_t)blocks * threads * 8));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        CK(cudaEventRecord(ctx->ev0, ctx->stream));
        svp::dadd_probe_kernel<<<blocks, threads, 0, ctx->stream>>>(o.as<double>(), iters, 1.0000001);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ctx->ev1, ctx->stream));
        CK(cudaEventSynchronize(ctx->ev1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        if (rep > 0 && ms < best) best = ms;
    }
    *dadd_lane_ops = (double)blocks * threads * iters * 8.0 / (best * 1e-3);
    o.release();
    return SVP_OK;
}

int svp_stats(svp_ctx *ctx, int32_t *n_launches, double *ms_by_kernel, int reset)
{
    if (!ctx) return SVP_ERR_ARG;
    // the lanes of a split polish batch keep their own counters: report the sums
    std::vector<svp_ctx *> all(1, ctx);
    all.insert(all.end(), ctx->lanes.begin(), ctx->lanes.end());
    if (n_launches) *n_launches = 0;
    if (ms_by_kernel) for (int i = 0; i < ST_N; ++i) ms_by_kernel[i] = 0;
    for (svp_ctx *c : all)
    {
        resolve_timers(c);
        if (n_launches) *n_launches += c->n_launches;
        if (ms_by_kernel) for (int i = 0; i < ST_N; ++i) ms_by_kernel[i] += c->ms[i];
        if (reset == 1) { c->n_launches = 0; for (int i = 0; i < ST_N; ++i) c->ms[i] = 0; }
        if (reset == 2) c->timing = true;    // enable per-kernel CUDA-event timing (events resolved lazily, no extra syncs)
        if (reset == 3) c->timing = false;
    }
    return SVP_OK;
}

int svp_k1_cells(svp_ctx *ctx, double *cells, int reset)
{
    if (!ctx || !cells) return SVP_ERR_ARG;
    std::vector<svp_ctx *> all(1, ctx);
    all.insert(all.end(), ctx->lanes.begin(), ctx->lanes.end());
    cells[0] = cells[1] = 0;
    for (svp_ctx *c : all)
    {
        if (!c->mixcnt.p || !c->mixcnt_zeroed) continue;
        unsigned long long v[2] = {0, 0};
        CK(cudaStreamSynchronize(c->stream));
        CK(cudaMemcpy(v, c->mixcnt.as<char>() + 32, 16, cudaMemcpyDeviceToHost));
        cells[0] += (double)v[0]; cells[1] += (double)v[1];
        if (reset) CK(cudaMemset(c->mixcnt.as<char>() + 32, 0, 16));
    }
    return SVP_OK;
}

int svp_k1_failed_certificates(svp_ctx *ctx, double *reads, int reset)
{
    if (!ctx || !reads) return SVP_ERR_ARG;
    std::vector<svp_ctx *> all(1, ctx);
    all.insert(all.end(), ctx->lanes.begin(), ctx->lanes.end());
    *reads = 0;
    for (svp_ctx *c : all)
    {
        if (!c->mixcnt.p || !c->mixcnt_zeroed) continue;
        unsigned long long v = 0;
        CK(cudaStreamSynchronize(c->stream));
        CK(cudaMemcpy(&v, c->mixcnt.as<char>() + 56, 8, cudaMemcpyDeviceToHost));
        *reads += (double)v;
        if (reset) CK(cudaMemset(c->mixcnt.as<char>() + 56, 0, 8));
    }
    return SVP_OK;
}

double svp_assign_quality(int32_t n_cigar, int32_t flanking_region)
{
    double error_rate = ((double)n_cigar - 1.0) / (flanking_region * 2.0);
    double fuzzyness = (1.0 - error_rate / 0.15) * 100.0;
    return std::max(fuzzyness, 0.0);
}

}  // extern "C"

// =====================================================================================================================
// svp_final_align_static: the final alignment (K5 + traceback) for callers that must not allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Same kernels as svp_final_align_batch, but every byte of scratch -- gap-step / edge tables, padded
// copies of the two sequences, the job descriptor, the band boundary rows and the 4-bit trace -- lives in ONE file-scope
// device array of fixed size, and inputs / outputs are DEVICE pointers of the caller. The entry point calls no
// cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine; everything is queued on the context's stream.
// =====================================================================================================================
namespace svp_static
{
constexpr int MAXD = SVP_STATIC_MAX_LEN;                       // longest sequence on either side
constexpr int NBANDS = (MAXD + svp::K5_BAND - 1) / svp::K5_BAND;
constexpr size_t al(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr int TBL = MAXD + 4;                                  // table entries 0..TBL
constexpr size_t OFF_GAPSTEP = 0;
constexpr size_t OFF_EDGE = OFF_GAPSTEP + al((size_t)(TBL + 1) * 8);
constexpr size_t OFF_REF = OFF_EDGE + al((size_t)(TBL + 1) * 8);
constexpr size_t OFF_SEQ = OFF_REF + al((size_t)MAXD + 64);
constexpr size_t OFF_JOB = OFF_SEQ + al((size_t)MAXD + 64);
constexpr size_t OFF_END = OFF_JOB + al(sizeof(svp::FinalJob));
constexpr size_t OFF_BNDS = OFF_END + al(sizeof(svp::FinalEnd));
constexpr size_t OFF_BNDV = OFF_BNDS + al((size_t)(MAXD + 8) * 8);
constexpr size_t OFF_BNDVL = OFF_BNDV + al((size_t)(MAXD + 8) * 8);
constexpr size_t OFF_TRACE = OFF_BNDVL + al((size_t)(MAXD + 8) * 4);
constexpr size_t TRACE_BYTES = (size_t)((MAXD + 1) / 2 + svp::K5_THREADS) * (size_t)(NBANDS * svp::K5_THREADS) * 4;
constexpr size_t TOTAL = OFF_TRACE + al(TRACE_BYTES);          // ~10.7 MB for MAXD = 4096
}  // namespace svp_static

// Held by every GPU that loads the library until the process exits. Shared by all calls and all contexts of a device:
// two calls that overlap on the GPU (different streams) race on it.
static __device__ __align__(256) unsigned char svp_static_scratch[svp_static::TOTAL];

namespace svp_static
{
// Copies the caller's sequences into the scratch (bases beyond N -> N, zero padding behind the end: the kernel reads one
// column past it), writes the job descriptor and clears what must start at zero (the end-cell record; the caller's result).
// Device code names the array directly.
__global__ void prepare_kernel(const uint8_t *__restrict__ ref, int n, const uint8_t *__restrict__ seq, int m,
                               svp::CigarOut *__restrict__ result)
{
    uint8_t *sref = svp_static_scratch + OFF_REF, *sseq = svp_static_scratch + OFF_SEQ;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
    for (int i = tid; i < n + 16; i += nt) { uint8_t b = i < n ? ref[i] : 0; sref[i] = b > 4 ? 4 : b; }
    for (int i = tid; i < m + 16; i += nt) { uint8_t b = i < m ? seq[i] : 0; sseq[i] = b > 4 ? 4 : b; }
    if (tid == 0)
    {
        svp::FinalJob j;
        j.ref = sref; j.seq = sseq; j.n = n; j.m = m;
        int nb = (m + svp::K5_BAND - 1) / svp::K5_BAND; if (nb < 1) nb = 1;
        j.row_threads = nb * svp::K5_THREADS;
        j.trace_off = 0; j.bnd_off = 0; j.valid = 1;
        *reinterpret_cast<svp::FinalJob *>(svp_static_scratch + OFF_JOB) = j;
        svp::FinalEnd e; e.S = 0; e.H = 0; e.V = 0;
        *reinterpret_cast<svp::FinalEnd *>(svp_static_scratch + OFF_END) = e;
        svp::CigarOut o; o.n_cigar = 0; o.stored = 0; o.score = 0;
        *result = o;
    }
}

// per device: address of the scratch array as the DEVICE sees it (the symbol itself is only a host-side stand-in),
// and whether the tables have been uploaded
struct PerDevice { unsigned char *base = nullptr; bool ready = false; };
static PerDevice g_dev[64];
static std::mutex g_mu;
}  // namespace svp_static

// Set-up, once per device, with that device current: the device address of the scratch array (cudaGetSymbolAddress -- in
// host code the array's name is NOT a device address) and the two tables every job shares, computed with the same IEEE
// operations as ensure_tables() (api_final.inc) into static host arrays and copied into the scratch. No allocation.
static int svp_static_setup(svp_ctx *ctx, unsigned char **base_out)
{
    using namespace svp_static;
    if (ctx->device < 0 || ctx->device >= 64) return svp_fail_msg(ctx, SVP_ERR_ARG, "svp_final_align_static: device index beyond 63");
    std::lock_guard<std::mutex> lk(g_mu);
    PerDevice &pd = g_dev[ctx->device];
    if (!pd.ready)
    {
        CK(cudaSetDevice(ctx->device));
        unsigned char *p = nullptr;
        CK(cudaGetSymbolAddress((void **)&p, svp_static_scratch));
        if (!p) return svp_fail_msg(ctx, SVP_ERR_CUDA, "svp_final_align_static: no device address for the scratch array");
        static double gs[TBL + 1], ed[TBL + 1];
        gs[0] = 0;
        for (int len = 1; len <= TBL; ++len)
        {
            volatile double q = F_GO / (double)len;
            gs[len] = (q < F_GE) ? (double)q : F_GE;
        }
        ed[0] = 0;
        volatile double V = F_INF, S = 0;
        int vl = 1;
        for (int i = 1; i <= TBL; ++i)
        {
            volatile double ext = V + gs[vl];
            volatile double opn = S + F_GO;
            if (ext < opn) { V = opn; vl = 1; } else { V = ext; vl = vl + 1; }
            S = V;
            ed[i] = S;
        }
        CK(cudaMemcpy(p + OFF_GAPSTEP, gs, sizeof gs, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(p + OFF_EDGE, ed, sizeof ed, cudaMemcpyHostToDevice));
        pd.base = p; pd.ready = true;
    }
    *base_out = pd.base;
    return SVP_OK;
}

extern "C" int svp_final_align_static(svp_ctx *ctx, const uint8_t *ref_part_dev, int32_t ref_part_len, const uint8_t *seq_dev,
                                      int32_t seq_len, char *cigar_op_dev, uint32_t *cigar_cnt_dev, int32_t cigar_cap,
                                      svp_final_static_result *result_dev)
{
    using namespace svp_static;
    static_assert(sizeof(svp_final_static_result) == sizeof(svp::CigarOut), "result layout");
    if (!ctx) return SVP_ERR_ARG;
    if (!ref_part_dev || !seq_dev || !cigar_op_dev || !cigar_cnt_dev || !result_dev || cigar_cap < 1)
        return svp_fail_msg(ctx, SVP_ERR_ARG, "svp_final_align_static: null pointer or empty CIGAR buffer");
    if (ref_part_len < 1 || seq_len < 1) return svp_fail_msg(ctx, SVP_ERR_ARG, "svp_final_align_static: empty sequence");
    if (ref_part_len > MAXD || seq_len > MAXD)
        return svp_fail_msg(ctx, SVP_ERR_CAPACITY, "svp_final_align_static: sequence longer than SVP_STATIC_MAX_LEN (" +
                                                       std::to_string(MAXD) + "); use svp_final_align_batch");
    unsigned char *base = nullptr;
    int rc = svp_static_setup(ctx, &base);
    if (rc) return rc;
    CK(cudaSetDevice(ctx->device));
    svp::CigarOut *result = reinterpret_cast<svp::CigarOut *>(result_dev);
    prepare_kernel<<<8, 256, 0, ctx->stream>>>(ref_part_dev, ref_part_len, seq_dev, seq_len, result);
    CK(cudaGetLastError());
    svp::K5Tables tb;
    tb.gapstep = reinterpret_cast<const double *>(base + OFF_GAPSTEP);
    tb.edge = reinterpret_cast<const double *>(base + OFF_EDGE);
    tb.size = TBL;
    const svp::FinalJob *job = reinterpret_cast<const svp::FinalJob *>(base + OFF_JOB);
    svp::FinalEnd *ends = reinterpret_cast<svp::FinalEnd *>(base + OFF_END);
    uint32_t *trace = reinterpret_cast<uint32_t *>(base + OFF_TRACE);
    svp::k5_final_kernel<true><<<1, svp::K5_THREADS, (size_t)(TBL + 1) * 8, ctx->stream>>>(
        job, tb, F_MM, F_MX, F_GO, F_INF, trace, reinterpret_cast<double *>(base + OFF_BNDS),
        reinterpret_cast<double *>(base + OFF_BNDV), reinterpret_cast<int32_t *>(base + OFF_BNDVL), ends);
    CK(cudaGetLastError());
    svp::k5_traceback_kernel<<<1, svp::K5TB_WARPS * 32, 0, ctx->stream>>>(job, 1, trace, ends, F_INF, cigar_op_dev, cigar_cnt_dev,
                                                                        cigar_cap, result);
    CK(cudaGetLastError());
    ctx->n_launches += 3;
    return SVP_OK;
}
