// Host-side context, growable device buffers and small utility kernels.
#pragma once
#include <sched.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/sviper_b200.h"
#include "common.cuh"
#include "k1_trace.cuh"
#include "k3_pileup.cuh"
#include "k5_final.cuh"
#include "k6_msa.cuh"

struct DevBuf
{
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaError_t e = cudaFree(p); if (e != cudaSuccess) return e; p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) return e;
        cap = want;
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *)p; }
};

// Pinned host staging: a bump allocator over page-locked blocks. Everything a batch call copies to or from
// the device goes through it (one cudaMemcpyAsync per array, no per-sequence copies, no pageable staging by
// the driver). reset() is called at the start of a batch call, after the stream has been synchronised.
struct HostStage
{
    struct Block { char *p; size_t cap; };
    std::vector<Block> blocks;
    size_t used = 0;       // bytes used in blocks.back()
    size_t total_req = 0;  // bytes requested since the last reset (sizes the consolidated block)
    void *alloc(size_t bytes)
    {
        bytes = (bytes + 255) & ~(size_t)255;
        total_req += bytes;
        if (blocks.empty() || used + bytes > blocks.back().cap)
        {
            size_t cap = bytes > ((size_t)8 << 20) ? bytes : ((size_t)8 << 20);
            if (!blocks.empty() && cap < blocks.back().cap * 2) cap = blocks.back().cap * 2;
            char *p = nullptr;
            if (cudaHostAlloc((void **)&p, cap, cudaHostAllocDefault) != cudaSuccess) return nullptr;
            blocks.push_back({p, cap});
            used = 0;
        }
        void *r = blocks.back().p + used;
        used += bytes;
        return r;
    }
    template <typename T> T *get(size_t n) { return (T *)alloc(n * sizeof(T) + sizeof(T)); }
    void reset()
    {
        if (blocks.size() > 1)
        {
            // consolidate: one block that holds everything the last call needed
            for (Block &b : blocks) cudaFreeHost(b.p);
            blocks.clear();
            size_t cap = total_req + total_req / 2;
            char *p = nullptr;
            if (cudaHostAlloc((void **)&p, cap, cudaHostAllocDefault) == cudaSuccess) blocks.push_back({p, cap});
        }
        used = 0; total_req = 0;
    }
    void release()
    {
        for (Block &b : blocks) cudaFreeHost(b.p);
        blocks.clear(); used = 0; total_req = 0;
    }
};

enum { ST_K1 = 0, ST_K1T, ST_K2, ST_K3, ST_K5, ST_K5TB, ST_PACK, ST_OTHER, ST_K1M /* score pass of mixed rounds */, ST_N };

// device-resident packed sequences (reads or windows)
struct DevSeqs
{
    DevBuf bytes, w2, wn, off32, len, unit_seq, byte_off;
    int n = 0;
    int64_t units = 0;
    svp::SeqStore store() const
    {
        svp::SeqStore s;
        s.w2 = w2.as<uint32_t>(); s.wn = wn.as<uint32_t>(); s.off32 = off32.as<int32_t>(); s.len = len.as<int32_t>();
        return s;
    }
};

struct PolishLayout;
struct FinalSpec;
struct svp_engine;
struct MsaSession;

constexpr int SVP_DEFAULT_LANES = 4, SVP_MAX_LANES = 16, SVP_MIN_PER_LANE = 4;

enum { SVP_SYNC_YIELD = 0, SVP_SYNC_SPIN = 1, SVP_SYNC_BLOCK = 2 };

struct svp_ctx
{
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaStream_t side[svp::TC_N] = {nullptr};        // traced-pass class kernels run side by side
    cudaEvent_t ev_fork = nullptr, ev_side[svp::TC_N] = {nullptr};
    // ---- lanes: a polish batch is split over n_lanes sub-contexts (own stream, arenas and host thread each), so that
    // the latency-bound kernels of one lane (traced pass, traceback, vote, final alignment) overlap the score
    // kernel of another ----
    int n_lanes = SVP_DEFAULT_LANES;
    bool is_lane = false;
    int share = 1;                          // lanes that share the GPU with this context (1: it has the GPU to itself)
    std::vector<svp_ctx *> lanes;
    std::vector<int> lane_range;            // variant ranges of the last split batch; empty: it ran on this context
    cudaEvent_t ev_call = nullptr;          // start (parent -> lanes) / end (lane -> parent) ordering of a split call
    int n_launches = 0;
    double ms[ST_N] = {0};
    bool timing = false;
    std::vector<cudaEvent_t> ev_pool;
    std::vector<int> ev_slot;
    size_t ev_used = 0;
    size_t arena_budget = (size_t)96 << 30;
    int full_window = 0;   // debug: trace the whole window instead of the provable sub-window
    const uint8_t *win_bytes = nullptr;   // windows as bytes 0..4 (+ wins.byte_off): align batches -> wins.bytes, polish -> cons0

    // ---- align pipeline state (also reused by the polish pipeline) ----
    DevSeqs reads, wins;
    DevBuf read_quals, qual_off;            // qualities (bytes) per read, offsets
    DevBuf task_read, task_win, task_idx19, task_idx32;
    DevBuf task_mate, win_ins;              // polish batches: mate task per task (-1 none), (insert mean, sd) per window
    bool prune_pairs = false;               // run_trace_pass: skip the traceback of pairs that cannot be proper
    int ntasks = 0, n19 = 0, n32 = 0, n_short = 0, n_long = 0;   // tasks; by score-kernel class (<= 152 / longer rows); by strip class (<= 160)
    DevBuf score_out, plans, units, E, tunits, TE, cls_count, cls_list, scan_tmp, trace, ops, alns;
    DevBuf rec_code, rec_qual, rec_inslen, rec_insq;
    int64_t E_total = 0;
    // ---- polish pipeline ----
    DevBuf descs, states, states_init, pair_task, cons0, cons1, inslen, voff, profile, cov, isgap, decision, proper, cons_init;
    // strand certificates (k1_score.cuh, K1Mix): per-task strand state, per-round task lists, per-window edit density
    DevBuf sstate, reg19, reg32, mix19, mix32, fail19, fail32, mixcnt, edit, emax;
    bool predict_certificates = true;       // SVP_NO_PREDICT=1: off (A/B switch)
    bool mix_reg_wide = false;              // SVP_MIX_REGULAR=wide: regular tasks of a mixed round in the wide shape (A/B switch)
    int k1_threads = 0;                     // 0: choose per launch; 64 / 128: force the block size of the throughput shape (SVP_K1_THREADS)
    bool mixcnt_zeroed = false;             // mixcnt: 4 list counters (reset every round) + the DP cell counter at byte 32
    int64_t pairs_total = 0, n_mateless = 0;   // pairs of the batch; pairs with a dummy mate
    bool mixed_rounds = true;               // SVP_NO_MIXED=1 turns the mixed tasks off (A/B measurements)
    struct PolishLayout *resident = nullptr;   // last svp_polish_batch, kept for svp_polish_rerun
    int resident_nv = 0;
    // ---- final alignment ----
    DevBuf fjobs, fref, fseq, ftrace, fbndS, fbndV, fbndVL, fends, fcig_op, fcig_cnt, fcig_out, gapstep, edge;
    int table_size = 0;
    bool k5_global_table = false;           // SVP_K5_GLOBAL_TABLE=1: K5 reads the gap-step table from global memory at every size (test switch)
    // ---- host staging / result gather ----
    HostStage stage;
    DevBuf win_done, gather_desc, gather_out;
    // ---- round control: device block + two pinned copies (read one round late) ----
    DevBuf ctl;
    svp::RoundCtl *h_ctl = nullptr;          // pinned, 2 slots
    cudaEvent_t ev_ctl[2] = {nullptr, nullptr};
    int64_t cap_units = 0, cap_words = 0;   // what the per-round arenas (records, paths, trace) currently hold
    int sync_mode = 0;                      // SVP_SYNC_* (svp_wait_event)
    cudaEvent_t ev_sync = nullptr;
    int n_regrow = 0;                       // rounds repeated because an arena had to grow
    bool sync_rounds = true;                // wait for the outcome of a round before queueing the next (SVP_SPECULATIVE_ROUNDS=1: one round ahead)
    int64_t est_units = 0, est_words = 0;   // first-round estimate of the current batch (polish_upload / upload_tasks)
    // ---- asynchronous batch scheduler (api_final.inc): one worker thread per lane ----
    struct svp_engine *engine = nullptr;
    // ---- pairwise fills of build_consensus (api_msa.inc) ----
    struct MsaSession *msa = nullptr;
};

// Host waits. Three ways (SVP_SYNC=yield|spin|block, default yield):
//   spin   the CUDA runtime's spinning wait (cudaStreamSynchronize / cudaEventSynchronize)
//   yield  polls the event and calls sched_yield() between polls: as prompt as spinning while cores are free, but a waiting
//          thread gives its core away when more threads than cores are runnable -- the case on an 8-GPU box, where a rank has
//          a 4-core slice for 4 lane threads + the submitting thread (pure spinning there: 151 k variants/s end to end in one
//          run, 102 k in the next, depending on which thread the scheduler starved)
//   block  waits on an event created with cudaEventBlockingSync (the thread sleeps): 2.8x slower at N = 8 (38,654 vs 109,913
//          variants/s) -- a round ends with a wait, and every wake-up adds tens of microseconds to a ~0.5 ms round
static inline cudaError_t svp_wait_event(svp_ctx *ctx, cudaEvent_t ev)
{
    if (ctx->sync_mode != SVP_SYNC_YIELD) return cudaEventSynchronize(ev);
    for (;;)
    {
        const cudaError_t e = cudaEventQuery(ev);
        if (e != cudaErrorNotReady) return e;
        sched_yield();
    }
}

// waits for everything queued on the context's stream
static inline cudaError_t svp_sync(svp_ctx *ctx)
{
    if (ctx->sync_mode == SVP_SYNC_SPIN) return cudaStreamSynchronize(ctx->stream);
    cudaError_t e = cudaEventRecord(ctx->ev_sync, ctx->stream);
    if (e != cudaSuccess) return e;
    return svp_wait_event(ctx, ctx->ev_sync);
}

static int svp_fail(svp_ctx *ctx, cudaError_t e, const char *what, const char *file, int line)
{
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
    if (ctx) ctx->err = buf;
    fprintf(stderr, "[sviper_b200] %s\n", buf);
    return SVP_ERR_CUDA;
}

static int svp_fail_msg(svp_ctx *ctx, int code, const std::string &msg)
{
    if (ctx) ctx->err = msg;
    return code;
}

namespace svp
{
// bytes (0..4) -> 2-bit words + N plane. One thread per 32-base unit.
// len_override (may be null): current length per sequence (polish rounds change the consensus length).
__global__ void pack_kernel(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ byte_off,
                            const int32_t *__restrict__ len, const int32_t *__restrict__ off32,
                            const int32_t *__restrict__ unit_seq, int64_t nunits, uint32_t *__restrict__ w2,
                            uint32_t *__restrict__ wn)
{
    int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= nunits) return;
    int s = unit_seq[u];
    int pos0 = (int)(u - off32[s]) * 32;
    int L = len[s];
    const uint8_t *b = bytes + byte_off[s];
    uint32_t lo = 0, hi = 0, nm = 0;
#pragma unroll 8
    for (int i = 0; i < 32; ++i)
    {
        int p = pos0 + i;
        uint32_t c = (p < L) ? b[p] : 0u;
        if (c > 3u) { nm |= 1u << i; c = 0u; }
        if (i < 16) lo |= c << (2 * i); else hi |= c << (2 * (i - 16));
    }
    w2[u * 2] = lo; w2[u * 2 + 1] = hi; wn[u] = nm;
}

// bases beyond N are N (seqan::Dna5 has 5 values); the score kernel indexes its profile with these bytes
__global__ void clamp_bases_kernel(uint8_t *__restrict__ b, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && b[i] > 4) b[i] = 4;
}

// polish rounds: refresh window lengths / byte offsets from the per-variant state
__global__ void window_refresh_kernel(const VarDesc *__restrict__ descs, const VarState *__restrict__ states, int nv,
                                      int32_t *__restrict__ win_len, int64_t *__restrict__ win_byte_off,
                                      int64_t buf1_delta, uint8_t *__restrict__ win_done)
{
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    win_len[v] = states[v].L;
    win_byte_off[v] = descs[v].cons_off + (states[v].cur ? buf1_delta : 0);
    win_done[v] = states[v].done ? 1 : 0;
}

// ctl->active = variants still being polished, ctl->live_pairs = their read pairs (an upper bound of half the live tasks);
// both were zeroed by fit_kernel earlier in the round
__global__ void count_active_kernel(const VarDesc *__restrict__ descs, const VarState *__restrict__ states, int nv,
                                    RoundCtl *__restrict__ ctl)
{
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    int a = (v < nv) ? !states[v].done : 0;
    if (a) atomicAdd(&ctl->live_pairs, descs[v].n_pairs);
    a = __syncthreads_count(a);
    if (threadIdx.x == 0 && a) atomicAdd(&ctl->active, a);
}

// DPX probe: independent dependent-chains per thread, enough threads to fill the chip
template <int MODE>
__global__ void dpx_probe_kernel(uint32_t *out, int iters)
{
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 0x00010001u + i;
    uint32_t c = blockIdx.x | 0x20002000u;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int i = 0; i < 8; ++i)
        {
            if (MODE == 0) a[i] = __viaddmax_s16x2(a[i], 0xFFFFFFFFu, c);
            else if (MODE == 1) a[i] = __vimax3_s16x2(a[i], c, a[(i + 1) & 7]);
            else
            {
                uint32_t x = a[i] + 0x00020007u;                   // IADD
                uint32_t hh = __viaddmax_s16x2(a[i], 0xFFFFFFFFu, c);
                uint32_t vv = __viaddmax_s16x2(c, 0xFFFFFFFFu, a[i]);
                a[i] = __vimax3_s16x2(x, hh, vv) + 0xFFF9FFFAu;      // VIMNMX3 + IADD
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// fp64 add probe: 8 independent chains per thread
__global__ void dadd_probe_kernel(double *out, int iters, double inc)
{
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = (double)threadIdx.x + i;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __dadd_rn(a[i], inc);
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
}  // namespace svp
