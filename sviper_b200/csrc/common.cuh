// Shared device-side definitions for the sviper_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef __CUDA_ARCH__
#define SVP_HOST_PASS 1
#endif

namespace svp
{
// BWA-MEM-like scores hard-coded by the reference (include/sviper/mapping.h:94-97)
constexpr int MATCH = 1, MISMATCH = -4, GAP_EXT = -1, GAP_OPEN = -6;

// Sequences live in HBM as 2-bit packed bases (16 per u32 word) plus a 1-bit N plane (32 per word).
// Every sequence starts on a 32-base boundary; off32 counts 32-base units.
struct SeqStore
{
    const uint32_t *w2;    // 2 words per unit
    const uint32_t *wn;    // 1 word per unit
    const int32_t *off32;  // per sequence
    const int32_t *len;    // per sequence
};

__device__ __forceinline__ int seq_base(const SeqStore &s, int unit, int pos)
{
    uint32_t w = __ldg(s.w2 + (size_t)unit * 2 + (pos >> 4));
    uint32_t n = __ldg(s.wn + (size_t)unit + (pos >> 5));
    int c = (w >> ((pos & 15) * 2)) & 3;
    return ((n >> (pos & 31)) & 1) ? 4 : c;
}

// K1 (score pass) output per task
struct ScoreOut
{
    int32_t score_f, score_r;   // forward / reverse-complement score
    int32_t col_f, col_r;       // leftmost DP column (1-based, 0 = initial column) of the last-row maximum
};

// per-alignment plan for the traced pass: winner strand + DP geometry
struct TracePlan
{
    int32_t read_id, win_id;
    int32_t is_rc;
    int32_t score;        // winner score
    int32_t end_col;      // DP column of the end cell (0..L)
    int32_t c0;           // first window column (0-based) of the traced sub-matrix; DP column 0 == window col c0
    int32_t ncols;        // number of window columns traced (end_col - c0)
    int32_t band_w;       // >= 0: banded traced pass over the diagonals |(j - i) - (end_col - R)| <= band_w; -1: strip kernel
    int32_t dlo;          // band: lowest diagonal j - i
    int32_t pad_;
};
// Device-resident control block of one pass (one per context): what the round needs from the per-round arenas and
// what is left to do. Written by fit_kernel / count_active_kernel, copied to pinned host memory once per round; the
// host reads it one round LATE (the next round is already queued), so no kernel launch waits for a host round trip.
struct RoundCtl
{
    int32_t ovf;                      // != 0: an arena is too small: traced pass, traceback and vote skip the round (no state changes)
    int32_t round;
    int64_t need_units, need_words;   // record / path units and trace words this round needs
    int32_t cls[4];                   // tasks per traced-pass class
    int32_t active, live_pairs;       // windows still being polished, their read pairs
};

// traced-pass classes (plan_kernel bins tasks by band width)
enum { TC_BAND16 = 0, TC_BAND32 = 1, TC_STRIP5 = 2, TC_STRIP8 = 3, TC_N = 4 };
}  // namespace svp

#define SVP_CHECK(call)                                                                              \
    do                                                                                               \
    {                                                                                                \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) return svp_fail(ctx, e_, #call, __FILE__, __LINE__);                  \
    } while (0)
