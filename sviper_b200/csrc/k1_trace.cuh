// K1t: TRACED pass of the semi-global affine DP for the winning strand of every task, restricted to the
// column window that provably contains every alignment scoring >= the winner (see plan_kernel), and
// K2: traceback + pileup records.
//
// Tie rules, trace bits and the end-cell correction follow the reference exactly
// (/root/reference/extern/seqan/include/seqan/align/dp_formula_affine.h:66-125, dp_formula.h:152-163,
// dp_algorithm_impl.h:1170-1185); the traceback follows dp_traceback_impl.h:391-420 / :243-259 / :321-336.
// Only 4 bits per cell are kept: bit0 = horizontal gap EXTENDED (else opened), bit1 = vertical gap
// EXTENDED, bits2-3 = 0 diagonal / 1 from the vertical matrix / 2 from the horizontal matrix.
//
// Mapping: one warp per alignment, lane l owns RPL consecutive read rows (rows aligned to the END of the
// warp, unused top rows behave as extra row 0s, see k1_score.cuh), wavefront over columns, int32 cells.
// A lane's RPL nibbles of one column form one 32-bit word: trace[(col-1)*32 + lane].
#pragma once
#include "common.cuh"
#include "../../include/sviper_b200.h"

namespace svp
{
constexpr int KT_WARPS = 4;
constexpr int KT_NEG = INT32_MIN / 2;   // seqan/align/dp_cell.h:144-145

// ---------------------------------------------------------------------------------------------
// plan: strand choice (mapping.h:121), mapQ, and the part of the matrix the traced pass has to cover.
//
// Column window: any alignment of the read ending in the last row at column e with score >= s spans at most
// R + max(0, R - s - 5) columns (m matches <= R, D deleted columns cost >= 5 + D), so the sub-matrix
// starting at c0 = e - R - max(0, R - s) - 8 contains the reference's traceback path and every path
// that ties with it at the end cell.
// Diagonal band: a cell (i, j) lies on a path that ends in (R, e) with score >= s only if
// |(e - j) - (R - i)| <= R - s - 5 (the path still has to absorb that many gap characters, costing at
// least 5 + that many, and can collect at most R matches). Cells outside the band w = max(0, R - s) can be
// replaced by -infinity: every cell ON a co-optimal path keeps its exact value (its optimal predecessor is
// in the band), and a candidate that is lowered was strictly smaller than the winner already (had it tied,
// it would itself be on a co-optimal path and therefore inside the band), so every trace decision the
// traceback reads is identical to the full matrix. full_window != 0 disables both restrictions.
// ---------------------------------------------------------------------------------------------
constexpr int KB_MAXW16 = 15;   // 2w+1 <= 32 diagonals: 8 lanes x 4 diagonals
constexpr int KB_MAXW32 = 31;   // 2w+1 <= 64 diagonals: 16 lanes x 4 diagonals

__host__ __device__ inline int kb_row_words(int R) { return (R >> 3) + 1; }   // trace words per diagonal (rows 0..R)

__global__ void plan_kernel(const int32_t *__restrict__ task_read, const int32_t *__restrict__ task_win, int ntasks,
                            const int32_t *__restrict__ read_len, const ScoreOut *__restrict__ sc,
                            const uint8_t *__restrict__ win_done, TracePlan *__restrict__ plan,
                            int64_t *__restrict__ units, int64_t *__restrict__ tunits,
                            int32_t *__restrict__ cls_count, int32_t *__restrict__ cls_list, int full_window,
                            const int32_t *__restrict__ task_mate, const double *__restrict__ win_ins)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntasks) return;
    ScoreOut o = sc[t];
    TracePlan p;
    p.read_id = task_read[t]; p.win_id = task_win[t]; p.pad_ = 0;
    if (win_done && win_done[p.win_id])
    {
        p.is_rc = 0; p.score = 0; p.end_col = 0; p.c0 = 0; p.ncols = -1; p.band_w = -1; p.dlo = 0;
        plan[t] = p; units[t] = 0; tunits[t] = 0;
        return;
    }
    p.is_rc = !(o.score_f > o.score_r);
    p.score = p.is_rc ? o.score_r : o.score_f;
    p.end_col = p.is_rc ? o.col_r : o.col_f;
    int R = read_len[p.read_id];
    int slack = R - p.score; if (slack < 0) slack = 0;
    // Pair pruning (polish batches: task_mate != null). Only proper pairs enter the pileup (polishing.h:135), and
    // whether a pair is proper (mapping.h:168-172) can often be decided from the score pass alone: the strands,
    // and bounds on gapsEndPos of the forward read / gapsBeginPos of the reverse one. An alignment of score s
    // holds at most (R-s-5)/2 inserted read bases (s <= R - 2I - 5) and R-s-5 deleted columns (s <= R - 5 - D), so
    //   endPos(F) = e_F + I_F  in [e_F, e_F + Imax_F],   beginPos(B) = e_B - R_B + I_B - D_B  in [.. - Dmax_B, .. + Imax_B].
    // A pair that cannot be proper needs no traceback and no pileup records: only its scores are used
    // (compute_mappQ_stats). Both mates evaluate the same predicate.
    if (task_mate)
    {
        bool prune;
        const int u = task_mate[t];
        if (u < 0) prune = true;                               // no mate: proper_pair stays false (mapping.h:34,162)
        else
        {
            const ScoreOut ou = sc[u];
            const int u_rc = !(ou.score_f > ou.score_r);
            if (u_rc == p.is_rc) prune = true;                 // same strand
            else
            {
                const int Ru = read_len[task_read[u]];
                const int su = u_rc ? ou.score_r : ou.score_f, eu = u_rc ? ou.col_r : ou.col_f;
                int slu = Ru - su; if (slu < 0) slu = 0;
                // F = forward read, B = reverse read
                const int eF = p.is_rc ? eu : p.end_col, slF = p.is_rc ? slu : slack;
                const int eB = p.is_rc ? p.end_col : eu, slB = p.is_rc ? slack : slu, RB = p.is_rc ? R : Ru;
                const int ImaxF = slF > 5 ? (slF - 5) / 2 : 0, ImaxB = slB > 5 ? (slB - 5) / 2 : 0;
                const int DmaxB = slB > 5 ? slB - 5 : 0;
                const int dd_max = eB - RB + ImaxB - eF, dd_min = eB - RB - DmaxB - eF - ImaxF;
                const double mean = win_ins[2 * p.win_id], sd = win_ins[2 * p.win_id + 1];
                const double rs = fmax(0.0, mean - 3 * sd);                       // config.h:101-109
                const double re = fmin(mean + 3 * mean, mean + 3 * sd);
                prune = !(dd_max > rs) || !(dd_min < re);
            }
        }
        if (prune)
        {
            p.c0 = 0; p.ncols = -2; p.band_w = -1; p.dlo = 0;
            plan[t] = p; units[t] = 0; tunits[t] = 0;
            return;
        }
    }
    int c0 = p.end_col - R - slack - 8;
    if (c0 < 0 || full_window) c0 = 0;
    p.c0 = c0; p.ncols = p.end_col - c0;
    int cls;
    if (!full_window && slack <= KB_MAXW32)
    {
        p.band_w = slack; p.dlo = p.end_col - R - slack;
        cls = slack <= KB_MAXW16 ? TC_BAND16 : TC_BAND32;
        tunits[t] = (int64_t)(slack <= KB_MAXW16 ? 32 : 64) * kb_row_words(R);   // every diagonal the lanes own
    }
    else
    {
        p.band_w = -1; p.dlo = 0;
        cls = R <= 160 ? TC_STRIP5 : TC_STRIP8;
        tunits[t] = (int64_t)(p.ncols > 0 ? p.ncols : 0) * 32;
    }
    plan[t] = p;
    units[t] = p.ncols + 2;   // pileup records / (+ read length) path ops reserved for this task
    int at = atomicAdd(cls_count + cls, 1);
    cls_list[(size_t)cls * ntasks + at] = t;
}

// After the scans: does the round fit its arenas? (E_end / TE_end: totals of the exclusive scans)
__global__ void fit_kernel(const int64_t *__restrict__ E_end, const int64_t *__restrict__ TE_end,
                           int32_t *__restrict__ cls_count, RoundCtl *__restrict__ ctl, int round, int64_t cap_units,
                           int64_t cap_words)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int64_t nu = *E_end, nw = *TE_end;
    ctl->need_units = nu; ctl->need_words = nw; ctl->round = round;
    for (int c = 0; c < TC_N; ++c) { ctl->cls[c] = cls_count[c]; cls_count[c] = 0; }
    ctl->active = 0; ctl->live_pairs = 0;
    ctl->ovf = (nu > cap_units || nw > cap_words) ? 1 : 0;
}

// Both exclusive scans of a round (record units -> E, trace words -> TE; n1 = tasks + 1 entries each) and the fit check in
// ONE single-block launch, for batches of up to SCAN2_MAX tasks: at that size the two library scans are four launches of
// a few microseconds each, i.e. mostly launch latency on the round's critical path. Also re-arms the class counters.
constexpr int SCAN2_THREADS = 1024, SCAN2_MAX = 128 * 1024;
__global__ void __launch_bounds__(SCAN2_THREADS) scan2_fit_kernel(const int64_t *__restrict__ units, int64_t *__restrict__ E,
                                                               const int64_t *__restrict__ tunits, int64_t *__restrict__ TE, int n1,
                                                               int32_t *__restrict__ cls_count, RoundCtl *__restrict__ ctl, int round,
                                                               int64_t cap_units, int64_t cap_words)
{
    // warp w owns the contiguous chunk [w * per, (w + 1) * per) (per: a multiple of 32); lanes read consecutive
    // entries, so every access is coalesced. Entry n1 - 1 only receives the totals: its own value is never read.
    __shared__ long long sh[2][SCAN2_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = n1 - 1;
    const int per = ((n1 + SCAN2_THREADS - 1) / SCAN2_THREADS) * 32;
    const int lo = min(n1, warp * per), hi = min(n1, lo + per);
    long long a = 0, b = 0;
    for (int i = lo + lane; i < hi; i += 32) if (i < n) { a += units[i]; b += tunits[i]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if (lane == 0) { sh[0][warp] = a; sh[1][warp] = b; }
    __syncthreads();
    if (warp == 0)
    {
        // exclusive prefix over the 32 warp totals; lane 31 also keeps the grand totals
        const long long wa = sh[0][lane], wb = sh[1][lane];
        long long pa = wa, pb = wb;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const long long xa = __shfl_up_sync(0xffffffffu, pa, o), xb = __shfl_up_sync(0xffffffffu, pb, o);
            if (lane >= o) { pa += xa; pb += xb; }
        }
        sh[0][lane] = pa - wa; sh[1][lane] = pb - wb;
        if (lane == 31)
        {
            ctl->need_units = pa; ctl->need_words = pb; ctl->round = round;
            for (int c = 0; c < TC_N; ++c) { ctl->cls[c] = cls_count[c]; cls_count[c] = 0; }
            ctl->active = 0; ctl->live_pairs = 0;
            ctl->ovf = (pa > cap_units || pb > cap_words) ? 1 : 0;
        }
    }
    __syncthreads();
    long long ca = sh[0][warp], cb = sh[1][warp];     // running exclusive prefix of this warp's chunk
    for (int base = lo; base < hi; base += 32)
    {
        const int i = base + lane;
        const long long u = (i < n && i < hi) ? units[i] : 0, w = (i < n && i < hi) ? tunits[i] : 0;
        long long pu = u, pw = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const long long xu = __shfl_up_sync(0xffffffffu, pu, o), xw = __shfl_up_sync(0xffffffffu, pw, o);
            if (lane >= o) { pu += xu; pw += xw; }
        }
        if (i < hi) { E[i] = ca + pu - u; TE[i] = cb + pw - w; }
        ca += __shfl_sync(0xffffffffu, pu, 31); cb += __shfl_sync(0xffffffffu, pw, 31);
    }
}

template <int RPL>
struct KTShape
{
    static constexpr int NSLOT = 32 * RPL;
    static constexpr int PROF_WORDS = 5 * NSLOT;
    static constexpr int SMEM_BYTES = KT_WARPS * PROF_WORDS * 4;
};

template <int RPL>
__global__ void __launch_bounds__(KT_WARPS * 32, 6) k1_trace_kernel(const TracePlan *__restrict__ plans,
                                                                const RoundCtl *__restrict__ ctl, int cls,
                                                                const int32_t *__restrict__ task_idx,
                                                                SeqStore reads, SeqStore wins,
                                                                const int64_t *__restrict__ E,
                                                                uint32_t *__restrict__ trace)
{
    using Sh = KTShape<RPL>;
    constexpr int NSLOT = Sh::NSLOT;
    extern __shared__ int32_t kt_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = blockIdx.x * KT_WARPS + warp;
    if (ctl->ovf || slot >= ctl->cls[cls]) return;     // the grid is sized for an upper bound of the class
    const int task = task_idx[slot];
    int32_t *prof = kt_smem + warp * Sh::PROF_WORDS;   // prof[c][k][lane]
    const TracePlan pl = plans[task];
    const int ncols = pl.ncols;
    if (ncols <= 0) return;                            // end cell in the initial column: no DP needed
    const int R = reads.len[pl.read_id], runit = reads.off32[pl.read_id];
    const int wunit = wins.off32[pl.win_id];
    const int pad = NSLOT - R;
    uint32_t *tr = trace + E[task];   // E: word offsets of the trace arena

    int s[RPL], h[RPL];
#pragma unroll
    for (int k = 0; k < RPL; ++k)
    {
        int row = lane * RPL + k - pad + 1;
        int rb = -1;
        if (row >= 1)
        {
            if (pl.is_rc) { int b = seq_base(reads, runit, R - row); rb = (b < 4) ? 3 - b : 4; }
            else rb = seq_base(reads, runit, row - 1);
        }
#pragma unroll
        for (int c = 0; c < 5; ++c)
            prof[c * NSLOT + k * 32 + lane] = (row < 1) ? 0 : ((c == rb) ? MATCH : MISMATCH);
        s[k] = (row >= 1) ? -(5 + row) : 0;   // initial column, not free (dp_formula_affine.h:291-325)
        h[k] = KT_NEG;
    }
    int prev_up;
    {
        int rowa = lane * RPL - pad;
        prev_up = (lane == 0 || rowa < 1) ? 0 : -(5 + rowa);
    }
    int out_s = s[RPL - 1], out_v = KT_NEG;
    __syncwarp();

    const int nsteps = ncols + 31;
    for (int t = 0; t < nsteps; ++t)
    {
        int r_s = __shfl_up_sync(0xffffffffu, out_s, 1);
        int r_v = __shfl_up_sync(0xffffffffu, out_v, 1);
        const int jj = t - lane;   // local window index; DP column jj + 1
        if (jj >= 0 && jj < ncols)
        {
            if (lane == 0) { r_s = 0; r_v = KT_NEG; }
            const int c = seq_base(wins, wunit, pl.c0 + jj);
            const int32_t *pr = prof + c * NSLOT + lane;
            int diag = prev_up;
            prev_up = r_s;
            int up_s = r_s, up_v = r_v;
            uint32_t tw = 0;
#pragma unroll
            for (int k = 0; k < RPL; ++k)
            {
                int inter = diag + pr[k * 32];
                diag = s[k];
                int hext = h[k] + GAP_EXT, hopn = s[k] + GAP_OPEN;
                uint32_t code = (hext < hopn) ? 0u : 1u;             // tie -> extension
                int hh = max(hext, hopn);
                h[k] = hh;
                int vext = up_v + GAP_EXT, vopn = up_s + GAP_OPEN;
                code |= (vext < vopn) ? 0u : 2u;
                up_v = max(vext, vopn);
                int gap = max(up_v, hh);
                uint32_t gsel = (up_v < hh) ? 8u : 4u;               // tie -> vertical
                int sn = max(inter, gap);
                if (inter < gap) code |= gsel;                       // tie -> diagonal
                if (k == RPL - 1 && lane == 31 && jj == ncols - 1)
                {
                    // end-cell correction (dp_algorithm_impl.h:1170-1185)
                    if (up_v == sn) code = (code & 3u) | 4u;
                    else if (hh == sn) code = (code & 3u) | 8u;
                }
                tw |= code << (4 * k);
                s[k] = sn; up_s = sn;
            }
            out_s = up_s; out_v = up_v;
            tr[(size_t)jj * 32 + lane] = tw;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K1b: BANDED traced pass. LPA lanes per alignment (128 / LPA alignments per block), lane l owns the DPL
// consecutive diagonals d = DPL*l + x (x = 0..DPL-1; d = (j - i) - dlo). One iteration covers two
// anti-diagonals: the even step computes the cells of the even slots x, the odd step those of the odd slots.
// Slots 2p and 2p+1 of a lane work on the same read row i = i0 - p in one iteration (i0 advances by one per
// iteration), so the read base is fetched once per slot pair. Predecessors:
//   diagonal (i-1, j-1): same diagonal, the slot's own previous value
//   up       (i-1, j)  : diagonal d+1 = slot x+1 (or lane l+1's slot 0 via shuffle), computed one step earlier
//   left     (i, j-1)  : diagonal d-1 = slot x-1 (or lane l-1's last slot via shuffle), computed one step earlier
// Row 0 (S = 0, free begin) and DP column 0 (S = go + (i-1) ge) take their initial values, anything else
// outside the matrix or the band is -infinity; all of that is select-based (no divergent early exits).
// Trace: 4 bits per cell as in the strip kernel; a slot collects the nibbles of 8 consecutive rows of its
// diagonal into a word: trace[d * RW + (i >> 3)], RW = kb_row_words(R).
// ---------------------------------------------------------------------------------------------
constexpr int KB_THREADS = 128;
constexpr int KB_DPL = 4;                                   // diagonals per lane
constexpr int KB_RMAX = SVP_MAX_READ_LEN + 8;               // staged read bytes
constexpr int KB_WMAX = SVP_MAX_READ_LEN + 2 * KB_MAXW32 + 8;   // staged window bytes (columns jbase .. e)
constexpr int KB_SEQ_BYTES = KB_RMAX + KB_WMAX + 8;

template <int LPA, int DPL>
__global__ void __launch_bounds__(KB_THREADS, 6) k1_band_kernel(const TracePlan *__restrict__ plans,
                                                            const RoundCtl *__restrict__ ctl, int cls,
                                                            const int32_t *__restrict__ task_idx, SeqStore reads,
                                                            SeqStore wins, const int64_t *__restrict__ TE,
                                                            uint32_t *__restrict__ trace)
{
    static_assert(DPL % 2 == 0 && LPA <= 32 && 32 % LPA == 0, "band kernel shape");
    constexpr int APB = KB_THREADS / LPA;   // alignments per block
    constexpr int NP = DPL / 2;             // slot pairs per lane
    __shared__ uint8_t kb_seq[APB][KB_SEQ_BYTES];
    const int l = threadIdx.x % LPA, grp = threadIdx.x / LPA;
    const int slot = blockIdx.x * APB + grp;
    const unsigned gmask = (LPA == 32) ? 0xffffffffu : (((1u << (LPA & 31)) - 1u) << (LPA * ((threadIdx.x & 31) / LPA)));
    const int ntasks = ctl->ovf ? 0 : ctl->cls[cls];    // the grid is sized for an upper bound of the class
    if (blockIdx.x * APB >= ntasks) return;             // (whole block)
    const bool live = slot < ntasks;
    int R = 0, e = 0, dlo = 0, rc = 0, runit = 0, wunit = 0, jbase = 0;
    uint32_t *tr = trace;
    if (live)
    {
        const int task = task_idx[slot];
        const TracePlan pl = plans[task];
        R = reads.len[pl.read_id]; runit = reads.off32[pl.read_id];
        wunit = wins.off32[pl.win_id];
        e = pl.end_col; dlo = pl.dlo; rc = pl.is_rc;
        tr = trace + TE[task];
        jbase = dlo > 0 ? dlo : 0;   // window columns jbase+1 .. e are staged at sW[j - jbase]
    }
    uint8_t *sR = kb_seq[grp], *sW = kb_seq[grp] + KB_RMAX;
    for (int i = l; i < R; i += LPA)
    {
        int b;
        if (rc) { b = seq_base(reads, runit, R - 1 - i); b = (b < 4) ? 3 - b : 4; }
        else b = seq_base(reads, runit, i);
        sR[i + 1] = (uint8_t)b;                       // sR[row], rows 1..R
    }
    for (int j = jbase + 1 + l; j <= e; j += LPA) sW[j - jbase] = (uint8_t)seq_base(wins, wunit, j - 1);
    __syncwarp(gmask);

    // values of the last cell computed on each own diagonal. A slot that has no cell inside the matrix in some
    // step keeps a stale value: a cell inside the matrix only ever reads neighbours that are inside too (or the
    // lane-edge -infinity), and those are by construction the most recent cells of their slots.
    int S[DPL], H[DPL], V[DPL];
    uint32_t acc[DPL];
#pragma unroll
    for (int x = 0; x < DPL; ++x) { S[x] = KT_NEG; H[x] = KT_NEG; V[x] = KT_NEG; acc[x] = 0; }
    const int RW = kb_row_words(R);
    // first step: the smallest anti-diagonal with a cell in the matrix, rounded down to even parity (even slots)
    int a_first = jbase;
    if ((a_first - dlo) & 1) --a_first;
    const int a_last = R + e;
    const int my_n = live ? (a_last - a_first + 2) / 2 : 0;                 // iterations of this alignment
    const int nsteps = __reduce_max_sync(0xffffffffu, my_n);
    int i0 = (a_first - dlo - DPL * l) >> 1;                                // row of slot pair 0 (exact: even)
    const int jd = dlo + DPL * l;                                           // j = i + jd + x
    const uint8_t *pW = sW + (jd - jbase);                                  // window byte of cell (i, slot x): pW[i + x]
    uint32_t *trl = tr + (size_t)(DPL * l) * RW;                            // trace words of slot x: trl[x * RW + (i >> 3)]
    // rows i0 for which all DPL cells of the lane are regular inner cells (1 <= i < R, 1 <= j < e) ...
    const int f_lo = max(NP, 1 - jd), f_hi = min(R - 1, e - 1 - jd - NP);
    // ... and rows for which at least one cell lies inside the matrix (0 <= i <= R, 0 <= j <= e)
    const int n_lo = max(0, -(jd + NP)), n_hi = live ? min(R + NP - 1, e - jd) : -1;

    // the recurrence of one inner cell; returns the 4 trace bits
    auto inner = [&](int sub, int diagS, int upS, int upV, int leftS, int leftH, int &So, int &Ho, int &Vo, int &gap_o) {
        const int inter = diagS + sub;
        const int hext = leftH + GAP_EXT, hopn = leftS + GAP_OPEN;
        uint32_t code = (hext < hopn) ? 0u : 1u;             // tie -> extension
        const int hh = max(hext, hopn);
        const int vext = upV + GAP_EXT, vopn = upS + GAP_OPEN;
        code |= (vext < vopn) ? 0u : 2u;
        const int vv = max(vext, vopn);
        const int gap = max(vv, hh);
        const uint32_t gsel = (vv < hh) ? 8u : 4u;           // tie -> vertical
        const int sn = max(inter, gap);
        if (inter < gap) code |= gsel;                       // tie -> diagonal
        So = sn; Ho = hh; Vo = vv; gap_o = gap;
        return code;
    };
    // general cell: boundaries, end cell, flush on every way out of the matrix
    auto cell = [&](int i, int x, int rb, int diagS, int upS, int upV, int leftS, int leftH, int &So, int &Ho, int &Vo) {
        const int j = i + jd + x;
        if ((unsigned)i > (unsigned)R || (unsigned)j > (unsigned)e) return;
        if (i == 0 || j == 0)
        {
            // free first row (S = 0), initial column (dp_formula_affine.h:291-325)
            So = (i == 0) ? 0 : -(5 + i); Ho = KT_NEG; Vo = KT_NEG;
            return;
        }
        int sn, hh, vv, gap;
        uint32_t code = inner((rb == pW[i + x]) ? MATCH : MISMATCH, diagS, upS, upV, leftS, leftH, sn, hh, vv, gap);
        if (i == R && j == e)
        {
            // end-cell correction (dp_algorithm_impl.h:1170-1185)
            if (vv == sn) code = (code & 3u) | 4u;
            else if (hh == sn) code = (code & 3u) | 8u;
        }
        So = sn; Ho = hh; Vo = vv;
        acc[x] |= code << (4 * (i & 7));
        if ((i & 7) == 7 || i == R || j == e) { trl[x * RW + (i >> 3)] = acc[x]; acc[x] = 0; }
    };

    for (int it = 0; it < nsteps; ++it, ++i0)
    {
        const bool reg_all = i0 >= f_lo && i0 <= f_hi;
        const bool none = i0 < n_lo || i0 > n_hi;
        const bool fast = __all_sync(0xffffffffu, reg_all || none);
        // even step. slot 0's left neighbour = lane l-1's last slot
        int lS = __shfl_up_sync(0xffffffffu, S[DPL - 1], 1, LPA);
        int lH = __shfl_up_sync(0xffffffffu, H[DPL - 1], 1, LPA);
        if (l == 0) { lS = KT_NEG; lH = KT_NEG; }
        if (fast)
        {
            // every lane of the warp is either fully inside (inner cells only, last row excluded) or has nothing to do
            int rbp[NP], shp[NP];
            if (reg_all)
            {
#pragma unroll
                for (int p = 0; p < NP; ++p)
                {
                    const int i = i0 - p, x = 2 * p;
                    int gap;
                    rbp[p] = sR[i]; shp[p] = 4 * (i & 7);
                    const uint32_t code = inner((rbp[p] == pW[i + x]) ? MATCH : MISMATCH, S[x], S[x + 1], V[x + 1],
                                                p == 0 ? lS : S[x - 1], p == 0 ? lH : H[x - 1], S[x], H[x], V[x], gap);
                    acc[x] |= code << shp[p];
                }
            }
            int uS = __shfl_down_sync(0xffffffffu, S[0], 1, LPA);
            int uV = __shfl_down_sync(0xffffffffu, V[0], 1, LPA);
            if (l == LPA - 1) { uS = KT_NEG; uV = KT_NEG; }
            if (reg_all)
            {
#pragma unroll
                for (int p = 0; p < NP; ++p)
                {
                    const int i = i0 - p, x = 2 * p + 1;
                    int gap;
                    const uint32_t code = inner((rbp[p] == pW[i + x]) ? MATCH : MISMATCH, S[x], p == NP - 1 ? uS : S[x + 1],
                                                p == NP - 1 ? uV : V[x + 1], S[x - 1], H[x - 1], S[x], H[x], V[x], gap);
                    acc[x] |= code << shp[p];
                    if (shp[p] == 28)
                    {
                        uint32_t *w = trl + (i >> 3);
                        w[(x - 1) * RW] = acc[x - 1]; w[x * RW] = acc[x];
                        acc[x - 1] = 0; acc[x] = 0;
                    }
                }
            }
            continue;
        }
        int rbp[NP];
#pragma unroll
        for (int p = 0; p < NP; ++p)
        {
            const int i = i0 - p, x = 2 * p;
            rbp[p] = ((unsigned)(i - 1) < (unsigned)R) ? sR[i] : 0xFF;
            cell(i, x, rbp[p], S[x], S[x + 1], V[x + 1], p == 0 ? lS : S[x - 1], p == 0 ? lH : H[x - 1], S[x], H[x], V[x]);
        }
        // odd step. the last slot's up neighbour = lane l+1's slot 0
        int uS = __shfl_down_sync(0xffffffffu, S[0], 1, LPA);
        int uV = __shfl_down_sync(0xffffffffu, V[0], 1, LPA);
        if (l == LPA - 1) { uS = KT_NEG; uV = KT_NEG; }
#pragma unroll
        for (int p = 0; p < NP; ++p)
        {
            const int i = i0 - p, x = 2 * p + 1;
            cell(i, x, rbp[p], S[x], p == NP - 1 ? uS : S[x + 1], p == NP - 1 ? uV : V[x + 1], S[x - 1], H[x - 1], S[x], H[x],
                 V[x]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2: traceback (one WARP per alignment) + the per-reference-position records the pileup consumes.
// Literal restatement of add_read_to_profile's bookkeeping (include/sviper/polishing.h:54-107) on the
// view columns of the alignment; see oracle/oracle_polish.c:orc_add_read for the same logic on the CPU.
// ---------------------------------------------------------------------------------------------
struct AlnOut
{
    int32_t lead;      // gapsBeginPos(gapsRead)
    int32_t endpos;    // gapsEndPos(gapsRead)
    int32_t npath;     // path ops stored (reverse order) in the ops buffer; -1: not traced (pair cannot be proper)
    int32_t lead_i;    // leading 'I' columns (path reached DP column 0)
    int32_t p0, n;     // records cover reference positions [p0, p0 + n)
    int32_t mapq;      // max(0, score)  (mapping.h:125,134)
    int32_t is_rc;
    int32_t err;       // != 0: internal invariant violated (window restriction / capacity)
};

struct RecArrays
{
    uint8_t *code;     // per record: 0x80 | noins<<3 | sym (0..4 base, 5 gap); 0 = no main update
    uint8_t *qual;
    uint16_t *inslen;  // inserted read bases in front of this reference position
    uint16_t *insq;    // oriented read index of the first inserted base
};

// One warp per alignment, both trace layouts (band: k1_band_kernel, strip: k1_trace_kernel<5|8>).
// Phase A follows the path in bulk: the 32 lanes read the next 32 cells along the current direction
// (diagonal, or the row / column of a gap being extended), a ballot finds where the run ends and the whole
// run is emitted at once, so the number of dependent trace loads is ~ the number of path segments, not
// the path length. All lanes carry the same (col, row, np) state.
// Phase B builds the pileup records for 32 view columns at a time from ballots / prefix popcounts; it
// is the data-parallel form of the sequential bookkeeping in add_read_to_profile (see the comments there).
constexpr int K2_WARPS = 4;

__global__ void __launch_bounds__(K2_WARPS * 32, 6) k2_traceback_kernel(
    const TracePlan *__restrict__ plans, const RoundCtl *__restrict__ ctl, int ntasks, SeqStore reads, SeqStore wins,
    const uint8_t *__restrict__ read_quals, const int64_t *__restrict__ qual_off,
    const uint32_t *__restrict__ trace, uint8_t *__restrict__ ops_buf, RecArrays rec,
    const int64_t *__restrict__ E, const int64_t *__restrict__ TE, AlnOut *__restrict__ out)
{
    const int t = blockIdx.x * K2_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (t >= ntasks || ctl->ovf) return;
    const TracePlan pl = plans[t];
    if (pl.ncols == -2)
    {
        // pair that cannot be proper (plan_kernel): only the score is used downstream
        if (lane == 0)
        {
            AlnOut o;
            o.lead = 0; o.endpos = 0; o.npath = -1; o.lead_i = 0; o.p0 = 0; o.n = 0;
            o.mapq = pl.score > 0 ? pl.score : 0; o.is_rc = pl.is_rc; o.err = 0;
            out[t] = o;
        }
        return;
    }
    if (pl.ncols < 0) return;   // window already converged
    const unsigned FULL = 0xffffffffu;
    const int R = reads.len[pl.read_id], L = wins.len[pl.win_id];
    const uint32_t *tr = trace + TE[t];
    const int RW = kb_row_words(R);
    const int band = pl.band_w >= 0;
    const int rpl = (R <= 160) ? 5 : 8;            // strip layout: rows per lane (plan_kernel's class rule)
    const int spad = 32 * rpl - R - 1;
    int bad_cell = 0;
    // trace nibble of the cell in local column col (DP column c0 + col), row
    auto codeat = [&](int col, int row) -> uint32_t {
        if (band)
        {
            const int d = pl.c0 + col - row - pl.dlo;
            if (d < 0 || d > 2 * pl.band_w) return 0x100u;   // outside the band: stops every probe (checked below)
            return (__ldg(tr + (size_t)d * RW + (row >> 3)) >> (4 * (row & 7))) & 15u;
        }
        const int slot = spad + row;
        const int sl = (rpl == 5) ? slot / 5 : slot >> 3, sk = (rpl == 5) ? slot % 5 : slot & 7;
        return (__ldg(tr + (size_t)(col - 1) * 32 + sl) >> (4 * sk)) & 15u;
    };
    uint8_t *ops = ops_buf + E[t] + (int64_t)t * SVP_MAX_READ_LEN;      // cap: ncols + 2 + max read length
    const int ops_cap = (int)(E[t + 1] - E[t]) + SVP_MAX_READ_LEN;
    AlnOut o;
    o.err = 0; o.is_rc = pl.is_rc; o.mapq = pl.score > 0 ? pl.score : 0;

    // ---- phase A: traceback (dp_traceback_impl.h:391-420, :243-259, :321-336) -----------------
    int col = pl.ncols, row = R, np = 0, path_tail_d = 0;
    bool only_d = true;
    auto emit = [&](char ch, int k) {   // k <= 32 ops of the same kind
        if (lane < k && np + lane < ops_cap) ops[np + lane] = (uint8_t)ch;
        np += k;
    };
    while (col > 0 && row > 0)
    {
        const bool in = (col - lane > 0) && (row - lane > 0);
        const uint32_t c = in ? codeat(col - lane, row - lane) : 0xffu;
        const unsigned stop = __ballot_sync(FULL, !in || (c >> 2) != 0);
        const int nd = stop ? (__ffs(stop) - 1) : 32;
        if (nd) { emit('M', nd); col -= nd; row -= nd; only_d = false; }
        if (nd == 32 || col <= 0 || row <= 0) continue;
        const uint32_t code = __shfl_sync(FULL, c, nd);   // first non-diagonal cell, now at (col, row)
        if (code == 0x100u) { bad_cell = 1; break; }         // must not happen: the path left the band
        if ((code >> 2) == 1)
        {
            only_d = false;
            if ((code & 2u) && row != 1)
            {
                for (;;)
                {
                    const int r = row - 1 - lane;
                    const uint32_t cc = (r >= 1) ? codeat(col, r) : 0u;
                    const unsigned st = __ballot_sync(FULL, r < 1 || !(cc & 2u) || r == 1);
                    const int k = st ? __ffs(st) : 32;
                    emit('I', k); row -= k;
                    if (st) break;
                }
            }
            --row; emit('I', 1);
        }
        else
        {
            int nD = 0;
            if ((code & 1u) && col != 1)
            {
                for (;;)
                {
                    const int cl = col - 1 - lane;
                    const uint32_t cc = (cl >= 1) ? codeat(cl, row) : 0u;
                    const unsigned st = __ballot_sync(FULL, cl < 1 || !(cc & 1u) || cl == 1);
                    const int k = st ? __ffs(st) : 32;
                    emit('D', k); col -= k; nD += k;
                    if (st) break;
                }
            }
            --col; emit('D', 1); ++nD;
            if (only_d) path_tail_d += nD;
        }
    }
    if (np > ops_cap) o.err = 3;
    if (bad_cell) o.err = 4;
    if (row > 0 && pl.c0 > 0) o.err = 1;   // must not happen: the window restriction was too tight
    const int lead_d = (row == 0) ? pl.c0 + col : 0;
    const int lead_i = row;
    const int tail_d = L - pl.end_col;
    const int lead = lead_d + lead_i;
    const int nops = lead + np + tail_d;
    o.npath = np; o.lead_i = lead_i;
    o.lead = lead_d;
    o.endpos = nops - (tail_d + path_tail_d);
    __syncwarp();
    // trailing 'I's of the ref row: only when nothing follows the path
    int trail_i = 0;
    if (tail_d == 0 && o.err == 0)
    {
        for (int base = 0; base < np; base += 32)
        {
            const int x = base + lane;
            const unsigned notI = __ballot_sync(FULL, x >= np || __ldcg(ops + x) != 'I');
            if (notI) { trail_i += __ffs(notI) - 1; break; }
            trail_i += 32;
        }
        if (trail_i > np) trail_i = np;
        if (trail_i == np) trail_i += lead_i;
    }
    const int ref_end = nops - trail_i;

    // ---- phase B: pileup records (polishing.h:54-107) -----------------------------------------
    // The reference walks the view columns [begin, end) sequentially; a run of consensus gaps ('I') is consumed
    // together with the column that follows it (even when that column is `end` itself), every other column
    // yields one record. Per record: the symbol, the length of the insertion run in front of it and the read
    // index where that run starts, and the "no insertion after this base" flag of the else-branch (:88-95).
    if (rec.code != nullptr && o.err == 0)
    {
        const int64_t ro = E[t];
        const int rcap = (int)(E[t + 1] - ro);
        const int runit = reads.off32[pl.read_id];
        const uint8_t *rq = read_quals + qual_off[pl.read_id];
        auto opat = [&](int idx) -> int {
            if (idx < lead) return (lead_i > 0) ? 'I' : 'D';
            if (idx < lead + np) return __ldcg(ops + (np - 1 - (idx - lead)));
            return 'D';
        };
        const int begin = lead;   // max(leading ref gaps, leading read gaps): one of them is zero
        const int end = (ref_end < o.endpos) ? ref_end : o.endpos;
        int i_carry = 0, q_carry = 0, run_carry = 0, n_rec = 0, over = 0;
        for (int base = begin; base <= end; base += 32)
        {
            const int c = base + lane;
            const bool valid = c <= end;
            const int op = valid ? opat(c) : 'M';
            const bool nextI = valid && (c + 1 < nops) && opat(c + 1) == 'I';
            const bool isI = valid && op == 'I', isD = valid && op == 'D';
            const unsigned mV = __ballot_sync(FULL, valid), mI = __ballot_sync(FULL, isI), mD = __ballot_sync(FULL, isD);
            const unsigned below = (1u << lane) - 1u;
            const int i_rec = i_carry + __popc(mV & ~mI & below);
            const int qc = q_carry + __popc(mV & ~mD & below);
            const unsigned x = ~mI & below;                      // non-'I' columns below this lane
            const int il = x ? (lane - 1 - (31 - __clz(x))) : (lane + run_carry);
            const bool emitrec = valid && !isI && (c < end || il > 0);
            if (emitrec)
            {
                uint32_t cd = 0x80u;
                if (il == 0 && c + 1 < nops && !nextI) cd |= 8u;
                uint8_t qual = 0;
                if (!isD)
                {
                    const int q = lead_i + qc;
                    const int src = pl.is_rc ? (R - 1 - q) : q;
                    int b = seq_base(reads, runit, src);
                    if (pl.is_rc) b = (b < 4) ? 3 - b : 4;
                    cd |= (uint32_t)b; qual = rq[src];
                }
                else
                    cd |= 5u;
                if (i_rec < rcap)
                {
                    rec.code[ro + i_rec] = (uint8_t)cd; rec.qual[ro + i_rec] = qual;
                    rec.inslen[ro + i_rec] = (uint16_t)il;
                    rec.insq[ro + i_rec] = (uint16_t)(il > 0 ? lead_i + qc - il : 0);
                }
                else
                    over = 1;
            }
            n_rec += __popc(__ballot_sync(FULL, emitrec));
            i_carry += __popc(mV & ~mI);
            q_carry += __popc(mV & ~mD);
            run_carry = (mI == FULL) ? run_carry + 32 : __clz(~mI);
        }
        if (__any_sync(FULL, over)) o.err = 2;
        o.p0 = lead_d; o.n = n_rec;
    }
    else { o.p0 = 0; o.n = 0; }
    if (lane == 0) out[t] = o;
}
}  // namespace svp
