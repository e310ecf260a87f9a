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
// plan: strand choice (mapping.h:121), mapQ, and the traced column window.
// Any alignment of the read ending in the last row at column e with score >= s spans at most
// R + max(0, R - s - 5) columns (m matches <= R, D deleted columns cost >= 5 + D), so the sub-matrix
// starting at c0 = e - R - max(0, R - s) - 8 contains the reference's traceback path and every path
// that ties with it at the end cell; cells on that path keep their values, hence all trace decisions
// on the path are identical to the full matrix. full_window != 0 disables the restriction.
// ---------------------------------------------------------------------------------------------
__global__ void plan_kernel(const int32_t *__restrict__ task_read, const int32_t *__restrict__ task_win, int ntasks,
                            const int32_t *__restrict__ read_len, const ScoreOut *__restrict__ sc,
                            const uint8_t *__restrict__ win_done, TracePlan *__restrict__ plan,
                            int64_t *__restrict__ units, int full_window)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntasks) return;
    ScoreOut o = sc[t];
    TracePlan p;
    p.read_id = task_read[t]; p.win_id = task_win[t];
    if (win_done && win_done[p.win_id])
    {
        p.is_rc = 0; p.score = 0; p.end_col = 0; p.c0 = 0; p.ncols = -1; p.trace_off = 0;
        plan[t] = p; units[t] = 0;
        return;
    }
    p.is_rc = !(o.score_f > o.score_r);
    p.score = p.is_rc ? o.score_r : o.score_f;
    p.end_col = p.is_rc ? o.col_r : o.col_f;
    int R = read_len[p.read_id];
    int slack = R - p.score; if (slack < 0) slack = 0;
    int c0 = p.end_col - R - slack - 8;
    if (c0 < 0 || full_window) c0 = 0;
    p.c0 = c0; p.ncols = p.end_col - c0;
    p.trace_off = 0;
    plan[t] = p;
    units[t] = p.ncols + 2;   // trace columns / pileup records / (+ read length) path ops reserved for this task
}

template <int RPL>
struct KTShape
{
    static constexpr int NSLOT = 32 * RPL;
    static constexpr int PROF_WORDS = 5 * NSLOT;
    static constexpr int SMEM_BYTES = KT_WARPS * PROF_WORDS * 4;
};

template <int RPL>
__global__ void __launch_bounds__(KT_WARPS * 32) k1_trace_kernel(const TracePlan *__restrict__ plans, int ntasks,
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
    if (slot >= ntasks) return;
    const int task = task_idx ? task_idx[slot] : slot;
    int32_t *prof = kt_smem + warp * Sh::PROF_WORDS;   // prof[c][k][lane]
    const TracePlan pl = plans[task];
    const int ncols = pl.ncols;
    if (ncols <= 0) return;                            // end cell in the initial column: no DP needed
    const int R = reads.len[pl.read_id], runit = reads.off32[pl.read_id];
    const int wunit = wins.off32[pl.win_id];
    const int pad = NSLOT - R;
    uint32_t *tr = trace + E[task] * 32;

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
// K2: traceback (one thread per alignment) + the per-reference-position records the pileup consumes.
// Literal restatement of add_read_to_profile's bookkeeping (include/sviper/polishing.h:54-107) on the
// view columns of the alignment; see oracle/oracle_polish.c:orc_add_read for the same logic on the CPU.
// ---------------------------------------------------------------------------------------------
struct AlnOut
{
    int32_t lead;      // gapsBeginPos(gapsRead)
    int32_t endpos;    // gapsEndPos(gapsRead)
    int32_t npath;     // path ops stored (reverse order) in the ops buffer
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

template <int RPL>
__device__ __forceinline__ uint32_t kt_code(const uint32_t *tr, int pad, int col, int row)
{
    int slot = pad + row - 1;
    uint32_t w = tr[(size_t)(col - 1) * 32 + slot / RPL];
    return (w >> (4 * (slot % RPL))) & 15u;
}

template <int RPL>
__global__ void k2_traceback_kernel(const TracePlan *__restrict__ plans, int ntasks,
                                    const int32_t *__restrict__ task_idx, SeqStore reads, SeqStore wins,
                                    const uint8_t *__restrict__ read_quals, const int64_t *__restrict__ qual_off,
                                    const uint32_t *__restrict__ trace, uint8_t *__restrict__ ops_buf,
                                    RecArrays rec, const int64_t *__restrict__ E, AlnOut *__restrict__ out)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntasks) return;
    if (task_idx) t = task_idx[t];
    const TracePlan pl = plans[t];
    if (pl.ncols < 0) return;   // window already converged
    const int R = reads.len[pl.read_id], L = wins.len[pl.win_id];
    const int pad = 32 * RPL - R;
    const uint32_t *tr = trace + E[t] * 32;
    uint8_t *ops = ops_buf + E[t] + (int64_t)t * SVP_MAX_READ_LEN;      // cap: ncols + 2 + max read length
    const int64_t ops_cap = E[t + 1] - E[t] + SVP_MAX_READ_LEN;
    AlnOut o;
    o.err = 0; o.is_rc = pl.is_rc; o.mapq = pl.score > 0 ? pl.score : 0;

    int col = pl.ncols, row = R, np = 0, path_tail_d = 0;
    bool only_d = true;
    while (col > 0 && row > 0)
    {
        uint32_t code = kt_code<RPL>(tr, pad, col, row);
        uint32_t sel = code >> 2;
        if (sel == 0)
        {
            if (np < ops_cap) ops[np] = 'M'; ++np; --col; --row; only_d = false;
        }
        else if (sel == 1)
        {
            only_d = false;
            if (code & 2u)
            {
                while ((code & 2u) && row != 1)
                {
                    --row; code = kt_code<RPL>(tr, pad, col, row);
                    if (np < ops_cap) ops[np] = 'I'; ++np;
                }
            }
            --row;
            if (np < ops_cap) ops[np] = 'I'; ++np;
        }
        else
        {
            if (code & 1u)
            {
                while ((code & 1u) && col != 1)
                {
                    --col; code = kt_code<RPL>(tr, pad, col, row);
                    if (np < ops_cap) ops[np] = 'D'; ++np;
                    if (only_d) ++path_tail_d;
                }
            }
            --col;
            if (np < ops_cap) ops[np] = 'D'; ++np;
            if (only_d) ++path_tail_d;
        }
    }
    if (np > ops_cap) o.err = 3;
    if (row > 0 && pl.c0 > 0) o.err = 1;   // must not happen: the window restriction was too tight
    const int lead_d = (row == 0) ? pl.c0 + col : 0;
    const int lead_i = row;
    const int tail_d = L - pl.end_col;
    const int lead = lead_d + lead_i;
    const int nops = lead + np + tail_d;
    o.npath = np; o.lead_i = lead_i;
    o.lead = lead_d;
    o.endpos = nops - (tail_d + path_tail_d);
    // trailing 'I's of the ref row: only when nothing follows the path
    int trail_i = 0;
    if (tail_d == 0) { while (trail_i < np && ops[trail_i] == 'I') ++trail_i; if (trail_i == np) trail_i += lead_i; }
    const int ref_end = nops - trail_i;

    if (rec.code != nullptr && o.err == 0)
    {
        // forward pass over the view columns [begin, end)
        const int64_t ro = E[t];
        const int64_t rcap = E[t + 1] - ro;
        const int runit = reads.off32[pl.read_id];
        const uint8_t *rq = read_quals + qual_off[pl.read_id];
        auto opat = [&](int idx) -> int {
            if (idx < lead) return (lead_i > 0) ? 'I' : 'D';
            if (idx < lead + np) return ops[np - 1 - (idx - lead)];
            return 'D';
        };
        auto refgap = [&](int idx) -> bool {
            if (idx < nops) return opat(idx) == 'I';
            return opat(nops - 1) != 'I';   // seqan isGap() at position == length (gaps_array.h:749-783)
        };
        const int begin = lead;   // max(leading ref gaps, leading read gaps): one of them is zero
        const int end = (ref_end < o.endpos) ? ref_end : o.endpos;
        int p = lead_d, q = lead_i;
        o.p0 = p;
        int i = 0;
        for (int idx = begin; idx < end; ++idx)
        {
            uint32_t cd = 0x80u;
            uint16_t il = 0, iq = 0;
            if (opat(idx) == 'I')
            {
                iq = (uint16_t)q;
                while (refgap(idx)) { ++il; ++q; ++idx; }
            }
            else if (!refgap(idx + 1))
                cd |= 8u;
            uint8_t qual = 0;
            if (opat(idx) != 'D')
            {
                int src = pl.is_rc ? (R - 1 - q) : q;
                int b = seq_base(reads, runit, src);
                if (pl.is_rc) b = (b < 4) ? 3 - b : 4;
                cd |= (uint32_t)b; qual = rq[src]; ++q;
            }
            else
                cd |= 5u;
            if (i < rcap)
            {
                rec.code[ro + i] = (uint8_t)cd; rec.qual[ro + i] = qual;
                rec.inslen[ro + i] = il; rec.insq[ro + i] = iq;
            }
            ++i; ++p;
        }
        if (i > rcap) o.err = 2;
        o.n = i;
    }
    else { o.p0 = 0; o.n = 0; }
    out[t] = o;
}
}  // namespace svp
