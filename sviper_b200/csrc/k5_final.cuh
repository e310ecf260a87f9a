// K5: final alignment of the polished sequence (V, rows) against the reference part (H, columns):
// seqan::globalAlignment(gapsRef, gapsSeq, Score<double>(MM,MX,GE,GO), AlignConfig<0,0,0,0>, ConvexGaps)
// + getCigarString (/root/reference/include/sviper/sviper.h:452-466).
//
// The recurrence is the vendored fork's convex-gap formula, replayed op for op in IEEE fp64
// (seqan/align/dp_formula_convex.h:62-141, :276-380; cell = {S,H,V,Hlen,Vlen}, dp_cell_convex.h:58-67):
//   ext = X + min(GE, GO / len_prev)   (X = H or V of the predecessor, len_prev its stored gap length)
//   opn = S_prev + GO;  tie -> extension;  len = extension ? len_prev + 1 : 1
// with "infinity" = numeric_limits<double>::min()/2, a POSITIVE denormal (dp_cell.h:137-145) that the
// first row/column and never-opened gaps start from. min(GE, GO/len) comes from a table computed with the
// same IEEE division on the host (len is an integer-valued double in the reference), and the first
// row/column (Horizontal/Vertical recursion) is the same host-computed chain for every job.
// Compile with -fmad=false; denormals are never flushed (no -ftz).
//
// Mapping: one block per job, thread T owns K5_RPL consecutive rows, block-wide wavefront over columns
// (thread T works on column t-T at step t), bottom rows handed down through shared memory, one
// __syncthreads per step. Sequences longer than K5_THREADS*K5_RPL rows run as successive row bands with
// the band's bottom row kept in global memory.
// Per cell: 4 DADD (diag+sub, H+step, V+step, S+GO -- S+GO serves both the cell to the right and the cell
// below), 4 fp64 compares and 2 loads of the gap step from a shared-memory copy of the table, indexed by the
// carried gap lengths (the kernel is bound by instruction issue, not by the latency of a step's dependency chain:
// a load where the step is used is cheaper than carrying the value through fp64 selects).
// Trace: 4 bits per cell as in k1_trace.cuh, one u16 per thread and step, stored anti-diagonal-major
// (index = step * row_threads + thread) so that a warp's stores of one step are contiguous.
#pragma once
#include "common.cuh"

namespace svp
{
constexpr int K5_THREADS = 512;
constexpr int K5_RPL = 4;
constexpr int K5_BAND = K5_THREADS * K5_RPL;

struct FinalJob
{
    const uint8_t *ref;   // H, length n (device)
    const uint8_t *seq;   // V, length m (device)
    int32_t n, m;
    int64_t trace_off;    // u32 units: ((n + 1) / 2 + K5_THREADS) * row_threads
    int32_t row_threads;  // nbands * K5_THREADS
    int64_t bnd_off;      // per-column band boundary (n + 1 entries)
    int32_t valid;
};

struct FinalEnd
{
    double S, H, V;
};

struct K5Tables
{
    const double *gapstep;   // gapstep[len] = min(GE, GO / len), len >= 1
    const double *edge;      // edge[i] = score of cell (i,0) == cell (0,i)
    int32_t size;
};

// SMEM_TABLE: the whole gap-step table (tb.size + 1 doubles) sits in dynamic shared memory; otherwise it is
// read through the read-only cache (only for tables beyond ~190 KB, i.e. sequences longer than ~24,000).
template <bool SMEM_TABLE>
__global__ void __launch_bounds__(K5_THREADS) k5_final_kernel(const FinalJob *__restrict__ jobs, K5Tables tb,
                                                             double MM, double MX, double GO, double INF,
                                                             uint32_t *__restrict__ trace, double *__restrict__ bndS,
                                                             double *__restrict__ bndV, int32_t *__restrict__ bndVL,
                                                             FinalEnd *__restrict__ ends)
{
    const FinalJob job = jobs[blockIdx.x];
    if (!job.valid) return;
    const int T = threadIdx.x, n = job.n, m = job.m;
    // bottom-row handoff to the thread below, double buffered by step parity, two columns (A, B) per step
    __shared__ double shS[2][2][K5_THREADS], shV[2][2][K5_THREADS];
    __shared__ int shVL[2][2][K5_THREADS];
    extern __shared__ double k5_sg[];
    if (SMEM_TABLE)
    {
        for (int i = T; i <= tb.size; i += K5_THREADS) k5_sg[i] = tb.gapstep[i];
        __syncthreads();
    }
    // gap lengths are carried as byte offsets into the table (len * 8): a lookup is one load, no index arithmetic
    auto G = [&](int off) -> double {
        return SMEM_TABLE ? *reinterpret_cast<const double *>(reinterpret_cast<const char *>(k5_sg) + off)
                          : __ldg(reinterpret_cast<const double *>(reinterpret_cast<const char *>(tb.gapstep) + off));
    };
    uint32_t *tr = trace + job.trace_off;
    double *bS = bndS + job.bnd_off, *bV = bndV + job.bnd_off;
    int32_t *bVL = bndVL + job.bnd_off;
    const int nbands = job.row_threads / K5_THREADS;

    for (int band = 0; band < nbands; ++band)
    {
        const int row0 = band * K5_BAND + T * K5_RPL;   // rows row0+1 .. row0+K5_RPL
        double S[K5_RPL], SO[K5_RPL], H[K5_RPL];
        int HL[K5_RPL], sv[K5_RPL];   // HL: horizontal gap length * 8
#pragma unroll
        for (int k = 0; k < K5_RPL; ++k)
        {
            int row = row0 + k + 1;
            S[k] = (row <= m) ? tb.edge[row] : 0.0;
            SO[k] = __dadd_rn(S[k], GO);
            H[k] = INF; HL[k] = 8;
            sv[k] = (row <= m) ? job.seq[row - 1] : 255;
        }
        double prev_up = (row0 <= m) ? tb.edge[row0] : 0.0;   // S[row0][0]
        const bool live = row0 < m;                            // this thread owns at least one real row
        const int endk = m - 1 - row0;                         // slot of the last row, if this thread owns it
        // every step advances TWO columns (A = 2(t-T)+1, B = A+1): the fp64 dependency chains of the two columns
        // are independent (cell (k, B) only needs cell (k, A)), which doubles the instruction-level parallelism
        // and halves the number of block-wide barriers. An odd n gets one harmless extra column.
        const int nsteps = (n + 1) / 2 + K5_THREADS - 1;
        const size_t tcol = (size_t)band * K5_THREADS + T;
        for (int t = 0; t < nsteps; ++t)
        {
            const int colA = 2 * (t - T) + 1, colB = colA + 1;
            const int par = t & 1;
            if (colA >= 1 && colA <= n && live)
            {
                double uSa, uVa, uSb, uVb; int uLa, uLb;   // uL*: vertical gap length * 8
                if (T == 0)
                {
                    if (band == 0)
                    {
                        uSa = tb.edge[colA]; uSb = tb.edge[colB]; uVa = INF; uVb = INF; uLa = 8; uLb = 8;
                    }
                    else
                    {
                        uSa = bS[colA]; uVa = bV[colA]; uLa = bVL[colA];
                        uSb = bS[colB]; uVb = bV[colB]; uLb = bVL[colB];
                    }
                }
                else
                {
                    uSa = shS[par ^ 1][0][T - 1]; uVa = shV[par ^ 1][0][T - 1]; uLa = shVL[par ^ 1][0][T - 1];
                    uSb = shS[par ^ 1][1][T - 1]; uVb = shV[par ^ 1][1][T - 1]; uLb = shVL[par ^ 1][1][T - 1];
                }
                const int hcA = job.ref[colA - 1], hcB = job.ref[colB - 1];   // ref is padded behind its end
                double dA = prev_up;      // S(row above, A - 1)
                prev_up = uSb;            // S(row above, B): the diagonal of the next step's column A
                double dB = uSa;          // S(row above, A)
                double uOa = __dadd_rn(uSa, GO), uOb = __dadd_rn(uSb, GO);
                double VkA[K5_RPL], VkB[K5_RPL], HkA[K5_RPL], SkA[K5_RPL];
                uint32_t twA = 0, twB = 0;
#pragma unroll
                for (int k = 0; k < K5_RPL; ++k)
                {
                    // ---- column A: left neighbour = the state arrays (column A - 1)
                    const double interA = __dadd_rn(dA, (hcA == sv[k]) ? MM : MX);
                    dA = S[k];
                    const double hextA = __dadd_rn(H[k], G(HL[k]));
                    const bool hxA = !(hextA < SO[k]);             // tie -> extension (dp_formula.h:152-163)
                    const double hhA = hxA ? hextA : SO[k];
                    const int hlA = hxA ? HL[k] + 8 : 8;
                    const double vextA = __dadd_rn(uVa, G(uLa));
                    const bool vxA = !(vextA < uOa);
                    uVa = vxA ? vextA : uOa;
                    uLa = vxA ? uLa + 8 : 8;
                    const bool gvA = !(uVa < hhA);                 // tie -> vertical
                    const double gapA = gvA ? uVa : hhA;
                    const bool dgA = !(interA < gapA);             // tie -> diagonal
                    const double snA = dgA ? interA : gapA;
                    twA |= ((hxA ? 1u : 0u) | (vxA ? 2u : 0u) | (dgA ? 0u : (gvA ? 4u : 8u))) << (4 * k);
                    uOa = __dadd_rn(snA, GO);
                    VkA[k] = uVa; HkA[k] = hhA; SkA[k] = snA;
                    // ---- column B: left neighbour = column A of this row
                    const double interB = __dadd_rn(dB, (hcB == sv[k]) ? MM : MX);
                    dB = snA;
                    const double hextB = __dadd_rn(hhA, G(hlA));
                    const bool hxB = !(hextB < uOa);
                    const double hhB = hxB ? hextB : uOa;
                    HL[k] = hxB ? hlA + 8 : 8;
                    H[k] = hhB;
                    const double vextB = __dadd_rn(uVb, G(uLb));
                    const bool vxB = !(vextB < uOb);
                    uVb = vxB ? vextB : uOb;
                    uLb = vxB ? uLb + 8 : 8;
                    const bool gvB = !(uVb < hhB);
                    const double gapB = gvB ? uVb : hhB;
                    const bool dgB = !(interB < gapB);
                    const double snB = dgB ? interB : gapB;
                    twB |= ((hxB ? 1u : 0u) | (vxB ? 2u : 0u) | (dgB ? 0u : (gvB ? 4u : 8u))) << (4 * k);
                    VkB[k] = uVb;
                    S[k] = snB;
                    uOb = __dadd_rn(snB, GO);
                    SO[k] = uOb;
                }
                if ((colA == n || colB == n) && (unsigned)endk < (unsigned)K5_RPL)
                {
                    // the end cell (m, n): report S/H/V and apply the end-cell correction (dp_algorithm_impl.h:1212-1229)
                    const bool inA = (colA == n);
#pragma unroll
                    for (int k = 0; k < K5_RPL; ++k)
                        if (k == endk)
                        {
                            FinalEnd e;
                            e.S = inA ? SkA[k] : S[k]; e.H = inA ? HkA[k] : H[k]; e.V = inA ? VkA[k] : VkB[k];
                            ends[blockIdx.x] = e;
                            uint32_t tw = inA ? twA : twB;
                            uint32_t c = (tw >> (4 * k)) & 15u;
                            if (e.V == e.S) c = (c & 3u) | 4u;
                            else if (e.H == e.S) c = (c & 3u) | 8u;
                            tw = (tw & ~(15u << (4 * k))) | (c << (4 * k));
                            if (inA) twA = tw; else twB = tw;
                        }
                }
                shS[par][0][T] = SkA[K5_RPL - 1]; shV[par][0][T] = uVa; shVL[par][0][T] = uLa;
                shS[par][1][T] = S[K5_RPL - 1]; shV[par][1][T] = uVb; shVL[par][1][T] = uLb;
                if (T == K5_THREADS - 1 && band + 1 < nbands)
                {
                    bS[colA] = SkA[K5_RPL - 1]; bV[colA] = uVa; bVL[colA] = uLa;
                    bS[colB] = S[K5_RPL - 1]; bV[colB] = uVb; bVL[colB] = uLb;
                }
                tr[(size_t)t * job.row_threads + tcol] = twA | (twB << 16);
            }
            __syncthreads();
        }
        __threadfence_block();
        __syncthreads();
    }
}

// traceback + run-length CIGAR (seqan/bam_io/cigar.h:354-420 with the D->N rule disabled, sviper.h:466)
struct CigarOut
{
    int32_t n_cigar;
    int32_t stored;      // elements actually written (<= cap)
    double score;
};

// One warp per job. The path is followed in bulk: the 32 lanes read the next 32 cells along the current
// direction (diagonal, or the row / column of a gap being extended), a ballot finds where the run ends, and
// the whole run is emitted at once -- the number of dependent trace loads drops from one per path step to
// about one per 32. All lanes carry the same (col, row, CIGAR) state; lane 0 writes.
constexpr int K5TB_WARPS = 4;
__global__ void __launch_bounds__(K5TB_WARPS * 32) k5_traceback_kernel(
    const FinalJob *__restrict__ jobs, int njobs, const uint32_t *__restrict__ trace, const FinalEnd *__restrict__ ends,
    double INF, char *__restrict__ cig_op, uint32_t *__restrict__ cig_cnt, int cap, CigarOut *__restrict__ out)
{
    const int j = blockIdx.x * K5TB_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (j >= njobs) return;
    const FinalJob job = jobs[j];
    if (!job.valid) return;
    char *op = cig_op + (size_t)j * cap;
    uint32_t *cnt = cig_cnt + (size_t)j * cap;
    const uint32_t *tr = trace + job.trace_off;
    int ne = 0;          // elements emitted (reverse order)
    char last = 0; uint32_t run = 0;
    auto emit = [&](char c, uint32_t k) {
        if (k == 0) return;
        if (c == last) { run += k; return; }
        if (run) { if (ne < cap && lane == 0) { op[ne] = last; cnt[ne] = run; } ++ne; }
        last = c; run = k;
    };
    CigarOut o;
    int col = job.n, row = job.m;
    if (!(ends[j].S > INF))
    {
        // the scout never leaves its default cell (dp_scout.h:168-180): n x 'D' then m x 'I'
        emit('I', (uint32_t)job.m); emit('D', (uint32_t)job.n);
        col = 0; row = 0;
        o.score = INF;
    }
    else
        o.score = ends[j].S;
    auto codeat = [&](int c, int r) -> uint32_t {
        const int slot = r - 1, tg = slot / K5_RPL;
        const int t = ((c - 1) >> 1) + (tg % K5_THREADS);
        uint32_t w = tr[(size_t)t * job.row_threads + tg] >> (16 * ((c - 1) & 1));
        return (w >> (4 * (slot % K5_RPL))) & 15u;
    };
    const unsigned FULL = 0xffffffffu;
    while (col > 0 && row > 0)
    {
        // ---- diagonal run ----
        const bool in = (col - lane > 0) && (row - lane > 0);
        const uint32_t c = in ? codeat(col - lane, row - lane) : 0xffu;
        const unsigned stop = __ballot_sync(FULL, !in || (c >> 2) != 0);
        const int nd = stop ? (__ffs(stop) - 1) : 32;
        if (nd) { emit('M', (uint32_t)nd); col -= nd; row -= nd; }
        if (nd == 32 || col <= 0 || row <= 0) continue;
        uint32_t code = __shfl_sync(FULL, c, nd);   // the first non-diagonal cell, now at (col, row)
        const uint32_t sel = code >> 2;
        if (sel == 1)
        {
            // vertical gap: while the cell says "extended" and row != 1 step up; then one more step
            if ((code & 2u) && row != 1)
            {
                for (;;)
                {
                    const int r = row - 1 - lane;
                    const uint32_t cc = (r >= 1) ? codeat(col, r) : 0u;
                    // lane i holds the code read after i+1 steps; the walk stops there if it is not an extension or r == 1
                    const unsigned st = __ballot_sync(FULL, r < 1 || !(cc & 2u) || r == 1);
                    const int k = st ? __ffs(st) : 32;      // steps taken in this probe (1..32)
                    emit('I', (uint32_t)k); row -= k;
                    if (st) break;
                }
            }
            --row; emit('I', 1);
        }
        else
        {
            if ((code & 1u) && col != 1)
            {
                for (;;)
                {
                    const int cl = col - 1 - lane;
                    const uint32_t cc = (cl >= 1) ? codeat(cl, row) : 0u;
                    const unsigned st = __ballot_sync(FULL, cl < 1 || !(cc & 1u) || cl == 1);
                    const int k = st ? __ffs(st) : 32;
                    emit('D', (uint32_t)k); col -= k;
                    if (st) break;
                }
            }
            --col; emit('D', 1);
        }
    }
    emit('I', (uint32_t)row);
    emit('D', (uint32_t)col);
    if (run) { if (ne < cap && lane == 0) { op[ne] = last; cnt[ne] = run; } ++ne; }
    o.n_cigar = ne;
    int st = ne < cap ? ne : cap;
    o.stored = st;
    __syncwarp();
    // reverse the stored prefix in place (if truncated the caller only gets n_cigar right)
    if (ne <= cap && lane == 0)
        for (int a = 0, b = ne - 1; a < b; ++a, --b)
        {
            char c = op[a]; op[a] = op[b]; op[b] = c;
            uint32_t x = cnt[a]; cnt[a] = cnt[b]; cnt[b] = x;
        }
    if (lane == 0) out[j] = o;
}
}  // namespace svp
