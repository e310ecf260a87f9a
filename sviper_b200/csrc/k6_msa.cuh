// K6: the two pairwise DP fills of the long-read consensus.
//
// sviper::build_consensus (/root/reference/include/sviper/basics.h:484-520) is seqan::globalMsaAlignment
// (graph_msa/graph_align_tcoffee_msa.h:323-548) + a column vote. Measured with the unmodified reference
// (profiles/r2_msa_profile.jsonl): ~80 % of its time is the LOCAL pairwise library -- up to four rounds of the legacy
// "clumping" Smith-Waterman per sequence pair (graph_align_tcoffee_library.h:252-390, :431-462) -- and ~10 % the GLOBAL
// pairwise library (Gotoh, global_alignment_unbanded.h:386-407). Both are O(n x m) matrix fills followed by an O(n + m)
// traceback. These kernels do the fills and produce the trace matrices in SeqAn's own byte layouts, so that the host can
// hand them to SeqAn's UNCHANGED traceback code (_alignSmithWatermanTrace, _computeTraceback); everything after that
// (alignment graph, guide tree, triplet extension, progressive alignment, vote) stays the reference's host code.
//
// Mapping: one block per (pair, fill). Thread t owns MSA_RPT consecutive rows (sequence v) and sweeps the columns
// (sequence h) as a block-wide wavefront: at step s it works on column s - t + 1 and hands its bottom row to thread t + 1
// through shared memory (one barrier per step). Sequences longer than MSA_BAND rows run as successive row bands, the band's
// bottom row kept in global memory. int32 scores (5 x 7,000 does not fit 16 bits). oracle/oracle_msa.c is the CPU
// restatement these are compared with byte for byte (tests/test_gpu_msa.py).
#pragma once
#include "common.cuh"

namespace svp
{
constexpr int MSA_THREADS = 256;
constexpr int MSA_RPT = 8;
constexpr int MSA_BAND = MSA_THREADS * MSA_RPT;
constexpr int MSA_NEG = INT32_MIN / 2;          // seqan/align/dp_cell.h:144-145

struct MsaJob
{
    const uint8_t *h;        // columns: str[0], length n
    const uint8_t *v;        // rows:    str[1], length m
    int32_t n, m;
    uint8_t *trace;          // local: n * ((m + 1) / 2) bytes; global: (n + 1) * (m + 1) bytes
    int32_t *bndS, *bndV;    // n + 1 entries each: bottom row of the previous band
    const uint32_t *forb;    // local: bit (col - 1) * m + (row - 1) set = cell lies on an earlier local alignment; may be null
    int32_t local;           // 1: clumping Smith-Waterman, 0: global Gotoh
    int32_t out;             // index of the result record
};

struct MsaOut
{
    int32_t score;           // local: best score; global: score of the bottom-right cell
    int32_t row, col, dir;   // local: end cell (indexPair[0], indexPair[1]) and its trace value (initialDir)
    int32_t S, H, V;         // global: the bottom-right cell (what _correctTraceValue looks at)
    int32_t pad_;
};

// SeqAn trace bits (align/dp_profile.h:125-134)
enum : uint32_t { TB_DIAG = 1, TB_H = 2, TB_V = 4, TB_HO = 8, TB_VO = 16, TB_MAXH = 32, TB_MAXV = 64 };

__global__ void __launch_bounds__(MSA_THREADS) msa_fill_kernel(const MsaJob *__restrict__ jobs, MsaOut *__restrict__ outs,
                                                              int mt, int mm, int ge, int go)
{
    const MsaJob job = jobs[blockIdx.x];
    const int T = threadIdx.x, n = job.n, m = job.m;
    const bool local = job.local != 0;
    __shared__ int shS[2][MSA_THREADS], shV[2][MSA_THREADS];
    __shared__ int shBest[MSA_THREADS / 32][4];
    const int nbands = (m + MSA_BAND - 1) / MSA_BAND;
    const size_t numRows = ((size_t)m + 1) >> 1;        // local: trace bytes per column
    const size_t ld = (size_t)m + 1;                    // global: trace bytes per column
    int best = 0, brow = 0, bcol = 0, bdir = 3;

    if (!local && T == 0) job.trace[0] = 0;             // cell (0, 0): NONE
    for (int band = 0; band < nbands; ++band)
    {
        const int row0 = band * MSA_BAND + T * MSA_RPT;  // this thread owns rows row0 + 1 .. row0 + MSA_RPT
        const bool live = row0 < m;
        int S[MSA_RPT], G[MSA_RPT];                      // S of the previous column; G = horizontal-gap matrix
        uint8_t vb[MSA_RPT];
#pragma unroll
        for (int k = 0; k < MSA_RPT; ++k)
        {
            const int row = row0 + k + 1;
            vb[k] = row <= m ? job.v[row - 1] : 255;
            if (local) { S[k] = 0; G[k] = go - ge; }     // library.h:304-312
            else
            {
                S[k] = go + (row - 1) * ge;              // Vertical recursion down column 0 (dp_formula_affine.h:291-325)
                G[k] = MSA_NEG;
                if (row <= m) job.trace[row] = (uint8_t)((row == 1 ? TB_VO : TB_V) | TB_MAXV);
            }
        }
        // S[row0][0]: the diagonal of this thread's first row at column 1
        int prev_up = local ? 0 : (row0 == 0 ? 0 : go + (row0 - 1) * ge);
        const int nsteps = n + MSA_THREADS - 1;
        for (int s = 0; s < nsteps; ++s)
        {
            const int col = s - T + 1, par = s & 1;
            if (col >= 1 && col <= n && live)
            {
                int upS, upV;
                if (T == 0)
                {
                    if (band == 0)
                    {
                        if (local) { upS = 0; upV = go - ge; }                       // mat[0] = 0, vert = open - extend (library.h:316-319)
                        else
                        {
                            upS = go + (col - 1) * ge; upV = MSA_NEG;                // Horizontal recursion along row 0
                            job.trace[(size_t)col * ld] = (uint8_t)((col == 1 ? TB_HO : TB_H) | TB_MAXH);
                        }
                    }
                    else { upS = job.bndS[col]; upV = job.bndV[col]; }
                }
                else { upS = shS[par ^ 1][T - 1]; upV = shV[par ^ 1][T - 1]; }
                const int hc = job.h[col - 1];
                int diag = prev_up;
                prev_up = upS;
                if (local)
                {
                    uint32_t nib[MSA_RPT];
#pragma unroll
                    for (int k = 0; k < MSA_RPT; ++k)
                    {
                        const int row = row0 + k + 1;
                        nib[k] = 0;
                        if (row > m) continue;
                        int max_val, tv;
                        bool forbidden = false;
                        if (job.forb)
                        {
                            const size_t bit = (size_t)(col - 1) * (size_t)m + (size_t)(row - 1);
                            forbidden = (job.forb[bit >> 5] >> (bit & 31)) & 1u;
                        }
                        if (forbidden) { nib[k] = 3; max_val = 0; tv = 3; upV = 0; G[k] = 0; }   // library.h:322-329
                        else
                        {
                            int a = upS + go, b = upV + ge;                                      // vertical: open only if strictly better
                            const uint32_t vo = a > b;
                            upV = vo ? a : b;
                            a = S[k] + go; b = G[k] + ge;
                            const uint32_t ho = a > b;
                            G[k] = ho ? a : b;
                            max_val = diag + (hc == vb[k] ? mt : mm); tv = 0;
                            if (upV > max_val) { max_val = upV; tv = 2; }
                            if (G[k] > max_val) { max_val = G[k]; tv = 1; }
                            if (0 >= max_val) { max_val = 0; tv = 3; }
                            nib[k] = (vo << 3) | (ho << 2) | (uint32_t)tv;
                        }
                        diag = S[k]; S[k] = max_val; upS = max_val;
                        // first maximum in the reference's visiting order: column by column, rows ascending (library.h:364-369)
                        if (max_val > best || (max_val == best && max_val > 0 && (col < bcol || (col == bcol && row < brow))))
                        { best = max_val; brow = row; bcol = col; bdir = tv; }
                    }
                    uint8_t *tc = job.trace + (size_t)(col - 1) * numRows + (size_t)(row0 >> 1);
#pragma unroll
                    for (int k = 0; k < MSA_RPT / 2; ++k)
                        if ((size_t)(row0 >> 1) + k < numRows) tc[k] = (uint8_t)((nib[2 * k] << 4) | nib[2 * k + 1]);
                }
                else
                {
                    uint8_t *tc = job.trace + (size_t)col * ld + row0 + 1;
#pragma unroll
                    for (int k = 0; k < MSA_RPT; ++k)
                    {
                        const int row = row0 + k + 1;
                        if (row > m) continue;
                        const int inter = diag + (hc == vb[k] ? mt : mm);
                        const int hext = G[k] + ge, hopn = S[k] + go;
                        uint32_t t = (hext < hopn) ? TB_HO : TB_H;                               // tie -> extension
                        const int hh = (hext < hopn) ? hopn : hext;
                        const int vext = upV + ge, vopn = upS + go;
                        t |= (vext < vopn) ? TB_VO : TB_V;
                        upV = (vext < vopn) ? vopn : vext;
                        const uint32_t t2 = (upV < hh) ? TB_MAXH : TB_MAXV;                      // tie -> vertical
                        const int gap = (upV < hh) ? hh : upV;
                        diag = S[k];
                        int sn;
                        if (inter < gap) { sn = gap; t |= t2; } else { sn = inter; t |= TB_DIAG; }   // tie -> diagonal
                        S[k] = sn; G[k] = hh; upS = sn;
                        tc[k] = (uint8_t)t;
                        if (row == m && col == n) { MsaOut o; o.score = sn; o.row = m; o.col = n; o.dir = 0; o.S = sn; o.H = hh; o.V = upV; o.pad_ = 0; outs[job.out] = o; }
                    }
                }
                shS[par][T] = upS; shV[par][T] = upV;
                if (T == MSA_THREADS - 1 && band + 1 < nbands) { job.bndS[col] = upS; job.bndV[col] = upV; }
            }
            __syncthreads();
        }
        __threadfence_block();
        __syncthreads();
    }
    if (local)
    {
        // block-wide: largest score, then smallest column, then smallest row
        auto better = [](int v1, int c1, int r1, int v2, int c2, int r2) { return v1 > v2 || (v1 == v2 && (c1 < c2 || (c1 == c2 && r1 < r2))); };
        if (best == 0) { bcol = INT32_MAX; brow = INT32_MAX; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            const int v2 = __shfl_xor_sync(0xffffffffu, best, o), c2 = __shfl_xor_sync(0xffffffffu, bcol, o);
            const int r2 = __shfl_xor_sync(0xffffffffu, brow, o), d2 = __shfl_xor_sync(0xffffffffu, bdir, o);
            if (better(v2, c2, r2, best, bcol, brow)) { best = v2; bcol = c2; brow = r2; bdir = d2; }
        }
        if ((T & 31) == 0) { shBest[T >> 5][0] = best; shBest[T >> 5][1] = bcol; shBest[T >> 5][2] = brow; shBest[T >> 5][3] = bdir; }
        __syncthreads();
        if (T == 0)
        {
            for (int w = 1; w < MSA_THREADS / 32; ++w)
                if (better(shBest[w][0], shBest[w][1], shBest[w][2], best, bcol, brow))
                { best = shBest[w][0]; bcol = shBest[w][1]; brow = shBest[w][2]; bdir = shBest[w][3]; }
            MsaOut o;
            o.score = best; o.row = best ? brow : 0; o.col = best ? bcol : 0; o.dir = best ? bdir : 3; o.S = o.H = o.V = 0; o.pad_ = 0;
            outs[job.out] = o;
        }
    }
}

// marks the cells of a local alignment found by the host's traceback as forbidden for the next round
// (_setForbiddenCell, align/local_alignment_banded_waterman_eggert_impl.h:177-186); cells = (col, row) pairs, 1-based
__global__ void msa_mark_kernel(uint32_t *__restrict__ forb, const int32_t *__restrict__ cells, int ncells, int m)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncells) return;
    const size_t bit = (size_t)(cells[2 * i] - 1) * (size_t)m + (size_t)(cells[2 * i + 1] - 1);
    atomicOr(forb + (bit >> 5), 1u << (bit & 31));
}
}  // namespace svp
