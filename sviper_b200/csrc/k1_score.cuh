// K1: batched semi-global affine DP, SCORE PASS (both strands of one read packed as s16x2).
//
// Replaces the two seqan::globalAlignment calls of sviper::map_single_read
// (/root/reference/include/sviper/mapping.h:111-119) as far as score, strand choice (mapping.h:121)
// and the end cell (leftmost last-row maximum, seqan/align/dp_scout.h:168-180) are concerned.
// The traceback of the winning strand is redone by the traced pass (k1_trace.cuh).
//
// Mapping: G = 8 lanes (one "group") per (read, window) task, 4 tasks per warp (G = 32: one task per warp). Lane g owns RPL
// consecutive read rows and sweeps the window columns as a wavefront: at step t it computes column
// t-g and hands its bottom row (S+go, V) to lane g+1 with one shuffle pair. The low s16 half carries
// the forward strand, the high half the reverse complement, so every DPX instruction advances two
// cells. Rows are aligned to the END of the group: unused rows at the top are "extra row 0s"
// (substitution score 0 keeps S == 0 there, which is exactly the free-begin row of
// AlignConfig<true,false,false,true>), so the read's last row is always the last slot of lane 7.
//
// Per cell pair: 1 LDS (query profile), IMAD (diag + sub), VIADDMNMX (H), VIADDMNMX (V), VIMNMX3 (S),
// IMAD (S + go) -- no trace, no compares. Three of the six sit on the 16-lane ALU pipe, which is the bound.
// The recurrence chains three dependent instructions down a lane's rows (V -> S -> S+go -> V of the next
// row), so every step advances TWO window columns: cell (k, a) and cell (k-1, b = a+1) are independent, the
// unrolled code interleaves the two chains, and per-step overhead (shuffles, range checks, window decode)
// is paid once per column pair. A window of odd length gets one harmless extra column that is never scored.
// Values are biased by +0x2000 so that both halves stay positive and plain 32-bit adds never carry between halves.
#pragma once
#include "common.cuh"

namespace svp
{
constexpr int K1_G = 8;                // lanes per task of the throughput shape (16 tasks per 128-thread block)
constexpr int K1_GW = 32;              // lanes per task of the WIDE shape: a third of the steps per task, more
                                       // per-step overhead. Used when few tasks are left (tail rounds), where the
                                       // kernel's duration is the latency of one task, not the chip's throughput.
constexpr int K1_THREADS = 128;         // default block size; launches of less than ~2 waves use 64-thread blocks (8 tasks),
                                        // which spread an uneven number of blocks more evenly over the 148 SMs
constexpr int K1_BIAS = 0x2000;
constexpr int K1_WEAK = 24;             // reads whose best strand scores below this (mostly reads hanging off a window edge,
                                        // score < 0: the certificate needs S >= 0) are not tried as mixed tasks in the next
                                        // round: they go to the regular list up front, which runs BESIDE the mixed launch

__host__ __device__ constexpr uint32_t k1_pk(int x) { return (uint32_t)((K1_BIAS + x) & 0xffff) * 0x00010001u; }

template <int G, int RPL, int THREADS = K1_THREADS>
struct K1Shape
{
    static constexpr int TASKS_PER_BLOCK = THREADS / G;
    static constexpr int NSLOT = G * RPL;
    // a lane's RPL profile entries of one window character are contiguous (fetched with LDS.128); the lane
    // stride KPAD is a multiple of 4 words with KPAD/4 odd, so the 8 lanes of a task (= one quarter-warp, the
    // conflict granularity of 128-bit loads) hit 8 distinct bank groups
    static constexpr int KPAD = ((RPL + 3) / 4 * 4) + ((((RPL + 3) / 4) & 1) ? 0 : 4);
    static constexpr int PST = G * KPAD;               // words per profile row (one window character); multiple of 32
    static constexpr int TASK_WORDS = 5 * PST + 8;
    static constexpr int SMEM_BYTES = TASKS_PER_BLOCK * TASK_WORDS * 4;
    static_assert(PST % 32 == 0 && TASK_WORDS % 4 == 0, "profile rows must keep bank phase and 16-byte alignment");
};

// Rounds >= 2 of a polish batch (MIXED): the strand of a read is known from an earlier round, together with an upper bound
// on the score of its OTHER strand, so the two halves of the s16x2 lanes carry the known strands of the two mates of a
// pair instead of the two strands of one read (half the tasks). Exactness: a consensus edit changes the score of an
// alignment that crosses it by at most 7 (substitution: 1 - (-4) = 5; deleted column: the path takes it as a new
// 1-column gap, 6; inserted column aligned to a read base: that base becomes a 1-base gap instead of a match, 7), so for
// every alignment P' of the other strand to the NEW consensus that spans at most W columns there is one to the OLD
// consensus with score >= score(P') - E, E = the largest sum of these costs over any W consecutive columns (k3 computes it
// per window for W = 320 and 512). An alignment with score >= S >= 0 has at most R - S - 5 <= R deleted columns, i.e.
// spans at most 2R <= W columns. Hence with bound' = max(bound + E, -6) (a path longer than W scores < -5):
// S >= 0 and S > bound' proves that no alignment of the other strand reaches S -- mapping.h:121 picks the same strand --
// and bound' is the other strand's bound for the next round. Reads that fail the test are appended to the regular
// (both strands) list of the same round.
struct K1Mix
{
    const int32_t *n_dev;      // number of tasks (pairs if MIXED) of this launch, on the device; null: `ntasks`
    int32_t *sstate;           // per task: bit 0 strand known, bit 1 is_rc, bit 2 weak (winner scored < K1_WEAK: its certificate
                               // would probably fail again), bits 3..12 the winner's last score (clamped to 0..1023) and
                               // bit 13 "that score comes from a mixed round" (both only used to PREDICT the certificate,
                               // never to decide it), bits 16..31 bound on the other strand's score
    const int32_t *emax;       // per window: edit cost E for W = 320, W = 512
    int32_t *fb19, *fb32;      // lists the failed certificates are appended to (both strands again, same round)
    int32_t *fb_cnt;           // their counters [2]
    unsigned long long *cells; // DP cells computed (statistics)
};

__device__ __forceinline__ uint32_t k1_pk2(int a, int b)
{
    return (uint32_t)((K1_BIAS + a) & 0xffff) | ((uint32_t)((K1_BIAS + b) & 0xffff) << 16);
}

template <int G, int RPL, bool MIXED, int THREADS>
__global__ void __launch_bounds__(THREADS, 512 / THREADS) k1_score_kernel(const int32_t *__restrict__ task_read,
                                                       const int32_t *__restrict__ task_win,
                                                       const int32_t *__restrict__ task_idx, int ntasks,
                                                       const uint8_t *__restrict__ win_done,
                                                       SeqStore reads, SeqStore wins,
                                                       const uint8_t *__restrict__ win_bytes,
                                                       const int64_t *__restrict__ win_byte_off,
                                                       ScoreOut *__restrict__ out,
                                                       uint32_t one /* always 1; opaque to the compiler */, K1Mix mix)
{
    using Sh = K1Shape<G, RPL, THREADS>;
    constexpr int NSLOT = Sh::NSLOT, PST = Sh::PST, KPAD = Sh::KPAD, NQ = (RPL + 3) / 4;
    extern __shared__ uint32_t k1_smem[];
    const int g = threadIdx.x & (G - 1), grp = threadIdx.x / G;
    const int slot = blockIdx.x * Sh::TASKS_PER_BLOCK + grp;
    uint32_t *prof = k1_smem + grp * Sh::TASK_WORDS;   // prof[c][g][k]

    if (mix.n_dev)
    {
        ntasks = *mix.n_dev;                                          // the grid was sized for an upper bound
        if (blockIdx.x * Sh::TASKS_PER_BLOCK >= ntasks) return;       // (whole block: no shuffle partner is left behind)
    }
    int R = 0, L = 0, runit = 0, task = -1;
    int R2 = 0, runit2 = 0, task2 = -1, rcA = 0, rcB = 0, wid = 0;   // MIXED: second read of the pair, known strands
    const uint8_t *wb = win_bytes;   // the window as bytes 0..4 (zero padded behind its end)
    if (slot < ntasks)
    {
        if (MIXED) { task = task_idx[2 * slot]; task2 = task_idx[2 * slot + 1]; }
        else task = task_idx ? task_idx[slot] : slot;
        int rid = task_read[task];
        wid = task_win[task];
        if (win_done && win_done[wid]) task = -1;       // window already converged: nothing to do
        else
        {
            R = reads.len[rid]; runit = reads.off32[rid];
            L = wins.len[wid]; wb = win_bytes + win_byte_off[wid];
            if (MIXED)
            {
                const int rid2 = task_read[task2];
                R2 = reads.len[rid2]; runit2 = reads.off32[rid2];
                rcA = (mix.sstate[task] >> 1) & 1; rcB = (mix.sstate[task2] >> 1) & 1;
            }
        }
    }
    const bool live = task >= 0;
    const int pad = NSLOT - R;   // extra "row 0"s at the top
    const int pad2 = NSLOT - R2;

    constexpr uint32_t GE2 = 0xFFFFFFFFu;           // (-1,-1)
    constexpr uint32_t GO2 = 0u - 0x00060006u;      // 32-bit add of this subtracts 6 from both halves (no borrow)
    constexpr uint32_t NEG2 = 0x00010001u;          // "-infinity": biased 1
    constexpr uint32_t ROW0_SO = k1_pk(GAP_OPEN);   // S == 0 plus gap open

    uint32_t so[RPL], h[RPL];
    // ---- query profile + initial column (non-free: S[i][0] = go + (i-1)*ge, dp_formula_affine.h:291-325)
#pragma unroll
    for (int k = 0; k < RPL; ++k)
    {
        int row = g * RPL + k - pad + 1;   // 1-based read row; <= 0 -> pad
        int row2 = MIXED ? g * RPL + k - pad2 + 1 : row;   // MIXED: the high half is another read
        int rf = -1, rr = -1;
        if (!MIXED)
        {
            if (row >= 1)
            {
                rf = seq_base(reads, runit, row - 1);
                int b = seq_base(reads, runit, R - row);
                rr = (b < 4) ? 3 - b : 4;
            }
        }
        else
        {
            // low half: read A on its known strand, high half: read B on its known strand
            if (row >= 1)
            {
                int b = seq_base(reads, runit, rcA ? R - row : row - 1);
                rf = rcA ? ((b < 4) ? 3 - b : 4) : b;
            }
            if (row2 >= 1)
            {
                int b = seq_base(reads, runit2, rcB ? R2 - row2 : row2 - 1);
                rr = rcB ? ((b < 4) ? 3 - b : 4) : b;
            }
        }
#pragma unroll
        for (int c = 0; c < 5; ++c)
        {
            uint32_t sf = (row < 1) ? 6u : ((c == rf) ? 7u : 2u);   // sub - go
            uint32_t sr = (row2 < 1) ? 6u : ((c == rr) ? 7u : 2u);
            prof[c * PST + g * KPAD + k] = sf | (sr << 16);
        }
        int s0 = (row >= 1) ? -(5 + row) : 0;
        int s02 = (row2 >= 1) ? -(5 + row2) : 0;
        so[k] = MIXED ? k1_pk2(s0 + GAP_OPEN, s02 + GAP_OPEN) : k1_pk(s0 + GAP_OPEN);
        h[k] = NEG2;
    }
    uint32_t prev_up;   // S[i-1][j-1] + go for this lane's first row
    {
        int rowa = g * RPL - pad;   // row above the first slot
        int sa = (rowa >= 1) ? -(5 + rowa) : 0;
        int rowa2 = g * RPL - pad2;
        int sa2 = (rowa2 >= 1) ? -(5 + rowa2) : 0;
        prev_up = (g == 0) ? ROW0_SO : (MIXED ? k1_pk2(sa + GAP_OPEN, sa2 + GAP_OPEN) : k1_pk(sa + GAP_OPEN));
    }
    uint32_t out_so0 = so[RPL - 1], out_v0 = NEG2, out_so1 = so[RPL - 1], out_v1 = NEG2;
    uint32_t best = so[RPL - 1];
    int colF = 0, colR = 0;
    // window characters of the lane's next column pair, fetched one step ahead (a lane's active steps are consecutive)
    uint32_t ca = 0, cb = 0;
    if (task >= 0) { ca = __ldg(wb); cb = __ldg(wb + 1); }
    // lane 0 of a task takes the free first row (S = 0, V = -inf) instead of a neighbour's bottom row; the blend
    // r * keep + row0 is an IMAD (FMA pipe) where a select would sit on the ALU pipe
    const uint32_t keep = min((uint32_t)g, one);   // 0 for lane 0, else 1 (through the opaque `one`, so that it stays a multiply)
    const uint32_t row0_so = (g == 0) ? ROW0_SO : 0u, row0_v = (g == 0) ? NEG2 : 0u;

    const int nsteps = __reduce_max_sync(0xffffffffu, live ? (L + 1) / 2 + G - 1 : 0);
    for (int t = 0; t < nsteps; ++t)
    {
        uint32_t r_so0 = __shfl_up_sync(0xffffffffu, out_so0, 1, G);
        uint32_t r_v0 = __shfl_up_sync(0xffffffffu, out_v0, 1, G);
        uint32_t r_so1 = __shfl_up_sync(0xffffffffu, out_so1, 1, G);
        uint32_t r_v1 = __shfl_up_sync(0xffffffffu, out_v1, 1, G);
        const int jj = 2 * (t - g);   // window columns jj (a) and jj + 1 (b)
        if (jj >= 0 && jj < L)
        {
            r_so0 = r_so0 * keep + row0_so; r_v0 = r_v0 * keep + row0_v;
            r_so1 = r_so1 * keep + row0_so; r_v1 = r_v1 * keep + row0_v;
            const uint32_t na = __ldg(wb + jj + 2), nb = __ldg(wb + jj + 3);   // next pair (zero padding behind the end)
            uint32_t pa[NQ * 4], pb[NQ * 4];
            {
                const uint4 *qa = reinterpret_cast<const uint4 *>(prof + ca * PST + g * KPAD);
                const uint4 *qb = reinterpret_cast<const uint4 *>(prof + cb * PST + g * KPAD);
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                {
                    const uint4 x = qa[q], y = qb[q];
                    pa[4 * q] = x.x; pa[4 * q + 1] = x.y; pa[4 * q + 2] = x.z; pa[4 * q + 3] = x.w;
                    pb[4 * q] = y.x; pb[4 * q + 1] = y.y; pb[4 * q + 2] = y.z; pb[4 * q + 3] = y.w;
                }
            }
            uint32_t dA = prev_up;        // So(row above, a - 1)
            prev_up = r_so1;              // So(row above, b): the diagonal of the next step's column a
            uint32_t dB = r_so0;          // So(row above, a)
            uint32_t uA = r_so0, vA = r_v0, uB = r_so1, vB = r_v1;
#pragma unroll
            for (int k = 0; k < RPL; ++k)
            {
                // column a
                const uint32_t interA = dA + pa[k];
                dA = so[k];
                const uint32_t hA = __viaddmax_s16x2(h[k], GE2, so[k]);
                vA = __viaddmax_s16x2(vA, GE2, uA);
                uA = __vimax3_s16x2(interA, hA, vA) * one + GO2;   // IMAD with a runtime 1: keeps this add off the ALU pipe
                // column b (its chain runs one row behind column a's)
                const uint32_t interB = dB + pb[k];
                dB = uA;
                const uint32_t hB = __viaddmax_s16x2(hA, GE2, uA);
                vB = __viaddmax_s16x2(vB, GE2, uB);
                uB = __vimax3_s16x2(interB, hB, vB) * one + GO2;
                h[k] = hB; so[k] = uB;
            }
            out_so0 = uA; out_v0 = vA; out_so1 = uB; out_v1 = vB;
            ca = na; cb = nb;
            if (g == G - 1)
            {
                // strict improvement only: the leftmost maximum wins (dp_scout.h:173). Improvements are rare
                // (a handful per task), so the per-half bookkeeping sits behind a branch.
                if (__vimax3_s16x2(best, uA, uB) != best)
                {
                    uint32_t m = __vmaxs2(best, uA);
                    if ((uA & 0xffffu) > (best & 0xffffu)) colF = jj + 1;   // halves are biased positive
                    if ((uA >> 16) > (best >> 16)) colR = jj + 1;
                    best = m;
                    if (jj + 1 < L)
                    {
                        m = __vmaxs2(best, uB);
                        if ((uB & 0xffffu) > (best & 0xffffu)) colF = jj + 2;
                        if ((uB >> 16) > (best >> 16)) colR = jj + 2;
                        best = m;
                    }
                }
            }
        }
    }
    if (g == G - 1 && live)
    {
        const int s_lo = (int)(best & 0xffff) - K1_BIAS - GAP_OPEN, s_hi = (int)(best >> 16) - K1_BIAS - GAP_OPEN;
        if (!MIXED)
        {
            ScoreOut o;
            o.score_f = s_lo; o.score_r = s_hi;
            o.col_f = colF; o.col_r = colR;
            out[task] = o;
            if (mix.sstate)
            {
                const int rc = !(s_lo > s_hi);                       // mapping.h:121
                int lose = rc ? s_lo : s_hi;
                lose = max(-32768, min(32767, lose));
                const int win = rc ? s_hi : s_lo;
                mix.sstate[task] = (int32_t)(((uint32_t)lose << 16) | (uint32_t)(rc << 1) | 1u | (win < K1_WEAK ? 4u : 0u) |
                                             ((uint32_t)max(0, min(1023, win)) << 3));
            }
            if (mix.cells) atomicAdd(mix.cells, 2ull * (unsigned long long)R * (unsigned long long)L);
        }
        else
        {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh)
            {
                const int t = hh ? task2 : task, sc = hh ? s_hi : s_lo, col = hh ? colR : colF, rc = hh ? rcB : rcA;
                const int Rt = hh ? R2 : R;
                const int old = mix.sstate[t] >> 16;
                const int nb = max(old + mix.emax[2 * wid + (Rt > 152)], -6);
                if (sc >= 0 && sc > nb && nb < 32000)
                {
                    ScoreOut o;
                    o.score_f = rc ? nb : sc; o.score_r = rc ? sc : nb;   // the loser's entry is only ever compared
                    o.col_f = rc ? 0 : col; o.col_r = rc ? col : 0;
                    out[t] = o;
                    mix.sstate[t] = (int32_t)(((uint32_t)nb << 16) | (uint32_t)(rc << 1) | 1u | (sc < K1_WEAK ? 4u : 0u) |
                                              ((uint32_t)max(0, min(1023, sc)) << 3) | (1u << 13));
                }
                else
                {
                    mix.sstate[t] = 0;
                    if (mix.cells) atomicAdd(mix.cells + 2, 1ull);                   // statistics: certificates that failed in a mixed launch
                    const int cls = Rt > 152;
                    const int at = atomicAdd(mix.fb_cnt + cls, 1);
                    (cls ? mix.fb32 : mix.fb19)[at] = t;
                }
            }
            if (mix.cells) atomicAdd(mix.cells, (unsigned long long)(R + R2) * (unsigned long long)L);   // (mix.cells points at the mixed-round counter)
        }
    }
}

// Rounds >= 2: splits the live pairs of a polish batch into mixed tasks (both strands known: one task per pair) and
// regular tasks (both strands of one read). cnt: [0] regular <= 152 rows, [1] regular <= 256, [2] mixed <= 152, [3] mixed <= 256
__global__ void mix_plan_kernel(const int32_t *__restrict__ pair_task, int64_t npairs, const int32_t *__restrict__ task_read,
                                const int32_t *__restrict__ task_win, const int32_t *__restrict__ read_len,
                                const uint8_t *__restrict__ win_done, const int32_t *__restrict__ sstate,
                                const int32_t *__restrict__ emax, int predict,
                                int32_t *__restrict__ reg19, int32_t *__restrict__ reg32, int32_t *__restrict__ mix19,
                                int32_t *__restrict__ mix32, int32_t *__restrict__ cnt)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const int t1 = pair_task[2 * p], t2 = pair_task[2 * p + 1];
    if (t1 < 0) return;
    if (win_done[task_win[t1]]) return;
    const int R1 = read_len[task_read[t1]], R2 = t2 >= 0 ? read_len[task_read[t2]] : 0;
    const int s1 = sstate[t1], s2 = t2 >= 0 ? sstate[t2] : 0;
    // Will the certificate hold? The mixed launch accepts a read iff its new score S satisfies S >= 0 and S > max(bound + E, -6)
    // (E = emax of its window). S is not known yet. After a REGULAR round the last score was taken against the unpolished
    // consensus and says little (the round's edits raise it by about what E charges for them): the read is tried. After a MIXED
    // round the consensus is nearly final, so a read whose last score (less a small margin) does not clear bound + E is expected
    // to fail and goes to the regular list right away -- that list runs BESIDE the mixed launch, whereas a failure inside it
    // costs another launch behind it (142 us when few tasks are live). Only a prediction: a wrong guess either way changes
    // where the read is aligned, never the result. (Measured: predicting after regular rounds too sends most reads of
    // windows with many edits to the regular list, configs[3] 889 vs 1,015 variants/s.) `predict` is set by the host only
    // while few tasks are live (api_polish.inc, K1_PREDICT_MAX_TASKS): on a full GPU a predicted read costs more than a failed one.
    auto shaky = [&](int st, int R) {
        const int e = emax[2 * task_win[t1] + (R > 152)];
        const int last = (st >> 3) & 1023, bound = st >> 16;
        return (st & 4) || (predict && ((st >> 13) & 1) && (last - 8) <= max(bound + e, -6));
    };
    if (t2 >= 0 && (s1 & 1) && (s2 & 1) && !shaky(s1, R1) && !shaky(s2, R2))
    {
        const int cls = (R1 > 152 || R2 > 152);
        const int at = atomicAdd(cnt + 2 + cls, 1);
        int32_t *m = cls ? mix32 : mix19;
        m[2 * at] = t1; m[2 * at + 1] = t2;
        return;
    }
    { const int cls = R1 > 152; (cls ? reg32 : reg19)[atomicAdd(cnt + cls, 1)] = t1; }
    if (t2 >= 0) { const int cls = R2 > 152; (cls ? reg32 : reg19)[atomicAdd(cnt + cls, 1)] = t2; }
}
}  // namespace svp
