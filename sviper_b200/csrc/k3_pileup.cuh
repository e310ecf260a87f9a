// K3/K4: quality-weighted pileup, vote, consensus rewrite and the polish_to_perfection state machine.
// One block per variant (window); atomic-free: thread p owns reference position p and adds the reads'
// contributions in pair order (read, then mate), which is the order of the reference's sequential loops
// (/root/reference/include/sviper/polishing.h:133-148), so every fp64 sum is bit-identical.
// fp64 products and sums use __dmul_rn/__dadd_rn: no FMA contraction (the reference build has no -march).
//
//   mappQ stats            polishing.h:174-195       (exact in any order: multiples of 0.5)
//   proper pair            mapping.h:164-173, config.h:101-109
//   fill_profile           polishing.h:54-172        (records come from k2_traceback_kernel)
//   consensus_from_profile polishing.h:236-293, basic/alphabet_profile.h:295-313, config.h:128-133
//   loop control           polishing.h:319-346
#pragma once
#include <cub/block/block_scan.cuh>
#include <cub/block/block_reduce.cuh>
#include "common.cuh"
#include "k1_trace.cuh"

namespace svp
{
constexpr int K3_THREADS = 1024;
constexpr int K3_UNROLL = 4;      // covering alignments whose records are fetched together
constexpr int K3_WARPS = K3_THREADS / 32;
constexpr int K3_LIST = 512;        // per-warp list of alignments overlapping the warp's 32 positions

// dynamic shared memory of k3_pileup_vote_kernel for a batch whose largest variant has max_pairs pairs
__host__ __device__ inline size_t k3_smem_bytes(int max_pairs)
{
    size_t nt = (size_t)2 * max_pairs + 2;
    return nt * (8 + 4 * 4) + (size_t)K3_WARPS * K3_LIST * 2 + 64;
}

struct VarDesc
{
    int32_t n_pairs;
    int32_t pair_base;      // first entry of this variant in pair_task (2 per pair: read, mate or -1)
    int64_t cons_off;       // byte offset of this variant's consensus buffers
    int32_t Lcap;           // capacity of a consensus buffer
    int32_t flank;
    double bq_mean, ins_mean, ins_sd, min_cov;
    int64_t prof_off;       // first view column of this variant in the profile buffers
    int32_t prof_cap;
    int32_t win_id;         // window slot in the packed store
};

struct VarState
{
    int32_t L;              // current consensus length
    int32_t fix_indels;
    int32_t rounds;
    int32_t done;
    int32_t err;
    uint32_t subst, ins, del;
    double mappq_mean;
    int32_t Lp;             // view columns of the last round
    int32_t cur;            // which consensus buffer is current (0/1)
    int32_t aligned;        // rounds that ran the alignment kernels (rounds minus the votes that reused a mapping)
    int32_t pad_;
};

struct K3Buffers
{
    uint8_t *cons[2];       // ping-pong consensus bytes (0..4)
    int32_t *inslen;        // per reference position (indexed like cons)
    int32_t *voff;
    double *profile;        // view columns x 6
    uint32_t *cov;
    uint8_t *isgap;
    uint8_t *decision;      // per view column: 0..4 emit that base, 255 emit nothing
    uint8_t *proper;        // per pair
    uint8_t *edit;          // per view column: by how much this column can change the score of an alignment that crosses it
                            // (0 unchanged, 5 substituted, 6 deleted consensus column, 7 emitted inserted column)
    int32_t *emax;          // per variant: largest sum of these costs over any 320 / 512 consecutive columns of the new
                            // consensus (upper bound; feeds the strand certificate of k1_score.cuh, K1Mix)
};
constexpr int K3_HB = 1024; // histogram bins of the edit cost, 32 new-consensus columns each (the last bin takes the rest)

__device__ __forceinline__ int oriented_base(const SeqStore &reads, int rid, int R, int is_rc, int q)
{
    int src = is_rc ? (R - 1 - q) : q;
    int b = seq_base(reads, reads.off32[rid], src);
    return is_rc ? ((b < 4) ? 3 - b : 4) : b;
}

__global__ void __launch_bounds__(K3_THREADS) k3_pileup_vote_kernel(
    const VarDesc *__restrict__ descs, VarState *__restrict__ states, const int32_t *__restrict__ pair_task,
    const AlnOut *__restrict__ alns, const TracePlan *__restrict__ plans, RecArrays rec,
    const int64_t *__restrict__ rec_off, SeqStore reads, const uint8_t *__restrict__ read_quals,
    const int64_t *__restrict__ qual_off, K3Buffers buf, int force_single_round, int max_pairs,
    const RoundCtl *__restrict__ ctl)
{
    const int v = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (ctl->ovf) return;   // the round's alignments were not produced (arena too small): it is repeated, nothing may change
    VarState st = states[v];
    if (st.done) return;
    const VarDesc d = descs[v];
    const int L = st.L, P = d.n_pairs, flank = d.flank;
    const uint8_t *cons = buf.cons[st.cur] + d.cons_off;
    uint8_t *next = buf.cons[st.cur ^ 1] + d.cons_off;
    int32_t *inslen = buf.inslen + d.cons_off;
    int32_t *voff = buf.voff + d.cons_off;
    double *profile = buf.profile + d.prof_off * 6;
    uint32_t *cov = buf.cov + d.prof_off;
    uint8_t *isgap = buf.isgap + d.prof_off;
    uint8_t *decision = buf.decision + d.prof_off;
    uint8_t *edit = buf.edit + d.prof_off;
    __shared__ int sh_hist[K3_HB + 32];
    __shared__ int sh_emax[2];
    uint8_t *proper = buf.proper + d.pair_base / 2;
    const int32_t *pt = pair_task + d.pair_base;

    // per-alignment metadata of this variant, staged once: entry e = 2 * pair + (0 read | 1 mate), in the
    // reference's accumulation order (polishing.h:133-148). Alignments that contribute nothing have n == 0.
    extern __shared__ unsigned char k3_dyn[];
    const int NTcap = 2 * max_pairs + 2, NT = 2 * P;
    int64_t *e_roff = (int64_t *)k3_dyn;
    int32_t *e_p0 = (int32_t *)(e_roff + NTcap);
    int32_t *e_n = e_p0 + NTcap;
    int32_t *e_mapq = e_n + NTcap;
    int32_t *e_task = e_mapq + NTcap;
    uint16_t *wlist = (uint16_t *)(e_task + NTcap) + (size_t)warp * K3_LIST;

    typedef cub::BlockReduce<long long, K3_THREADS> BR;
    typedef cub::BlockScan<int, K3_THREADS> BS;
    __shared__ union { typename BR::TempStorage r; typename BS::TempStorage s; } tmp;
    __shared__ long long sh_ll;
    __shared__ int sh_i[4];

    // ---- 1. proper pairs + mappQ mean -------------------------------------------------------
    const double rs = fmax(0.0, d.ins_mean - 3 * d.ins_sd);
    const double re = fmin(d.ins_mean + 3 * d.ins_mean, d.ins_mean + 3 * d.ins_sd);
    long long msum = 0;
    int aerr = 0;
    for (int i = tid; i < P; i += K3_THREADS)
    {
        int t1 = pt[2 * i], t2 = pt[2 * i + 1];
        AlnOut a = alns[t1], b;
        aerr |= a.err;
        uint8_t pr = 0;
        if (t2 >= 0)
        {
            b = alns[t2];
            aerr |= b.err;
            msum += a.mapq + b.mapq;
            if (a.npath < 0 || b.npath < 0) pr = 0;     // plan_kernel proved that the pair is not proper
            else if (!a.is_rc && b.is_rc) { int dd = b.lead - a.endpos; pr = (dd > rs) && (dd < re); }
            else if (a.is_rc && !b.is_rc) { int dd = a.lead - b.endpos; pr = (dd > rs) && (dd < re); }
        }
        else
            msum += 2 * a.mapq;
        proper[i] = pr;
        e_task[2 * i] = t1; e_task[2 * i + 1] = t2;
        e_p0[2 * i] = a.p0; e_n[2 * i] = pr ? a.n : 0; e_mapq[2 * i] = a.mapq; e_roff[2 * i] = rec_off[t1];
        if (pr) { e_p0[2 * i + 1] = b.p0; e_n[2 * i + 1] = b.n; e_mapq[2 * i + 1] = b.mapq; e_roff[2 * i + 1] = rec_off[t2]; }
        else { e_p0[2 * i + 1] = 0; e_n[2 * i + 1] = 0; e_mapq[2 * i + 1] = 0; e_roff[2 * i + 1] = 0; }
    }
    msum = BR(tmp.r).Sum(msum);
    if (tid == 0) sh_ll = msum;
    aerr = __syncthreads_or(aerr);
    const double mappq = (P > 0) ? ((double)sh_ll * 0.5) / (double)P : st.mappq_mean;
    if (aerr) { if (tid == 0) { states[v].err = 10 + aerr; states[v].done = 1; } return; }
    __syncthreads();

    // the warp's 32 consecutive positions share (almost) the same covering alignments: compact them once,
    // order preserved. Returns the list length, or -1 when it does not fit (then every entry is visited).
    auto build_list = [&](int lo, int hi) -> int {
        int cnt = 0;
        for (int base = 0; base < NT; base += 32)
        {
            int e = base + lane;
            bool hit = false;
            if (e < NT) { int n = e_n[e], p0 = e_p0[e]; hit = n > 0 && p0 <= hi && p0 + n > lo; }
            unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) { int at = cnt + __popc(m & ((1u << lane) - 1)); if (at < K3_LIST) wlist[at] = (uint16_t)e; }
            cnt += __popc(m);
        }
        __syncwarp();
        return (cnt <= K3_LIST && NT <= 65535) ? cnt : -1;
    };

    // ---- 2. insertion sub-profile lengths per reference position ---------------------------
    for (int base = 0; base < L; base += K3_THREADS)
    {
        const int p = base + tid;
        const int lo = base + warp * 32;
        if (lo >= L) continue;                       // whole warp out of range (warp-uniform)
        const int nl = build_list(lo, min(lo + 31, L - 1));
        const int cnt = nl < 0 ? NT : nl;
        if (p < L)
        {
            // covering alignments in chunks of K3_UNROLL: all record loads of a chunk are issued before the first
            // use (the loop is bound by load latency, not by bandwidth); the entries are then applied in order
            int il = 0;
            for (int x0 = 0; x0 < cnt; x0 += K3_UNROLL)
            {
                bool ok[K3_UNROLL]; int64_t rr[K3_UNROLL]; uint8_t cdv[K3_UNROLL]; uint16_t ilv[K3_UNROLL];
#pragma unroll
                for (int u = 0; u < K3_UNROLL; ++u)
                {
                    const int x = x0 + u;
                    const int e = (x < cnt) ? (nl < 0 ? x : wlist[x]) : 0;
                    const unsigned dd = (unsigned)(p - e_p0[e]);
                    ok[u] = (x < cnt) && dd < (unsigned)e_n[e];
                    rr[u] = e_roff[e] + dd;
                }
#pragma unroll
                for (int u = 0; u < K3_UNROLL; ++u)
                {
                    cdv[u] = 0; ilv[u] = 0;
                    if (ok[u]) { cdv[u] = rec.code[rr[u]]; ilv[u] = rec.inslen[rr[u]]; }
                }
#pragma unroll
                for (int u = 0; u < K3_UNROLL; ++u)
                {
                    if (!ok[u]) continue;
                    const int l = ilv[u];
                    if (l > il) il = l;
                    if ((cdv[u] & 8u) && il < 1) il = 1;
                }
            }
            inslen[p] = il;
        }
        __syncwarp();
    }
    __syncthreads();
    // ---- 3. view offsets: exclusive scan of inslen[p] + 1 ------------------------------------
    {
        int carry = 0;
        for (int base = 0; base < L; base += K3_THREADS)
        {
            int p = base + tid;
            int x = (p < L) ? inslen[p] + 1 : 0, ex, tot;
            BS(tmp.s).ExclusiveSum(x, ex, tot);
            if (p < L) voff[p] = carry + ex;
            carry += tot;
            __syncthreads();
        }
        if (tid == 0) sh_i[0] = carry;
    }
    __syncthreads();
    const int Lp = sh_i[0];
    if (Lp > d.prof_cap) { if (tid == 0) { states[v].err = 4; states[v].done = 1; } return; }

    // ---- 4. pileup + vote per reference position ---------------------------------------------
    const double qq = __dmul_rn(d.bq_mean, mappq);
    int fix = st.fix_indels;
    unsigned n_sub = 0, n_ins = 0, n_del = 0;
    auto vote = [&](const double *c, unsigned cv, int gap, int base) -> int {
        // returns the base to emit (0..4) or 255; updates the counters
        if (!fix && gap) return 255;
        int idx = 0; double mx = c[0];
#pragma unroll
        for (int k = 1; k < 6; ++k) if (c[k] > mx) { idx = k; mx = c[k]; }
        double half = (double)cv / 2;
        double thr = __dmul_rn((d.min_cov < half) ? half : d.min_cov, qq);
        if (!(mx >= thr)) return gap ? 255 : base;
        if (idx < 5) { if (gap) ++n_ins; else if (base != idx) ++n_sub; return idx; }
        if (!fix) return gap ? 255 : base;
        if (!gap) ++n_del;
        return 255;
    };
    for (int base = 0; base < L; base += K3_THREADS)
    {
        const int p = base + tid;
        const int lo = base + warp * 32;
        if (lo >= L) continue;                       // warp-uniform
        const int nl = build_list(lo, min(lo + 31, L - 1));
        const int cnt = nl < 0 ? NT : nl;
        if (p < L)
        {
            double m[6] = {0, 0, 0, 0, 0, 0};
            double gap0 = 0.0;
            unsigned cv = 0, noins = 0, runs = 0;
            const int il = inslen[p];
            for (int x0 = 0; x0 < cnt; x0 += K3_UNROLL)
            {
                bool ok[K3_UNROLL]; int64_t rr[K3_UNROLL]; int mqv[K3_UNROLL];
                uint8_t cdv[K3_UNROLL], qv[K3_UNROLL]; uint16_t ilv[K3_UNROLL];
#pragma unroll
                for (int u = 0; u < K3_UNROLL; ++u)
                {
                    const int x = x0 + u;
                    const int e = (x < cnt) ? (nl < 0 ? x : wlist[x]) : 0;
                    const unsigned dd = (unsigned)(p - e_p0[e]);
                    ok[u] = (x < cnt) && dd < (unsigned)e_n[e];
                    rr[u] = e_roff[e] + dd; mqv[u] = e_mapq[e];
                }
#pragma unroll
                for (int u = 0; u < K3_UNROLL; ++u)
                {
                    cdv[u] = 0; qv[u] = 0; ilv[u] = 0;
                    if (ok[u]) { cdv[u] = rec.code[rr[u]]; qv[u] = rec.qual[rr[u]]; ilv[u] = rec.inslen[rr[u]]; }
                }
#pragma unroll
                for (int u = 0; u < K3_UNROLL; ++u)
                {
                    if (!ok[u]) continue;                       // entries are applied in the reference's order
                    const uint8_t cd = cdv[u];
                    const double mq = (double)mqv[u];
                    const double gw = __dmul_rn(d.bq_mean, mq);
                    if (ilv[u]) ++runs;
                    if (cd & 8u) { gap0 = __dadd_rn(gap0, gw); ++noins; }
                    const int sym = cd & 7u;
                    const double wgt = (sym < 5) ? __dmul_rn((double)qv[u], mq) : gw;
                    // m[sym] += wgt without a dynamically indexed local array
#pragma unroll
                    for (int q = 0; q < 6; ++q) if (q == sym) m[q] = __dadd_rn(m[q], wgt);
                    ++cv;
                }
            }
            const int col0 = voff[p];
            const bool ins_in = (p > flank) && (p <= L - flank);
            // inserted columns in front of p
            for (int k = 0; k < il; ++k)
            {
                double c[6] = {0, 0, 0, 0, 0, gap0};
                unsigned ck = (k == 0) ? noins : 0;   // ins_cov[k]; noins bookings count in column 0
                if (runs)
                {
                    for (int x = 0; x < cnt; ++x)
                    {
                        const int e = nl < 0 ? x : wlist[x];
                        const unsigned dd = (unsigned)(p - e_p0[e]);
                        if (dd >= (unsigned)e_n[e]) continue;
                        const int64_t r = e_roff[e] + dd;
                        if (rec.inslen[r] <= k) continue;
                        const int t = e_task[e];
                        const int rid = plans[t].read_id, R = reads.len[rid], rc = alns[t].is_rc;
                        const int q = rec.insq[r] + k;
                        const int b = oriented_base(reads, rid, R, rc, q);
                        const int ql = read_quals[qual_off[rid] + (rc ? R - 1 - q : q)];
                        const double wgt = __dmul_rn((double)ql, (double)e_mapq[e]);
#pragma unroll
                        for (int qx = 0; qx < 5; ++qx) if (qx == b) c[qx] = __dadd_rn(c[qx], wgt);
                        ++ck;
                    }
                }
                if (k >= 1) ck += noins;   // polishing.h:159-160
                int col = col0 + k;
#pragma unroll
                for (int s = 0; s < 6; ++s) profile[(size_t)col * 6 + s] = c[s];
                cov[col] = ck; isgap[col] = 1;
                decision[col] = ins_in ? (uint8_t)vote(c, ck, 1, -1) : 255;
                edit[col] = decision[col] != 255 ? 7 : 0;
            }
            {
                int col = col0 + il;
#pragma unroll
                for (int s = 0; s < 6; ++s) profile[(size_t)col * 6 + s] = m[s];
                cov[col] = cv; isgap[col] = 0;
                int bs = cons[p];
                decision[col] = (p >= flank && p < L - flank) ? (uint8_t)vote(m, cv, 0, bs) : 255;
                edit[col] = ((p >= flank && p < L - flank) && decision[col] != bs) ? (decision[col] == 255 ? 6 : 5) : 0;
            }
        }
        __syncwarp();
    }
    // A phase-1 round that changes nothing is followed by the first indel round on the SAME consensus
    // (polishing.h:327-343): its mapping, profile and coverage are identical, only consensus_from_profile runs
    // with fix_indels set. That round is done right here as a second vote over the stored profile (pass 1) instead
    // of a full alignment round.
    for (int pass = 0;; ++pass)
    {
        if (pass == 1)
        {
            fix = 1; n_sub = 0; n_ins = 0; n_del = 0;
            for (int p = tid; p < L; p += K3_THREADS)
            {
                const int il = inslen[p], col0 = voff[p];
                const bool ins_in = (p > flank) && (p <= L - flank);
                for (int k = 0; k <= il; ++k)
                {
                    const int col = col0 + k;
                    double c[6];
#pragma unroll
                    for (int q = 0; q < 6; ++q) c[q] = profile[(size_t)col * 6 + q];
                    if (k < il) { decision[col] = ins_in ? (uint8_t)vote(c, cov[col], 1, -1) : 255; edit[col] = decision[col] != 255 ? 7 : 0; }
                    else
                    {
                        decision[col] = (p >= flank && p < L - flank) ? (uint8_t)vote(c, cov[col], 0, cons[p]) : 255;
                        edit[col] = ((p >= flank && p < L - flank) && decision[col] != cons[p]) ? (decision[col] == 255 ? 6 : 5) : 0;
                    }
                }
            }
        }
        {
            long long pk = ((long long)n_sub) | ((long long)n_ins << 20) | ((long long)n_del << 40);
            __syncthreads();
            pk = BR(tmp.r).Sum(pk);
            if (tid == 0) sh_ll = pk;
        }
        __syncthreads();
        // ---- 5. compaction: new consensus = prefix + emitted + suffix ----------------------------
        const int vb = voff[flank] + inslen[flank];                 // column of base `flank`
        const int ve = voff[L - flank] + inslen[L - flank];         // column of base L - flank
        int carry = flank;
        for (int i = tid; i < flank; i += K3_THREADS) next[i] = cons[i];
        for (int i = tid; i < K3_HB + 32; i += K3_THREADS) sh_hist[i] = 0;
        if (tid < 2) sh_emax[tid] = 0;
        __syncthreads();
        for (int base = vb; base < ve; base += K3_THREADS)
        {
            int col = base + tid;
            int dcs = (col < ve) ? decision[col] : 255;
            int x = (dcs != 255), ex, tot;
            BS(tmp.s).ExclusiveSum(x, ex, tot);
            if (x && carry + ex < d.Lcap) next[carry + ex] = (uint8_t)dcs;
            // edit cost in NEW-consensus coordinates: an emitted column sits at carry + ex, a deleted one is booked on
            // the column that follows it
            if (col < ve && edit[col]) atomicAdd(&sh_hist[min((carry + ex) >> 5, K3_HB - 1)], (int)edit[col]);
            carry += tot;
            __syncthreads();
        }
        {
            // W consecutive columns touch at most W / 32 + 1 bins
            int m0 = 0, m1 = 0;
            for (int i = tid; i < K3_HB; i += K3_THREADS)
            {
                int a0 = 0;
                for (int j = 0; j < 11; ++j) a0 += sh_hist[i + j];
                int a1 = a0;
                for (int j = 11; j < 17; ++j) a1 += sh_hist[i + j];
                m0 = max(m0, a0); m1 = max(m1, a1);
            }
            m0 = __reduce_max_sync(0xffffffffu, m0); m1 = __reduce_max_sync(0xffffffffu, m1);
            if ((tid & 31) == 0) { atomicMax(&sh_emax[0], m0); atomicMax(&sh_emax[1], m1); }
            __syncthreads();
            if (tid < 2) buf.emax[2 * v + tid] = sh_emax[tid];
        }
        const int newL = carry + flank;
        if (newL > d.Lcap) { if (tid == 0) { states[v].err = 5; states[v].done = 1; } return; }
        for (int i = tid; i < flank; i += K3_THREADS) next[carry + i] = cons[L - flank + i];
        __syncthreads();
        // ---- 6. convergence + loop control (polishing.h:327-343) ---------------------------------
        int diff = (newL != L);
        if (!diff)
            for (int i = tid; i < L; i += K3_THREADS) diff |= (next[i] != cons[i]);
        diff = __syncthreads_or(diff);
        // every thread keeps its copy of the state current (pass 1 needs it), thread 0 publishes it
        const long long pk = sh_ll;
        st.subst += (unsigned)(pk & 0xfffff); st.ins += (unsigned)((pk >> 20) & 0xfffff); st.del += (unsigned)(pk >> 40);
        st.mappq_mean = mappq;
        st.rounds += 1;
        if (pass == 0) st.aligned += 1;
        st.Lp = Lp;
        st.L = newL;
        if (pass == 0) st.cur ^= 1;    // pass 1 rewrites the buffer pass 0 made current (its input is the identical old one)
        bool again = false;
        if (force_single_round) st.done = 1;
        else if (!st.fix_indels)
        {
            if (!diff || st.rounds >= 20) st.fix_indels = 1;   // phase 2 always starts with one more round
            if (st.rounds >= 30) st.done = 1;
            again = !diff && !st.done;
        }
        else if (!diff || st.rounds >= 30) st.done = 1;
        if (!again)
        {
            if (tid == 0) states[v] = st;
            break;
        }
        __syncthreads();
    }
}
}  // namespace svp
