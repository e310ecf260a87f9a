/* TEST INFRASTRUCTURE ONLY. Unity build of the CPU oracle -> oracle/liboracle.so (see Makefile).
 * Flat accessors at the bottom are for ctypes (oracle/oracle.py). */
#include "sviper_oracle.c"
#include "oracle_polish.c"
#include "oracle_final.c"
#include "oracle_eval.c"
#include "oracle_msa.c"
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- map_single_read for ctypes: returns score via out params, ops into caller buffer ---- */
int orc_map_read(const uint8_t *read, const uint8_t *qual, int m, const uint8_t *ref, int n,
                 int *score, int *score_rc, int *is_rc, int *lead, int *endpos, char *ops, int *nops)
{
    orc_aln a;
    orc_map_single_read(read, qual, m, ref, n, &a);
    *score = a.score; *score_rc = a.score_rc; *is_rc = a.is_rc;
    *lead = read_begin(&a); *endpos = read_end(&a);
    memcpy(ops, a.ops, (size_t)a.nops); *nops = a.nops;
    orc_aln_free(&a);
    return 0;
}

/* ---- orc_round accessors ---- */
int orc_round_Lp(const orc_round *r) { return r->Lp; }
const double *orc_round_profile(const orc_round *r) { return r->profile; }
const unsigned *orc_round_cov(const orc_round *r) { return r->cov; }
const uint8_t *orc_round_isgap(const orc_round *r) { return r->isgap; }
const uint8_t *orc_round_out(const orc_round *r) { return r->out; }
int orc_round_out_len(const orc_round *r) { return r->out_len; }
int orc_round_proper(const orc_round *r, int i) { return r->mobs[i].proper; }
/* which = 0 read, 1 mate */
int orc_round_aln(const orc_round *r, int i, int which, int *score, int *score_rc, int *is_rc, char *ops)
{
    const orc_aln *a = which ? &r->mobs[i].m : &r->mobs[i].r;
    *score = a->score; *score_rc = a->score_rc; *is_rc = a->is_rc;
    if (ops && a->nops) memcpy(ops, a->ops, (size_t)a->nops);
    return a->nops;
}

/* ---- score-only batch of map_single_read for the alignment microbench CPU baseline.
 * tasks: read t (bases at rb+roff[t] .. roff[t+1]) against window win[t] (wb+woff[w]..).
 * Uses all OpenMP threads over tasks (the reference itself only parallelises over variants). ---- */
int orc_align_batch(int ntasks, const uint8_t *rb, const uint8_t *rq, const int32_t *roff,
                    const int32_t *win, const uint8_t *wb, const int32_t *woff,
                    int *score, int *score_rc, int *is_rc, int *lead, int *endpos)
{
    int t;
#pragma omp parallel for schedule(dynamic, 16)
    for (t = 0; t < ntasks; ++t)
    {
        orc_aln a; int w = win[t];
        orc_map_single_read(rb + roff[t], rq + roff[t], roff[t + 1] - roff[t], wb + woff[w], woff[w + 1] - woff[w], &a);
        score[t] = a.score; score_rc[t] = a.score_rc; is_rc[t] = a.is_rc;
        lead[t] = read_begin(&a); endpos[t] = read_end(&a);
        orc_aln_free(&a);
    }
    return 0;
}

/* ---- whole hot path for nv variants (polish_to_perfection + final alignment), one OpenMP thread per
 * variant like the reference's loop at src/sviper.cpp:43-47. Per-variant inputs are arrays of pointers.
 * n_cigar[v] receives the number of CIGAR elements, rounds[v] the rounds. CPU baseline for bench.py. ---- */
int orc_polish_many(int nv, const orc_config *cfgs, const int *np,
                    const uint8_t **b1, const uint8_t **q1, const int32_t **off1,
                    const uint8_t **b2, const uint8_t **q2, const int32_t **off2,
                    const uint8_t **cons, const int *L, const uint8_t **refp, const int *refL,
                    int *n_cigar, int *rounds)
{
    int v;
#pragma omp parallel for schedule(dynamic, 1)
    for (v = 0; v < nv; ++v)
    {
        orc_config c = cfgs[v];
        int cap = 2 * L[v] + 4096, covlen = 0, n, nops = 0, k;
        uint8_t *out = (uint8_t *)malloc((size_t)cap);
        n = orc_polish_to_perfection(&c, np[v], b1[v], q1[v], off1[v], b2[v], q2[v], off2[v], cons[v], L[v], out, cap,
                                     NULL, 0, &covlen);
        rounds[v] = (int)c.rounds;
        n_cigar[v] = 0;
        if (n > 0 && refp && refp[v])
        {
            char *ops = (char *)malloc((size_t)n + refL[v] + 2);
            char last = 0;
            orc_align_convex(refp[v], refL[v], out, n, 90000.0, -80000.0, -1.0, -100000.0, ops, &nops);
            for (k = 0; k < nops; ++k) if (ops[k] != last) { ++n_cigar[v]; last = ops[k]; }
            free(ops);
        }
        free(out);
    }
    return 0;
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
