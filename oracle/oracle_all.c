/* TEST INFRASTRUCTURE ONLY. Unity build of the CPU oracle -> oracle/liboracle.so (see Makefile).
 * Flat accessors at the bottom are for ctypes (oracle/oracle.py). */
#include "sviper_oracle.c"
#include "oracle_polish.c"
#include "oracle_final.c"
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

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
