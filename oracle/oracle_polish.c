/* TEST INFRASTRUCTURE ONLY (see sviper_oracle.c). Part of the unity build oracle_all.c.
 * Pileup, vote and the polishing loop: include/sviper/polishing.h. */

/* one polishing round's observable state (what ref_harness ROUND prints) */
typedef struct
{
    int np;
    orc_mob *mobs;
    int Lp;              /* profile length after splicing inserted columns */
    double *profile;     /* Lp x 6 */
    unsigned *cov;       /* Lp */
    uint8_t *isgap;      /* Lp: 1 = inserted column (gap in contigGaps) */
    uint8_t *out; int out_len;
} orc_round;

void orc_round_free(orc_round *r)
{
    int i;
    if (!r) return;
    for (i = 0; i < r->np; ++i) { orc_aln_free(&r->mobs[i].r); orc_aln_free(&r->mobs[i].m); }
    free(r->mobs); free(r->profile); free(r->cov); free(r->isgap); free(r->out); free(r);
}

typedef struct { double *cnt; unsigned *cov; int len; } orc_insprof;

static void ins_grow(orc_insprof *p, int pos)
{
    if (p->len <= pos)
    {
        int nl = pos + 1, k;
        p->cnt = (double *)realloc(p->cnt, (size_t)nl * 6 * sizeof(double));
        p->cov = (unsigned *)realloc(p->cov, (size_t)nl * sizeof(unsigned));
        for (k = p->len; k < nl; ++k) { memset(p->cnt + (size_t)k * 6, 0, 6 * sizeof(double)); p->cov[k] = 0; }
        p->len = nl;
    }
}

/* seqan isGap on the ref row incl. the position == length quirk (seqan/align/gaps_array.h:749-783) */
static int ref_isgap(const orc_aln *a, int idx)
{
    if (idx < a->nops) return a->ops[idx] == 'I';
    return a->ops[a->nops - 1] != 'I';
}

/* polishing.h:54-107, literal */
static void orc_add_read(double *profile, unsigned *cov, orc_insprof *ins, unsigned *no_ins_cov,
                         const orc_aln *a, double mapQ, const orc_config *c)
{
    int rb = ref_begin(a), qb = read_begin(a), re = ref_end(a), qe = read_end(a);
    unsigned begin = (unsigned)(rb > qb ? rb : qb), end = (unsigned)(re < qe ? re : qe), idx;
    /* running source positions at column 0 */
    int *refpos = (int *)malloc(((size_t)a->nops + 2) * sizeof(int));
    int *readpos = (int *)malloc(((size_t)a->nops + 2) * sizeof(int));
    int k, rp = 0, qp = 0;
    for (k = 0; k < a->nops; ++k)
    {
        refpos[k] = rp; readpos[k] = qp;          /* toSourcePosition, RightOfViewPos (gaps_array.h:477-511) */
        if (a->ops[k] != 'I') ++rp;
        if (a->ops[k] != 'D') ++qp;
    }
    refpos[a->nops] = rp; readpos[a->nops] = qp;
    for (idx = begin; idx < end; ++idx)
    {
        unsigned p = (unsigned)refpos[idx];
        if (ref_isgap(a, (int)idx))
        {
            unsigned ins_pos = 0;
            while (ref_isgap(a, (int)idx))
            {
                int b = a->bases[readpos[idx]], q = a->quals[readpos[idx]];
                ins_grow(&ins[p], (int)ins_pos);
                ins[p].cnt[(size_t)ins_pos * 6 + b] += (double)q * mapQ;   /* config.h:115-122, paired */
                ++ins[p].cov[ins_pos];
                ++idx; ++ins_pos;
            }
        }
        else if (!ref_isgap(a, (int)idx + 1))
        {
            ins_grow(&ins[p], 0);
            ins[p].cnt[5] += c->baseQ_mean * mapQ;
            ++ins[p].cov[0];
            ++no_ins_cov[p];
        }
        if ((int)idx < a->nops && a->ops[idx] != 'D')
        {
            int b = a->bases[readpos[idx]], q = a->quals[readpos[idx]];
            profile[(size_t)p * 6 + b] += (double)q * mapQ;
        }
        else
            profile[(size_t)p * 6 + 5] += c->baseQ_mean * mapQ;
        ++cov[p];
    }
    free(refpos); free(readpos);
}

/* one call of sviper::polish (polishing.h:295-317): mapping, mappQ stats, fill_profile, consensus */
orc_round *orc_polish_round(orc_config *c, int np,
                            const uint8_t *b1, const uint8_t *q1, const int32_t *off1,
                            const uint8_t *b2, const uint8_t *q2, const int32_t *off2,
                            const uint8_t *ref, int L)
{
    orc_round *R = (orc_round *)calloc(1, sizeof *R);
    double *profile = (double *)calloc((size_t)L * 6 + 6, sizeof(double));
    unsigned *cov = (unsigned *)calloc((size_t)L + 1, sizeof(unsigned));
    unsigned *no_ins = (unsigned *)calloc((size_t)L + 1, sizeof(unsigned));
    orc_insprof *ins = (orc_insprof *)calloc((size_t)L + 1, sizeof *ins);
    int i, p, Lp, pos;
    R->np = np; R->mobs = (orc_mob *)calloc((size_t)np + 1, sizeof(orc_mob));
    /* mapping.h:148-179. The pairs are independent (each writes its own mob), so large windows (configs[3]: 13 kb x 1,300
     * pairs) are spread over the host cores; nested inside orc_polish_many's loop over variants this stays sequential. */
#pragma omp parallel for schedule(dynamic, 4) if (np >= 64 && L >= 3000)
    for (i = 0; i < np; ++i)
    {
        orc_mob *mb = &R->mobs[i];
        orc_map_single_read(b1 + off1[i], q1 + off1[i], off1[i + 1] - off1[i], ref, L, &mb->r);
        if (off2[i + 1] - off2[i] != 0)
        {
            orc_map_single_read(b2 + off2[i], q2 + off2[i], off2[i + 1] - off2[i], ref, L, &mb->m);
            if ((!mb->r.is_rc && mb->m.is_rc && orc_insert_in_range(c, read_end(&mb->r), read_begin(&mb->m))) ||
                (mb->r.is_rc && !mb->m.is_rc && orc_insert_in_range(c, read_end(&mb->m), read_begin(&mb->r))))
                mb->proper = 1;
        }
    }
    /* polishing.h:174-195 */
    if (np != 0)
    {
        double avg = 0.0;
        for (i = 0; i < np; ++i)
        {
            orc_mob *mb = &R->mobs[i];
            avg += (mb->m.len != 0) ? (mb->r.mapQ + mb->m.mapQ) / 2 : mb->r.mapQ;
        }
        c->mappQ_mean = avg / (double)np;
    }
    /* polishing.h:109-148 */
    for (i = 0; i < np; ++i)
    {
        orc_mob *mb = &R->mobs[i];
        if (!mb->proper) continue;
        orc_add_read(profile, cov, ins, no_ins, &mb->r, mb->r.mapQ, c);
        if (mb->m.len != 0) orc_add_read(profile, cov, ins, no_ins, &mb->m, mb->m.mapQ, c);
    }
    /* polishing.h:150-171: splice insertion sub-profiles in front of their base */
    Lp = L; for (p = 0; p < L; ++p) Lp += ins[p].len;
    R->Lp = Lp;
    R->profile = (double *)calloc((size_t)Lp * 6 + 6, sizeof(double));
    R->cov = (unsigned *)calloc((size_t)Lp + 1, sizeof(unsigned));
    R->isgap = (uint8_t *)calloc((size_t)Lp + 1, 1);
    pos = 0;
    for (p = 0; p < L; ++p)
    {
        int k;
        for (k = 0; k < ins[p].len; ++k)
        {
            memcpy(R->profile + (size_t)pos * 6, ins[p].cnt + (size_t)k * 6, 6 * sizeof(double));
            R->profile[(size_t)pos * 6 + 5] = ins[p].cnt[5];
            R->cov[pos] = ins[p].cov[k] + (k >= 1 ? no_ins[p] : 0);
            R->isgap[pos] = 1;
            ++pos;
        }
        memcpy(R->profile + (size_t)pos * 6, profile + (size_t)p * 6, 6 * sizeof(double));
        R->cov[pos] = cov[p]; R->isgap[pos] = 0;
        ++pos;
    }
    /* polishing.h:236-293 consensus_from_profile */
    {
        uint8_t *out = (uint8_t *)malloc((size_t)Lp + 1);
        int o = 0, begin = -1, end = -1, src = 0, v;
        int sb = c->flank, se = L - c->flank;
        double qq = c->baseQ_mean * c->mappQ_mean;
        for (v = 0; v < Lp; ++v)            /* toViewPosition of a source position = column of that base */
            if (!R->isgap[v]) { if (src == sb) begin = v; if (src == se) end = v; ++src; }
        if (begin < 0) begin = Lp;
        if (end < 0) end = Lp;
        memcpy(out, ref, (size_t)c->flank); o = c->flank;
        src = sb; /* source index of the next base column at or after 'begin' */
        for (v = begin; v < end; ++v)
        {
            int gap = R->isgap[v], idx = 0, k, base = gap ? -1 : ref[src];
            const double *pc = R->profile + (size_t)v * 6;
            double mx = pc[0], half, thr;
            if (!gap) ++src;
            if (!c->fix_indels && gap) continue;
            for (k = 1; k < 6; ++k) if (pc[k] > mx) { idx = k; mx = pc[k]; }   /* alphabet_profile.h:295-313 */
            half = (double)R->cov[v] / 2;
            thr = ((c->min_coverage < half) ? half : c->min_coverage) * qq;    /* config.h:128-133 */
            if (!(pc[idx] >= thr)) { if (!gap) out[o++] = (uint8_t)base; continue; }
            if (idx < 5)
            {
                out[o++] = (uint8_t)idx;
                if (gap) ++c->ins; else if (base != idx) ++c->subst;
            }
            else if (!c->fix_indels) { if (!gap) out[o++] = (uint8_t)base; }
            else { if (!gap) ++c->del; }
        }
        memcpy(out + o, ref + (L - c->flank), (size_t)c->flank); o += c->flank;
        R->out = out; R->out_len = o;
    }
    for (p = 0; p < L; ++p) { free(ins[p].cnt); free(ins[p].cov); }
    free(ins); free(no_ins); free(profile); free(cov);
    return R;
}

/* polishing.h:319-346. out must hold cap bytes; cov_out (may be NULL) receives the last round's
 * coverage profile (config.cov_profile), *cov_len its length. Returns the polished length or -1. */
int orc_polish_to_perfection(orc_config *c, int np,
                             const uint8_t *b1, const uint8_t *q1, const int32_t *off1,
                             const uint8_t *b2, const uint8_t *q2, const int32_t *off2,
                             const uint8_t *ref0, int L0, uint8_t *out, int cap,
                             unsigned *cov_out, int cov_cap, int *cov_len)
{
    uint8_t *ref = (uint8_t *)malloc((size_t)L0 + 1), *old = NULL;
    int L = L0, oldL = 0, phase;
    memcpy(ref, ref0, (size_t)L0);
    c->fix_indels = 0;
    for (phase = 0; phase < 2; ++phase)
    {
        unsigned limit = phase ? 30u : 20u;
        if (phase) { c->fix_indels = 1; oldL = 0; }      /* clear(old_ref): empty string */
        while (!(L == oldL && (L == 0 || memcmp(ref, old, (size_t)L) == 0)) && c->rounds < limit)
        {
            orc_round *R;
            free(old); old = ref; oldL = L;
            R = orc_polish_round(c, np, b1, q1, off1, b2, q2, off2, old, oldL);
            ref = R->out; L = R->out_len; R->out = NULL;
            if (cov_out) { int k, n = R->Lp < cov_cap ? R->Lp : cov_cap; for (k = 0; k < n; ++k) cov_out[k] = R->cov[k]; if (cov_len) *cov_len = R->Lp; }
            orc_round_free(R);
            ++c->rounds;
        }
    }
    if (L > cap) { free(ref); free(old); return -1; }
    memcpy(out, ref, (size_t)L);
    free(ref); free(old);
    return L;
}
