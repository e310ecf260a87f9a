/* TEST INFRASTRUCTURE ONLY -- CPU restatement of SViper's per-variant polishing hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker. The product path
 * (sviper_b200/csrc) never links or calls it.
 *
 * Parity status: PINNED. Every function here is checked (tests/test_oracle_vs_ref.py,
 * tests/golden/) against (a) SeqAn's own golden vectors for the semi-global Gotoh
 * alignment (extern/seqan/tests/align/test_alignment_algorithms_global.h:2084-2151)
 * and (b) outputs of the unmodified reference headers run through oracle/_ref/ref_harness.
 *
 * File:line citations are into /root/reference. "seqan/" = extern/seqan/include/seqan/.
 * Bases are 0..4 = A,C,G,T,N (seqan::Dna5 ordValue); qualities 0..62, N has quality 0
 * (seqan/basic/alphabet_residue.h:331-365).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <limits.h>
#include <float.h>
#include <math.h>

/* trace bits: seqan/align/dp_profile.h:125-134 */
enum { T_NONE = 0, T_DIAG = 1, T_H = 2, T_V = 4, T_HO = 8, T_VO = 16, T_MFH = 32, T_MFV = 64 };

/* ------------------------------------------------------------------------------------------
 * Traceback (seqan/align/dp_traceback_impl.h:391-420 priority, :243-259 / :321-336 gap walks,
 * :501-549 head/tail segments, :436-460 initial direction with PreferGapsAtEnd) followed by
 * the Gaps adaptor (dp_traceback_adaptor.h:60-118). The result is one op per view column:
 *   'M' both rows have a base, 'D' gap in the V row (HORIZONTAL), 'I' gap in the H row (VERTICAL).
 * T is column-major: T[col*(m+1)+row]. Returns the number of ops.
 * ---------------------------------------------------------------------------------------- */
static int orc_traceback(const uint8_t *T, int n, int m, int col, int row, char *ops)
{
    int k = 0, i;
    size_t ld = (size_t)m + 1;
    unsigned tv;
    /* TraceTail (:525-533) */
    for (i = row; i < m; ++i) ops[k++] = 'I';
    for (i = col; i < n; ++i) ops[k++] = 'D';
    tv = T[(size_t)col * ld + row];
    if (tv & T_MFV) tv &= (T_V | T_VO | T_MFV);
    else if (tv & T_MFH) tv &= (T_H | T_HO | T_MFH);
    while (col > 0 && row > 0 && tv != T_NONE)
    {
        if (tv & T_DIAG)
        {
            ops[k++] = 'M'; --col; --row; tv = T[(size_t)col * ld + row];
        }
        else if ((tv & T_MFV) && (tv & T_V))
        {
            while ((!(tv & T_VO) || (tv & T_V)) && row != 1)
            {
                --row; tv = T[(size_t)col * ld + row]; ops[k++] = 'I';
            }
            --row; tv = T[(size_t)col * ld + row]; ops[k++] = 'I';
        }
        else if ((tv & T_MFV) && (tv & T_VO))
        {
            --row; tv = T[(size_t)col * ld + row]; ops[k++] = 'I';
        }
        else if ((tv & T_MFH) && (tv & T_H))
        {
            while ((!(tv & T_HO) || (tv & T_H)) && col != 1)
            {
                --col; tv = T[(size_t)col * ld + row]; ops[k++] = 'D';
            }
            --col; tv = T[(size_t)col * ld + row]; ops[k++] = 'D';
        }
        else if ((tv & T_MFH) && (tv & T_HO))
        {
            --col; tv = T[(size_t)col * ld + row]; ops[k++] = 'D';
        }
        else
            break; /* undefined trace value: cannot happen with SingleTrace */
    }
    /* TraceHead (:541-548) */
    for (i = 0; i < row; ++i) ops[k++] = 'I';
    for (i = 0; i < col; ++i) ops[k++] = 'D';
    for (i = 0; i < k / 2; ++i) { char c = ops[i]; ops[i] = ops[k - 1 - i]; ops[k - 1 - i] = c; }
    return k;
}

/* ------------------------------------------------------------------------------------------
 * Semi-global affine alignment exactly as mapping.h:111-114 calls it:
 * globalAlignment(gapsH, gapsV, Score<int,Simple>(mm,mx,ge,go), AlignConfig<true,false,false,true>, AffineGaps)
 *  - recurrence + tie rules: seqan/align/dp_formula_affine.h:66-125 (All), :291-325 (Vertical),
 *    dp_formula.h:152-163 (_maxScore: left wins ties), :275-289 (Zero)
 *  - cell kinds: dp_meta_info.h:178-221; scout: dp_scout.h:168-180 (strict >, last row only)
 *  - end-cell correction: dp_algorithm_impl.h:1170-1185
 * H = ref (columns, length n), V = read (rows, length m). go is the cost of the FIRST gap char.
 * ops needs room for n+m chars. Returns the score; *end_col = column of the chosen end cell.
 * ---------------------------------------------------------------------------------------- */
int orc_align_semiglobal(const uint8_t *ref, int n, const uint8_t *read, int m,
                         int mm, int mx, int ge, int go, char *ops, int *nops, int *end_col)
{
    const int NEG = INT_MIN / 2; /* seqan/align/dp_cell.h:144-145 */
    size_t ld = (size_t)m + 1;
    int *S = (int *)malloc(ld * sizeof(int));
    int *H = (int *)malloc(ld * sizeof(int));
    uint8_t *T = (uint8_t *)malloc(ld * ((size_t)n + 1));
    int i, j, best = NEG, bestH = NEG, bestV = NEG, bestCol = 0, first = 1;
    if (n <= 0 || m <= 0) { free(S); free(H); free(T); *nops = 0; *end_col = 0; return INT_MIN; }
    /* initial column */
    {
        int cvS = 0, cvV = NEG;
        S[0] = 0; H[0] = NEG; T[0] = T_NONE;
        for (i = 1; i <= m; ++i)
        {
            int ext = cvV + ge, opn = cvS + go;
            unsigned tv;
            if (ext < opn) { cvV = opn; tv = T_VO; } else { cvV = ext; tv = T_V; }
            H[i] = NEG; S[i] = cvV; cvS = S[i];
            T[i] = (uint8_t)(tv | T_MFV);
        }
        best = S[m]; bestH = H[m]; bestV = cvV; bestCol = 0; first = 0;
    }
    for (j = 1; j <= n; ++j)
    {
        uint8_t *Tc = T + (size_t)j * ld;
        int diag, cvS, cvV = NEG;
        uint8_t h = ref[j - 1];
        /* row 0: Zero */
        diag = S[0]; /* value before overwrite is also 0 */
        S[0] = 0; diag = 0; cvS = 0; Tc[0] = T_NONE;
        for (i = 1; i <= m; ++i)
        {
            int inter = diag + ((h == read[i - 1]) ? mm : mx);
            int ext, opn, g, s;
            unsigned tv, tv2;
            diag = S[i];
            ext = H[i] + ge; opn = S[i] + go;
            if (ext < opn) { H[i] = opn; tv = T_HO; } else { H[i] = ext; tv = T_H; }
            ext = cvV + ge; opn = cvS + go;
            if (ext < opn) { cvV = opn; tv |= T_VO; } else { cvV = ext; tv |= T_V; }
            if (cvV < H[i]) { g = H[i]; tv2 = T_MFH; } else { g = cvV; tv2 = T_MFV; }
            if (inter < g) { s = g; tv = tv2 | tv; } else { s = inter; tv = T_DIAG | tv; }
            S[i] = s; cvS = s;
            Tc[i] = (uint8_t)tv;
        }
        if (first || S[m] > best) { best = S[m]; bestH = H[m]; bestV = cvV; bestCol = j; first = 0; }
    }
    {
        uint8_t *tp = T + (size_t)bestCol * ld + m;
        if (bestV == best) { *tp &= (uint8_t)~T_DIAG; *tp |= T_MFV; }
        else if (bestH == best) { *tp &= (uint8_t)~T_DIAG; *tp |= T_MFH; }
    }
    *nops = orc_traceback(T, n, m, bestCol, m, ops);
    *end_col = bestCol;
    free(S); free(H); free(T);
    return best;
}

/* ==========================================================================================
 * map_single_read / mapping  (include/sviper/mapping.h:84-137, :148-179)
 * ======================================================================================== */
typedef struct
{
    int len;          /* read length (0 = dummy mate) */
    uint8_t *bases;   /* oriented as aligned (reverse-complemented when is_rc) */
    uint8_t *quals;
    char *ops;        /* one op per view column */
    int nops;
    int is_rc;
    int score, score_rc;
    double mapQ;
} orc_aln;

static void orc_aln_free(orc_aln *a) { free(a->bases); free(a->quals); free(a->ops); memset(a, 0, sizeof *a); }

/* seqan/modifier/modifier_functors.h:275-284: complement keeps the quality; N stays N */
static void orc_revcomp(const uint8_t *b, const uint8_t *q, int m, uint8_t *ob, uint8_t *oq)
{
    int i;
    for (i = 0; i < m; ++i)
    {
        uint8_t c = b[m - 1 - i];
        ob[i] = (c < 4) ? (uint8_t)(3 - c) : c;
        oq[i] = q[m - 1 - i];
    }
}

static int lead_of(const char *ops, int n, char g) { int k = 0; while (k < n && ops[k] == g) ++k; return k; }
static int trail_of(const char *ops, int n, char g) { int k = 0; while (k < n && ops[n - 1 - k] == g) ++k; return k; }
/* mapping.h:59-70 on the read row ('D' = gap in read row) and the ref row ('I' = gap in ref row) */
static int read_begin(const orc_aln *a) { return lead_of(a->ops, a->nops, 'D'); }
static int read_end(const orc_aln *a) { return a->nops - trail_of(a->ops, a->nops, 'D'); }
static int ref_begin(const orc_aln *a) { return lead_of(a->ops, a->nops, 'I'); }
static int ref_end(const orc_aln *a) { return a->nops - trail_of(a->ops, a->nops, 'I'); }

void orc_map_single_read(const uint8_t *read, const uint8_t *qual, int m, const uint8_t *ref, int n, orc_aln *out)
{
    uint8_t *rb, *rq; char *ops1, *ops2; int n1, n2, e1, e2, s, src;
    memset(out, 0, sizeof *out);
    if (m == 0) return;                       /* mapping.h:99-100 */
    rb = (uint8_t *)malloc(m); rq = (uint8_t *)malloc(m);
    orc_revcomp(read, qual, m, rb, rq);
    ops1 = (char *)malloc((size_t)n + m + 1); ops2 = (char *)malloc((size_t)n + m + 1);
    s = orc_align_semiglobal(ref, n, read, m, 1, -4, -1, -6, ops1, &n1, &e1);
    src = orc_align_semiglobal(ref, n, rb, m, 1, -4, -1, -6, ops2, &n2, &e2);
    out->len = m; out->score = s; out->score_rc = src;
    if (s > src)                              /* mapping.h:121 */
    {
        out->bases = (uint8_t *)malloc(m); out->quals = (uint8_t *)malloc(m);
        memcpy(out->bases, read, m); memcpy(out->quals, qual, m);
        out->ops = ops1; out->nops = n1; out->is_rc = 0; out->mapQ = (double)(s > 0 ? s : 0);
        free(rb); free(rq); free(ops2);
    }
    else
    {
        out->bases = rb; out->quals = rq;
        out->ops = ops2; out->nops = n2; out->is_rc = 1; out->mapQ = (double)(src > 0 ? src : 0);
        free(ops1);
    }
}

typedef struct
{
    int flank;             /* SViperConfig::ref_flank_length (sviper.h:363 sets 500) */
    double ins_mean, ins_sd;
    double min_coverage;   /* config.h:79, 4 */
    double baseQ_mean;     /* polishing.h:197-234, computed once per variant */
    int fix_indels;
    double mappQ_mean;
    unsigned subst, ins, del, rounds;
} orc_config;

/* config.h:101-109 */
static int orc_insert_in_range(const orc_config *c, int end_fwd, int begin_rev)
{
    double a = c->ins_mean - 3 * c->ins_sd, b1 = c->ins_mean + 3 * c->ins_mean, b2 = c->ins_mean + 3 * c->ins_sd;
    double rs = (0.0 < a) ? a : 0.0;           /* std::max(0.0, a) */
    double re = (b2 < b1) ? b2 : b1;           /* std::min(b1, b2) */
    return (begin_rev - end_fwd > rs) && (begin_rev - end_fwd < re);
}

typedef struct { orc_aln r, m; int proper; } orc_mob;

/* polishing.h:197-234 (std is never used downstream and is not reproduced) */
double orc_baseq_mean(const uint8_t *q1, const int32_t *off1, const uint8_t *q2, const int32_t *off2, int np)
{
    double avg = 0.0; int i, k;
    if (np == 0) return 0.0;
    for (i = 0; i < np; ++i) for (k = off1[i]; k < off1[i + 1]; ++k) avg += q1[k];
    for (i = 0; i < np; ++i) for (k = off2[i]; k < off2[i + 1]; ++k) avg += q2[k];
    return avg / (double)((unsigned long)np * (unsigned long)(off1[1] - off1[0]) +
                          (unsigned long)np * (unsigned long)(off2[1] - off2[0]));
}
