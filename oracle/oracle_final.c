/* TEST INFRASTRUCTURE ONLY (see sviper_oracle.c). Part of the unity build oracle_all.c.
 *
 * Final alignment of the polished sequence against the reference flank (sviper.h:452-466):
 * globalAlignment(gapsRef, gapsSeq, Score<double,Simple>(MM,MX,GE,GO), AlignConfig<0,0,0,0>, ConvexGaps)
 * + getCigarString. ConvexGaps exists only in the vendored SeqAn fork:
 *   recurrences  seqan/align/dp_formula_convex.h:62-141 (All), :276-325 (Horizontal), :331-380 (Vertical)
 *   cell         seqan/align/dp_cell_convex.h:58-67  {S, H, V, Hlen=1, Vlen=1}
 *   "infinity"   seqan/align/dp_cell.h:137-145 = numeric_limits<double>::min()/2 (a POSITIVE denormal)
 *   end cell     only (n,m) is tracked (dp_meta_info.h:204-216); correction dp_algorithm_impl.h:1212-1229
 * Compiled with -ffp-contract=off: every add/div below is one IEEE-754 operation, as in the
 * reference build (-O3, no -march, no fast-math).
 */

/* std::min(a, b) */
static double orc_min(double a, double b) { return (b < a) ? b : a; }

/* ops needs n+m chars. Returns the score of cell (n,m). */
double orc_align_convex(const uint8_t *ref, int n, const uint8_t *seq, int m,
                        double MM, double MX, double GE, double GO, char *ops, int *nops)
{
    const double INF = DBL_MIN / 2;
    size_t ld = (size_t)m + 1;
    double *S = (double *)malloc(ld * sizeof(double));
    double *H = (double *)malloc(ld * sizeof(double));
    double *HL = (double *)malloc(ld * sizeof(double));
    uint8_t *T = (uint8_t *)malloc(ld * ((size_t)n + 1));
    double cvS, cvV, cvVL, diag, endS, endH, endV;
    int i, j;
    if (n <= 0 || m <= 0) { free(S); free(H); free(HL); free(T); *nops = 0; return -DBL_MAX; }
    for (i = 0; i <= m; ++i) { S[i] = INF; H[i] = INF; HL[i] = 1; }
    /* column 0: Zero, then Vertical */
    S[0] = 0; cvS = 0; cvV = INF; cvVL = 1; T[0] = T_NONE;
    for (i = 1; i <= m; ++i)
    {
        double ext = cvV + orc_min(GE, GO / cvVL), opn = cvS + GO, vl;
        unsigned tv;
        if (ext < opn) { cvV = opn; tv = T_VO; vl = 1; } else { cvV = ext; tv = T_V; vl = cvVL + 1; }
        H[i] = INF; S[i] = cvV;
        cvS = S[i]; cvVL = vl;
        T[i] = (uint8_t)(tv | T_MFV);
    }
    endS = S[m]; endH = H[m]; endV = cvV;
    for (j = 1; j <= n; ++j)
    {
        uint8_t *Tc = T + (size_t)j * ld;
        uint8_t h = ref[j - 1];
        /* row 0: Horizontal */
        {
            double ext, opn; unsigned tv;
            diag = S[0];
            ext = H[0] + orc_min(GE, GO / HL[0]); opn = S[0] + GO;
            if (ext < opn) { H[0] = opn; tv = T_HO; HL[0] = 1; } else { H[0] = ext; tv = T_H; HL[0] = HL[0] + 1; }
            cvV = INF; S[0] = H[0];
            cvS = S[0]; cvVL = 1;      /* the row-0 column cell's vertical length is never written: stays 1 */
            Tc[0] = (uint8_t)(tv | T_MFH);
        }
        for (i = 1; i <= m; ++i)
        {
            double inter = diag + ((h == seq[i - 1]) ? MM : MX);
            double ext, opn, g, s, vl;
            unsigned tv, tv2;
            diag = S[i];
            ext = H[i] + orc_min(GE, GO / HL[i]); opn = S[i] + GO;
            if (ext < opn) { H[i] = opn; tv = T_HO; HL[i] = 1; } else { H[i] = ext; tv = T_H; HL[i] = HL[i] + 1; }
            ext = cvV + orc_min(GE, GO / cvVL); opn = cvS + GO;
            if (ext < opn) { cvV = opn; tv |= T_VO; vl = 1; } else { cvV = ext; tv |= T_V; vl = cvVL + 1; }
            if (cvV < H[i]) { g = H[i]; tv2 = T_MFH; } else { g = cvV; tv2 = T_MFV; }
            if (inter < g) { s = g; tv = tv2 | tv; } else { s = inter; tv = T_DIAG | tv; }
            S[i] = s; cvS = s; cvVL = vl;
            Tc[i] = (uint8_t)tv;
        }
        endS = S[m]; endH = H[m]; endV = cvV;
    }
    /* dp_scout.h:168-180: the scout starts from a default cell (S = INF, host position 0) and only takes
     * the tracked cell (n,m) if its score is strictly greater. A non-positive final score therefore leaves
     * the end cell at (0,0): the traceback degenerates to the tail segments (n x 'D' then m x 'I') and the
     * returned score is INF. Verified against the compiled reference. */
    if (!(endS > INF))
    {
        for (i = 0; i < n; ++i) ops[i] = 'D';
        for (i = 0; i < m; ++i) ops[n + i] = 'I';
        *nops = n + m;
        free(S); free(H); free(HL); free(T);
        return INF;
    }
    {
        uint8_t *tp = T + (size_t)n * ld + m;
        if (endV == endS) { *tp &= (uint8_t)~T_DIAG; *tp |= T_MFV; }
        else if (endH == endS) { *tp &= (uint8_t)~T_DIAG; *tp |= T_MFH; }
    }
    *nops = orc_traceback(T, n, m, n, m, ops);
    free(S); free(H); free(HL); free(T);
    return endS;
}

/* seqan/bam_io/cigar.h:354-420 with the D->N threshold disabled (sviper.h:466): plain run-length
 * of M/I/D. Returns the number of CIGAR elements. */
int orc_cigar(const char *ops, int nops, char *op_out, uint32_t *cnt_out, int cap)
{
    int k = 0, i;
    for (i = 0; i < nops; ++i)
    {
        if (k > 0 && op_out[k - 1] == ops[i]) { ++cnt_out[k - 1]; continue; }
        if (k == cap) return -1;
        op_out[k] = ops[i]; cnt_out[k] = 1; ++k;
    }
    return k;
}

/* include/sviper/evaluate_final_mapping.h:17-28 */
double orc_assign_quality(int n_cigar, int flanking_region)
{
    double error_rate = ((double)n_cigar - 1.0) / (flanking_region * 2.0);
    double fuzzyness = (1.0 - error_rate / 0.15) * 100.0;
    return (fuzzyness < 0.0) ? 0.0 : fuzzyness;   /* std::max(fuzzyness, 0.0) */
}
