/* TEST INFRASTRUCTURE ONLY (see sviper_oracle.c). Part of the unity build oracle_all.c.
 *
 * The two pairwise DP fills inside sviper::build_consensus (include/sviper/basics.h:484-520 ->
 * seqan::globalMsaAlignment, graph_msa/graph_align_tcoffee_msa.h:323-548): SURVEY 8f row f2. Measured
 * (profiles/r2_msa_profile.jsonl): ~80 % of build_consensus is the local pairwise library, ~10 % the global one.
 * Restated here as plain C so that the CUDA fills (sviper_b200/csrc/k6_msa.cuh) can be compared byte for byte;
 * pinned by fixtures the reference harness printed (tests/golden/msa_*.txt).
 *
 *   orc_msa_sw_fill      graph_msa/graph_align_tcoffee_library.h:252-390  _alignSmithWaterman (legacy "clumping"
 *                        Smith-Waterman: affine gaps, cells on earlier local alignments are forbidden)
 *   orc_msa_global_fill  align/dp_formula_affine.h:66-125, :243-325 + dp_algorithm_impl.h:489-573: the trace matrix of
 *                        globalAlignment(String<Fragment>, ..., AlignConfig<>, Gotoh()) (global_alignment_unbanded.h:386-407)
 * seq h = str[0] = columns (len1 / length n), seq v = str[1] = rows (len2 / length m); bases 0..4, equal == match
 * (score_base.h:333-340). */

/* Legacy trace nibble (library.h:262, :331-360): bit 3 vertical gap OPENED here, bit 2 horizontal gap OPENED here,
 * bits 0-1: 0 diagonal, 1 horizontal, 2 vertical, 3 stop. Two rows per byte, odd rows (1, 3, ...) in the high
 * nibble; a column holds (m + 1) / 2 bytes (library.h:299, :372-375).
 * forbidden: n * m bytes, [(col - 1) * m + (row - 1)] (library.h:128-136), may be NULL.
 * out4: score_max, row, col (indexPair[0], indexPair[1]), initialDir. */
void orc_msa_sw_fill(const uint8_t *h, int n, const uint8_t *v, int m, int mt, int mm, int ge, int go,
                     const uint8_t *forbidden, uint8_t *trace, int *out4)
{
    int *mat = (int *)malloc(((size_t)m + 1) * sizeof(int)), *hor = (int *)malloc(((size_t)m + 1) * sizeof(int));
    int score_max = 0, brow = 0, bcol = 0, bdir = 3, row, col;
    size_t nrows = ((size_t)m >> 1) + (m & 1), it = 0;
    unsigned char cur;
    memset(trace, 0, (size_t)n * nrows);
    mat[0] = 0;
    for (row = 1; row <= m; ++row) { mat[row] = 0; hor[row] = go - ge; }
    for (col = 1; col <= n; ++col)
    {
        int diag = mat[0], vert = go - ge, max_val = 0, tv = 0;
        mat[0] = 0;
        cur = 0;
        for (row = 1; row <= m; ++row)
        {
            if (forbidden && forbidden[(size_t)(col - 1) * m + (row - 1)])
            {
                cur = (unsigned char)(cur << 3); cur |= 3;
                max_val = 0; vert = 0; hor[row] = 0; tv = 3;
            }
            else
            {
                int a = mat[row - 1] + go, b = vert + ge;         /* mat[row-1] already holds this column's value */
                if (a > b) { vert = a; cur |= 1; } else vert = b;
                cur = (unsigned char)(cur << 1);
                a = mat[row] + go; b = hor[row] + ge;
                if (a > b) { hor[row] = a; cur |= 1; } else hor[row] = b;
                cur = (unsigned char)(cur << 2);
                max_val = diag + (h[col - 1] == v[row - 1] ? mt : mm);
                tv = 0;
                if (vert > max_val) { max_val = vert; tv = 2; }
                if (hor[row] > max_val) { max_val = hor[row]; tv = 1; }
                if (0 >= max_val) { max_val = 0; tv = 3; }
                cur |= (unsigned char)tv;
            }
            diag = mat[row];
            mat[row] = max_val;
            if (max_val > score_max) { brow = row; bcol = col; score_max = max_val; bdir = tv; }
            if (row & 1) cur = (unsigned char)(cur << 1);
            else { trace[it++] = cur; cur = 0; }
        }
        if (m & 1) { cur = (unsigned char)(cur << 3); trace[it++] = cur; }
    }
    /* (the forbidden branch leaves tvMat untouched in the reference; initialDir is only taken where max_val > score_max,
     * which a forbidden cell (max_val = 0) never is, so the value of tv there is never observed) */
    out4[0] = score_max; out4[1] = brow; out4[2] = bcol; out4[3] = bdir;
    free(mat); free(hor);
}

/* SeqAn trace bits (align/dp_profile.h:125-134) */
enum { TB_NONE = 0, TB_DIAG = 1, TB_H = 2, TB_V = 4, TB_HO = 8, TB_VO = 16, TB_MAXH = 32, TB_MAXV = 64 };

/* Trace matrix as DPMatrix_<TraceValue, FullDPMatrix> stores it: (m + 1) x (n + 1) bytes, element (row i, column j)
 * at j * (m + 1) + i. Tie rules of _maxScore, SingleTrace: (a < b) ? b : a (dp_formula.h:152-163) -- extension,
 * vertical and diagonal win ties. end3: S, H, V of the bottom-right cell (the scout's _maxScore; _correctTraceValue,
 * dp_algorithm_impl.h:1170-1185, is applied by the caller). Returns the alignment score. */
int orc_msa_global_fill(const uint8_t *h, int n, const uint8_t *v, int m, int mt, int mm, int ge, int go,
                        uint8_t *trace, int *end3)
{
    const int NEG = INT_MIN / 2;                      /* dp_cell.h:144-145 */
    int *S = (int *)malloc(((size_t)m + 1) * sizeof(int)), *H = (int *)malloc(((size_t)m + 1) * sizeof(int));
    int i, j, vlast = NEG;
    size_t ld = (size_t)m + 1;
    /* initial column: (0,0) Zero, below it the Vertical recursion (dp_formula_affine.h:291-325) */
    S[0] = 0; H[0] = NEG; trace[0] = TB_NONE;
    {
        int V = NEG;
        for (i = 1; i <= m; ++i)
        {
            int ext = V + ge, opn = S[i - 1] + go;
            uint8_t t = (ext < opn) ? TB_VO : TB_V;
            V = (ext < opn) ? opn : ext;
            S[i] = V; H[i] = NEG;
            trace[i] = (uint8_t)(t | TB_MAXV);
            vlast = V;
        }
    }
    if (n == 0) { end3[0] = S[m]; end3[1] = NEG; end3[2] = (m ? vlast : NEG); j = S[m]; free(S); free(H); return j; }
    for (j = 1; j <= n; ++j)
    {
        uint8_t *tc = trace + (size_t)j * ld;
        int diag = S[0], V = NEG;
        /* first row: Horizontal recursion (dp_formula_affine.h:243-285) */
        {
            int ext = H[0] + ge, opn = S[0] + go;
            uint8_t t = (ext < opn) ? TB_HO : TB_H;
            H[0] = (ext < opn) ? opn : ext;
            S[0] = H[0];
            tc[0] = (uint8_t)(t | TB_MAXH);
        }
        for (i = 1; i <= m; ++i)
        {
            int inter = diag + (h[j - 1] == v[i - 1] ? mt : mm);
            int hext = H[i] + ge, hopn = S[i] + go, vext = V + ge, vopn = S[i - 1] + go, hh, gap, sn;
            uint8_t t = (hext < hopn) ? TB_HO : TB_H, t2;
            hh = (hext < hopn) ? hopn : hext;
            t |= (vext < vopn) ? TB_VO : TB_V;
            V = (vext < vopn) ? vopn : vext;
            t2 = (V < hh) ? TB_MAXH : TB_MAXV;
            gap = (V < hh) ? hh : V;
            diag = S[i];
            if (inter < gap) { sn = gap; t |= t2; } else { sn = inter; t |= TB_DIAG; }
            S[i] = sn; H[i] = hh;
            tc[i] = t;
            vlast = V;
        }
    }
    end3[0] = S[m]; end3[1] = H[m]; end3[2] = (m ? vlast : NEG);
    j = S[m];
    free(S); free(H);
    return j;
}
