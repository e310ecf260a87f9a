/* TEST INFRASTRUCTURE ONLY -- CPU restatement of what the reference does with the final CIGAR:
 * refine_variant (include/sviper/variant.h:277-353) on the record built at include/sviper/sviper.h:460-466.
 * (assign_quality, evaluate_final_mapping.h:17-28, is orc_assign_quality in oracle_final.c.)
 * Pinned by tests/golden/eval_expected.txt (outputs of the unmodified reference, oracle/ref_harness.cpp EVAL). */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORC_DEV_SIZE 0.6 /* variant.h:10 */
#define ORC_DEV_POS 0.5  /* variant.h:11 */

/* One element of advance_in_cigar's loop body (include/sviper/basics.h:148-170). -1: unknown operation. */
static int orc_cigar_step(char op, unsigned cnt, int *ref_pos, int *read_pos)
{
    if (op == 'M' || op == 'X' || op == '=') { *read_pos += (int)cnt; *ref_pos += (int)cnt; }
    else if (op == 'I' || op == 'S') *read_pos += (int)cnt;
    else if (op == 'D') *ref_pos += (int)cnt;
    else if (op == 'H') {}
    else return -1;
    return 0;
}

/* refine_variant, coordinates only (variant.h:277-319). sv_type 1 = DEL, 2 = INS (variant.h:13-21).
 * Returns 1 and updates pos / end / svlen (and *read_pos = start of the inserted bases in the polished sequence)
 * when the CIGAR still holds the variant, 0 when it was "polished away" (FAIL5, sviper.h:474-478). */
int orc_refine_variant(int begin_pos, const char *op, const unsigned *cnt, int n, int sv_type,
                       int *pos, int *end, int *svlen, int *read_pos_out)
{
    int i = 0, ref_pos = begin_pos, read_pos = 0, has = 0;
    const int vpos = *pos, vlen = *svlen;
    /* first advance (variant.h:285-289): stop once ref >= ref_pos - DEV_POS * sv_length (double compare) */
    while (i < n)
    {
        if ((double)ref_pos >= (double)vpos - ORC_DEV_POS * (double)vlen) break;
        if (orc_cigar_step(op[i], cnt[i], &ref_pos, &read_pos)) return -1;
        ++i;
    }
    /* second advance (variant.h:292-310): the criterion looks at the element about to be consumed */
    while (i < n)
    {
        int same = (op[i] == 'D' && sv_type == 1) || (op[i] == 'I' && sv_type == 2);
        if (same && (double)cnt[i] >= (double)vlen - ORC_DEV_SIZE * (double)vlen &&
            (double)cnt[i] <= (double)vlen + ORC_DEV_SIZE * (double)vlen)
        {
            has = 1;
            break;
        }
        if ((double)ref_pos > (double)(vpos + vlen) + ORC_DEV_POS * (double)vlen) break;
        if (orc_cigar_step(op[i], cnt[i], &ref_pos, &read_pos)) return -1;
        ++i;
    }
    if (has) /* variant.h:312-319 */
    {
        *pos = ref_pos;
        *svlen = (int)cnt[i];
        *end = sv_type == 2 ? ref_pos : ref_pos + *svlen;
        *read_pos_out = read_pos;
    }
    return has;
}

/* std::regex_replace(info, "<lit>[class]*", repl) with the ECMAScript defaults: every non-overlapping match,
 * leftmost, greedy. Returns the number of matches. cls: 0 = [0-9], 1 = [A-Za-z]. */
static int orc_replace_all(const char *s, const char *lit, int cls, const char *repl, char *out, size_t cap)
{
    size_t ll = strlen(lit), o = 0, rl = strlen(repl);
    int hits = 0;
    while (*s)
    {
        if (strncmp(s, lit, ll) == 0)
        {
            s += ll;
            while (cls == 0 ? (*s >= '0' && *s <= '9') : ((*s >= 'A' && *s <= 'Z') || (*s >= 'a' && *s <= 'z'))) ++s;
            if (o + rl < cap) memcpy(out + o, repl, rl);
            o += rl;
            ++hits;
        }
        else
        {
            if (o + 1 < cap) out[o] = *s;
            ++o; ++s;
        }
    }
    if (o < cap) out[o] = 0;
    return hits;
}

/* The INFO rewrite of refine_variant (variant.h:321-349). ins_seq: the inserted bases as text (INS only, else NULL).
 * out must hold strlen(info) + strlen(ins_seq) + 64 bytes per possible match; returns 0. */
int orc_refine_info(const char *info, int sv_type, int svlen, int end, const char *ins_seq, char *out, size_t cap)
{
    char *a = (char *)malloc(cap), *b = (char *)malloc(cap), num[64];
    char *rep;
    snprintf(a, cap, "%s", info);
    /* SVLEN (variant.h:327-332): negative form first, then the plain one, else append */
    snprintf(num, sizeof num, "SVLEN=-%d", svlen);
    if (orc_replace_all(a, "SVLEN=-", 0, num, b, cap)) { char *t = a; a = b; b = t; }
    else
    {
        snprintf(num, sizeof num, "SVLEN=%d", svlen);
        if (orc_replace_all(a, "SVLEN=", 0, num, b, cap)) { char *t = a; a = b; b = t; }
        else { snprintf(b, cap, "%s;SVLEN=%d", a, svlen); { char *t = a; a = b; b = t; } }
    }
    /* END (variant.h:334-337); note that the pattern also hits the tail of CIEND= */
    snprintf(num, sizeof num, "END=%d", end);
    if (orc_replace_all(a, "END=", 0, num, b, cap)) { char *t = a; a = b; b = t; }
    else { snprintf(b, cap, "%s;END=%d", a, end); { char *t = a; a = b; b = t; } }
    /* SEQ (variant.h:339-348), insertions only */
    if (sv_type == 2)
    {
        size_t n = strlen(ins_seq ? ins_seq : "") + 8;
        rep = (char *)malloc(n);
        snprintf(rep, n, "SEQ=%s", ins_seq ? ins_seq : "");
        if (orc_replace_all(a, "SEQ=", 1, rep, b, cap)) { char *t = a; a = b; b = t; }
        else { snprintf(b, cap, "%s;%s", a, rep); { char *t = a; a = b; b = t; } }
        free(rep);
    }
    snprintf(out, cap, "%s", a);
    free(a); free(b);
    return 0;
}
