/* sviper_b200 -- C ABI of the B200 (sm_100a) implementation of SViper's per-variant polishing hot path.
 *
 * The reference (smehringer/SViper) has no plugin/FFI interface: the hot path is three call seams inside
 * one translation unit. Each entry point below replaces one of them:
 *
 *   svp_polish_batch        include/sviper/polishing.h:319-346  polish_to_perfection(reads1, reads2, ref, config),
 *                           called at include/sviper/sviper.h:431, plus (when ref_part is given) the final
 *                           alignment + CIGAR at include/sviper/sviper.h:452-466
 *   svp_final_align_batch   include/sviper/sviper.h:457-466     globalAlignment(..., ConvexGaps()) + getCigarString
 *   svp_align_batch         include/sviper/mapping.h:84-137     map_single_read (fwd + RC semi-global affine
 *                           alignment, strand pick, mapQ) -- also the config-5 GCUPS microbenchmark entry
 *   svp_submit_* / svp_wait_any   src/sviper.cpp:43-47          the per-variant OpenMP loop -> device batch scheduler
 *
 * Conventions: plain C types; all pointers are HOST memory owned by the caller; device memory is owned by
 * the context; bases are uint8 0..4 = A,C,G,T,N (seqan::Dna5 ordValue); qualities uint8 0..62 with N carrying
 * 0 (seqan::Dna5Q, extern/seqan/include/seqan/basic/alphabet_residue.h:331-365). Every function returns
 * SVP_OK (0) or a negative error; any CUDA error is fatal for the call -- there is NO CPU fallback.
 * One context per GPU; a context is used by one thread at a time.
 */
#ifndef SVIPER_B200_H
#define SVIPER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVP_OK 0
#define SVP_ERR_CUDA (-1)        /* a CUDA call or kernel failed; svp_last_error() has the text */
#define SVP_ERR_ARG (-2)         /* invalid argument (empty read1, window shorter than 2*flank, ...) */
#define SVP_ERR_CAPACITY (-3)    /* a window outgrew its buffers (polished sequence longer than svp_window_capacity(len)),
                                    a caller buffer is too small, or too many read pairs in one variant (~4,000) */
#define SVP_ERR_UNSUPPORTED (-4) /* read longer than SVP_MAX_READ_LEN */

#define SVP_MAX_READ_LEN 256

typedef struct svp_ctx svp_ctx;

/* Creates a context on cuda_device. cuda_stream: a cudaStream_t (as void*) every kernel and copy of this
 * context is issued on; NULL = the context creates its own stream. */
svp_ctx *svp_create(int cuda_device, void *cuda_stream);
void svp_destroy(svp_ctx *ctx);
const char *svp_last_error(const svp_ctx *ctx);
const char *svp_version(void);

/* ------------------------------------------------------------------------------------------------
 * map_single_read batch (include/sviper/mapping.h:84-137).
 * Task t aligns read read_id[t] against window win_id[t], forward and reverse-complement, full
 * unbanded semi-global DP, Score<int>(1,-4,-1,-6) (mapping.h:94-97).
 * ---------------------------------------------------------------------------------------------- */
typedef struct
{
    int32_t n_reads;
    const uint8_t *read_bases;   /* concatenated */
    const int32_t *read_off;     /* n_reads + 1 offsets into read_bases */
    int32_t n_windows;
    const uint8_t *win_bases;    /* concatenated */
    const int32_t *win_off;      /* n_windows + 1 */
    int32_t n_tasks;
    const int32_t *task_read;    /* n_tasks */
    const int32_t *task_win;     /* n_tasks */
} svp_align_in;

typedef struct
{
    /* each n_tasks long; any pointer may be NULL to skip that output */
    int32_t *score;      /* forward-strand score */
    int32_t *score_rc;   /* reverse-complement score */
    int32_t *is_rc;      /* 1 iff !(score > score_rc)           (mapping.h:121) */
    int32_t *lead;       /* gapsBeginPos(gapsRead) of the winner (mapping.h:66-70) */
    int32_t *endpos;     /* gapsEndPos(gapsRead) of the winner   (mapping.h:59-63) */
    /* optional alignment of the winner, one op per view column between the read row's leading and
     * trailing gaps: 'M' base/base, 'D' gap in read row, 'I' gap in consensus row. ops_cap bytes per
     * task; ops_len[t] = number of ops (ops beyond ops_cap are dropped). May be NULL. */
    char *ops;
    int32_t ops_cap;
    int32_t *ops_len;
} svp_align_out;

/* with_traceback = 0: scores / strand only (lead, endpos, ops are not written). */
int svp_align_batch(svp_ctx *ctx, const svp_align_in *in, svp_align_out *out, int with_traceback);

/* Same work as svp_align_batch(with_traceback=0) but inputs already resident on the device from the last
 * svp_align_batch call: re-runs the kernels only (used by bench.py for the HBM-resident `value`).
 * kernel_ms (may be NULL) receives the CUDA-event time of the dominant (score) kernel alone. */
int svp_align_rerun(svp_ctx *ctx, int with_traceback, float *kernel_ms);

/* ------------------------------------------------------------------------------------------------
 * polish_to_perfection (+ final alignment) for a batch of independent variants.
 * ---------------------------------------------------------------------------------------------- */
typedef struct
{
    /* flanked consensus (include/sviper/sviper.h:419-422) */
    const uint8_t *consensus;
    int32_t consensus_len;
    int32_t ref_flank_length;           /* SViperConfig::ref_flank_length, 500 (sviper.h:363) */
    /* read pairs: reads2[i] is the mate of reads1[i]; a zero-length mate is the dummy (basics.h:317,398) */
    int32_t n_pairs;
    const uint8_t *reads1_bases, *reads1_quals;
    const int32_t *reads1_off;          /* n_pairs + 1 */
    const uint8_t *reads2_bases, *reads2_quals;
    const int32_t *reads2_off;          /* n_pairs + 1 */
    /* SViperConfig scalars (include/sviper/config.h) */
    double baseQ_mean;                  /* compute_baseQ_stats, polishing.h:197-234 (host, sviper.h:426) */
    double insert_mean, insert_sd;      /* CmdOptions, config.h:16-17 */
    double min_coverage;                /* config.h:79 (4) */
    /* optional reference part for the final alignment (sviper.h:447-450); NULL = polish only */
    const uint8_t *ref_part;
    int32_t ref_part_len;
} svp_window_in;

typedef struct
{
    int32_t status;                     /* SVP_OK or a per-variant error */
    /* polished sequence (caller buffer of polished_cap bytes) */
    uint8_t *polished;
    int32_t polished_cap;
    int32_t polished_len;
    /* SViperConfig side effects (config.h:72-90) */
    uint32_t rounds, substituted_bases, inserted_bases, deleted_bases;
    double mappQ_mean;
    /* coverage profile of the last round (config.cov_profile; logged at sviper.h:441-443). Optional. */
    uint32_t *cov_profile;
    int32_t cov_cap;
    int32_t cov_len;
    /* final alignment CIGAR (sviper.h:466); filled when ref_part was given. Optional buffers. */
    char *cigar_op;
    uint32_t *cigar_cnt;
    int32_t cigar_cap;
    int32_t n_cigar;                    /* number of CIGAR elements (assign_quality uses only this) */
    double final_score;
    /* rounds for which the reads were re-aligned: `rounds` minus the first indel round when it follows a round
     * that changed nothing (same consensus, same mapping; only consensus_from_profile differs, polishing.h:327-343) */
    uint32_t alignment_rounds;
    uint32_t reserved_;
} svp_result;

/* Per-variant limits (read longer than SVP_MAX_READ_LEN, empty read1, window shorter than 2 x flank, more read pairs than
 * the pileup kernel stages) are reported in that variant's svp_result.status; such a variant is left out of the device
 * batch and the OTHER variants of the call are polished normally. The return value is non-zero only when the call as a
 * whole failed (CUDA error, out of device memory). */
int svp_polish_batch(svp_ctx *ctx, int32_t n_variants, const svp_window_in *in, svp_result *out);

/* Longest polished sequence a window of consensus_len bases can grow to on the device (consensus_len + consensus_len / 4 +
 * 512): size svp_result.polished with it; 3 x that + 1024 entries always hold the coverage profile. The reference has no
 * limit (polishing.h:236-293 appends to a string); a window that outgrows it ends with status SVP_ERR_CAPACITY. */
int32_t svp_window_capacity(int32_t consensus_len);

/* Re-runs the last svp_polish_batch (every round and the final alignment) from its HBM-resident inputs;
 * nothing is copied from or to the host. Used by bench.py for the device-resident throughput. */
int svp_polish_rerun(svp_ctx *ctx);

/* Final alignment only: ref_part[i] (H) vs seq[i] (V), Score<double>(90000,-80000,-1,-100000), ConvexGaps. */
typedef struct
{
    const uint8_t *ref_part;
    int32_t ref_part_len;
    const uint8_t *seq;
    int32_t seq_len;
} svp_final_in;

int svp_final_align_batch(svp_ctx *ctx, int32_t n, const svp_final_in *in, svp_result *out);

/* Final alignment of ONE pair without any allocation -- NO COUNTERPART IN THE REFERENCE; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same kernels and results as svp_final_align_batch, but all scratch (tables, padded sequence copies, job
 * descriptor, band rows, trace) lives in one file-scope device array of fixed size (about 10.7 MB, held by every GPU that
 * loads the library until the process exits), so ref_part_len and seq_len are limited to SVP_STATIC_MAX_LEN: longer
 * sequences are refused with SVP_ERR_CAPACITY (svp_last_error names the limit). Unlike the rest of this header the sequence,
 * CIGAR and result pointers are DEVICE memory owned by the caller (bases uint8 0..4; cigar buffers of cigar_cap elements).
 * The call only queues three kernels on the context's stream and returns; it calls no cudaMalloc / cudaFree / cudaHostAlloc
 * and no allocating library routine. The first call on a device does a one-time synchronous set-up (symbol address, 64 KB of
 * tables copied into the scratch): make it outside a graph capture. Every call on a device shares the scratch, which keeps
 * the previous call's contents (what must start at zero is cleared by the call itself): calls that overlap on the GPU --
 * different streams or contexts -- race; calls on one stream are ordered and safe. result->n_cigar > cigar_cap means the
 * CIGAR was truncated (result->stored elements were written). */
#define SVP_STATIC_MAX_LEN 4096
typedef struct
{
    int32_t n_cigar;     /* CIGAR elements of the alignment */
    int32_t stored;      /* elements written to cigar_op_dev / cigar_cnt_dev (<= cigar_cap) */
    double score;        /* the alignment score the reference discards (sviper.h:457) */
} svp_final_static_result;
int svp_final_align_static(svp_ctx *ctx, const uint8_t *ref_part_dev, int32_t ref_part_len, const uint8_t *seq_dev,
                           int32_t seq_len, char *cigar_op_dev, uint32_t *cigar_cnt_dev, int32_t cigar_cap,
                           svp_final_static_result *result_dev);

/* include/sviper/evaluate_final_mapping.h:17-28 (host arithmetic, provided for completeness) */
double svp_assign_quality(int32_t n_cigar, int32_t flanking_region);

/* ------------------------------------------------------------------------------------------------
 * What polish_variant does with the final CIGAR (include/sviper/sviper.h:468-489). Host arithmetic over
 * the few dozen CIGAR elements of svp_result; together with svp_assign_quality these define the VCF
 * fields the parity contract names: POS, END, SVLEN, SEQ, QUAL, FILTER (FAIL5).
 * ---------------------------------------------------------------------------------------------- */
#define SVP_SV_DEL 1   /* SV_TYPE::DEL, include/sviper/variant.h:13-21 */
#define SVP_SV_INS 2   /* SV_TYPE::INS */

typedef struct
{
    int32_t sv_type;       /* SVP_SV_DEL or SVP_SV_INS */
    int32_t ref_pos;       /* Variant::ref_pos */
    int32_t ref_pos_end;   /* Variant::ref_pos_end */
    int32_t sv_length;     /* Variant::sv_length */
} svp_variant_pos;

/* refine_variant, coordinates (include/sviper/variant.h:277-319). begin_pos = final_record.beginPos
 * (sviper.h:461), cigar = svp_result.cigar_op / cigar_cnt / n_cigar. Returns 1 and updates *v (and *ins_begin =
 * offset of the inserted bases in the polished sequence, may be NULL) when the alignment still holds the variant;
 * 0 when it was "polished away" (caller sets FILTER=FAIL5, sviper.h:474-478; *v unchanged); SVP_ERR_ARG for an
 * unknown CIGAR operation (the reference throws) or an sv_type other than DEL / INS. */
int svp_refine_variant(int32_t begin_pos, const char *cigar_op, const uint32_t *cigar_cnt, int32_t n_cigar,
                       svp_variant_pos *v, int32_t *ins_begin);

/* refine_variant, INFO rewrite (include/sviper/variant.h:321-349): SVLEN= / END= / SEQ= replaced where present
 * (every occurrence, with std::regex_replace's semantics: END= also hits the tail of CIEND=), appended otherwise.
 * *v is the refined variant, polished the polished sequence (uint8 0..4), ins_begin from svp_refine_variant
 * (both used for insertions only). Writes at most out_cap - 1 characters + NUL; returns the full length. */
int svp_refine_info(const char *info, const svp_variant_pos *v, const uint8_t *polished, int32_t polished_len,
                    int32_t ins_begin, char *out, int32_t out_cap);

/* QUAL as Variant::write prints it (default ostream format of a double, include/sviper/variant.h:142). */
int svp_format_quality(double quality, char *out, int32_t out_cap);

/* ------------------------------------------------------------------------------------------------
 * What polish_variant does with BAM records BEFORE the hot path. Host arithmetic over
 * (op, count) CIGAR arrays, no SeqAn types; reading the BAM / FASTA files stays the reference's own code.
 * ---------------------------------------------------------------------------------------------- */
/* record_supports_variant (include/sviper/variant.h:239-275, called for every long read at sviper.h:236-249):
 * 1 if the record's CIGAR holds a D (DEL) / I (INS) element of 0.4..1.6 x sv_length inside
 * [ref_pos - sv_length/2, ref_pos + 1.5 x sv_length], else 0 (also for any other sv_type); SVP_ERR_ARG for an unknown
 * CIGAR operation (the reference throws std::logic_error). */
int svp_record_supports_variant(int32_t begin_pos, const char *cigar_op, const uint32_t *cigar_cnt, int32_t n_cigar,
                                const svp_variant_pos *v);

/* get_read_region_boundaries (include/sviper/basics.h:189-241, the long-read crop of sviper.h:318-340): the read
 * positions that correspond to the reference region [ref_region_begin, ref_region_end) under the record's CIGAR.
 * Returns 0; 1 with (0, 0) when begin >= end (the reference prints a message and returns (0, 0)); SVP_ERR_ARG for an
 * unknown operation or a region that ends before the record starts (the reference reads cigar[-1] there). */
int svp_read_region_boundaries(int32_t begin_pos, const char *cigar_op, const uint32_t *cigar_cnt, int32_t n_cigar,
                               int32_t ref_region_begin, int32_t ref_region_end, int32_t *read_region_begin,
                               int32_t *read_region_end);

/* seqan::getAlignmentLengthInRef (extern/seqan/include/seqan/bam_io/cigar.h:598-610): everything but S, H, I. */
int32_t svp_alignment_length_in_ref(const char *cigar_op, const uint32_t *cigar_cnt, int32_t n_cigar);

/* cut_down_high_coverage (include/sviper/basics.h:403-442, sviper.h:405): the short-read records of a region, sorted by
 * (position ascending, mapping quality descending) and thinned so that no more than mean_coverage kept reads start
 * while that many are still open. keep[] (capacity n) receives the indices of the kept records in output order;
 * returns their number, or SVP_ERR_ARG where the reference would run out of its `ends` array (a record that reaches
 * beyond the last record's end + 500, or a gap of more than 32,767 bases). */
int svp_cut_down_high_coverage(int32_t n, const int32_t *begin_pos, const uint8_t *mapq, const int32_t *len_in_ref,
                               int32_t mean_coverage, int32_t *keep);

/* merge_record_group (include/sviper/merge_split_alignments.h:378-534; merge_alignments, :536-561, calls it once per read
 * name, sviper.h:262-266): the supplementary alignments of a read are folded into its primary alignment, so that a split
 * long read shows the DEL / INS as one CIGAR element. recs = all records of ONE read name in input order. *chosen = index
 * of the record the result is based on (first after the reference's sort: primary first, then by mapping quality; its
 * flag / mapQ / name carry over), begin_pos / cigar / seq = the merged alignment. SVP_ERR_ARG for an unknown CIGAR
 * operation or where the reference indexes out of bounds (empty CIGAR, clip longer than the other record's sequence);
 * SVP_ERR_CAPACITY if the output buffers are too small (n_cigar / seq_len then hold the needed sizes). */
typedef struct
{
    int32_t begin_pos;          /* BamAlignmentRecord::beginPos (0-based) */
    int32_t rid;                /* rID */
    uint16_t flag;              /* BAM flag: 0x10 reverse, 0x100 secondary, 0x800 supplementary */
    uint8_t mapq;
    const char *cigar_op;
    const uint32_t *cigar_cnt;
    int32_t n_cigar;
    const char *seq;            /* as stored in the record (hard-clipped bases absent) */
    int32_t seq_len;
} svp_bam_record;

int svp_merge_record_group(int32_t n, const svp_bam_record *recs, int32_t *chosen, int32_t *begin_pos, char *cigar_op,
                           uint32_t *cigar_cnt, int32_t cigar_cap, int32_t *n_cigar, char *seq, int32_t seq_cap,
                           int32_t *seq_len);

/* ------------------------------------------------------------------------------------------------
 * The pairwise DP fills of build_consensus (include/sviper/basics.h:484-520 ->
 * seqan::globalMsaAlignment, extern/seqan/include/seqan/graph_msa/graph_align_tcoffee_msa.h:323-548). ~90 % of
 * build_consensus is spent filling pairwise DP matrices for the T-Coffee library; these entry points do the fills on the
 * device and return the TRACE MATRICES in SeqAn's own byte layouts, so that the host side hands them to SeqAn's unchanged
 * traceback code and everything after it (alignment graph, guide tree, progressive alignment, column vote) stays the
 * reference's host code. One session = one globalMsaAlignment call on one context.
 *   svp_msa_begin        the sequence pairs selectPairs() chose (graph_align_tcoffee_msa.h:346-347): h = str[0] = matrix
 *                        columns, v = str[1] = rows; scores = SimpleScore(match, mismatch, gap_extend, gap_open) (5,-3,-1,-3
 *                        at basics.h:494)
 *   svp_msa_global_fill  global pairwise library (graph_msa/graph_align_tcoffee_library.h:979-1008 -> globalAlignment(...,
 *                        Gotoh()), align/global_alignment_unbanded.h:386-407): trace_out[i] receives the (n + 1) x (m + 1)
 *                        TraceBitMap_ bytes of DPMatrix_<.., FullDPMatrix>, element (row i, column j) at j * (m + 1) + i
 *                        (align/dp_formula_affine.h:66-125, :243-325), res[i].S/H/V the bottom-right cell
 *   svp_msa_local_fill   one round of the local pairwise library's clumping Smith-Waterman (library.h:252-390
 *                        _alignSmithWaterman): trace_out[i] receives n * ((m + 1) / 2) bytes (two rows per byte, odd rows in
 *                        the high nibble), res[i] score / row / col / dir = score_max, indexPair, initialDir.
 *                        new_forbidden[i] = n_new[i] (col, row) pairs (1-based) the caller's traceback of the round before
 *                        marked with _setForbiddenCell; they stay forbidden for the rest of the session.
 * ---------------------------------------------------------------------------------------------- */
typedef struct
{
    const uint8_t *h; int32_t n;     /* bases 0..4 */
    const uint8_t *v; int32_t m;
} svp_msa_pair;

typedef struct
{
    int32_t score, row, col, dir;
    int32_t S, H, V;
} svp_msa_fill_result;

int svp_msa_begin(svp_ctx *ctx, int32_t n_pairs, const svp_msa_pair *pairs, int32_t match, int32_t mismatch,
                  int32_t gap_extend, int32_t gap_open);
int svp_msa_global_fill(svp_ctx *ctx, int32_t n, const int32_t *pair_idx, uint8_t *const *trace_out, svp_msa_fill_result *res);
int svp_msa_local_fill(svp_ctx *ctx, int32_t n, const int32_t *pair_idx, const int32_t *const *new_forbidden,
                       const int32_t *n_new, uint8_t *const *trace_out, svp_msa_fill_result *res);

/* ------------------------------------------------------------------------------------------------
 * Asynchronous device batch scheduler (replaces the loop body at src/sviper.cpp:43-47; SURVEY 8b "submissions
 * asynchronous with an explicit wait"). svp_submit_polish queues one variant and returns at once; one worker thread per
 * lane (svp_set_lanes) forms device batches from what is queued, runs them (staging, H2D, every round, final alignment,
 * D2H into the caller's buffers) and hands the results to svp_wait_any in completion order. The lanes run out of phase,
 * so the copies of one batch overlap the kernels of the others. Thread-safe: any thread may submit, any thread may wait.
 * Lifetime: *w is copied, but everything it points to, *r and everything *r points to must stay valid until svp_wait_any
 * has returned that result. A variant outside the per-variant limits completes at once with its status set.
 * A lane starts a batch when batch_target variants are queued (default 256, env SVP_BATCH_TARGET), on svp_flush, or when
 * a thread blocks in svp_wait_any with nothing running. Do not call the synchronous batch functions on the same context
 * while submissions are outstanding.
 * ---------------------------------------------------------------------------------------------- */
int svp_submit_polish(svp_ctx *ctx, uint64_t tag, const svp_window_in *w, svp_result *r);
int svp_flush(svp_ctx *ctx);                                     /* start on whatever is queued now */
int svp_wait_any(svp_ctx *ctx, uint64_t *tag, svp_result **r);   /* blocks; returns 1 when nothing is outstanding */
int svp_engine_config(svp_ctx *ctx, int32_t batch_target, int32_t max_batch);   /* <= 0: keep */
int svp_engine_stats(svp_ctx *ctx, uint64_t *batches, uint64_t *variants);      /* device batches formed so far */

/* ------------------------------------------------------------------------------------------------
 * Debug / measurement hooks
 * ---------------------------------------------------------------------------------------------- */
/* One polishing round (sviper::polish, polishing.h:295-317) on a single window with the profile exposed:
 * profile = Lp x 6 doubles, cov = Lp, isgap = Lp (1 = inserted column), proper = n_pairs flags. */
int svp_debug_round(svp_ctx *ctx, const svp_window_in *w, int fix_indels, double *profile, uint32_t *cov,
                    uint8_t *isgap, int32_t cap_cols, int32_t *n_cols, uint8_t *proper, uint8_t *out_seq,
                    int32_t out_cap, int32_t *out_len, double *mappQ_mean, uint32_t *counters /* subst, ins, del */);

/* DP cells the score kernel has computed for polish batches on this context since the last reset (both halves of every
 * packed lane count): cells[0] in regular rounds (both strands of every read, the reference's 2 x |read| x |window|),
 * cells[1] in mixed rounds (a pair's two known strands as one task: half of that). Synchronises the context's streams. */
int svp_k1_cells(svp_ctx *ctx, double *cells /* 2 entries */, int reset);

/* Reads whose strand certificate failed INSIDE a mixed launch since the last reset (they are re-aligned on both strands by a
 * further launch behind it; reads expected to fail are sent to the regular list up front and are not counted). */
int svp_k1_failed_certificates(svp_ctx *ctx, double *reads, int reset);

/* Measures the sustained rate of the DPX-class packed 16x2 instructions the score kernel is made of
 * (VIADDMNMX / VIMNMX3 / IADD), in lane-operations per second over the whole chip. */
int svp_probe_dpx(svp_ctx *ctx, double *viaddmax_lane_ops, double *vimax3_lane_ops, double *mix_lane_ops);

/* Measures the sustained fp64 add rate of the chip (independent DADD chains, all SMs), in lane-operations per second:
 * the yardstick of the final-alignment kernel, whose recurrence is 4 DADD + 4 fp64 compares per cell. */
int svp_probe_fp64(svp_ctx *ctx, double *dadd_lane_ops);

/* Per-kernel launch counts and CUDA-event milliseconds accumulated since the last reset.
 * names: k1_score (regular rounds), k1_trace, k2_traceback, k3_pileup_vote, k5_final, k5_traceback, pack, other,
 * k1_score_mixed (score pass of the mixed rounds: list building + mixed + regular launches) */
int svp_stats(svp_ctx *ctx, int32_t *n_launches, double *ms_by_kernel /* 9 entries */, int reset);

/* A polish batch of >= 8 variants is split into n_lanes contiguous parts that run concurrently on the same GPU
 * (own stream, arenas and host thread per lane), so that the latency-bound stages of one part overlap the
 * alignment kernel of another. Results do not depend on the setting. Default 4 (environment: SVP_LANES);
 * 1 = the whole batch as one part. The reference's counterpart is the OpenMP loop over variants
 * (src/sviper.cpp:43-47), whose iterations are independent in the same way. */
int svp_set_lanes(svp_ctx *ctx, int32_t n_lanes);

#ifdef __cplusplus
}
#endif
#endif
