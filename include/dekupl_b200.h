/*
 * dekupl_b200.h — C ABI of libdekupl_b200.so: the B200-native count-matrix hot
 * path of DE-kupl (joinCounts -> computeNF -> TtestFilter).
 *
 * The reference (Transipedia/dekupl-run) has no FFI: its boundary is three
 * single-threaded CLIs over gz/TSV text. Each entry point below cites the
 * reference interface it replaces; the drop-in CLIs (same names, argv, stdout)
 * in dekupl-run_b200/host/ are thin main()s over this ABI.
 *
 * Conventions: plain C, int return codes (0 = ok, <0 = DK_ERR_*), message via
 * dk_last_error(); no exceptions; all array arguments are CALLER-OWNED; pointers
 * are HOST pointers unless the name ends in _dev; library-owned objects
 * (dk_matrix, dk_ttest) are released with their *_free. A context is bound to
 * one GPU and one caller thread (the reference tools are single-threaded).
 * There is NO CPU fallback: without a CUDA device every call fails loudly.
 *
 * Packed formats: k-mer key = uint64, 2 bits/base A=0 C=1 G=2 T=3, first base in
 * the most significant used bits (k <= 31; as share/mergeTags/dna.h:49-59), so
 * unsigned order == strcmp order of the reference. Count = int32 (reference:
 * `int` via atoi). Matrix row = key + S counts, row-major.
 */
#ifndef DEKUPL_B200_H
#define DEKUPL_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define DK_OK               0
#define DK_ERR_CUDA        -1   /* CUDA runtime/driver error (message has details) */
#define DK_ERR_ARG         -2   /* bad argument (S out of range, misaligned pointer, ...) */
#define DK_ERR_NOMEM       -3
#define DK_ERR_UNSUPPORTED -4   /* reference behaviour we refuse to imitate (e.g. group of 1 => Boost abort) */
#define DK_ERR_NODEVICE    -5

#define DK_MAX_SAMPLES 256

#define DK_NF_REFERENCE 0  /* computeNF.c:138-168 prefix / early-stop semantics */
#define DK_NF_FULL      1  /* median of ratios over ALL rows (README semantics) */

typedef struct dk_ctx dk_ctx;
typedef struct dk_matrix dk_matrix;   /* device-resident filtered count matrix */
typedef struct dk_ttest dk_ttest;     /* device-resident per-row test results  */

/* ---- context ---------------------------------------------------------------*/
int  dk_init(dk_ctx **out, int device);           /* device = CUDA ordinal */
void dk_free(dk_ctx *ctx);
const char *dk_last_error(const dk_ctx *ctx);     /* never freed by the caller; ctx may be NULL */
const char *dk_version(void);
/* work is enqueued on this CUDA stream (a cudaStream_t passed as void*); default: a NON-BLOCKING stream the ctx owns.
 * A caller that passes _dev buffers written or read by its own kernels / collectives (the multi-GPU building blocks
 * below) MUST either put the context on the stream those run on with dk_set_stream, or order the two streams itself
 * (event record + cudaStreamWaitEvent around every such call): nothing else orders a non-blocking stream against them. */
int  dk_set_stream(dk_ctx *ctx, void *cuda_stream);
int  dk_sync(dk_ctx *ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
uint64_t dk_launch_count(const dk_ctx *ctx);
/* device time (ms, CUDA events on the ctx stream) of the last call of each stage:
 * which = 0 join, 1 nf, 2 ttest rows, 3 BH+filter, 4 merge kernel alone */
double dk_stage_ms(const dk_ctx *ctx, int which);

/* ---- joinCounts (replaces share/joinCounts/joinCounts.c:72-115 main loop) ----
 * keys[s]/cnts[s]: n[s] records of sample s, keys ascending. Row kept iff
 * #{s : present && cnt >= min_abund} >= min_rec (joinCounts.c:81-83,104).
 * emit_stale_tail != 0 reproduces the extra all-zero row printed when
 * min_rec <= 0 (joinCounts.c:104-114). Device pointers must be 16-byte aligned. */
int dk_join_counts(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts,
                   const uint64_t *n, int min_rec, int min_abund, int emit_stale_tail,
                   dk_matrix **out);
int dk_join_counts_dev(dk_ctx *ctx, int S, const uint64_t *const *keys_dev, const int32_t *const *cnts_dev,
                       const uint64_t *n, int min_rec, int min_abund, int emit_stale_tail,
                       dk_matrix **out);

/* ---- shard boundaries for the multi-GPU path ("equal-count splitters"; the reference is one process and
 * has no counterpart, its order of output rows — ascending k-mers, joinCounts.c:72-115 — is what sharding by k-mer
 * range preserves). Writes G-1 ascending k-mers: shard g takes the records with splitters[g-1] <= key < splitters[g]
 * (shard 0 from the smallest key, shard G-1 to the largest). The boundaries are chosen so that every shard holds about
 * 1/G of ALL records of all dumps, whatever the distribution of the k-mers (every dump is sampled at 256 * G regular
 * positions, at least 1024 and at most 16384: a shard is off by < 0.4 %). Host pointers unless device_inputs. */
int dk_key_splitters(dk_ctx *ctx, int S, const uint64_t *const *keys, const uint64_t *n, int device_inputs, int G,
                     uint64_t *splitters /* out [G-1] */);

/* ---- k-mer counting of one sample (replaces `jellyfish count -L L -m K [-C]` on the two mates with
 * bin/revCompFastq.pl:33-44 on one of them, `jellyfish dump -c` and `sort -k1`: Snakefile:523-582) ----
 * bases: the sequence lines of all reads back to back (host); read r is bases[offsets[r] .. offsets[r+1]).
 * rc_flag (may be NULL): 1 for reads the pipeline reverse-complements first (mate 1 of an "rf" library, mate 2 of
 * "fr"). canonical != 0 is -C (unstranded). Windows holding a byte other than ACGTacgt are skipped. The result is
 * the sorted (k-mer, count >= min_count) table on the device, in the layout dk_join_counts_dev takes.
 * One call handles < 2^32-2 windows (about 55 M reads of 100 bases at k = 31). */
typedef struct dk_dump dk_dump;
int dk_count_kmers(dk_ctx *ctx, const char *bases, const uint64_t *offsets, uint64_t n_reads, const uint8_t *rc_flag,
                   int k, int canonical, uint32_t min_count, dk_dump **out);
uint64_t dk_dump_size(const dk_dump *d);
const uint64_t *dk_dump_keys_dev(const dk_dump *d);       /* 16-byte aligned device pointers, valid until dk_dump_free */
const int32_t  *dk_dump_cnts_dev(const dk_dump *d);
void dk_dump_stats(const dk_dump *d, uint64_t *windows, uint64_t *valid, uint64_t *distinct);
int dk_dump_export(dk_ctx *ctx, const dk_dump *d, uint64_t *keys, int32_t *cnts);   /* host copies; either may be NULL */
void dk_dump_free(dk_ctx *ctx, dk_dump *d);
/* a sample too large for one call: count its reads in batches with min_count = 1, then add the tables up
 * (k-way merge = dk_join_counts, row sums = the sample's counts, min_count applied to the sums); <= DK_MAX_SAMPLES tables */
int dk_dump_merge(dk_ctx *ctx, dk_dump *const *parts, int n_parts, uint32_t min_count, dk_dump **out);

/* ---- matrix ----------------------------------------------------------------*/
/* wrap an existing row-major matrix (what computeNF / TtestFilter read from raw-counts.tsv) */
int dk_matrix_from_host(dk_ctx *ctx, int S, uint64_t R, const uint64_t *keys /* may be NULL */,
                        const int32_t *cnts, dk_matrix **out);
uint64_t dk_matrix_rows(const dk_matrix *m);      /* R: rows that passed the filter */
uint64_t dk_matrix_union(const dk_matrix *m);     /* U: union k-mers merged (0 for from_host) */
int      dk_matrix_samples(const dk_matrix *m);
/* gather rows [row0,row0+nrows) in reference order into caller buffers (either may be NULL) */
int dk_matrix_export(dk_ctx *ctx, const dk_matrix *m, uint64_t row0, uint64_t nrows,
                     uint64_t *keys, int32_t *cnts);
int dk_matrix_export_dev(dk_ctx *ctx, const dk_matrix *m, uint64_t row0, uint64_t nrows,
                         uint64_t *keys_dev, int32_t *cnts_dev);
/* text of rows [row0,row0+nrows) exactly as joinCounts prints them (joinCounts.c:105-109: "%s" k-mer of k bases,
 * "\t%d" per sample, "\n"), formatted on the device into the host buffer out (capacity cap). *len = bytes written;
 * when cap is too small nothing is copied, *len = bytes needed and DK_ERR_ARG is returned. */
int dk_matrix_format_rows(dk_ctx *ctx, const dk_matrix *m, uint64_t row0, uint64_t nrows, int k,
                          char *out, uint64_t cap, uint64_t *len);
/* order-independent checksums over all rows: xor of keys, sum of all counts (mod 2^64) */
int dk_matrix_checksum(dk_ctx *ctx, const dk_matrix *m, uint64_t *key_xor, uint64_t *cnt_sum);
void dk_matrix_free(dk_ctx *ctx, dk_matrix *m);

/* ---- reference-transcript mask (replaces bin/diffFilter.pl:50-69, rule Snakefile:634-644) ----
 * Sorted anti-join: *out = the rows of m whose k-mer does not occur in mask_keys (ascending,
 * duplicates allowed; host buffer unless mask_on_device). m is left untouched: computeNF reads
 * the raw matrix (Snakefile:366), TtestFilter the masked one (Snakefile:654). An unsorted mask
 * or matrix is refused (the Perl script silently mis-joins). n_removed may be NULL. */
int dk_matrix_mask(dk_ctx *ctx, const dk_matrix *m, const uint64_t *mask_keys, uint64_t n_mask,
                   int mask_on_device, dk_matrix **out, uint64_t *n_removed);
/* keys-only form for a caller that keeps the table rows itself (the diffFilter CLI prints the
 * surviving text lines byte for byte): keep[i] = 1 iff keys[i] is absent from mask_keys. Host buffers. */
int dk_keys_antijoin(dk_ctx *ctx, const uint64_t *keys, uint64_t n, const uint64_t *mask_keys,
                     uint64_t n_mask, uint8_t *keep);

/* ---- computeNF (replaces share/computeNF/computeNF.c:109-168) ---------------
 * mode DK_NF_REFERENCE: step_rows = rows per step (0 => the reference's 1,000,000).
 * info (may be NULL): info[0] = rows consumed, info[1] = steps evaluated,
 * info[2] = 1 if R fell in the reference's undefined zone step<=R<2*step (we
 * return the step-1 factors there), info[3] = all-positive rows used. */
int dk_compute_nf(dk_ctx *ctx, const dk_matrix *m, int mode, uint64_t step_rows,
                  double *nf_out, uint64_t *info);
/* multi-GPU DK_NF_FULL building blocks: exact global median of the per-sample ratios by
 * MSB-first radix select on the order-preserving uint64 image of the fp64 ratio, 8 rounds
 * of 8 bits. Per round the caller all-reduces (sum) S*256 uint64 bins over the ranks:
 *   dk_nf_begin -> all-reduce n_pos -> dk_nf_set_rank(total/2)
 *   8x { dk_nf_hist_dev(hist) -> all-reduce(hist) -> dk_nf_advance_dev(hist) }
 *   dk_nf_result -> dk_nf_end                                                     */
typedef struct dk_nf_state dk_nf_state;
int dk_nf_begin(dk_ctx *ctx, const dk_matrix *m, dk_nf_state **out, uint64_t *n_pos /* all-positive rows of this shard */);
int dk_nf_set_rank(dk_ctx *ctx, dk_nf_state *st, uint64_t k /* 0-based rank, reference: size/2 */);
int dk_nf_hist_dev(dk_ctx *ctx, dk_nf_state *st, uint64_t *hist_dev /* [S*256] device, overwritten */);
int dk_nf_advance_dev(dk_ctx *ctx, dk_nf_state *st, uint64_t *hist_dev /* reduced bins; zeroed on return */);
int dk_nf_result(dk_ctx *ctx, dk_nf_state *st, double *nf_out /* exp(median ratio), S values */);
void dk_nf_end(dk_ctx *ctx, dk_nf_state *st);

/* ---- TtestFilter (replaces share/TtestFilter/TtestFilter.c:373-497, t-test mode) --
 * idxA/idxB: column indices of the two groups in conditions-file order
 * (TtestFilter.c:251-284); nf[S]: normalisation factors as TtestFilter sees them
 * (i.e. after the %f round trip). Groups of size < 2 => DK_ERR_UNSUPPORTED (the
 * reference aborts inside Boost).
 * log2FC (TtestFilter.c:72-75,433) is read by the reference for the rows it prints only: for groups of <= 8 samples in
 * column order it is computed when first read (by dk_ttest_filter at the candidate rows, by dk_ttest_export_rows at every
 * row), from the matrix `m`, which must therefore outlive the test object, as dk_ttest_export_survivors already requires. */
int dk_ttest_rows(dk_ctx *ctx, const dk_matrix *m, const double *nf,
                  const int *idxA, int nA, const int *idxB, int nB, dk_ttest **out);
/* TtestFilter -l: Poisson likelihood ratio of the two group means of (count+1) against the pooled rate
 * (TtestFilter.c:152-185), group totals N1/N2 over all rows first (:340-362); counts are not normalised. The result
 * object feeds the same BH / filter / export calls (exported "normalised" counts are the raw counts, as the
 * reference prints them). info (may be NULL): [0]=N1, [1]=N2, [2]=rows whose statistic was negative (Boost throws
 * there and the reference aborts; p is set to 1). */
int dk_poisson_rows(dk_ctx *ctx, const dk_matrix *m, const int *idxA, int nA, const int *idxB, int nB, dk_ttest **out,
                    uint64_t *info);
/* BH over candidates p <= p_cut (exact for every row whose adjusted p <= p_cut; p_cut >= 1 => all rows).
 * Single GPU: dk_ttest_adjust(ctx,t,p_cut). Multi GPU: _candidates -> caller all-gathers
 * the candidate p-values and all-reduces R -> _adjust_global. */
int dk_ttest_candidates(dk_ctx *ctx, dk_ttest *t, double p_cut, uint64_t *n_cand, const double **cand_p_dev);
int dk_ttest_adjust_global(dk_ctx *ctx, dk_ttest *t, const double *all_cand_p_dev, uint64_t n_all, uint64_t R_global);
int dk_ttest_adjust(dk_ctx *ctx, dk_ttest *t, double p_cut);
/* Candidate narrowing (what dk_ttest_adjust does internally; the multi-GPU caller does it over global counts): with
 * n = #{p <= c} over ALL rows of all ranks, no row above c' = dk_bh_next_cut(p_cut, n, R_global) = p_cut * n / R_global
 * can survive or lower an adjusted value to <= p_cut, so the cut may fall from p_cut to c', then again from the new
 * count, ... (a few counting passes), and the candidate list shrinks from |{p <= p_cut}| to about the rows that matter.
 * dk_ttest_count_le counts this object's rows; dk_ttest_candidates_narrowed builds the list at the narrowed cut while
 * keeping p_cut as the bound up to which adjusted values are exact (what dk_ttest_filter checks). */
int dk_ttest_count_le(dk_ctx *ctx, const dk_ttest *t, double cut, uint64_t *n);
double dk_bh_next_cut(double p_cut, uint64_t n_le_cut_global, uint64_t R_global);
int dk_ttest_candidates_narrowed(dk_ctx *ctx, dk_ttest *t, double cut, double p_cut, uint64_t *n_cand, const double **cand_p_dev);
/* Scalable multi-GPU BH: after the all-gather every rank takes ONE p-value range of the candidates (equal p-values
 * always share a range), ranks it with dk_bh_rank_slice (n_below = candidates below the range), lowers F by the
 * minimum F of all higher ranges, all-gathers (sorted p, F) in range order and applies them to its own rows. */
int dk_bh_rank_slice(dk_ctx *ctx, const double *p_dev, uint64_t n, uint64_t n_below, uint64_t R_global,
                     double *sorted_p_dev /* out [n] ascending */, double *F_dev /* out [n] running minimum inside the slice */);
int dk_ttest_apply_adjusted(dk_ctx *ctx, dk_ttest *t, const double *sorted_p_dev, const double *F_dev, uint64_t n_all);
/* dk_bh_rank_slice for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO counterpart in the reference. Same arguments and same
 * results as dk_bh_rank_slice, for n <= DK_BH_NOALLOC_MAX p-values (a larger n is refused with DK_ERR_ARG). It calls no
 * cudaMalloc / cudaFree and no allocating routine: its scratch is one fixed device array inside the library (about
 * 4.3 MB, held by every GPU that loads the library until exit). The context's device must be the current device.
 * The work is only ENQUEUED on the context's stream (no synchronisation): dk_sync before reading the outputs.
 * All calls on one GPU share the scratch, so calls that overlap on that GPU (two contexts or two streams) race;
 * calls on one stream are ordered and need nothing cleared between them. */
#define DK_BH_NOALLOC_MAX (1 << 18)
int dk_bh_rank_slice_noalloc(dk_ctx *ctx, const double *p_dev, uint64_t n, uint64_t n_below, uint64_t R_global,
                             double *sorted_p_dev /* out [n] ascending */, double *F_dev /* out [n] */);
/* TtestFilter.c:486: survivors = adj_p <= pmax && fabs(log2FC) >= fmin; returns D in *n_keep.
 * pmax must not exceed the p_cut the adjusted values were computed for (dk_ttest_adjust / _candidates): they are exact
 * only up to it and absent beyond it, so a larger pmax is refused with DK_ERR_ARG.
 * ties (may be NULL): ties[0] = rows with adj_p within 1e-9 rel of pmax, ties[1] = rows with
 * fabs(log2FC) within 1e-9 rel of fmin, ties[2] = rows whose sd was within 1e-13 of the sd==0 test */
int dk_ttest_filter(dk_ctx *ctx, dk_ttest *t, double pmax, double fmin, uint64_t *n_keep, uint64_t *ties);
/* survivors in reference order; any output may be NULL. norm_cnts = S doubles per row (count/NF, TtestFilter.c:382) */
int dk_ttest_export_survivors(dk_ctx *ctx, const dk_ttest *t, uint64_t *row_idx, uint64_t *keys, double *p_adj,
                              double *meanA, double *meanB, double *log2fc, double *norm_cnts);
/* per-row vectors (R entries, reference order); p_adj is NaN where not computed (p_raw > p_cut) */
int dk_ttest_export_rows(dk_ctx *ctx, const dk_ttest *t, double *p_raw, double *log2fc, double *p_adj);
uint64_t dk_ttest_rows_count(const dk_ttest *t);
void dk_ttest_free(dk_ctx *ctx, dk_ttest *t);

/* ---- mergeTags (replaces share/mergeTags/mergeTags.c:68-212, stranded mode) -----------------
 * Contigs of the k-mers of ONE regulation direction (the caller splits by log2FC > 0, mergeTags.c:296-301): for
 * every overlap i = k-1 .. min_k, contigs whose i-base right end equals another contig's i-base left end are joined,
 * an end shared by two contigs on the same side being unusable (:56-66). seqs: n k-mers of k upper-case ACGT bases
 * back to back; rep_id[n]: the caller's id of each k-mer (its line number); a contig reports the id of its member
 * with the lowest p-value (the rightmost such member on ties, :147-153) and the number of k-mers merged into it.
 * Contigs linked in a ring are left unmerged at that level (the reference's result there depends on hash order);
 * dk_contigs_rings counts them. The unstranded procedure (-n) is not provided. */
typedef struct dk_contigs dk_contigs;
int dk_merge_tags(dk_ctx *ctx, const char *seqs, uint64_t n, int k, const double *pvalue, const uint32_t *rep_id,
                  int min_k, dk_contigs **out);
uint64_t dk_contigs_count(const dk_contigs *c);
uint64_t dk_contigs_chars(const dk_contigs *c);
uint64_t dk_contigs_rings(const dk_contigs *c);
/* rep[n], nb[n], off[n+1] (contig j = chars[off[j] .. off[j+1])), chars[dk_contigs_chars]; any may be NULL */
int dk_contigs_export(const dk_contigs *c, uint32_t *rep, uint32_t *nb, uint64_t *off, char *chars);
void dk_contigs_free(dk_contigs *c);

/* ---- whole path, host buffers in, survivors out (what bench.py's e2e times) ---- */
typedef struct {
    uint64_t U, R, R_pos, D;          /* union rows, filtered rows, all-positive rows used by NF, survivors */
    uint64_t ties[3];
    double   nf[DK_MAX_SAMPLES];
    double   ms_join, ms_nf, ms_ttest, ms_bh;
} dk_path_stats;
int dk_run_path(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                int device_inputs, int min_rec, int min_abund, int nf_mode,
                const int *idxA, int nA, const int *idxB, int nB, double pmax, double fmin,
                int nf_text_roundtrip /* round NFs through "%f" like the pipeline does */,
                dk_path_stats *stats, dk_matrix **out_matrix /* may be NULL */, dk_ttest **out_test /* may be NULL */);
/* dk_run_path that also RETURNS the filtered count matrix (the rows joinCounts prints, joinCounts.c:104-110; the
 * pipeline's raw-counts.tsv.gz, Snakefile:594-596) in caller-owned host buffers, reference order: matrix_keys_host
 * [cap_rows], matrix_cnts_host [cap_rows * S] row-major; pinned memory makes the copy asynchronous. With host dumps the
 * rows of every key-range slice leave on a third stream while later slices are still being copied in (PCIe is full
 * duplex). *matrix_rows = rows written; if they exceed cap_rows the call fails with DK_ERR_ARG and *matrix_rows = the need. */
int dk_run_path_matrix_out(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                           int device_inputs, int min_rec, int min_abund, int nf_mode,
                           const int *idxA, int nA, const int *idxB, int nB, double pmax, double fmin,
                           int nf_text_roundtrip, uint64_t *matrix_keys_host, int32_t *matrix_cnts_host,
                           uint64_t matrix_cap_rows, uint64_t *matrix_rows /* may be NULL */,
                           dk_path_stats *stats, dk_matrix **out_matrix /* may be NULL */, dk_ttest **out_test /* may be NULL */);
/* RNA-Seq mode (Snakefile:654): the same path with the mask between computeNF and TtestFilter.
 * NFs come from the raw matrix; the t-test, BH and the survivors from the masked one, which is
 * what *out_matrix receives. stats->R stays the raw row count; *rows_masked = rows the mask removed. */
int dk_run_path_masked(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                       int device_inputs, int min_rec, int min_abund, int nf_mode,
                       const int *idxA, int nA, const int *idxB, int nB, double pmax, double fmin,
                       int nf_text_roundtrip, const uint64_t *mask_keys, uint64_t n_mask, int mask_on_device,
                       dk_path_stats *stats, uint64_t *rows_masked /* may be NULL */,
                       dk_matrix **out_matrix /* may be NULL */, dk_ttest **out_test /* may be NULL */);

/* ---- several GPUs, one process per GPU ("NCCL over NVLink only for the global pieces") -----------
 * The reference is one process; these calls have no counterpart there. The context can own an NCCL communicator:
 * rank 0 calls dk_comm_unique_id and hands the DK_COMM_ID_BYTES to every rank by any means (a file, MPI, a socket,
 * torch.distributed); every rank then calls dk_comm_init (collective). NCCL is resolved with dlopen("libnccl.so.2") at
 * that moment; single-GPU callers never need it. Rank order must be k-mer order: rank g owns the slice of every sorted
 * dump between two boundaries of dk_key_splitters.
 * dk_run_path_sharded is dk_run_path on this rank's shard with the global pieces done by NCCL inside: R and U (one
 * all-reduce), computeNF's factors (reference mode: rank 0's prefix + one broadcast, DK_ERR_UNSUPPORTED on every rank if
 * its early stop does not fire inside rank 0's shard; full mode: one all-reduce of S*256 bins per radix round), BH
 * (candidate narrowing over global counts, then an all-gather of the few candidates left) and D. stats holds the GLOBAL
 * U, R, D, ties and the NFs on every rank; local_URD (may be NULL) this rank's U, R, D; *out_matrix / *out_test this
 * rank's shard (row indices are local). Without a communicator it is dk_run_path. */
#define DK_COMM_ID_BYTES 128
int dk_comm_unique_id(dk_ctx *ctx, unsigned char *id /* out [DK_COMM_ID_BYTES] */);
int dk_comm_init(dk_ctx *ctx, const unsigned char *id, int rank, int world);
void dk_comm_free(dk_ctx *ctx);
int dk_comm_rank(const dk_ctx *ctx);
int dk_comm_world(const dk_ctx *ctx);
int dk_run_path_sharded(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                        int device_inputs, int min_rec, int min_abund, int nf_mode,
                        const int *idxA, int nA, const int *idxB, int nB, double pmax, double fmin,
                        int nf_text_roundtrip, dk_path_stats *stats, uint64_t *local_URD /* [3], may be NULL */,
                        dk_matrix **out_matrix /* may be NULL */, dk_ttest **out_test /* may be NULL */);

/* ---- chunk-streamed path: dumps larger than device memory, kept in (pinned) HOST memory -----------------------
 * The reference streams too: joinCounts holds one record per input (joinCounts.c:43-69), TtestFilter reads the table
 * row by row (TtestFilter.c:340-444). Here the key space is cut into chunks of about chunk_records records (all dumps
 * together; equal-count k-mer boundaries); chunk c+1 is copied in on a second stream while chunk c is merged, filtered,
 * tested and reduced to its candidate rows (raw p <= p_cut), in k-mer order; the chunk's matrix is then dropped.
 * computeNF's factors (computeNF.c:138-168) come from the head of the table when its early stop fires there, else from
 * ALL all-positive rows of the table, which the reference also reads to the end in that case:
 * dk_stream_nf_reference: the head = the first chunk; *stop_fired says whether the stop fired inside it. With widen != 0
 *   the head grows to the first 2, 4, 8 ... chunks until it fires or the dumps are exhausted — exact when these dumps are
 *   the whole table; a rank of a sharded run passes widen = 0 and, when the stop did not fire, uses:
 * dk_stream_positive_rows: one sweep that keeps only the rows computeNF can use (every count > 0; with many samples a
 *   vanishing share of the table). *pos_rows is a dense matrix whose KEY column holds each row's number in this rank's
 *   filtered table; stats->R / ->U are this rank's totals, stats->n_cand the rows kept.
 * dk_nf_reference_sparse: computeNF's step / early-stop replay over such rows of a table of R_total rows (row numbers
 *   ascending, made global by the caller: + the filtered rows of the ranks before). Host arrays. info as dk_compute_nf.
 * dk_stream_path: the test sweep. nf = the factors as TtestFilter sees them (after the "%f" round trip). On return
 *   *cand_matrix / *cand_test hold the candidate rows only (both NULL when there is none) and take the ordinary calls
 *   dk_ttest_candidates -> dk_ttest_adjust_global(..., R_global = stats->R summed over all ranks) -> dk_ttest_filter ->
 *   dk_ttest_export_survivors; row indices they report count candidate rows. Free both with their *_free. */
typedef struct {
    uint64_t U, R, n_cand, chunks, h2d_bytes;      /* union rows, filtered rows, candidate rows, chunks walked, bytes copied in */
    double   h2d_seconds;                          /* time the copy stream spent copying (sum over chunks) */
    double   ms_join, ms_merge, ms_ttest, ms_total;
} dk_stream_stats;
int dk_stream_nf_reference(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                           int min_rec, int min_abund, uint64_t chunk_records, uint64_t step_rows /* 0 = 1,000,000 */,
                           int widen, double *nf_out, uint64_t *info /* may be NULL */, int *stop_fired /* may be NULL */);
int dk_stream_positive_rows(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                            int min_rec, int min_abund, uint64_t chunk_records, dk_stream_stats *stats /* may be NULL */,
                            dk_matrix **pos_rows);
int dk_nf_reference_sparse(dk_ctx *ctx, int S, uint64_t n_pos, const uint64_t *row_idx, const int32_t *cnts /* [n_pos][S] */,
                           uint64_t R_total, uint64_t step_rows, double *nf_out, uint64_t *info /* may be NULL */);
int dk_stream_path(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                   int min_rec, int min_abund, uint64_t chunk_records, const double *nf,
                   const int *idxA, int nA, const int *idxB, int nB, double p_cut,
                   dk_stream_stats *stats /* may be NULL */, dk_matrix **cand_matrix, dk_ttest **cand_test);

#ifdef __cplusplus
}
#endif
#endif
