/*
 * dekupl_oracle.h — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C restatement of the reference's joinCounts -> computeNF -> TtestFilter
 * algorithms on packed (uint64 k-mer, int32 count) arrays. Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may link or call this. The product (libdekupl_b200.so) never does.
 *
 * Pinning status:
 *   - orc_join / orc_nf: pinned against the REAL reference binaries built from
 *     /root/reference/share/{joinCounts,computeNF} into oracle/_ref/ (see
 *     oracle/Makefile, tests/test_oracle_vs_ref.py) and against goldens captured
 *     from them (tests/golden/).
 *   - orc_ttest: PARITY UNPINNED by the reference: TtestFilter needs Boost.Math
 *     (absent here, unvendored, version unpinned) and the reference ships no
 *     expected outputs. Pins we add: scipy cross-check of the Student-t tail
 *     (tests/golden/ttest_scipy_*.json, made by tests/golden/make_golden.py).
 *
 * Reference files followed (relative to /root/reference):
 *   share/joinCounts/joinCounts.c:72-115   (merge + recurrence filter)
 *   share/computeNF/computeNF.c:18-47,109-168 (logmean, ratios, median, steps)
 *   share/TtestFilter/TtestFilter.c:72-150,373-476,483-497 (moments, Welch, BH, filter)
 */
#ifndef DEKUPL_ORACLE_H
#define DEKUPL_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* joinCounts.c:72-115. keys[s] sorted ascending (duplicates allowed, emitted as
 * separate rows like the reference). Counts are C ints (atoi). Writes up to cap
 * rows: out_keys[row], out_cnts[row*S+s]. emit_stale_tail reproduces the extra
 * all-zero row the reference prints when min_rec <= 0 (joinCounts.c:104-114).
 * Returns the number of rows the reference would print (may exceed cap; rows
 * beyond cap are counted but not stored). *n_union receives the union size. */
size_t orc_join(int S, const uint64_t *const *keys, const int32_t *const *cnts,
                const size_t *n, int min_rec, int min_abund, int emit_stale_tail,
                uint64_t *out_keys, int32_t *out_cnts, size_t cap, size_t *n_union);

/* computeNF.c:18-47,109-168 on a row-major int matrix cnts[R*S].
 * step = rows per step (reference: 1,000,000; parameterised so tests can reach
 * the step logic with small inputs). Returns the number of rows consumed.
 * For step <= R < 2*step the reference prints uninitialised memory; the oracle
 * (and the product) return the step-1 factors there and set *undefined_zone=1. */
size_t orc_nf_reference(const int32_t *cnts, size_t R, int S, size_t step,
                        double *nf_out, int *undefined_zone);

/* median-of-ratios over ALL rows (README semantics; the multi-GPU "full" mode). */
void orc_nf_full(const int32_t *cnts, size_t R, int S, double *nf_out);

/* Student-t two-sided tail 2*Q(|t|; df) — replaces the Boost call at
 * TtestFilter.c:146-147 (local regularised incomplete beta, fp64). */
double orc_student_t_2tail(double t, double df);
/* the same tail through Boost.Math's branch structure (v > 2t^2: ibetac(1/2, v/2, t^2/(v+t^2)); else ibeta(v/2, 1/2, v/(v+t^2))) */
double orc_student_t_2tail_boost_branches(double t, double df);

/* TtestFilter.c:373-440 per row (t-test mode): outputs have R entries each. */
void orc_ttest_rows(const int32_t *cnts, size_t R, int S, const double *nf,
                    const int *idxA, int nA, const int *idxB, int nB,
                    double *p_raw, double *meanA, double *meanB, double *log2fc);

/* Poisson likelihood-ratio mode (-l): TtestFilter.c:152-185 per row, totals of :340-362 in N_out[2]. */
double orc_poisson_lr(double K1, double K2, int64_t N1, int64_t N2);
void orc_poisson_rows(const int32_t *cnts, size_t R, int S, const int *idxA, int nA, const int *idxB, int nB,
                      double *p_raw, double *meanA, double *meanB, double *log2fc, int64_t *N_out);

/* bin/diffFilter.pl:50-69 (reference-transcript mask): keep[i] = 1 iff keys[i] is not in mask.
 * Both ascending; returns the number of rows kept. */
size_t orc_antijoin(const uint64_t *keys, size_t n, const uint64_t *mask, size_t n_mask, uint8_t *keep);

/* Snakefile:523-582 (jellyfish count -L -m [-C] | dump -c | sort -k1, revCompFastq.pl on flagged reads) on packed k-mers.
 * Returns the number of pairs written, -1 when cap is too small. Not pinned by the binary (jellyfish absent). */
long long orc_count_kmers(const char *bases, const uint64_t *offsets, size_t n_reads, const uint8_t *rc_flag, int k, int canonical,
                          uint32_t min_count, uint64_t *out_keys, int32_t *out_cnts, size_t cap);

/* TtestFilter.c:459-476: Benjamini-Hochberg, in place semantic on a copy. */
void orc_bh(const double *p_raw, size_t R, double *p_adj);

/* TtestFilter.c:486: keep[i] = adj<=pmax && fabs(l2fc)>=fmin. Returns D. */
size_t orc_filter(const double *p_adj, const double *log2fc, size_t R,
                  double pmax, double fmin, uint8_t *keep);

/* text I/O used by bench.py's CPU baseline only: dumps as joinCounts reads them, rows as it prints them */
long orc_write_dump_text(const char *path, const uint64_t *keys, const int32_t *cnts, size_t n, int k);
long orc_parse_matrix_text(const char *path, int S, int32_t *cnts, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
