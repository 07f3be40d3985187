/* CPU ORACLE — test infrastructure only (see dekupl_oracle.h). */
#include "dekupl_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- joinCounts (reference: share/joinCounts/joinCounts.c:72-115) ----------
 * The reference keeps one cursor per file and, per output row, scans all
 * cursors twice: once to collect counts of cursors equal to the current minimum
 * (advancing them) and once to find the next minimum. Packed keys compare like
 * strcmp on equal-length ACGT strings (A<C<G<T == 0<1<2<3, MSB first). */
size_t orc_join(int S, const uint64_t *const *keys, const int32_t *const *cnts,
                const size_t *n, int min_rec, int min_abund, int emit_stale_tail,
                uint64_t *out_keys, int32_t *out_cnts, size_t cap, size_t *n_union)
{
    size_t *pos = (size_t *)calloc((size_t)S, sizeof(size_t));
    int32_t *row = (int32_t *)calloc((size_t)S, sizeof(int32_t));
    size_t R = 0, U = 0;
    uint64_t last = 0;
    for (;;) {
        int any = 0;
        uint64_t mn = 0;
        for (int s = 0; s < S; s++)
            if (pos[s] < n[s] && (!any || keys[s][pos[s]] < mn)) { mn = keys[s][pos[s]]; any = 1; }
        if (!any) break;
        int rec = 0;
        for (int s = 0; s < S; s++) {
            if (pos[s] < n[s] && keys[s][pos[s]] == mn) {      /* joinCounts.c:78-83 */
                row[s] = cnts[s][pos[s]];
                if (row[s] >= min_abund) rec++;
                pos[s]++;
            } else row[s] = 0;                                   /* joinCounts.c:96 */
        }
        U++; last = mn;
        if (rec >= min_rec) {                                    /* joinCounts.c:104 */
            if (R < cap) { out_keys[R] = mn; memcpy(out_cnts + R * S, row, sizeof(int32_t) * S); }
            R++;
        }
    }
    /* joinCounts.c:104-114: one more iteration runs on the stale min_kmer with
     * all counts 0 and rec 0; it prints iff 0 >= min_rec. */
    if (emit_stale_tail && U > 0 && 0 >= min_rec) {
        if (R < cap) { out_keys[R] = last; memset(out_cnts + R * S, 0, sizeof(int32_t) * S); }
        R++;
    }
    if (n_union) *n_union = U;
    free(pos); free(row);
    return R;
}

/* ---- computeNF (reference: share/computeNF/computeNF.c) -------------------- */
typedef struct { double *a; size_t n, m; } dvec;
static void dpush(dvec *v, double x) {
    if (v->n == v->m) { v->m = v->m ? v->m * 2 : 1024; v->a = (double *)realloc(v->a, v->m * sizeof(double)); }
    v->a[v->n++] = x;
}
static int cmp_double(const void *a, const void *b) {            /* computeNF.c:27-36 */
    double x = *(const double *)a, y = *(const double *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}
static void compute_nf(dvec *v, int S, double *nf) {             /* computeNF.c:38-47 */
    for (int j = 0; j < S; j++) {
        qsort(v[j].a, v[j].n, sizeof(double), cmp_double);
        nf[j] = v[j].n ? exp(v[j].a[v[j].n / 2]) : NAN;          /* reference: UB on empty */
    }
}
/* computeNF.c:18-25 + 122-130: returns 1 and pushes when the row is all-positive */
static void push_row(const int32_t *c, int S, dvec *v) {
    double sum = 0;
    for (int j = 0; j < S; j++) sum += log((double)c[j]);        /* log(int) promotes to double */
    double m = sum / S;
    for (int j = 0; j < S; j++)
        if (c[j] > 0 && isfinite(m)) dpush(&v[j], log((double)c[j]) - m);
}

size_t orc_nf_reference(const int32_t *cnts, size_t R, int S, size_t step,
                        double *nf_out, int *undefined_zone)
{
    dvec *v = (dvec *)calloc((size_t)S, sizeof(dvec));
    double *nf = (double *)calloc((size_t)S, sizeof(double));
    double *nw = (double *)calloc((size_t)S, sizeof(double));
    double min_error = 0.001 * S;                                 /* computeNF.c:53,97 */
    size_t nb_step = 0, line = 0;
    int have_nf = 0;
    if (undefined_zone) *undefined_zone = 0;
    for (size_t i = 0; i < R; i++) {
        line++;
        push_row(cnts + i * S, S, v);
        if (line % step == 0) {                                   /* computeNF.c:138 */
            compute_nf(v, S, nw);
            if (nb_step > 0) {
                double sum = 0;
                /* at nb_step==1 the reference compares against uninitialised
                 * memory (computeNF.c:95,145); we model "never converges there". */
                for (int j = 0; j < S; j++) sum += fabs(nf[j] - nw[j]);
                int stop = (nb_step >= 2) && (sum < min_error);
                memcpy(nf, nw, sizeof(double) * S); have_nf = 1;
                if (stop) break;
            } else {
                /* step 0 result is discarded by the reference (leaked); we keep it
                 * only to give the 1*step <= R < 2*step zone a defined answer. */
                memcpy(nf, nw, sizeof(double) * S);
            }
            nb_step++;
        }
    }
    if (nb_step == 0) compute_nf(v, S, nf);                       /* computeNF.c:166-168 */
    else if (!have_nf && undefined_zone) *undefined_zone = 1;
    memcpy(nf_out, nf, sizeof(double) * S);
    for (int j = 0; j < S; j++) free(v[j].a);
    free(v); free(nf); free(nw);
    return line;
}

void orc_nf_full(const int32_t *cnts, size_t R, int S, double *nf_out)
{
    dvec *v = (dvec *)calloc((size_t)S, sizeof(dvec));
    for (size_t i = 0; i < R; i++) push_row(cnts + i * S, S, v);
    compute_nf(v, S, nf_out);
    for (int j = 0; j < S; j++) free(v[j].a);
    free(v);
}

/* bin/diffFilter.pl:50-69 restated on packed keys: one cursor over the (sorted) mask, advanced
 * while it is below the current table k-mer (:59-66); the row is printed unless the cursor then
 * sits on an equal k-mer (:67-69). For equal-length ACGT strings Perl's `lt`/`ne` are the integer
 * compares used here. The cursor never moves back, so an unsorted input mis-joins exactly as the
 * script does. */
size_t orc_antijoin(const uint64_t *keys, size_t n, const uint64_t *mask, size_t n_mask, uint8_t *keep)
{
    size_t d = 0, kept = 0;
    for (size_t i = 0; i < n; i++) {
        while (d < n_mask && mask[d] < keys[i]) d++;
        keep[i] = !(d < n_mask && mask[d] == keys[i]);
        kept += keep[i];
    }
    return kept;
}

/* k-mer counting of one sample as the pipeline does it (Snakefile:523-582): `jellyfish count -L min_count -m k [-C]`
 * over the reads (those flagged rc first pass through bin/revCompFastq.pl:33-44: reversed and complemented),
 * `jellyfish dump -c`, `sort -k1`. Restated on packed k-mers: every window of k bases without a byte outside
 * ACGTacgt is one occurrence; -C counts min(k-mer, reverse complement). jellyfish is not in this image, so this
 * restatement follows its documented behaviour and is not pinned by the binary.
 * Returns the number of (k-mer, count) pairs written (ascending k-mers); -1 when cap is too small. */
static int cmp_u64(const void *a, const void *b)
{
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}
static uint64_t revcomp_k(uint64_t v, int k)
{
    uint64_t r = 0;
    for (int i = 0; i < k; i++) { r = (r << 2) | (3 - (v & 3)); v >>= 2; }
    return r;
}
long long orc_count_kmers(const char *bases, const uint64_t *offsets, size_t n_reads, const uint8_t *rc_flag, int k, int canonical,
                          uint32_t min_count, uint64_t *out_keys, int32_t *out_cnts, size_t cap)
{
    size_t W = 0, n = 0;
    for (size_t r = 0; r < n_reads; r++) { size_t len = offsets[r + 1] - offsets[r]; if (len >= (size_t)k) W += len - k + 1; }
    uint64_t *v = (uint64_t *)malloc(sizeof(uint64_t) * (W ? W : 1));
    for (size_t r = 0; r < n_reads; r++) {
        const char *s = bases + offsets[r];
        size_t len = offsets[r + 1] - offsets[r];
        for (size_t i = 0; i + k <= len; i++) {
            uint64_t key = 0;
            int ok = 1;
            for (int j = 0; j < k && ok; j++) {
                int c;
                switch (s[i + j]) {
                    case 'A': case 'a': c = 0; break;
                    case 'C': case 'c': c = 1; break;
                    case 'G': case 'g': c = 2; break;
                    case 'T': case 't': c = 3; break;
                    default: c = -1;
                }
                if (c < 0) ok = 0; else key = (key << 2) | (uint64_t)c;
            }
            if (!ok) continue;
            uint64_t rc = revcomp_k(key, k);
            if (canonical) key = key < rc ? key : rc;
            else if (rc_flag && rc_flag[r]) key = rc;            /* the k-mers of revcomp(read) */
            v[n++] = key;
        }
    }
    qsort(v, n, sizeof(uint64_t), cmp_u64);
    size_t o = 0;
    for (size_t i = 0; i < n;) {
        size_t j = i;
        while (j < n && v[j] == v[i]) j++;
        if (j - i >= min_count) {
            if (o >= cap) { free(v); return -1; }
            out_keys[o] = v[i]; out_cnts[o] = (int32_t)(j - i); o++;
        }
        i = j;
    }
    free(v);
    return (long long)o;
}

/* ---- text I/O for the CPU baseline of bench.py (test infrastructure) ---------------------------------------
 * orc_write_dump_text: one sample's dump as `jellyfish dump -c | sort` leaves it ("KMER COUNT\n", the format
 * joinCounts.c:55-68 reads), from the packed arrays. orc_parse_matrix_text: the rows joinCounts prints
 * ("KMER(\tCOUNT)xS\n", joinCounts.c:105-109) back into an int matrix, tokens as kseq's ks_getuntil / atoi see them. */
#include <stdio.h>
long orc_write_dump_text(const char *path, const uint64_t *keys, const int32_t *cnts, size_t n, int k)
{
    FILE *f = fopen(path, "w");
    if (!f) return -1;
    static const char B[4] = {'A', 'C', 'G', 'T'};
    char line[96];
    long bytes = 0;
    for (size_t i = 0; i < n; i++) {
        for (int j = 0; j < k; j++) line[j] = B[(keys[i] >> (2 * (k - 1 - j))) & 3];
        int len = k + snprintf(line + k, sizeof line - k, "\t%d\n", cnts[i]);
        fwrite(line, 1, (size_t)len, f);
        bytes += len;
    }
    fclose(f);
    return bytes;
}
/* returns the number of rows parsed (<= cap); a first line that starts with a non-ACGT token is skipped as a header */
long orc_parse_matrix_text(const char *path, int S, int32_t *cnts, size_t cap)
{
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    size_t bufn = 1 << 20;
    char *line = (char *)malloc(bufn);
    long R = 0;
    while ((size_t)R < cap && fgets(line, (int)bufn, f)) {
        char *p = line;
        while (*p && *p != '\t' && *p != ' ' && *p != '\n') p++;          /* the k-mer token */
        if (R == 0 && !(line[0] == 'A' || line[0] == 'C' || line[0] == 'G' || line[0] == 'T')) continue;
        int s = 0;
        while (s < S && *p && *p != '\n') {
            while (*p == '\t' || *p == ' ') p++;
            cnts[(size_t)R * S + s++] = atoi(p);                          /* joinCounts.c:68, TtestFilter.c:380 */
            while (*p && *p != '\t' && *p != ' ' && *p != '\n') p++;
        }
        for (; s < S; s++) cnts[(size_t)R * S + s] = 0;
        R++;
    }
    free(line);
    fclose(f);
    return R;
}
