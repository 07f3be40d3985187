// joinCounts — drop-in replacement of share/joinCounts/joinCounts.c (same argv, stdout, exit codes),
// merge + recurrence filter on the GPU through libdekupl_b200.so. Invoked by Snakefile:584-597 as
//   joinCounts -r MIN_REC -a MIN_REC_AB s1.txt.gz ... sN.txt.gz | gzip -c >> raw-counts.tsv.gz
// Deviations (documented in DESIGN.md): k-mers must be <= 31 ACGT bases of one length (the
// reference accepts any string); an empty input file is an empty sample (the reference
// segfaults, joinCounts.c:59-61). No CPU fallback: without a GPU the tool exits non-zero.
#include <unistd.h>
#include "../../include/dekupl_b200.h"
#include "dk_text.h"

int main(int argc, char *argv[])
{
    int min_recurrence = 1, min_recurrence_abundance = 1, c;
    while ((c = getopt(argc, argv, "r:a:")) >= 0) {
        switch (c) {
            case 'r': min_recurrence = atoi(optarg); break;
            case 'a': min_recurrence_abundance = atoi(optarg); break;
        }
    }
    if (optind == argc) {   // joinCounts.c:31-38
        fprintf(stderr, "\n");
        fprintf(stderr, "Usage:   joinCounts [options] <counts.tsv> [<counts.tsv> ...]\n\n");
        fprintf(stderr, "Options: -r INT    min recurrence [%d]\n", min_recurrence);
        fprintf(stderr, "Options: -a INT    min recurrence abundance [%d]\n", min_recurrence_abundance);
        fprintf(stderr, "\n");
        return 1;
    }
    int S = argc - optind;
    if (S > DK_MAX_SAMPLES) { fprintf(stderr, "joinCounts: at most %d input files are supported\n", DK_MAX_SAMPLES); return EXIT_FAILURE; }
    std::vector<std::vector<uint64_t>> keys(S);
    std::vector<std::vector<int32_t>> cnts(S);
    int k = 0;
    for (int i = 0; i < S; i++) {
        const char *path = argv[optind + i];
        DkFile f;
        if (!f.load(path)) { fprintf(stderr, "Failed to open %s\n", path); exit(EXIT_FAILURE); }   // joinCounts.c:52
        DkCursor cur(f);
        const char *s;
        size_t len;
        bool first = true;
        keys[i].reserve(f.size() / 34 + 16);
        cnts[i].reserve(f.size() / 34 + 16);
        // joinCounts.c:55-68 (first record: skip the rest of its line) and :86-92 (then strictly token pairs)
        while (cur.token(s, len)) {
            if (len == 0 && cur.eof()) break;
            if (len == 0) { if (!first) break; }        // ks_getuntil(...) > 0 fails on an empty token: input ends (:86)
            uint64_t key;
            if (!dk_encode(s, len, key)) {
                fprintf(stderr, "joinCounts: %s: record %zu: k-mer '%.*s' is not 1..31 ACGT bases (unsupported by the GPU path)\n", path,
                        keys[i].size() + 1, (int)len, s);
                return EXIT_FAILURE;
            }
            if (k == 0) k = (int)len;
            if ((int)len != k) { fprintf(stderr, "joinCounts: %s: mixed k-mer lengths (%d and %zu)\n", path, k, len); return EXIT_FAILURE; }
            const char *cs = "";
            size_t cl = 0;
            if (cur.dret != 0) cur.token(cs, cl);
            keys[i].push_back(key);
            cnts[i].push_back(dk_atoi(cs, cl));
            if (first) { cur.skip_line(); first = false; }
        }
    }
    dk_ctx *ctx = nullptr;
    if (dk_init(&ctx, 0) != DK_OK) { fprintf(stderr, "joinCounts: %s\n", dk_last_error(nullptr)); return EXIT_FAILURE; }
    std::vector<const uint64_t *> kp(S);
    std::vector<const int32_t *> cp(S);
    std::vector<uint64_t> n(S);
    for (int i = 0; i < S; i++) { kp[i] = keys[i].data(); cp[i] = cnts[i].data(); n[i] = keys[i].size(); }
    dk_matrix *m = nullptr;
    if (dk_join_counts(ctx, S, kp.data(), cp.data(), n.data(), min_recurrence, min_recurrence_abundance, 1, &m) != DK_OK) {
        fprintf(stderr, "joinCounts: %s\n", dk_last_error(ctx));
        return EXIT_FAILURE;
    }
    uint64_t R = dk_matrix_rows(m);
    DkOut out;
    out.fp = stdout;
    const uint64_t BLK = 1 << 20;
    std::vector<uint64_t> rk(BLK);
    std::vector<int32_t> rc(BLK * (size_t)S);
    char kbuf[32];
    for (uint64_t r0 = 0; r0 < R; r0 += BLK) {
        uint64_t nr = R - r0 < BLK ? R - r0 : BLK;
        if (dk_matrix_export(ctx, m, r0, nr, rk.data(), rc.data()) != DK_OK) { fprintf(stderr, "joinCounts: %s\n", dk_last_error(ctx)); return EXIT_FAILURE; }
        for (uint64_t r = 0; r < nr; r++) {               // joinCounts.c:105-109
            dk_decode(rk[r], k, kbuf);
            out.put(kbuf, (size_t)k);
            for (int s = 0; s < S; s++) { out.putc_('\t'); out.put_int(rc[r * S + s]); }
            out.putc_('\n');
            if (out.b.size() > (1 << 22) - 8192) out.flush();
        }
    }
    out.flush();
    dk_matrix_free(ctx, m);
    dk_free(ctx);
    return 0;
}
