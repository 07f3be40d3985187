// joinCounts — drop-in replacement of share/joinCounts/joinCounts.c (same argv, stdout, exit codes),
// merge + recurrence filter on the GPU through libdekupl_b200.so. Invoked by Snakefile:584-597 as
//   joinCounts -r MIN_REC -a MIN_REC_AB s1.txt.gz ... sN.txt.gz | gzip -c >> raw-counts.tsv.gz
// Deviations (documented in DESIGN.md): k-mers must be <= 31 ACGT bases of one length (the
// reference accepts any string); an empty input file is an empty sample (the reference
// segfaults, joinCounts.c:59-61). No CPU fallback: without a GPU the tool exits non-zero.
#include <unistd.h>
#include <algorithm>
#include <atomic>
#include <thread>
#include "../../include/dekupl_b200.h"
#include <sys/stat.h>
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
    dk_async_init_start();                      // CUDA context comes up while the dumps are inflated and parsed
    if (S > DK_MAX_SAMPLES) { fprintf(stderr, "joinCounts: at most %d input files are supported\n", DK_MAX_SAMPLES); return EXIT_FAILURE; }
    std::vector<std::vector<uint64_t>> keys(S);
    std::vector<std::vector<int32_t>> cnts(S);
    // every dump is inflated and parsed on its own thread (the text side is the CLI's bottleneck);
    // diagnostics are reported afterwards for the first failing file in argv order, as a sequential reader would
    std::vector<int> kk(S, 0), status(S, 0);              // status: 0 ok, 1 cannot open, 2 parse error (message in err)
    std::vector<std::string> err(S);
    auto parse_one = [&](int i) {
        const char *path = argv[optind + i];
        DkStreamCursor cur;                                // inflated and tokenised block by block, never held as a whole
        if (!cur.open(path)) { status[i] = 1; return; }   // joinCounts.c:52
        const char *s;
        size_t len;
        bool first = true;
        int k = 0;
        {   // one allocation instead of a doubling series: ~34 bytes of text per record, gzip shrinks a dump about 4x
            struct stat sb;
            if (stat(path, &sb) == 0 && S_ISREG(sb.st_mode)) {
                const size_t n = strlen(path);
                const bool gz = n > 3 && strcmp(path + n - 3, ".gz") == 0;
                const size_t guess = (size_t)sb.st_size * (gz ? 4 : 1) / 34 + 16;
                keys[i].reserve(guess);
                cnts[i].reserve(guess);
            }
        }
        // joinCounts.c:55-68 (first record: skip the rest of its line) and :86-92 (then strictly token pairs)
        while (cur.token(s, len)) {
            if (len == 0 && cur.eof()) break;
            if (len == 0) { if (!first) break; }          // ks_getuntil(...) > 0 fails on an empty token: input ends (:86)
            uint64_t key;
            char msg[256];
            if (!dk_encode(s, len, key)) {
                snprintf(msg, sizeof msg, "joinCounts: %s: record %zu: k-mer '%.*s' is not 1..31 ACGT bases (unsupported by the GPU path)",
                         path, keys[i].size() + 1, (int)(len > 64 ? 64 : len), s);
                status[i] = 2; err[i] = msg;
                return;
            }
            if (k == 0) k = (int)len;
            if ((int)len != k) {
                snprintf(msg, sizeof msg, "joinCounts: %s: mixed k-mer lengths (%d and %zu)", path, k, len);
                status[i] = 2; err[i] = msg;
                return;
            }
            const char *cs = "";
            size_t cl = 0;
            if (cur.dret != 0) cur.token(cs, cl);
            keys[i].push_back(key);
            cnts[i].push_back(dk_atoi(cs, cl));
            if (first) { cur.skip_line(); first = false; }
        }
        kk[i] = k;
    };
    {
        unsigned hw = std::thread::hardware_concurrency();
        const int nthr = std::max(1, std::min(S, (int)(hw ? hw : 4)));
        std::atomic<int> next(0);
        std::vector<std::thread> pool;
        for (int t = 0; t < nthr; t++)
            pool.emplace_back([&] { for (int i; (i = next.fetch_add(1)) < S;) parse_one(i); });
        for (auto &t : pool) t.join();
    }
    int k = 0;
    for (int i = 0; i < S; i++) {
        if (status[i] == 1) { fprintf(stderr, "Failed to open %s\n", argv[optind + i]); exit(EXIT_FAILURE); }
        if (status[i] == 2) { fprintf(stderr, "%s\n", err[i].c_str()); return EXIT_FAILURE; }
        if (kk[i] && k == 0) k = kk[i];
        if (kk[i] && kk[i] != k) { fprintf(stderr, "joinCounts: %s: mixed k-mer lengths (%d and %d)\n", argv[optind + i], k, kk[i]); return EXIT_FAILURE; }
    }
    dk_ctx *ctx = nullptr;
    if (!(ctx = dk_async_init_obj.wait())) { fprintf(stderr, "joinCounts: %s\n", dk_async_init_obj.err.c_str()); return EXIT_FAILURE; }
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
    // rows are formatted on the device (dk_matrix_format_rows), one block at a time; the host only writes bytes
    const uint64_t BLK = 1 << 20;
    std::vector<char> text((size_t)BLK * ((size_t)k + 2 + 12 * (size_t)S) + 64);
    for (uint64_t r0 = 0; r0 < R; r0 += BLK) {
        uint64_t nr = R - r0 < BLK ? R - r0 : BLK, len = 0;
        if (dk_matrix_format_rows(ctx, m, r0, nr, k, text.data(), text.size(), &len) != DK_OK) {   // joinCounts.c:105-109
            fprintf(stderr, "joinCounts: %s\n", dk_last_error(ctx));
            return EXIT_FAILURE;
        }
        out.flush();
        fwrite(text.data(), 1, (size_t)len, stdout);
    }
    out.flush();
    dk_matrix_free(ctx, m);
    dk_free(ctx);
    return 0;
}
