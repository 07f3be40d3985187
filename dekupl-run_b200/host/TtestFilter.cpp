// TtestFilter — drop-in replacement of share/TtestFilter/TtestFilter.c in t-test mode (same argv,
// stdout, stderr banners), per-k-mer Welch t-test + BH + filter on the GPU through libdekupl_b200.so.
// Invoked by bin/Ttest_diff_method.R:60 as
//   TtestFilter -p P -f F -r raw_pvals.txt.gz counts.tsv.gz sample_conditions_full.tsv A B | gzip -c > diff-counts.tsv.gz
// Not reproduced on purpose: the fixed temp file ./.TtestFilter.tmp (TtestFilter.c:22,331,497 — two
// runs in one directory corrupt each other) and the unused extra pass over the matrix (:340-362).
// -l runs the Poisson likelihood ratio of :152-185 (dk_poisson_rows); where the reference aborts (negative statistic:
// Boost throws) the row gets p = 1 and a warning is printed.
#include <math.h>
#include <unistd.h>
#include "../../include/dekupl_b200.h"
#include "dk_matrix_text.h"

int main(int argc, char *argv[])
{
    double pvalue_threshold = 0.05, log2fc_threshold = 0;
    char *raw_pval_file = NULL;
    bool use_t_test = true;
    int c;
    while ((c = getopt(argc, argv, "p:f:r:l")) >= 0) {
        switch (c) {
            case 'p': pvalue_threshold = atof(optarg); break;
            case 'f': log2fc_threshold = atof(optarg); break;
            case 'r': raw_pval_file = optarg; break;
            case 'l': use_t_test = false; break;
        }
    }
    if ((optind + 3) >= argc) {   // TtestFilter.c:207-216
        fprintf(stderr, "\n");
        fprintf(stderr, "Usage:   ttestFilter [options] <counts.tsv> <conditions.tsv> <conditionA> <conditionB>\n\n");
        fprintf(stderr, "Options: -p FLOAT  max pvalue [%.2f]\n", pvalue_threshold);
        fprintf(stderr, "         -f FLOAT  min log2 fold change (absolute) [%.2f]\n", log2fc_threshold);
        fprintf(stderr, "         -r STR    file were raw pvalues will be printed\n");
        fprintf(stderr, "         -l        disregard the program name, do a Poisson likelihood ratio (HAWK model)\n");
        fprintf(stderr, "\n");
        return 1;
    }
    const char *counts_file = argv[optind++], *conditions_file = argv[optind++];
    const char *conditionA = argv[optind++], *conditionB = argv[optind++];

    dk_async_init_start();                      // CUDA context comes up while the table is inflated and parsed
    DkMatrixText mt;                            // read in blocks: only the packed counts and the tags stay in memory
    bool opened = false;
    size_t bad = mt.load(counts_file, true, opened);
    if (!opened) { fprintf(stderr, "Failed to open %s\n", counts_file); exit(EXIT_FAILURE); }
    const int S = mt.S;

    // conditions: skip header; per line sample, condition, NF (TtestFilter.c:251-284)
    DkFile cf;
    if (!cf.load(conditions_file)) { fprintf(stderr, "Failed to open %s\n", conditions_file); exit(EXIT_FAILURE); }
    std::vector<double> nf(S > 0 ? S : 1, 0.0);
    std::vector<int> idxA, idxB;
    {
        DkCursor cur(cf);
        const char *s;
        size_t len;
        cur.line(s, len);
        while (cur.token(s, len)) {
            if (len == 0 && cur.eof()) break;
            std::string sample(s, len);
            if (cur.dret != '\n' && cur.dret != 0) {
                const char *cs; size_t cl;
                if (cur.token(cs, cl) && cl > 0) {
                    std::string cond(cs, cl);
                    if (cur.dret != '\n' && cur.dret != 0) {
                        const char *ns; size_t nl;
                        if (cur.token(ns, nl) && nl > 0) {
                            double norm_fact = atof(std::string(ns, nl).c_str());
                            int j = 0;
                            while (j < S && mt.names[j] != sample) j++;
                            if (j < S) {
                                if (cond == conditionA) idxA.push_back(j);
                                else if (cond == conditionB) idxB.push_back(j);
                                nf[j] = norm_fact;
                            }
                        }
                    }
                }
            }
            cur.skip_line();
        }
    }
    // friendly print (TtestFilter.c:301-318)
    fprintf(stderr, "Condition A: %s\n", conditionA);
    fprintf(stderr, "Condition B: %s\n", conditionB);
    fprintf(stderr, "Group A: ");
    for (size_t j = 0; j < idxA.size(); j++) fprintf(stderr, "%s%s", j ? ", " : "", mt.names[idxA[j]].c_str());
    fprintf(stderr, "\nGroup B: ");
    for (size_t j = 0; j < idxB.size(); j++) fprintf(stderr, "%s%s", j ? ", " : "", mt.names[idxB[j]].c_str());
    fprintf(stderr, "\n");
    fprintf(stderr, "Computing t-test for all k-mers\n");
    fprintf(stdout, "tag\tpvalue\tmeanA\tmeanB\tlog2FC");       // :320-323
    for (int j = 0; j < S; j++) fprintf(stdout, "\t%s", mt.names[j].c_str());
    fprintf(stdout, "\n");
    fflush(stdout);
    if (bad) { fprintf(stderr, "inconsistent number of column (line %zu)\n", bad); exit(EXIT_FAILURE); }   // :386-389
    if (S < 1 || S > DK_MAX_SAMPLES) { fprintf(stderr, "TtestFilter: %d samples unsupported (1..%d)\n", S, DK_MAX_SAMPLES); return EXIT_FAILURE; }

    dk_ctx *ctx = nullptr;
    if (!(ctx = dk_async_init_obj.wait())) { fprintf(stderr, "TtestFilter: %s\n", dk_async_init_obj.err.c_str()); return EXIT_FAILURE; }
    dk_matrix *m = nullptr;
    dk_ttest *t = nullptr;
#define CHECK(call) if ((call) != DK_OK) { fprintf(stderr, "TtestFilter: %s\n", dk_last_error(ctx)); return EXIT_FAILURE; }
    CHECK(dk_matrix_from_host(ctx, S, mt.R, nullptr, mt.cnts.data(), &m));
    if (use_t_test) {
        CHECK(dk_ttest_rows(ctx, m, nf.data(), idxA.data(), (int)idxA.size(), idxB.data(), (int)idxB.size(), &t));
    } else {
        uint64_t pinfo[3] = {0, 0, 0};
        CHECK(dk_poisson_rows(ctx, m, idxA.data(), (int)idxA.size(), idxB.data(), (int)idxB.size(), &t, pinfo));
        if (pinfo[2])
            fprintf(stderr, "TtestFilter: %llu k-mers had a negative likelihood-ratio statistic (the reference aborts inside Boost); p set to 1\n",
                    (unsigned long long)pinfo[2]);
    }
    if (raw_pval_file) {   // :335-338,426-430: gz text "kmer\t%f\n" with the RAW p of every row
        gzFile pf = gzopen(raw_pval_file, "w");
        if (!pf) { fprintf(stderr, "Failed to open %s\n", raw_pval_file); exit(EXIT_FAILURE); }
        std::vector<double> p(mt.R + 1);
        CHECK(dk_ttest_export_rows(ctx, t, p.data(), nullptr, nullptr));
        DkOut o;
        o.gz = pf;
        for (uint64_t i = 0; i < mt.R; i++) { o.put(mt.tag(i), mt.tag_len(i)); o.printf_("\t%f\n", p[i]); }
        o.flush();
        gzclose(pf);
    }
    fprintf(stderr, "Adjusting pvalues with Benjamini–Hochberg\n");
    CHECK(dk_ttest_adjust(ctx, t, pvalue_threshold));
    uint64_t D = 0, ties[3] = {0, 0, 0};
    CHECK(dk_ttest_filter(ctx, t, pvalue_threshold, log2fc_threshold, &D, ties));
    fprintf(stderr, "Print the final output\n");
    if (ties[0] || ties[1] || ties[2])
        fprintf(stderr, "TtestFilter: ties within 1e-9 of a threshold: adj_p %llu, log2FC %llu, sd==0 test %llu\n",
                (unsigned long long)ties[0], (unsigned long long)ties[1], (unsigned long long)ties[2]);
    std::vector<uint64_t> row(D + 1);
    std::vector<double> pa(D + 1), mA(D + 1), mB(D + 1), fc(D + 1), norm((D + 1) * (size_t)S);
    CHECK(dk_ttest_export_survivors(ctx, t, row.data(), nullptr, pa.data(), mA.data(), mB.data(), fc.data(), norm.data()));
    DkOut out;
    out.fp = stdout;
    for (uint64_t i = 0; i < D; i++) {   // :487-489
        out.put(mt.tag(row[i]), mt.tag_len(row[i]));
        out.printf_("\t%f", pa[i]); out.printf_("\t%f", mA[i]); out.printf_("\t%f", mB[i]); out.printf_("\t%f", fc[i]);
        for (int j = 0; j < S; j++) out.printf_("\t%.2f", norm[i * (size_t)S + j]);
        out.putc_('\n');
    }
    out.flush();
    dk_ttest_free(ctx, t);
    dk_matrix_free(ctx, m);
    dk_free(ctx);
    return 0;
}
