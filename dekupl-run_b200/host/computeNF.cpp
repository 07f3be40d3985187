// computeNF — drop-in replacement of share/computeNF/computeNF.c (same argv and stdout),
// median-of-ratios normalisation factors on the GPU through libdekupl_b200.so.
// Invoked by Snakefile:364-371 as `computeNF raw-counts.tsv.gz > normalization_factors.tsv`.
// DK_NF_MODE=full in the environment selects the all-rows median instead of the reference's
// step / early-stop prefix semantics (computeNF.c:138-168), which is the default.
#include <unistd.h>
#include "../../include/dekupl_b200.h"
#include "dk_matrix_text.h"

int main(int argc, char *argv[])
{
    if (optind >= argc) {   // computeNF.c:55-59
        fprintf(stderr, "\n");
        fprintf(stderr, "Usage:   computeNF [options] <counts.tsv>\n\n");
        return 1;
    }
    const char *path = argv[optind];
    dk_async_init_start();                      // CUDA context comes up while the table is inflated and parsed
    DkMatrixText mt;                            // read in blocks: only the packed counts stay in memory
    bool opened = false;
    size_t bad = mt.load(path, false, opened);
    if (!opened) { fprintf(stderr, "Failed to open %s\n", path); exit(EXIT_FAILURE); }
    if (bad) { fprintf(stderr, "inconsistent number of column (line %zu)\n", bad); exit(EXIT_FAILURE); }
    if (mt.S < 1 || mt.S > DK_MAX_SAMPLES) { fprintf(stderr, "computeNF: %d samples unsupported (1..%d)\n", mt.S, DK_MAX_SAMPLES); return EXIT_FAILURE; }
    dk_ctx *ctx = nullptr;
    if (!(ctx = dk_async_init_obj.wait())) { fprintf(stderr, "computeNF: %s\n", dk_async_init_obj.err.c_str()); return EXIT_FAILURE; }
    dk_matrix *m = nullptr;
    if (dk_matrix_from_host(ctx, mt.S, mt.R, nullptr, mt.cnts.data(), &m) != DK_OK) { fprintf(stderr, "computeNF: %s\n", dk_last_error(ctx)); return EXIT_FAILURE; }
    const char *mode_env = getenv("DK_NF_MODE");
    int mode = (mode_env && strcmp(mode_env, "full") == 0) ? DK_NF_FULL : DK_NF_REFERENCE;
    std::vector<double> nf(mt.S);
    uint64_t info[4] = {0, 0, 0, 0};
    if (dk_compute_nf(ctx, m, mode, 0, nf.data(), info) != DK_OK) { fprintf(stderr, "computeNF: %s\n", dk_last_error(ctx)); return EXIT_FAILURE; }
    for (uint64_t i = 0; i < info[1] && mode == DK_NF_REFERENCE && mt.R >= 1000000; i++) fprintf(stderr, "compute normalization_factors\n");
    if (info[2])
        fprintf(stderr, "computeNF: %llu rows lie in [1M,2M): the reference prints uninitialised memory here; "
                        "printing the factors of the first 1M rows\n", (unsigned long long)mt.R);
    fprintf(stdout, "sample\tnormalization_factor\n");          // computeNF.c:170-176
    for (int j = 0; j < mt.S; j++) fprintf(stdout, "%s\t%f\n", mt.names[j].c_str(), nf[j]);
    dk_matrix_free(ctx, m);
    dk_free(ctx);
    return 0;
}
