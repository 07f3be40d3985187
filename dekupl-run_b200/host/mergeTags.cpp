// mergeTags — drop-in replacement of share/mergeTags/mergeTags.c for the stranded procedure (same argv and stdout):
// assembles the differentially expressed k-mers of diff-counts.tsv.gz into contigs on the GPU (dk_merge_tags).
// Invoked by Snakefile:691-707 as `mergeTags -k K [-n] diff-counts.tsv.gz | gzip -c > merged-diff-counts.tsv.gz`.
// -n (unstranded libraries) is refused: the reference's result there depends on its hash-table iteration order.
#include <unistd.h>
#include <algorithm>
#include "../../include/dekupl_b200.h"
#include "dk_text.h"

int main(int argc, char *argv[])
{
    int k_length = 31, stranded = 1, min_assembly_k = 15, c;
    while ((c = getopt(argc, argv, "nk:m:")) >= 0) {
        switch (c) {
            case 'n': stranded = 0; break;
            case 'k': k_length = atoi(optarg); break;
            case 'm': min_assembly_k = atoi(optarg); break;
        }
    }
    if (optind == argc) {   // mergeTags.c:225-233
        fprintf(stderr, "\n");
        fprintf(stderr, "Usage:   mergeTags [options] <counts.tsv>\n\n");
        fprintf(stderr, "Options: -k INT    length of k-mers (max_value: 32) [%d]\n", k_length);
        fprintf(stderr, "         -m INT    min assembly overlap (max_value: k) [%d]\n", min_assembly_k);
        fprintf(stderr, "         -n        Unstranded merging procedure\n");
        fprintf(stderr, "\n");
        return 1;
    }
    const char *path = argv[optind];
    if (!stranded) { fprintf(stderr, "mergeTags: the unstranded procedure (-n) is not provided by the GPU build\n"); return EXIT_FAILURE; }
    dk_async_init_start();
    fprintf(stderr, "Loading k-mers into memory\n");
    DkFile f;
    if (!f.load(path)) { fprintf(stderr, "Failed to open %s\n", path); exit(EXIT_FAILURE); }
    DkCursor cur(f);
    const char *s;
    size_t len;
    // header: every token but the last one is counted (mergeTags.c:259-262) = the columns that follow the tag
    int nb_cols = 0;
    const char *hdr = cur.p;
    while (cur.token(s, len) && len > 0 && cur.dret != '\n') nb_cols++;
    size_t hdr_len = (size_t)(cur.p - hdr);
    if (hdr_len && hdr[hdr_len - 1] == '\n') hdr_len--;
    std::vector<std::pair<const char *, size_t>> lines;       // whole data lines, printed back unchanged
    std::vector<char> seqs[2];
    std::vector<double> pv[2];
    std::vector<uint32_t> ids[2];
    uint32_t nb_kmers = 0;
    while (!cur.eof()) {
        const char *l0 = cur.p;
        if (!cur.token(s, len)) break;
        if (len == 0 && cur.eof()) break;
        if ((int)len != k_length) { fprintf(stderr, "mergeTags: line %u: k-mer of %zu bases with -k %d\n", nb_kmers + 2, len, k_length); return EXIT_FAILURE; }
        const char *tag = s;
        double pvalue = 0, log2fc = 0;
        for (int i = 0; i < nb_cols; i++) {
            const char *cs;
            size_t cl;
            if (cur.dret == '\n' || cur.dret == 0 || !cur.token(cs, cl) || cl == 0) {
                fprintf(stderr, "Invalid number of columns at line %u\n", nb_kmers + 2);      // :275-278
                exit(1);
            }
            double v = atof(std::string(cs, cl).c_str());
            if (i == 0) pvalue = v;
            if (i == 3) log2fc = v;
        }
        if (cur.dret != '\n' && !(cur.dret == 0 && cur.eof())) { fprintf(stderr, "Invalid number of columns at line %u\n", nb_kmers + 2); exit(1); }
        size_t ll = (size_t)(cur.p - l0);
        if (ll && l0[ll - 1] == '\n') ll--;
        lines.emplace_back(l0, ll);
        const int g = log2fc > 0 ? 0 : 1;                      // :283-288 up- and down-regulated k-mers are assembled apart
        seqs[g].insert(seqs[g].end(), tag, tag + k_length);
        pv[g].push_back(pvalue);
        ids[g].push_back(nb_kmers);
        nb_kmers++;
    }
    fprintf(stderr, "%u k-mers loaded\n", nb_kmers);
    fprintf(stderr, "Assembling k-mers\n");
    dk_ctx *ctx = nullptr;
    if (!(ctx = dk_async_init_obj.wait())) { fprintf(stderr, "mergeTags: %s\n", dk_async_init_obj.err.c_str()); return EXIT_FAILURE; }
    struct Contig { uint32_t rep, nb; std::string seq; };
    std::vector<Contig> all;
    for (int g = 0; g < 2; g++) {
        fprintf(stderr, "Assembling %zu %s regulated k-mers\n", pv[g].size(), g == 0 ? "up" : "down");
        dk_contigs *cg = nullptr;
        if (dk_merge_tags(ctx, seqs[g].data(), pv[g].size(), k_length, pv[g].data(), ids[g].data(), min_assembly_k, &cg) != DK_OK) {
            fprintf(stderr, "mergeTags: %s\n", dk_last_error(ctx));
            return EXIT_FAILURE;
        }
        const uint64_t m = dk_contigs_count(cg), nch = dk_contigs_chars(cg);
        std::vector<uint32_t> rep(m), nb(m);
        std::vector<uint64_t> off(m + 1);
        std::vector<char> ch(nch + 1);
        dk_contigs_export(cg, rep.data(), nb.data(), off.data(), ch.data());
        if (dk_contigs_rings(cg))
            fprintf(stderr, "mergeTags: %llu contigs were linked in rings and left unmerged at their level\n",
                    (unsigned long long)dk_contigs_rings(cg));
        for (uint64_t j = 0; j < m; j++) all.push_back({rep[j], nb[j], std::string(ch.data() + off[j], ch.data() + off[j + 1])});
        dk_contigs_free(cg);
        fprintf(stderr, "%llu assemblies after merging\n", (unsigned long long)m);
    }
    fprintf(stderr, "Merging up and down assemblies\n");
    std::sort(all.begin(), all.end(), [](const Contig &a, const Contig &b) { return a.rep < b.rep; });      // :307
    DkOut out;
    out.fp = stdout;
    out.put("nb_merged_kmers\tcontig\t", 23);                 // :316
    out.put(hdr, hdr_len);
    out.putc_('\n');
    for (const Contig &cg : all) {                             // :317-326
        out.put_int((int)cg.nb);
        out.putc_('\t');
        out.put(cg.seq.data(), cg.seq.size());
        out.putc_('\t');
        out.put(lines[cg.rep].first, lines[cg.rep].second);
        out.putc_('\n');
    }
    out.flush();
    dk_free(ctx);
    return 0;
}
