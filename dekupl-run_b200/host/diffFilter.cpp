// diffFilter — drop-in replacement of bin/diffFilter.pl (same argv and stdout): prints the header and
// every row of the count table whose k-mer is absent from the sorted reference-transcript dump.
// Invoked by Snakefile:634-644 as `diffFilter.pl ref_transcripts.tsv.gz raw-counts.tsv.gz | gzip -c`.
// The anti-join runs on the GPU (dk_keys_antijoin); the text rows stay here and the survivors are
// printed byte for byte, so any count formatting in the table passes through untouched.
#include <unistd.h>
#include "../../include/dekupl_b200.h"
#include "dk_text.h"

static bool first_token(const char *s, size_t len, const char *&tok, size_t &tl)
{
    size_t i = 0;
    while (i < len && !isspace((unsigned char)s[i])) i++;
    tok = s; tl = i;
    return i > 0;
}

// first column of every line, 2-bit packed; dies on anything that is not one fixed-length ACGT word
static void pack_column(const DkFile &f, bool skip_header, const char *what, std::vector<uint64_t> &keys,
                        std::vector<std::pair<const char *, size_t>> *lines, std::pair<const char *, size_t> *header, size_t &k)
{
    DkCursor cur(f);
    const char *s;
    size_t len;
    if (skip_header) {
        if (!cur.line(s, len)) { if (header) *header = {nullptr, 0}; return; }
        if (header) *header = {s, len + (cur.dret == '\n' ? 1 : 0)};
    }
    while (cur.line(s, len)) {
        size_t full = len + (cur.dret == '\n' ? 1 : 0);
        const char *tok;
        size_t tl;
        uint64_t v;
        if (!first_token(s, len, tok, tl) || !dk_encode(tok, tl, v) || (k && tl != k)) {
            fprintf(stderr, "diffFilter: %s line %zu: k-mers must be one length of at most 31 ACGT bases\n", what, keys.size() + 1);
            exit(EXIT_FAILURE);
        }
        k = tl;
        keys.push_back(v);
        if (lines) lines->emplace_back(s, full);
    }
}

int main(int argc, char *argv[])
{
    if (argc < 3) {
        fprintf(stderr, "Usage:   diffFilter <ref_kmers.tsv[.gz]> <counts.tsv[.gz]>\n");
        return 1;
    }
    dk_async_init_start();                      // CUDA context comes up while the two files are inflated and parsed
    DkFile diff, counts;
    if (!diff.load(argv[1])) { fprintf(stderr, "Cannot open %s\n", argv[1]); return 2; }       // diffFilter.pl:35,37
    if (!counts.load(argv[2])) { fprintf(stderr, "Cannot open %s\n", argv[2]); return 2; }
    std::vector<uint64_t> mask, keys;
    std::vector<std::pair<const char *, size_t>> lines;
    std::pair<const char *, size_t> header{nullptr, 0};
    size_t k = 0;
    pack_column(counts, true, argv[2], keys, &lines, &header, k);
    pack_column(diff, false, argv[1], mask, nullptr, nullptr, k);
    std::vector<uint8_t> keep(keys.size());
    dk_ctx *ctx = nullptr;
    if (!(ctx = dk_async_init_obj.wait())) { fprintf(stderr, "diffFilter: %s\n", dk_async_init_obj.err.c_str()); return EXIT_FAILURE; }
    if (dk_keys_antijoin(ctx, keys.data(), keys.size(), mask.data(), mask.size(), keep.data()) != DK_OK) {
        fprintf(stderr, "diffFilter: %s\n", dk_last_error(ctx));
        return EXIT_FAILURE;
    }
    DkOut out;
    out.fp = stdout;
    if (header.first) out.put(header.first, header.second);                                    // diffFilter.pl:53-54
    for (size_t i = 0; i < lines.size(); i++)
        if (keep[i]) out.put(lines[i].first, lines[i].second);                                  // diffFilter.pl:67-69
    out.flush();
    dk_free(ctx);
    return 0;
}
