// CPU-only check of the block-streamed table reader (DkMatrixText::load) against the whole-file one (parse): names, rows,
// counts, tags and the reported inconsistent line must agree for any block size.
// usage: test_matrix_text <file> <block bytes>  -> prints "S R bad checksum"; exit 0 when both readers agree
#include "dk_matrix_text.h"

static uint64_t mix(uint64_t h, uint64_t v) { h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2); return h; }

int main(int argc, char **argv)
{
    if (argc < 3) return 2;
    DkFile f;
    if (!f.load(argv[1])) return 3;
    DkMatrixText a, b;
    const size_t bad_a = a.parse(f, true);
    bool opened = false;
    const size_t bad_b = b.load(argv[1], true, opened, (size_t)atol(argv[2]));
    if (!opened) return 3;
    bool same = bad_a == bad_b && a.S == b.S && a.names == b.names;
    if (!bad_a) {                                   // after an inconsistent line the reference exits: nothing else is defined
        same = same && a.R == b.R && a.cnts == b.cnts && a.tags.size() == b.tag_at.size();
        for (size_t i = 0; same && i < a.tags.size(); i++)
            same = a.tags[i].second == b.tag_len(i) && memcmp(a.tags[i].first, b.tag(i), a.tags[i].second) == 0;
    }
    uint64_t h = 0;
    for (int32_t c : b.cnts) h = mix(h, (uint32_t)c);
    for (size_t i = 0; i < b.tag_at.size(); i++) for (uint32_t j = 0; j < b.tag_len(i); j++) h = mix(h, (unsigned char)b.tag(i)[j]);
    printf("%d %llu %zu %016llx\n", b.S, (unsigned long long)b.R, bad_b, (unsigned long long)h);
    return same ? 0 : 1;
}
