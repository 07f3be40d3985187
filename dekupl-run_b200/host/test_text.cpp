// CPU-only check of the streaming tokeniser (DkStreamCursor) against the in-memory one (DkCursor): the same record
// stream joinCounts' parser sees, for buffer sizes down to a few bytes. usage: test_text <file> <bufsize>  -> prints records
#include "dk_text.h"

template <class C>
static void walk(C &cur, std::string &out)
{
    const char *s;
    size_t len;
    bool first = true;
    while (cur.token(s, len)) {                         // the loop of joinCounts.cpp (joinCounts.c:55-68, 86-92)
        if (len == 0 && cur.eof()) break;
        if (len == 0) { if (!first) break; }
        out.append(s, len);
        out += ' ';
        const char *cs = "";
        size_t cl = 0;
        if (cur.dret != 0) cur.token(cs, cl);
        out += std::to_string(dk_atoi(cs, cl));
        out += '\n';
        if (first) { cur.skip_line(); first = false; }
    }
}

int main(int argc, char **argv)
{
    if (argc < 3) return 2;
    DkFile f;
    if (!f.load(argv[1])) return 3;
    DkCursor a(f);
    std::string ra, rb;
    walk(a, ra);
    DkStreamCursor b;
    if (!b.open(argv[1], (size_t)atol(argv[2]))) return 3;
    walk(b, rb);
    fwrite(rb.data(), 1, rb.size(), stdout);
    return ra == rb ? 0 : 1;
}
