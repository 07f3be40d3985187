// dk_matrix_text.h — parser of the count-matrix text ("tag<ws>s1<ws>...sN\n" header + rows) as
// computeNF.c:79-135 and TtestFilter.c:237-247,373-389 read it.
#pragma once
#include "dk_text.h"

struct DkMatrixText {
    std::vector<std::string> names;
    std::vector<int32_t> cnts;                       // row-major R x S
    std::vector<std::pair<const char *, uint32_t>> tags;   // views into the file buffer (when keep_tags)
    uint64_t R = 0;
    int S = 0;
    // returns 0 ok, else the 1-based data line whose column count is inconsistent
    size_t parse(const DkFile &f, bool keep_tags)
    {
        DkCursor cur(f);
        const char *s;
        size_t len;
        cur.token(s, len);                            // first header token is ignored (computeNF.c:83)
        while (cur.dret != '\n' && cur.token(s, len)) {
            names.emplace_back(s, len);
            if (cur.dret == '\n') break;
        }
        S = (int)names.size();
        cnts.reserve(f.size() / 3);
        size_t line = 0;
        while (cur.token(s, len)) {
            if (len == 0 && cur.eof()) break;
            line++;
            if (keep_tags) tags.emplace_back(s, (uint32_t)len);
            int j = 0;
            while (j < S && cur.dret != 0) {
                const char *cs;
                size_t cl;
                if (!cur.token(cs, cl)) break;
                cnts.push_back(dk_atoi(cs, cl));
                j++;
            }
            while (j < S) { cnts.push_back(0); j++; }
            if (cur.dret != '\n' && !(cur.dret == 0 && cur.eof() && j == S)) return line;   // computeNF.c:132-135
            R++;
        }
        return 0;
    }
};
