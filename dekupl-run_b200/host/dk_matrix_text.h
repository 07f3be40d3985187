// dk_matrix_text.h — parser of the count-matrix text ("tag<ws>s1<ws>...sN\n" header + rows) as
// computeNF.c:79-135 and TtestFilter.c:237-247,373-389 read it. Large tables are parsed by several threads, each on
// a run of whole lines; the result is accepted only when every thread consumed exactly one line per row (true for
// every table the pipeline writes), otherwise the sequential reader, which follows the reference token by token,
// is used.
#pragma once
#include <thread>
#include "dk_text.h"

struct DkMatrixText {
    std::vector<std::string> names;
    std::vector<int32_t> cnts;                       // row-major R x S
    std::vector<std::pair<const char *, uint32_t>> tags;   // views into the file buffer (when keep_tags)
    uint64_t R = 0;
    int S = 0;

    struct Part {
        std::vector<int32_t> cnts;
        std::vector<std::pair<const char *, uint32_t>> tags;
        uint64_t rows = 0;
        size_t bad = 0;                              // 1-based row (within the part) with an inconsistent column count
    };
    // rows of [b,e) in the reference's token order; stops at the first inconsistent row
    static void parse_range(const char *b, const char *e, int S, bool keep_tags, Part &o)
    {
        DkCursor cur(b, e);
        const char *s;
        size_t len;
        o.cnts.reserve((size_t)(e - b) / 3);
        size_t line = 0;
        while (cur.token(s, len)) {
            if (len == 0 && cur.eof()) break;
            line++;
            if (keep_tags) o.tags.emplace_back(s, (uint32_t)len);
            int j = 0;
            while (j < S && cur.dret != 0) {
                const char *cs;
                size_t cl;
                if (!cur.token(cs, cl)) break;
                o.cnts.push_back(dk_atoi(cs, cl));
                j++;
            }
            while (j < S) { o.cnts.push_back(0); j++; }
            if (cur.dret != '\n' && !(cur.dret == 0 && cur.eof() && j == S)) { o.bad = line; return; }   // computeNF.c:132-135
            o.rows++;
        }
    }
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
        const char *b = cur.p, *e = cur.end;
        unsigned hw = std::thread::hardware_concurrency();
        const char *env = getenv("DK_PARSE_THREADS");
        int T = env ? atoi(env) : (int)(hw ? hw : 1);
        if (T > 32) T = 32;
        const size_t min_part = env ? 1 : (size_t)4 << 20;
        if (T > 1 && (size_t)(e - b) >= 2 * min_part && S > 0) {
            if ((size_t)(e - b) / min_part < (size_t)T) T = (int)((size_t)(e - b) / min_part);
            std::vector<const char *> cut(T + 1);
            cut[0] = b; cut[T] = e;
            for (int t = 1; t < T; t++) {             // cut after a newline
                const char *q = b + (size_t)(e - b) * t / T;
                if (q < cut[t - 1]) q = cut[t - 1];
                while (q < e && *q != '\n') q++;
                cut[t] = q < e ? q + 1 : e;
            }
            std::vector<Part> part(T);
            std::vector<std::thread> pool;
            for (int t = 0; t < T; t++) pool.emplace_back([&, t] { parse_range(cut[t], cut[t + 1], S, keep_tags, part[t]); });
            for (auto &th : pool) th.join();
            bool ok = true;
            for (int t = 0; t < T && ok; t++) {
                if (part[t].bad) { ok = false; break; }
                uint64_t nl = 0;                      // one line per row, the last row of the file may lack its newline
                for (const char *q = cut[t]; q < cut[t + 1]; q++) nl += *q == '\n';
                const bool open_tail = cut[t + 1] > cut[t] && cut[t + 1][-1] != '\n';
                if (part[t].rows != nl + (open_tail ? 1 : 0)) ok = false;
            }
            if (ok) {
                size_t tot = 0;
                for (auto &p : part) { R += p.rows; tot += p.cnts.size(); }
                cnts.resize(tot);
                size_t at = 0;
                for (auto &p : part) {
                    memcpy(cnts.data() + at, p.cnts.data(), p.cnts.size() * sizeof(int32_t));
                    at += p.cnts.size();
                    if (keep_tags) tags.insert(tags.end(), p.tags.begin(), p.tags.end());
                    std::vector<int32_t>().swap(p.cnts);
                }
                return 0;
            }
        }
        Part all;
        parse_range(b, e, S, keep_tags, all);
        cnts.swap(all.cnts);
        tags.swap(all.tags);
        R = all.rows;
        return all.bad;
    }
};
