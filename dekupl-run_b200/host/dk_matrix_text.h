// dk_matrix_text.h — parser of the count-matrix text ("tag<ws>s1<ws>...sN\n" header + rows) as
// computeNF.c:79-135 and TtestFilter.c:237-247,373-389 read it. Large tables are parsed by several threads, each on
// a run of whole lines; the result is accepted only when every thread consumed exactly one line per row (true for
// every table the pipeline writes), otherwise the sequential reader, which follows the reference token by token,
// is used. load() reads the file in blocks cut at line ends (the table of a 16-sample cohort is tens of GB of text; only
// the packed counts and the tags stay in memory) and hands anything that is not one line per row to the whole-file reader.
#pragma once
#include <thread>
#include "dk_text.h"

struct DkMatrixText {
    std::vector<std::string> names;
    std::vector<int32_t> cnts;                       // row-major R x S
    std::vector<std::pair<const char *, uint32_t>> tags;   // parse(): views into the file buffer (when keep_tags)
    std::vector<char> tag_pool;                      // load(): the tags' bytes, back to back
    std::vector<std::pair<uint64_t, uint32_t>> tag_at;     // load(): (offset into tag_pool, length) per row
    const char *tag(uint64_t i) const { return tag_pool.data() + tag_at[i].first; }
    uint32_t tag_len(uint64_t i) const { return tag_at[i].second; }
    uint64_t R = 0;
    int S = 0;

    struct Part {
        std::vector<int32_t> cnts;
        std::vector<std::pair<const char *, uint32_t>> tags;
        uint64_t rows = 0;
        size_t bad = 0;                              // 1-based row (within the part) with an inconsistent column count
    };
    // rows of [b,e) in the reference's token order; stops at the first inconsistent row. Only at the end of the FILE may a
    // row run out of tokens (the reference pads it with zeros there); a row that runs into the end of a range cut out of
    // the middle of the file is reported as inconsistent, which sends the caller to the whole-file reader
    static void parse_range(const char *b, const char *e, int S, bool keep_tags, Part &o, bool file_end = true)
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
            if (cur.dret != '\n' && !(file_end && cur.dret == 0 && cur.eof() && j == S)) { o.bad = line; return; }   // computeNF.c:132-135
            o.rows++;
        }
    }
    static int n_threads()
    {
        unsigned hw = std::thread::hardware_concurrency();
        const char *env = getenv("DK_PARSE_THREADS");
        int T = env ? atoi(env) : (int)(hw ? hw : 1);
        return T > 32 ? 32 : (T < 1 ? 1 : T);
    }
    void pool_tags(const std::vector<std::pair<const char *, uint32_t>> &v)
    {
        for (auto &t : v) {
            tag_at.emplace_back((uint64_t)tag_pool.size(), t.second);
            tag_pool.insert(tag_pool.end(), t.first, t.first + t.second);
        }
    }
    // rows of the whole lines [b,e) appended to cnts (and to tags, or to the pool when `pooled`), parsed by several threads
    // when the range is large; false (nothing appended) unless every thread consumed exactly one line per row
    bool parse_lines(const char *b, const char *e, bool keep_tags, bool pooled, size_t min_part, bool file_end)
    {
        int T = n_threads();
        if (getenv("DK_PARSE_THREADS")) min_part = 1;
        if (min_part < 1) min_part = 1;
        if ((size_t)(e - b) / min_part < (size_t)T) T = (int)((size_t)(e - b) / min_part);
        if (T < 1) T = 1;
        std::vector<const char *> cut(T + 1);
        cut[0] = b; cut[T] = e;
        for (int t = 1; t < T; t++) {                 // cut after a newline
            const char *q = b + (size_t)(e - b) * t / T;
            if (q < cut[t - 1]) q = cut[t - 1];
            while (q < e && *q != '\n') q++;
            cut[t] = q < e ? q + 1 : e;
        }
        std::vector<Part> part(T);
        if (T == 1) parse_range(cut[0], cut[1], S, keep_tags, part[0], file_end);
        else {
            std::vector<std::thread> pool;
            for (int t = 0; t < T; t++) pool.emplace_back([&, t] { parse_range(cut[t], cut[t + 1], S, keep_tags, part[t], file_end && t == T - 1); });
            for (auto &th : pool) th.join();
        }
        for (int t = 0; t < T; t++) {
            if (part[t].bad) return false;
            uint64_t nl = 0;                          // one line per row, the last row of the file may lack its newline
            for (const char *q = cut[t]; q < cut[t + 1]; q++) nl += *q == '\n';
            const bool open_tail = cut[t + 1] > cut[t] && cut[t + 1][-1] != '\n';
            if (part[t].rows != nl + (open_tail ? 1 : 0)) return false;
        }
        size_t tot = cnts.size();
        for (auto &p : part) tot += p.cnts.size();
        size_t at = cnts.size();
        cnts.resize(tot);
        for (auto &p : part) {
            R += p.rows;
            if (!p.cnts.empty()) memcpy(cnts.data() + at, p.cnts.data(), p.cnts.size() * sizeof(int32_t));
            at += p.cnts.size();
            if (keep_tags) { if (pooled) pool_tags(p.tags); else tags.insert(tags.end(), p.tags.begin(), p.tags.end()); }
            std::vector<int32_t>().swap(p.cnts);
        }
        return true;
    }
    // header line: the first token is ignored (computeNF.c:83), the others name the samples; returns the cursor position
    const char *parse_header(const char *b, const char *e)
    {
        DkCursor cur(b, e);
        const char *s;
        size_t len;
        cur.token(s, len);
        while (cur.dret != '\n' && cur.token(s, len)) {
            names.emplace_back(s, len);
            if (cur.dret == '\n') break;
        }
        S = (int)names.size();
        return cur.p;
    }
    // the table of `path`, read in blocks of about `block` bytes cut at line ends. Returns 0, or the 1-based data line whose
    // column count is inconsistent; opened = false when the file cannot be read. Tags (keep_tags) end up in the pool.
    size_t load(const char *path, bool keep_tags, bool &opened, size_t block = (size_t)64 << 20)
    {
        gzFile fp = gzopen(path, "r");
        opened = fp != nullptr;
        if (!fp) return 0;
        gzbuffer(fp, 1 << 20);
        std::vector<char> buf(block < 64 ? 64 : block);
        size_t len = 0;
        bool eof = false;
        auto fill = [&] {
            while (len < buf.size() && !eof) {
                const size_t want = buf.size() - len;
                const int n = gzread(fp, buf.data() + len, (unsigned)(want > (1u << 30) ? (1u << 30) : want));
                if (n <= 0) eof = true; else len += (size_t)n;
            }
        };
        auto has_nl = [&](size_t from) { return memchr(buf.data() + from, '\n', len - from) != nullptr; };
        fill();
        while (!eof && !has_nl(0)) { buf.resize(buf.size() * 2); fill(); }     // the header line, whole
        size_t start = (size_t)(parse_header(buf.data(), buf.data() + len) - buf.data());
        bool regular = S > 0;
        while (regular) {
            size_t stop = len;
            if (!eof) {                                                      // up to the last line end of the block
                while (stop > start && buf[stop - 1] != '\n') stop--;
                if (stop == start) {                                         // a line longer than the block
                    if (start) { memmove(buf.data(), buf.data() + start, len - start); len -= start; start = 0; }
                    else buf.resize(buf.size() * 2);
                    fill();
                    continue;
                }
            }
            if (stop > start && !parse_lines(buf.data() + start, buf.data() + stop, keep_tags, true, (size_t)4 << 20, eof)) { regular = false; break; }
            if (eof) break;
            memmove(buf.data(), buf.data() + stop, len - stop);
            len -= stop; start = 0;
            fill();
        }
        gzclose(fp);
        if (regular) return 0;
        // not one line per row somewhere: the whole-file reader follows the reference token by token
        names.clear(); cnts.clear(); tags.clear(); tag_pool.clear(); tag_at.clear(); R = 0; S = 0;
        std::vector<char>().swap(buf);
        DkFile f;
        if (!f.load(path)) { opened = false; return 0; }
        const size_t bad = parse(f, keep_tags);
        pool_tags(tags);
        tags.clear();
        return bad;
    }
    // returns 0 ok, else the 1-based data line whose column count is inconsistent
    size_t parse(const DkFile &f, bool keep_tags)
    {
        const char *e = f.buf.data() + f.size();
        const char *b = parse_header(f.buf.data(), e);
        if (S > 0 && parse_lines(b, e, keep_tags, false, (size_t)4 << 20, true)) return 0;
        cnts.clear(); tags.clear(); R = 0;
        Part all;
        parse_range(b, e, S, keep_tags, all);
        cnts.swap(all.cnts);
        tags.swap(all.tags);
        R = all.rows;
        return all.bad;
    }
};
