// dk_text.h — text side of the drop-in CLIs: gz/plain file slurp, kseq-compatible tokenisation,
// 2-bit k-mer codec, buffered formatting. The reference tools read through klib's kstream
// (share/*/kseq.h:90-141): tokens end at ANY isspace byte, files may be gzip (multi-member) or
// plain, the last line may lack '\n', numbers are parsed with atoi/atof.
#pragma once
#include <ctype.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>
#include <string>
#include <thread>
#include <vector>

struct DkFile {
    std::vector<char> buf;   // whole decompressed content + trailing '\0'
    bool load(const char *path)
    {
        gzFile fp = gzopen(path, "r");            // transparent for plain files, concatenated members handled by zlib
        if (!fp) return false;
        gzbuffer(fp, 1 << 20);
        size_t cap = 1 << 22, len = 0;
        buf.resize(cap);
        for (;;) {
            if (cap - len < (1 << 20)) { cap *= 2; buf.resize(cap); }
            int n = gzread(fp, buf.data() + len, (unsigned)(cap - len - 1 > (1u << 30) ? (1u << 30) : cap - len - 1));
            if (n <= 0) break;
            len += (size_t)n;
        }
        gzclose(fp);
        buf.resize(len + 1);
        buf[len] = '\0';
        return true;
    }
    size_t size() const { return buf.size() - 1; }
};

// cursor over a DkFile with ks_getuntil(KS_SEP_SPACE) semantics: a token is the (possibly empty)
// run of bytes up to the next isspace byte; dret = that delimiter (0 at EOF).
struct DkCursor {
    const char *p, *end;
    int dret = 0;
    DkCursor(const DkFile &f) : p(f.buf.data()), end(f.buf.data() + f.size()) {}
    DkCursor(const char *b, const char *e) : p(b), end(e) {}
    bool eof() const { return p >= end; }
    // returns false when no token could be read (EOF before any byte)
    bool token(const char *&s, size_t &len)
    {
        if (p >= end) { dret = 0; return false; }
        s = p;
        while (p < end && !isspace((unsigned char)*p)) p++;
        len = (size_t)(p - s);
        if (p < end) { dret = (unsigned char)*p; p++; } else dret = 0;
        return true;
    }
    void skip_line()   // as `while ((dret = ks_getc(ks)) > 0 && dret != '\n')`
    {
        if (dret == '\n') return;
        while (p < end && *p != '\n') p++;
        if (p < end) { dret = '\n'; p++; } else dret = 0;
    }
    bool line(const char *&s, size_t &len)   // KS_SEP_LINE
    {
        if (p >= end) return false;
        s = p;
        while (p < end && *p != '\n') p++;
        len = (size_t)(p - s);
        if (p < end) { dret = '\n'; p++; } else dret = 0;
        return true;
    }
};

// The same cursor over a file that is inflated in blocks instead of being held in memory as a whole (a dump of 300 M
// records is ~10 GB of text): the buffer is refilled when a token or a line runs into its end, the unfinished bytes moving
// to the front. A token pointer stays valid until the next call. `bufsize` is a parameter for the tokeniser's own test.
struct DkStreamCursor {
    gzFile fp = nullptr;
    std::vector<char> buf;
    size_t p = 0, end = 0;
    bool file_eof = false;
    int dret = 0;
    bool open(const char *path, size_t bufsize = (size_t)8 << 20)
    {
        fp = gzopen(path, "r");                   // transparent for plain files, concatenated members handled by zlib
        if (!fp) return false;
        gzbuffer(fp, 1 << 20);
        buf.resize(bufsize < 16 ? 16 : bufsize);
        return true;
    }
    ~DkStreamCursor() { if (fp) gzclose(fp); }
    // keep [from, end) at the front of the buffer (growing it when a single token is longer than the buffer) and read on
    bool refill(size_t from)
    {
        if (file_eof) return false;
        const size_t keep = end - from;
        if (keep && from) memmove(buf.data(), buf.data() + from, keep);
        if (keep == buf.size()) buf.resize(buf.size() * 2);
        p -= from; end = keep;
        const int n = gzread(fp, buf.data() + end, (unsigned)(buf.size() - end));
        if (n <= 0) { file_eof = true; return false; }
        end += (size_t)n;
        return true;
    }
    bool eof()
    {
        if (p < end) return false;
        p = end = 0;
        return !refill(0);
    }
    bool token(const char *&s, size_t &len)
    {
        if (eof()) { dret = 0; return false; }
        size_t start = p;
        for (;;) {
            while (p < end && !isspace((unsigned char)buf[p])) p++;
            if (p < end) break;                   // delimiter found
            const size_t from = start;
            if (!refill(from)) break;             // end of file: the token ends with the data
            start = 0;
        }
        s = buf.data() + start;
        len = p - start;
        if (p < end) { dret = (unsigned char)buf[p]; p++; } else dret = 0;
        return true;
    }
    void skip_line()
    {
        if (dret == '\n') return;
        for (;;) {
            while (p < end && buf[p] != '\n') p++;
            if (p < end) { dret = '\n'; p++; return; }
            p = end = 0;
            if (!refill(0)) { dret = 0; return; }
        }
    }
};

static inline int dk_atoi(const char *s, size_t len)   // atoi on a non-terminated token
{
    size_t i = 0;
    bool neg = false;
    if (i < len && (s[i] == '+' || s[i] == '-')) { neg = s[i] == '-'; i++; }
    long long v = 0;
    for (; i < len && s[i] >= '0' && s[i] <= '9'; i++) { v = v * 10 + (s[i] - '0'); if (v > 0x7fffffffLL + 1) v = 0x7fffffffLL + 1; }
    v = neg ? -v : v;
    return (int)v;
}

// 2-bit packing, first base most significant; returns false on k > 31 or a non-ACGT byte
static inline bool dk_encode(const char *s, size_t len, uint64_t &out)
{
    if (len == 0 || len > 31) return false;
    uint64_t v = 0;
    for (size_t i = 0; i < len; i++) {
        unsigned c;
        switch (s[i]) {
            case 'A': c = 0; break;
            case 'C': c = 1; break;
            case 'G': c = 2; break;
            case 'T': c = 3; break;
            default: return false;
        }
        v = (v << 2) | c;
    }
    out = v;
    return true;
}
static inline void dk_decode(uint64_t v, int k, char *out)
{
    for (int i = k - 1; i >= 0; i--) { out[i] = "ACGT"[v & 3]; v >>= 2; }
}

struct DkOut {   // buffered stdout/gz writer
    FILE *fp = nullptr;
    gzFile gz = nullptr;
    std::vector<char> b;
    DkOut() { b.reserve(1 << 22); }
    void flush()
    {
        if (b.empty()) return;
        if (gz) gzwrite(gz, b.data(), (unsigned)b.size()); else fwrite(b.data(), 1, b.size(), fp);
        b.clear();
    }
    void put(const char *s, size_t n) { b.insert(b.end(), s, s + n); if (b.size() > (1 << 22) - 4096) flush(); }
    void putc_(char c) { b.push_back(c); }
    void put_int(int v)   // "%d"
    {
        char t[16];
        int n = 0;
        long long x = v;
        bool neg = x < 0;
        if (neg) x = -x;
        do { t[n++] = (char)('0' + x % 10); x /= 10; } while (x);
        if (neg) t[n++] = '-';
        while (n) b.push_back(t[--n]);
    }
    void printf_(const char *fmt, double v)
    {
        char t[512];
        int n = snprintf(t, sizeof t, fmt, v);
        put(t, (size_t)n);
    }
};

// Creating the CUDA context takes about a second on a fresh process, as long as inflating and parsing a mid-sized
// input: the CLIs start it on a helper thread first thing and pick the context up when the text is ready.
#ifdef DEKUPL_B200_H
struct DkAsyncInit {
    dk_ctx *ctx = nullptr;
    int rc = DK_OK;
    std::string err;
    std::thread th;
    void start() { th = std::thread([this] { rc = dk_init(&ctx, 0); if (rc != DK_OK) err = dk_last_error(nullptr); }); }
    dk_ctx *wait() { if (th.joinable()) th.join(); return rc == DK_OK ? ctx : nullptr; }
    ~DkAsyncInit() { if (th.joinable()) th.join(); }
};
static DkAsyncInit dk_async_init_obj;
// exit() from any depth stays safe: the helper thread is joined before the C++ runtime tears the process down
static inline void dk_async_init_start()
{
    dk_async_init_obj.start();
    atexit([] { dk_async_init_obj.wait(); });
}
#endif
