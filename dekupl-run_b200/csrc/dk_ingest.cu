// dk_ingest.cu — K9, k-mer counting of one sample on the device: what the pipeline does with
// `jellyfish count -L 2 -m K [-C] <(zcat R1 | revCompFastq.pl) <(zcat R2)`, `jellyfish dump -c` and `sort -k1`
// (Snakefile:523-582, bin/revCompFastq.pl:33-44), producing directly the sorted (k-mer, count) table that
// dk_join_counts_dev reads — no .jf file, no text dump, no sort(1).
//   emit   one thread per read rolls the 2-bit forward and reverse-complement k-mers over the read; every window
//          without a non-ACGT base emits one 64-bit key: the forward k-mer, its reverse complement for a read
//          flagged "reverse-complemented" (the k-mers of revcomp(read) are the reverse complements of the read's),
//          or the smaller of the two in canonical mode (-C); other windows emit the all-ones sentinel;
//   sort   LSD radix sort of the keys (dk_sort.cu), sentinels last;
//   count  run heads -> scan -> run lengths; runs with length >= min_count (-L) are compacted into the dump.
// Lower-case acgt count like upper case (jellyfish's code table); any other byte resets the window.
#include "dk_common.cuh"

struct dk_dump {
    uint64_t n = 0;
    uint64_t *keys = nullptr;     // ascending, distinct
    int32_t *cnts = nullptr;
    uint64_t windows = 0, valid = 0, distinct = 0;
};

#define IG_T 128

__device__ __forceinline__ int ig_code(unsigned char c)
{
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return -1;
    }
}

__global__ void __launch_bounds__(IG_T) k_ingest_emit(const unsigned char *bases, const uint64_t *off, const uint64_t *woff,
                                                      const unsigned char *rc_flag, uint64_t n_reads, int k, int canonical,
                                                      uint64_t *out, unsigned long long *n_invalid)
{
    const uint64_t r = (uint64_t)blockIdx.x * IG_T + threadIdx.x;
    unsigned long long bad = 0;
    if (r < n_reads) {
        const uint64_t b = off[r], e = off[r + 1];
        uint64_t w = woff[r];
        const uint64_t mask = k == 32 ? ~0ull : ((1ull << (2 * k)) - 1ull);
        const int rc = rc_flag ? rc_flag[r] : 0, hs = 2 * (k - 1);
        uint64_t fw = 0, rv = 0;
        int run = 0;                                          // consecutive ACGT bases ending here
        for (uint64_t i = b; i < e; i++) {
            const int c = ig_code(bases[i]);
            if (c < 0) { run = 0; fw = rv = 0; }
            else {
                fw = ((fw << 2) | (uint64_t)c) & mask;
                rv = (rv >> 2) | ((uint64_t)(3 - c) << hs);
                run++;
            }
            if (i + 1 >= b + (uint64_t)k) {                   // window [i-k+1, i] exists
                uint64_t key = DK_SENTINEL;
                if (run >= k) key = canonical ? (fw < rv ? fw : rv) : (rc ? rv : fw);
                else bad++;
                out[w++] = key;
            }
        }
    }
    for (int o = 16; o; o >>= 1) bad += __shfl_down_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(n_invalid, bad);
}

__global__ void k_ingest_heads(const uint64_t *keys, uint64_t n_valid, uint32_t *head)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_valid) head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}
__global__ void k_ingest_upos(const uint32_t *head, const uint32_t *uid, uint64_t n_valid, uint32_t n_unique, uint32_t *upos)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_valid && head[i]) upos[uid[i]] = (uint32_t)i;
    if (i == 0) upos[n_unique] = (uint32_t)n_valid;
}
__global__ void k_ingest_keep(const uint32_t *upos, uint32_t n_unique, uint32_t min_count, uint32_t *keep)
{
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n_unique) keep[u] = (upos[u + 1] - upos[u]) >= min_count ? 1u : 0u;
}
__global__ void k_ingest_write(const uint64_t *keys, const uint32_t *upos, const uint32_t *keep, const uint32_t *kidx, uint32_t n_unique,
                               uint64_t *okeys, int32_t *ocnts)
{
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n_unique && keep[u]) {
        const uint32_t o = kidx[u], p = upos[u];
        okeys[o] = keys[p];
        ocnts[o] = (int32_t)(upos[u + 1] - p);
    }
}

static void dump_release(dk_ctx *ctx, dk_dump *d)
{
    if (!d) return;
    dk_dev_release(ctx, d->keys); dk_dev_release(ctx, d->cnts);
    delete d;
}

// bases: the reads' sequence lines back to back (host memory), read r = bases[offsets[r] .. offsets[r+1]).
// rc_flag (may be NULL): 1 for reads the pipeline passes through revCompFastq.pl (mate 1 of an "rf" library, mate 2 of
// "fr"). canonical != 0 is jellyfish -C. min_count is jellyfish -L. One call handles up to 2^32-2 windows.
extern "C" int dk_count_kmers(dk_ctx *ctx, const char *bases, const uint64_t *offsets, uint64_t n_reads, const uint8_t *rc_flag,
                              int k, int canonical, uint32_t min_count, dk_dump **out)
{
    if (!ctx || !out || (n_reads && (!bases || !offsets))) return dk_fail(ctx, DK_ERR_ARG, "count_kmers: NULL argument");
    if (k < 1 || k > 31) return dk_fail(ctx, DK_ERR_ARG, "count_kmers: k=%d outside [1,31]", k);
    *out = nullptr;
    cudaStream_t st = ctx->stream;
    dk_dump *d = new dk_dump();
    std::vector<uint64_t> woff(n_reads + 1);
    uint64_t W = 0;
    for (uint64_t r = 0; r < n_reads; r++) {
        if (offsets[r + 1] < offsets[r]) { delete d; return dk_fail(ctx, DK_ERR_ARG, "count_kmers: offsets not ascending at read %llu", (unsigned long long)r); }
        const uint64_t len = offsets[r + 1] - offsets[r];
        woff[r] = W;
        W += len >= (uint64_t)k ? len - k + 1 : 0;
    }
    woff[n_reads] = W;
    d->windows = W;
    if (W >= 0xFFFFFFFEull) { delete d; return dk_fail(ctx, DK_ERR_UNSUPPORTED, "count_kmers: %llu windows in one call; split the sample into read batches", (unsigned long long)W); }
    if (W == 0) { *out = d; return DK_OK; }
    const uint64_t nb = offsets[n_reads] - offsets[0];
    unsigned char *b_d = (unsigned char *)dk_dev_alloc(ctx, nb + 16);
    uint64_t *off_d = (uint64_t *)dk_dev_alloc(ctx, 8 * (n_reads + 1)), *woff_d = (uint64_t *)dk_dev_alloc(ctx, 8 * (n_reads + 1));
    unsigned char *rc_d = rc_flag ? (unsigned char *)dk_dev_alloc(ctx, n_reads + 16) : nullptr;
    uint64_t *k0 = (uint64_t *)dk_dev_alloc(ctx, 8 * W), *k1 = (uint64_t *)dk_dev_alloc(ctx, 8 * W);
    unsigned long long *ninv = (unsigned long long *)dk_dev_alloc(ctx, 64);
    int rc = DK_OK;
    uint32_t *head = nullptr, *uid = nullptr, *upos = nullptr, *keep = nullptr, *kidx = nullptr;
    auto cleanup = [&]() {
        dk_dev_release(ctx, b_d); dk_dev_release(ctx, off_d); dk_dev_release(ctx, woff_d); dk_dev_release(ctx, rc_d);
        dk_dev_release(ctx, k0); dk_dev_release(ctx, k1); dk_dev_release(ctx, ninv); dk_dev_release(ctx, head);
        dk_dev_release(ctx, uid); dk_dev_release(ctx, upos); dk_dev_release(ctx, keep); dk_dev_release(ctx, kidx);
    };
    if (!b_d || !off_d || !woff_d || (rc_flag && !rc_d) || !k0 || !k1 || !ninv) { cleanup(); dump_release(ctx, d); return DK_ERR_NOMEM; }
    std::vector<uint64_t> rel(n_reads + 1);
    for (uint64_t r = 0; r <= n_reads; r++) rel[r] = offsets[r] - offsets[0];
#define IG_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess && rc == DK_OK) rc = dk_fail(ctx, DK_ERR_CUDA, "count_kmers: %s: %s", #call, cudaGetErrorString(e__)); } while (0)
    IG_CUDA(cudaMemcpyAsync(b_d, bases + offsets[0], nb, cudaMemcpyHostToDevice, st));
    IG_CUDA(cudaMemcpyAsync(off_d, rel.data(), 8 * (n_reads + 1), cudaMemcpyHostToDevice, st));
    IG_CUDA(cudaMemcpyAsync(woff_d, woff.data(), 8 * (n_reads + 1), cudaMemcpyHostToDevice, st));
    if (rc_flag) IG_CUDA(cudaMemcpyAsync(rc_d, rc_flag, n_reads, cudaMemcpyHostToDevice, st));
    IG_CUDA(cudaMemsetAsync(ninv, 0, 64, st));
    if (rc == DK_OK) {
        k_ingest_emit<<<(unsigned)((n_reads + IG_T - 1) / IG_T), IG_T, 0, st>>>(b_d, off_d, woff_d, rc_d, n_reads, k, canonical, k0, ninv);
        ctx->launches++;
        IG_CUDA(cudaGetLastError());
    }
    uint64_t *sk = k0;
    uint32_t *sv = nullptr;
    if (rc == DK_OK) rc = dk_sort_pairs_u64(ctx, k0, nullptr, k1, nullptr, W, &sk, &sv);
    unsigned long long inv = 0;
    IG_CUDA(cudaMemcpyAsync(&inv, ninv, 8, cudaMemcpyDeviceToHost, st));
    IG_CUDA(cudaStreamSynchronize(st));                      // also keeps rel/woff alive until the copies are done
    const uint64_t V = W - inv;
    d->valid = V;
    uint32_t U = 0, K = 0;
    if (rc == DK_OK && V) {
        head = (uint32_t *)dk_dev_alloc(ctx, 4 * (V + 1)); uid = (uint32_t *)dk_dev_alloc(ctx, 4 * (V + 2));
        if (!head || !uid) rc = DK_ERR_NOMEM;
        if (rc == DK_OK) {
            k_ingest_heads<<<(unsigned)((V + 255) / 256), 256, 0, st>>>(sk, V, head); ctx->launches++;
            rc = dk_scan_excl_u32(ctx, head, uid, V);
        }
        IG_CUDA(cudaMemcpyAsync(&U, uid + V, 4, cudaMemcpyDeviceToHost, st));
        IG_CUDA(cudaStreamSynchronize(st));
        d->distinct = U;
    }
    if (rc == DK_OK && U) {
        upos = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)U + 2)); keep = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)U + 1));
        kidx = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)U + 2));
        if (!upos || !keep || !kidx) rc = DK_ERR_NOMEM;
        if (rc == DK_OK) {
            k_ingest_upos<<<(unsigned)((V + 255) / 256), 256, 0, st>>>(head, uid, V, U, upos); ctx->launches++;
            k_ingest_keep<<<(U + 255) / 256, 256, 0, st>>>(upos, U, min_count, keep); ctx->launches++;
            rc = dk_scan_excl_u32(ctx, keep, kidx, U);
        }
        IG_CUDA(cudaMemcpyAsync(&K, kidx + U, 4, cudaMemcpyDeviceToHost, st));
        IG_CUDA(cudaStreamSynchronize(st));
    }
    if (rc == DK_OK && K) {
        d->keys = (uint64_t *)dk_dev_alloc(ctx, 8 * ((size_t)K + 4));          // +4: dk_join_counts_dev stages whole 32-byte groups
        d->cnts = (int32_t *)dk_dev_alloc(ctx, 4 * ((size_t)K + 4));
        if (!d->keys || !d->cnts) rc = DK_ERR_NOMEM;
        if (rc == DK_OK) {
            k_ingest_write<<<(U + 255) / 256, 256, 0, st>>>(sk, upos, keep, kidx, U, d->keys, d->cnts); ctx->launches++;
            IG_CUDA(cudaGetLastError());
            IG_CUDA(cudaStreamSynchronize(st));
        }
        d->n = K;
    }
#undef IG_CUDA
    cleanup();
    if (rc != DK_OK) { dump_release(ctx, d); return rc; }
    *out = d;
    return DK_OK;
}

extern "C" uint64_t dk_dump_size(const dk_dump *d) { return d ? d->n : 0; }
extern "C" const uint64_t *dk_dump_keys_dev(const dk_dump *d) { return d ? d->keys : nullptr; }
extern "C" const int32_t *dk_dump_cnts_dev(const dk_dump *d) { return d ? d->cnts : nullptr; }
extern "C" void dk_dump_stats(const dk_dump *d, uint64_t *windows, uint64_t *valid, uint64_t *distinct)
{
    if (windows) *windows = d ? d->windows : 0;
    if (valid) *valid = d ? d->valid : 0;
    if (distinct) *distinct = d ? d->distinct : 0;
}
extern "C" int dk_dump_export(dk_ctx *ctx, const dk_dump *d, uint64_t *keys, int32_t *cnts)
{
    if (!ctx || !d) return dk_fail(ctx, DK_ERR_ARG, "dump_export: NULL argument");
    if (!d->n) return DK_OK;
    if (keys) DK_CUDA(ctx, cudaMemcpyAsync(keys, d->keys, 8 * d->n, cudaMemcpyDeviceToHost, ctx->stream));
    if (cnts) DK_CUDA(ctx, cudaMemcpyAsync(cnts, d->cnts, 4 * d->n, cudaMemcpyDeviceToHost, ctx->stream));
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DK_OK;
}
extern "C" void dk_dump_free(dk_ctx *ctx, dk_dump *d) { if (ctx) dump_release(ctx, d); }

// ---- samples larger than one call: count read batches with min_count = 1, then add the tables up ----------------
// one thread per row of the joined batch tables: total count, keep flag
__global__ void k_dump_rowsum(const int32_t *cnts, const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, int S, uint64_t R,
                              uint32_t min_count, uint32_t *total, uint32_t *keep)
{
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (lbase[mid] <= r) lo = mid; else hi = mid; }
    const int32_t *c = cnts + ((size_t)order[lo] * DK_CH + (size_t)(r - lbase[lo])) * (size_t)S;
    unsigned long long s = 0;
    for (int j = 0; j < S; j++) s += (unsigned long long)(uint32_t)c[j];
    const uint32_t t = s > 0x7fffffffull ? 0x7fffffffu : (uint32_t)s;
    total[r] = t;
    keep[r] = t >= min_count ? 1u : 0u;
}
__global__ void k_dump_compact(const uint64_t *keys, const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, uint64_t R,
                               const uint32_t *total, const uint32_t *keep, const uint32_t *kidx, uint64_t *okeys, int32_t *ocnts)
{
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R || !keep[r]) return;
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (lbase[mid] <= r) lo = mid; else hi = mid; }
    okeys[kidx[r]] = keys[(size_t)order[lo] * DK_CH + (size_t)(r - lbase[lo])];
    ocnts[kidx[r]] = (int32_t)total[r];
}

void dk_matrix_release(dk_ctx *ctx, dk_matrix *m);

// Sum of n_parts tables counted with min_count = 1 (read batches of one sample): the k-way merge is dk_join_counts_dev
// (every k-mer kept), the row sums are the sample's counts, -L is applied to the sums.
extern "C" int dk_dump_merge(dk_ctx *ctx, dk_dump *const *parts, int n_parts, uint32_t min_count, dk_dump **out)
{
    if (!ctx || !parts || !out || n_parts < 1) return dk_fail(ctx, DK_ERR_ARG, "dump_merge: bad argument");
    if (n_parts > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "dump_merge: at most %d tables per call", DK_MAX_SAMPLES);
    *out = nullptr;
    std::vector<const uint64_t *> kp(n_parts);
    std::vector<const int32_t *> cp(n_parts);
    std::vector<uint64_t> nn(n_parts);
    dk_dump *d = new dk_dump();
    for (int i = 0; i < n_parts; i++) {
        if (!parts[i]) { delete d; return dk_fail(ctx, DK_ERR_ARG, "dump_merge: NULL table %d", i); }
        kp[i] = parts[i]->keys; cp[i] = parts[i]->cnts; nn[i] = parts[i]->n;
        d->windows += parts[i]->windows; d->valid += parts[i]->valid;
    }
    dk_matrix *m = nullptr;
    int rc = dk_join_impl(ctx, n_parts, kp.data(), cp.data(), nn.data(), 1, 1, 0, &m);
    if (rc != DK_OK) { delete d; return rc; }
    const uint64_t R = m->R;
    d->distinct = R;
    cudaStream_t st = ctx->stream;
    uint32_t K = 0;
    uint32_t *total = nullptr, *keep = nullptr, *kidx = nullptr;
    if (R) {
        total = (uint32_t *)dk_dev_alloc(ctx, 4 * (R + 1)); keep = (uint32_t *)dk_dev_alloc(ctx, 4 * (R + 1));
        kidx = (uint32_t *)dk_dev_alloc(ctx, 4 * (R + 2));
        if (!total || !keep || !kidx) rc = DK_ERR_NOMEM;
        if (rc == DK_OK) {
            k_dump_rowsum<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks, m->S, R, min_count,
                                                                     total, keep);
            ctx->launches++;
            rc = dk_scan_excl_u32(ctx, keep, kidx, R);
        }
        if (rc == DK_OK) {
            cudaError_t e = cudaMemcpyAsync(&K, kidx + R, 4, cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) rc = dk_fail(ctx, DK_ERR_CUDA, "dump_merge: %s", cudaGetErrorString(e));
        }
        if (rc == DK_OK && K) {
            d->keys = (uint64_t *)dk_dev_alloc(ctx, 8 * ((size_t)K + 4));
            d->cnts = (int32_t *)dk_dev_alloc(ctx, 4 * ((size_t)K + 4));
            if (!d->keys || !d->cnts) rc = DK_ERR_NOMEM;
            if (rc == DK_OK) {
                k_dump_compact<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(m->keys, m->lbase, m->order, (uint32_t)m->n_chunks, R, total, keep,
                                                                          kidx, d->keys, d->cnts);
                ctx->launches++;
                cudaError_t e = cudaStreamSynchronize(st);
                if (e != cudaSuccess) rc = dk_fail(ctx, DK_ERR_CUDA, "dump_merge: %s", cudaGetErrorString(e));
            }
            d->n = K;
        }
    }
    dk_dev_release(ctx, total); dk_dev_release(ctx, keep); dk_dev_release(ctx, kidx);
    dk_matrix_release(ctx, m);
    if (rc != DK_OK) { dump_release(ctx, d); return rc; }
    *out = d;
    return DK_OK;
}
