// dk_sort.cu — stable LSD radix sort of (uint64 key, uint32 value) pairs, 8 bits per pass (vals may be NULL: keys only).
// Used for the BH candidate p-values (positive doubles order like their bit patterns)
// and for ordering survivor row indices. Passes whose digit is constant are skipped.
#include "dk_common.cuh"

#define RS_THREADS 256
#define RS_ITEMS 16
#define RS_TILE (RS_THREADS * RS_ITEMS)

// all 8 digit histograms in one read of the keys
__global__ void k_rs_hist8(const uint64_t *keys, size_t n, unsigned long long *hist /*[8][256]*/)
{
    __shared__ unsigned int sh[8 * 256];
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint64_t k = keys[i];
#pragma unroll
        for (int d = 0; d < 8; d++) atomicAdd(&sh[d * 256 + ((k >> (8 * d)) & 255u)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// per-tile digit counts, digit-major: cnt[digit*ntiles + tile]
__global__ void k_rs_tile_hist(const uint64_t *keys, size_t n, int shift, uint32_t ntiles, uint32_t *cnt)
{
    __shared__ unsigned int sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    size_t base = (size_t)blockIdx.x * RS_TILE;
    for (int i = 0; i < RS_ITEMS; i++) {
        size_t k = base + (size_t)i * RS_THREADS + threadIdx.x;
        if (k < n) atomicAdd(&sh[(keys[k] >> shift) & 255u], 1u);
    }
    __syncthreads();
    cnt[(size_t)threadIdx.x * ntiles + blockIdx.x] = sh[threadIdx.x];
}

// stable scatter: rounds of RS_THREADS consecutive elements; in-round rank by match_any
__global__ void k_rs_scatter(const uint64_t *keys, const uint32_t *vals, size_t n, int shift, uint32_t ntiles,
                             const uint32_t *off /*digit-major exclusive scan of cnt*/, uint64_t *okeys, uint32_t *ovals)
{
    __shared__ unsigned int run[256];                 // next output slot per digit for this tile
    __shared__ unsigned short wcnt[RS_THREADS / 32][256];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    run[threadIdx.x] = off[(size_t)threadIdx.x * ntiles + blockIdx.x];
    size_t base = (size_t)blockIdx.x * RS_TILE;
    for (int i = 0; i < RS_ITEMS; i++) {
        for (int q = w; q < RS_THREADS / 32; q += RS_THREADS / 32)
            for (int d = lane; d < 256; d += 32) wcnt[q][d] = 0;
        __syncthreads();
        size_t k = base + (size_t)i * RS_THREADS + threadIdx.x;
        bool valid = k < n;
        uint64_t key = valid ? keys[k] : 0;
        uint32_t val = (valid && vals) ? vals[k] : 0;
        unsigned dg = valid ? (unsigned)((key >> shift) & 255u) : 256u;   // 256 = "invalid" class
        unsigned peers = __match_any_sync(0xffffffffu, dg);
        unsigned rank = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank == 0) wcnt[w][dg] = (unsigned short)__popc(peers);
        __syncthreads();
        unsigned pos = 0;
        if (valid) {
            pos = run[dg] + rank;
            for (int q = 0; q < w; q++) pos += wcnt[q][dg];
        }
        __syncthreads();
        {   // thread d advances digit d's cursor by the round's total
            unsigned t = 0;
            for (int q = 0; q < RS_THREADS / 32; q++) t += wcnt[q][threadIdx.x];
            run[threadIdx.x] += t;
        }
        if (valid) { okeys[pos] = key; if (ovals) ovals[pos] = val; }
        __syncthreads();
    }
}

int dk_sort_pairs_u64(dk_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t *keys_tmp, uint32_t *vals_tmp, size_t n,
                      uint64_t **keys_sorted, uint32_t **vals_sorted)
{
    *keys_sorted = keys; *vals_sorted = vals;
    if (n < 2) return DK_OK;
    if (n >= 0xFFFFFFFFull) return dk_fail(ctx, DK_ERR_ARG, "sort: n too large");
    cudaStream_t st = ctx->stream;
    uint32_t ntiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
    unsigned long long *hist = (unsigned long long *)dk_dev_alloc(ctx, sizeof(unsigned long long) * 8 * 256);
    uint32_t *cnt = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (256 * (size_t)ntiles + 1));
    if (!hist || !cnt) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(hist, 0, sizeof(unsigned long long) * 8 * 256, st));
    int g = (int)((n + 4095) / 4096);
    if (g > ctx->sm_count * 8) g = ctx->sm_count * 8;
    k_rs_hist8<<<g, 256, 0, st>>>(keys, n, hist); DK_LAUNCHED(ctx);
    std::vector<unsigned long long> h(8 * 256);
    DK_CUDA(ctx, cudaMemcpyAsync(h.data(), hist, sizeof(unsigned long long) * 8 * 256, cudaMemcpyDeviceToHost, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    uint64_t *src_k = keys, *dst_k = keys_tmp;
    uint32_t *src_v = vals, *dst_v = vals_tmp;
    for (int d = 0; d < 8; d++) {
        bool constant = false;
        for (int b = 0; b < 256; b++) if (h[d * 256 + b] == n) constant = true;
        if (constant) continue;
        k_rs_tile_hist<<<ntiles, RS_THREADS, 0, st>>>(src_k, n, 8 * d, ntiles, cnt); DK_LAUNCHED(ctx);
        DK_TRY(dk_scan_excl_u32(ctx, cnt, cnt, 256 * (size_t)ntiles));
        k_rs_scatter<<<ntiles, RS_THREADS, 0, st>>>(src_k, src_v, n, 8 * d, ntiles, cnt, dst_k, dst_v); DK_LAUNCHED(ctx);
        uint64_t *tk = src_k; src_k = dst_k; dst_k = tk;
        uint32_t *tv = src_v; src_v = dst_v; dst_v = tv;
    }
    *keys_sorted = src_k; *vals_sorted = src_v;
    dk_dev_release(ctx, hist); dk_dev_release(ctx, cnt);
    return DK_OK;
}

// ---- key-only sort on caller-supplied scratch: no allocation, no host round trip (dk_bh_rank_slice_noalloc) -----
// exclusive scan of n <= a few 10^4 counters by ONE block (the allocating dk_scan_excl_u32 is not used here)
__global__ void k_rs_scan_small(uint32_t *v, uint32_t n)
{
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t b = 0; b < n; b += blockDim.x) {
        const uint32_t i = b + threadIdx.x;
        const uint32_t x = i < n ? v[i] : 0u;
        uint32_t incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) wsum[w] = incl;
        __syncthreads();
        uint32_t woff = 0, tot = 0;
        for (int q = 0; q < nw; q++) { const uint32_t c = wsum[q]; if (q < w) woff += c; tot += c; }
        const uint32_t base = carry;
        if (i < n) v[i] = base + woff + incl - x;
        __syncthreads();
        if (threadIdx.x == 0) carry = base + tot;
        __syncthreads();
    }
}
// all 8 digit passes (no constant-digit skip: that needs a host read-back); 8 is even, so the result lands in `keys`.
// cnt: 256 * ceil(n / RS_TILE) counters, fully rewritten by every pass (nothing has to start at zero).
int dk_sort_keys_u64_scratch(dk_ctx *ctx, uint64_t *keys, uint64_t *keys_tmp, size_t n, uint32_t *cnt)
{
    if (n < 2) return DK_OK;
    if (n >= 0xFFFFFFFFull) return dk_fail(ctx, DK_ERR_ARG, "sort: n too large");
    cudaStream_t st = ctx->stream;
    const uint32_t ntiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
    uint64_t *src = keys, *dst = keys_tmp;
    for (int d = 0; d < 8; d++) {
        k_rs_tile_hist<<<ntiles, RS_THREADS, 0, st>>>(src, n, 8 * d, ntiles, cnt); DK_LAUNCHED(ctx);
        k_rs_scan_small<<<1, 1024, 0, st>>>(cnt, 256u * ntiles); DK_LAUNCHED(ctx);
        k_rs_scatter<<<ntiles, RS_THREADS, 0, st>>>(src, nullptr, n, 8 * d, ntiles, cnt, dst, nullptr); DK_LAUNCHED(ctx);
        uint64_t *t = src; src = dst; dst = t;
    }
    return DK_OK;
}
