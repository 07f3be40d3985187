// dk_mask.cu — K7, the reference-transcript mask: a sorted anti-join that drops every
// matrix row whose k-mer occurs in the sorted dump of the reference transcripts
// (bin/diffFilter.pl:50-69, rule filter_transcript_counts Snakefile:634-644). TtestFilter
// reads the masked matrix (Snakefile:654) while computeNF keeps reading the raw one
// (Snakefile:366), so the mask produces a second matrix instead of editing the first.
//
// Both sides are ascending, so a block of MK_ROWS consecutive rows only needs the slice of
// the mask between its first and last key: two block-wide searches narrow the range, then
// every thread finishes with a few probes that stay in L1/L2. HBM traffic is one read of
// the matrix keys, one read of the mask, and one read + one write of the surviving rows.
#include "dk_common.cuh"

void dk_matrix_release(dk_ctx *ctx, dk_matrix *m);

#define MK_T 256
#define MK_V 4
#define MK_ROWS (MK_T * MK_V)

__device__ __forceinline__ size_t mk_phys_row(const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, uint64_t row)
{
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (lbase[mid] <= row) lo = mid; else hi = mid;
    }
    return (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
}

// first index in [lo,hi) with a[i] >= k
__device__ __forceinline__ uint64_t mk_lower(const uint64_t *a, uint64_t lo, uint64_t hi, uint64_t k)
{
    while (lo < hi) {
        uint64_t mid = lo + ((hi - lo) >> 1);
        if (a[mid] < k) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// flags[0] |= 1 when a[] is not ascending (duplicates allowed)
__global__ void k_mask_sorted(const uint64_t *a, uint64_t n, uint32_t *flags, uint32_t bit)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = i + 1 < n && a[i] > a[i + 1];
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flags, bit);
}

// keep flag per row of a dense ascending key list (the CLI keeps the text lines on the host)
__global__ void __launch_bounds__(MK_T) k_mask_keys(const uint64_t *keys, uint64_t n, const uint64_t *mask, uint64_t n_mask,
                                                    uint8_t *keep, uint32_t *flags)
{
    __shared__ uint64_t rng[2];
    uint64_t r0 = (uint64_t)blockIdx.x * MK_ROWS, r1 = r0 + MK_ROWS < n ? r0 + MK_ROWS : n;
    if (threadIdx.x == 0) {
        uint64_t kl = keys[r1 - 1], lo = mk_lower(mask, 0, n_mask, keys[r0]);
        rng[0] = lo;
        rng[1] = kl == ~0ull ? n_mask : mk_lower(mask, lo, n_mask, kl + 1);
    }
    __syncthreads();
    uint64_t mlo = rng[0], mhi = rng[1];
    bool bad = false;
    for (int v = 0; v < MK_V; v++) {
        uint64_t r = r0 + (uint64_t)v * MK_T + threadIdx.x;
        if (r >= r1) break;
        uint64_t k = keys[r];
        if (r + 1 < n && k > keys[r + 1]) bad = true;
        uint64_t p = mk_lower(mask, mlo, mhi, k);
        keep[r] = !(p < mhi && mask[p] == k);                     // diffFilter.pl:67 ($diff_tag ne $tag)
    }
    if (bad) atomicOr(flags, 2u);
}

// pass 1 over a chunked matrix: keep flags + kept rows per block
__global__ void __launch_bounds__(MK_T) k_mask_rows(const uint64_t *keys, const uint64_t *lbase, const uint32_t *order,
                                                    uint32_t n_chunks, uint64_t R, const uint64_t *mask, uint64_t n_mask,
                                                    uint8_t *keep, uint32_t *blk_kept, uint32_t *flags)
{
    __shared__ uint64_t rng[2];
    __shared__ uint32_t wsum[MK_T / 32];
    uint64_t r0 = (uint64_t)blockIdx.x * MK_ROWS, r1 = r0 + MK_ROWS < R ? r0 + MK_ROWS : R;
    if (threadIdx.x == 0) {
        uint64_t kf = keys[mk_phys_row(lbase, order, n_chunks, r0)], kl = keys[mk_phys_row(lbase, order, n_chunks, r1 - 1)];
        uint64_t lo = mk_lower(mask, 0, n_mask, kf);
        rng[0] = lo;
        rng[1] = kl == ~0ull ? n_mask : mk_lower(mask, lo, n_mask, kl + 1);
    }
    __syncthreads();
    uint64_t mlo = rng[0], mhi = rng[1];
    uint32_t kept = 0;
    bool bad = false;
    for (int v = 0; v < MK_V; v++) {
        uint64_t r = r0 + (uint64_t)v * MK_T + threadIdx.x;
        if (r >= r1) break;
        uint64_t k = keys[mk_phys_row(lbase, order, n_chunks, r)];
        if (r + 1 < R && k > keys[mk_phys_row(lbase, order, n_chunks, r + 1)]) bad = true;
        uint64_t p = mk_lower(mask, mlo, mhi, k);
        uint8_t kp = !(p < mhi && mask[p] == k);
        keep[r] = kp;
        kept += kp;
    }
    if (bad) atomicOr(flags, 2u);
    for (int o = 16; o; o >>= 1) kept += __shfl_down_sync(0xffffffffu, kept, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = kept;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < MK_T / 32; w++) t += wsum[w];
        blk_kept[blockIdx.x] = t;
    }
}

// pass 2: surviving rows of block b go to dense rows [blk_base[b], +kept) of the new matrix
__global__ void __launch_bounds__(MK_T) k_mask_copy(const uint64_t *keys, const int32_t *cnts, const uint64_t *lbase,
                                                    const uint32_t *order, uint32_t n_chunks, uint64_t R, int S, int has_keys,
                                                    const uint8_t *keep, const uint32_t *blk_base, uint64_t *okeys, int32_t *ocnts)
{
    __shared__ uint32_t src[MK_ROWS];            // physical row (low 32 bits are enough: < 2^32 rows per GPU)
    __shared__ uint32_t wsum[MK_T / 32];
    __shared__ uint32_t total;
    uint64_t r0 = (uint64_t)blockIdx.x * MK_ROWS;
    // thread t owns rows r0 + t*MK_V .. +MK_V (consecutive, so the output stays in row order)
    uint32_t mine = 0;
    uint8_t kp[MK_V];
    for (int v = 0; v < MK_V; v++) {
        uint64_t r = r0 + (uint64_t)threadIdx.x * MK_V + v;
        kp[v] = r < R ? keep[r] : 0;
        mine += kp[v];
    }
    uint32_t incl = mine;
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < MK_T / 32; i++) { uint32_t x = wsum[i]; wsum[i] = t; t += x; }
        total = t;
    }
    __syncthreads();
    uint32_t pos = wsum[w] + incl - mine;
    for (int v = 0; v < MK_V; v++)
        if (kp[v]) src[pos++] = (uint32_t)mk_phys_row(lbase, order, n_chunks, r0 + (uint64_t)threadIdx.x * MK_V + v);
    __syncthreads();
    uint32_t nk = total;
    size_t base = blk_base[blockIdx.x];
    if (has_keys) for (uint32_t i = threadIdx.x; i < nk; i += MK_T) okeys[base + i] = keys[src[i]];
    uint32_t words = nk * (uint32_t)S;
    for (uint32_t i = threadIdx.x; i < words; i += MK_T) {
        uint32_t r = i / (uint32_t)S, j = i - r * (uint32_t)S;
        ocnts[base * S + i] = cnts[(size_t)src[r] * S + j];
    }
}

__global__ void k_mask_identity(uint32_t *order, uint64_t *lbase, uint32_t *chunk_rows, uint32_t n_chunks, uint64_t R)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_chunks) {
        order[i] = i;
        lbase[i] = (uint64_t)i * DK_CH;
        uint64_t left = R - (uint64_t)i * DK_CH;
        chunk_rows[i] = left < DK_CH ? (uint32_t)left : DK_CH;
    }
    if (i == 0) lbase[n_chunks] = R;
}

static int mask_to_device(dk_ctx *ctx, const uint64_t *mask_keys, uint64_t n_mask, int on_device, const uint64_t **dev, uint64_t **owned)
{
    *owned = nullptr;
    *dev = mask_keys;
    if (on_device || !n_mask) return DK_OK;
    uint64_t *d = (uint64_t *)dk_dev_alloc(ctx, 8 * n_mask);
    if (!d) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemcpyAsync(d, mask_keys, 8 * n_mask, cudaMemcpyHostToDevice, ctx->stream));
    *dev = d; *owned = d;
    return DK_OK;
}

static int mask_check_flags(dk_ctx *ctx, uint32_t *flags_dev, const char *who)
{
    uint32_t f = 0;
    DK_CUDA(ctx, cudaMemcpyAsync(&f, flags_dev, 4, cudaMemcpyDeviceToHost, ctx->stream));
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (f & 1u) return dk_fail(ctx, DK_ERR_ARG, "%s: the mask dump is not sorted (diffFilter.pl needs `sort -k1` output)", who);
    if (f & 2u) return dk_fail(ctx, DK_ERR_ARG, "%s: the count table is not sorted by k-mer", who);
    return DK_OK;
}

// keep[i] = 1 when keys[i] is absent from mask_keys. Host buffers; used by the diffFilter CLI,
// which keeps the text lines and prints the survivors byte for byte.
extern "C" int dk_keys_antijoin(dk_ctx *ctx, const uint64_t *keys, uint64_t n, const uint64_t *mask_keys, uint64_t n_mask,
                                uint8_t *keep)
{
    if (!ctx || (n && (!keys || !keep)) || (n_mask && !mask_keys)) return dk_fail(ctx, DK_ERR_ARG, "keys_antijoin: NULL argument");
    if (!n) return DK_OK;
    cudaStream_t st = ctx->stream;
    uint64_t *dk = (uint64_t *)dk_dev_alloc(ctx, 8 * n), *dm = nullptr;
    uint8_t *dkeep = (uint8_t *)dk_dev_alloc(ctx, n);
    uint32_t *flags = (uint32_t *)dk_dev_alloc(ctx, 64);
    const uint64_t *mask_dev = nullptr;
    if (!dk || !dkeep || !flags) return DK_ERR_NOMEM;
    DK_TRY(mask_to_device(ctx, mask_keys, n_mask, 0, &mask_dev, &dm));
    DK_CUDA(ctx, cudaMemcpyAsync(dk, keys, 8 * n, cudaMemcpyHostToDevice, st));
    DK_CUDA(ctx, cudaMemsetAsync(flags, 0, 64, st));
    if (n_mask > 1) { k_mask_sorted<<<(unsigned)((n_mask + 255) / 256), 256, 0, st>>>(mask_dev, n_mask, flags, 1u); DK_LAUNCHED(ctx); }
    k_mask_keys<<<(unsigned)((n + MK_ROWS - 1) / MK_ROWS), MK_T, 0, st>>>(dk, n, mask_dev, n_mask, dkeep, flags); DK_LAUNCHED(ctx);
    int rc = mask_check_flags(ctx, flags, "keys_antijoin");
    if (rc == DK_OK) {
        cudaError_t e = cudaMemcpyAsync(keep, dkeep, n, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = dk_fail(ctx, DK_ERR_CUDA, "keys_antijoin: %s", cudaGetErrorString(e));
    }
    dk_dev_release(ctx, dk); dk_dev_release(ctx, dkeep); dk_dev_release(ctx, flags); dk_dev_release(ctx, dm);
    return rc;
}

// New matrix = rows of m whose k-mer is not in mask_keys (ascending, duplicates allowed). m is left untouched.
extern "C" int dk_matrix_mask(dk_ctx *ctx, const dk_matrix *m, const uint64_t *mask_keys, uint64_t n_mask, int mask_on_device,
                              dk_matrix **out, uint64_t *n_removed)
{
    if (!ctx || !m || !out || (n_mask && !mask_keys)) return dk_fail(ctx, DK_ERR_ARG, "matrix_mask: NULL argument");
    if (!m->has_keys) return dk_fail(ctx, DK_ERR_ARG, "matrix_mask: matrix has no k-mers");
    *out = nullptr;
    cudaStream_t st = ctx->stream;
    uint64_t R = m->R;
    unsigned nb = (unsigned)((R + MK_ROWS - 1) / MK_ROWS);
    const uint64_t *mask_dev = nullptr;
    uint64_t *dm = nullptr;
    DK_TRY(mask_to_device(ctx, mask_keys, n_mask, mask_on_device, &mask_dev, &dm));
    uint8_t *keep = (uint8_t *)dk_dev_alloc(ctx, R + 1);
    uint32_t *blk = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)nb + 1)), *base = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)nb + 2));
    uint32_t *flags = (uint32_t *)dk_dev_alloc(ctx, 64);
    if (!keep || !blk || !base || !flags) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(flags, 0, 64, st));
    uint32_t kept = 0;
    int rc = DK_OK;
    if (R) {
        if (n_mask > 1) { k_mask_sorted<<<(unsigned)((n_mask + 255) / 256), 256, 0, st>>>(mask_dev, n_mask, flags, 1u); DK_LAUNCHED(ctx); }
        k_mask_rows<<<nb, MK_T, 0, st>>>(m->keys, m->lbase, m->order, (uint32_t)m->n_chunks, R, mask_dev, n_mask, keep, blk, flags);
        DK_LAUNCHED(ctx);
        DK_TRY(dk_scan_excl_u32(ctx, blk, base, nb));
        DK_CUDA(ctx, cudaMemcpyAsync(&kept, base + nb, 4, cudaMemcpyDeviceToHost, st));
        rc = mask_check_flags(ctx, flags, "matrix_mask");
    }
    dk_matrix *o = nullptr;
    if (rc == DK_OK) {
        o = new dk_matrix();
        o->S = m->S; o->R = kept; o->U = m->U; o->has_keys = true;
        o->n_chunks = ((uint64_t)kept + DK_CH - 1) / DK_CH;
        o->cap_chunks = o->n_chunks + 1;
        o->keys = (uint64_t *)dk_dev_alloc(ctx, 8 * o->cap_chunks * DK_CH);
        o->cnts = (int32_t *)dk_dev_alloc(ctx, 4 * o->cap_chunks * DK_CH * o->S);
        o->chunk_rows = (uint32_t *)dk_dev_alloc(ctx, 4 * o->cap_chunks);
        o->order = (uint32_t *)dk_dev_alloc(ctx, 4 * (o->n_chunks + 2));
        o->lbase = (uint64_t *)dk_dev_alloc(ctx, 8 * (o->n_chunks + 3));
        if (!o->keys || !o->cnts || !o->chunk_rows || !o->order || !o->lbase) { dk_matrix_release(ctx, o); o = nullptr; rc = DK_ERR_NOMEM; }
    }
    if (rc == DK_OK) {
        if (R) {
            k_mask_copy<<<nb, MK_T, 0, st>>>(m->keys, m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks, R, m->S, 1, keep, base,
                                             o->keys, o->cnts);
            DK_LAUNCHED(ctx);
        }
        unsigned g = (unsigned)((o->n_chunks + 255) / 256);
        k_mask_identity<<<g ? g : 1, 256, 0, st>>>(o->order, o->lbase, o->chunk_rows, (uint32_t)o->n_chunks, o->R); DK_LAUNCHED(ctx);
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { dk_matrix_release(ctx, o); o = nullptr; rc = dk_fail(ctx, DK_ERR_CUDA, "matrix_mask: %s", cudaGetErrorString(e)); }
    }
    dk_dev_release(ctx, keep); dk_dev_release(ctx, blk); dk_dev_release(ctx, base); dk_dev_release(ctx, flags); dk_dev_release(ctx, dm);
    if (rc != DK_OK) return rc;
    *out = o;
    if (n_removed) *n_removed = R - kept;
    return DK_OK;
}
