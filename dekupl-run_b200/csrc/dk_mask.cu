// dk_mask.cu — K7, the reference-transcript mask: a sorted anti-join that drops every
// matrix row whose k-mer occurs in the sorted dump of the reference transcripts
// (bin/diffFilter.pl:50-69, rule filter_transcript_counts Snakefile:634-644). TtestFilter
// reads the masked matrix (Snakefile:654) while computeNF keeps reading the raw one
// (Snakefile:366), so the mask produces a second matrix instead of editing the first.
//
// Both sides are ascending, so a block of MK_ROWS consecutive rows only needs the slice of
// the mask between its first and last key: two block-wide searches narrow the range, then
// every thread finishes with a few probes that stay in L1/L2. HBM traffic is one read of
// the matrix keys, one read of the mask, a 4-byte row map written and read back, and one
// read + one write of the surviving rows.
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

// chunk holding logical row `row`, searched inside [clo, chi] (ordered chunk positions)
__device__ __forceinline__ uint32_t mk_chunk_of(const uint64_t *lbase, uint32_t clo, uint32_t chi, uint64_t row)
{
    uint32_t lo = clo, hi = chi + 1;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (lbase[mid] <= row) lo = mid; else hi = mid;
    }
    return lo;
}

// pass 0, one thread per block of MK_ROWS rows: the chunk holding the block's first row and the position of its
// first k-mer in the mask (first entry >= it, and first entry > it). Done for all blocks at once, so the long
// dependent probe chains of these whole-array searches overlap instead of stalling every block of pass 1.
__global__ void k_mask_bounds(const uint64_t *keys, const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, uint64_t R,
                              uint32_t nb, const uint64_t *mask, uint64_t n_mask, uint32_t *cb, uint32_t *mlo, uint32_t *mup,
                              uint32_t *flags)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nb) return;
    if (b == nb) { cb[nb] = n_chunks - 1; mlo[nb] = (uint32_t)n_mask; mup[nb] = (uint32_t)n_mask; return; }
    const uint64_t r0 = (uint64_t)b * MK_ROWS;
    const uint32_t c = mk_chunk_of(lbase, 0, n_chunks - 1, r0);
    const uint64_t kf = keys[(size_t)order[c] * DK_CH + (size_t)(r0 - lbase[c])];
    if (r0 > 0 && keys[mk_phys_row(lbase, order, n_chunks, r0 - 1)] > kf) atomicOr(flags, 2u);   // order across the block boundary
    uint64_t lo = mk_lower(mask, 0, n_mask, kf), up = lo;
    while (up < n_mask && mask[up] == kf) up++;              // duplicates in the mask are rare and short
    cb[b] = c; mlo[b] = (uint32_t)lo; mup[b] = (uint32_t)up;
}

// pass 1 over a chunked matrix: src[r] = physical row of logical row r, or MK_DROP when the mask holds its k-mer;
// kept rows per block.
#define MK_DROP 0xFFFFFFFFu
__global__ void __launch_bounds__(MK_T) k_mask_rows(const uint64_t *keys, const uint64_t *lbase, const uint32_t *order,
                                                    uint64_t R, const uint64_t *mask, const uint32_t *cb, const uint32_t *mlo,
                                                    const uint32_t *mup, uint32_t *src, uint32_t *blk_kept, uint32_t *flags)
{
    __shared__ uint64_t sk[MK_ROWS];
    __shared__ uint32_t wsum[MK_T / 32];
    const uint64_t r0 = (uint64_t)blockIdx.x * MK_ROWS, r1 = r0 + MK_ROWS < R ? r0 + MK_ROWS : R;
    const uint32_t nrows = (uint32_t)(r1 - r0);
    bool bad = false;
    // every k-mer of this block is >= its first one and <= the next block's first one
    const uint32_t c0 = cb[blockIdx.x], c1 = cb[blockIdx.x + 1];
    const uint64_t *mk = mask + mlo[blockIdx.x];
    const uint32_t mhi = mup[blockIdx.x + 1], mn = mhi > mlo[blockIdx.x] ? mhi - mlo[blockIdx.x] : 0u;   // (0 on unsorted input)
    // thread t owns MK_V consecutive rows: one chunk search and one mask search, then both cursors only walk forward
    const uint32_t i0 = threadIdx.x * MK_V;
    uint32_t ph[MK_V];
    uint64_t kk[MK_V];
    if (i0 < nrows) {
        uint32_t c = mk_chunk_of(lbase, c0, c1, r0 + i0);
        uint64_t cend = lbase[c + 1], cbeg = lbase[c];
        uint32_t pbase = order[c] * (uint32_t)DK_CH;
#pragma unroll
        for (int v = 0; v < MK_V; v++) {
            const uint64_t r = r0 + i0 + v;
            ph[v] = 0; kk[v] = 0;
            if (i0 + v < nrows) {
                while (r >= cend) { c++; cbeg = cend; cend = lbase[c + 1]; pbase = order[c] * (uint32_t)DK_CH; }   // skips empty chunks
                ph[v] = pbase + (uint32_t)(r - cbeg);
                kk[v] = keys[ph[v]];
                sk[i0 + v] = kk[v];
            }
        }
    }
    __syncthreads();
    uint32_t kept = 0;
    if (i0 < nrows) {
        uint32_t lo = 0, hi = mn;
        while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (mk[mid] < kk[0]) lo = mid + 1; else hi = mid; }
        uint32_t p = lo, o[MK_V];
#pragma unroll
        for (int v = 0; v < MK_V; v++) {
            o[v] = MK_DROP;
            if (i0 + v < nrows) {
                const uint64_t k = kk[v];
                if (i0 + v + 1 < nrows && k > sk[i0 + v + 1]) bad = true;
                // the cursor of diffFilter.pl:59-66 only moves forward: gallop from where the previous row left it
                uint32_t step = 1;
                while (p + step < mn && mk[p + step - 1] < k) { p += step; step <<= 1; }
                uint32_t ph_ = p + step < mn ? p + step : mn;
                while (p < ph_) { const uint32_t mid = (p + ph_) >> 1; if (mk[mid] < k) p = mid + 1; else ph_ = mid; }
                const bool kp = !(p < mn && mk[p] == k);                     // diffFilter.pl:67 ($diff_tag ne $tag)
                o[v] = kp ? ph[v] : MK_DROP;
                kept += kp;
            }
        }
        if (i0 + MK_V <= nrows) *reinterpret_cast<uint4 *>(src + r0 + i0) = make_uint4(o[0], o[1], o[2], o[3]);
        else for (int v = 0; v < MK_V && i0 + v < nrows; v++) src[r0 + i0 + v] = o[v];
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

// pass 2: surviving rows of block b go to dense rows [blk_base[b], +kept) of the new matrix. Rows are copied as
// 16-byte words when S % 4 == 0 (W = 4, Sw = S/4), else as 4-byte words (W = 1, Sw = S); divSw = ceil(2^32 / Sw).
template <typename VEC>
__global__ void __launch_bounds__(MK_T) k_mask_copy(const uint64_t *keys, const VEC *cnts, uint64_t R, uint32_t Sw, uint32_t divSw,
                                                    const uint32_t *src, const uint32_t *blk_base, uint64_t *okeys, VEC *ocnts)
{
    __shared__ uint32_t srow[MK_ROWS];
    __shared__ uint32_t wsum[MK_T / 32];
    __shared__ uint32_t total;
    const uint64_t r0 = (uint64_t)blockIdx.x * MK_ROWS;
    // thread t owns rows r0 + t*MK_V .. +MK_V (consecutive, so the output stays in row order)
    uint32_t mine = 0, sp[MK_V];
    if (r0 + (uint64_t)threadIdx.x * MK_V + MK_V <= R) {
        const uint4 q = *reinterpret_cast<const uint4 *>(src + r0 + (uint64_t)threadIdx.x * MK_V);
        sp[0] = q.x; sp[1] = q.y; sp[2] = q.z; sp[3] = q.w;
    } else {
#pragma unroll
        for (int v = 0; v < MK_V; v++) { const uint64_t r = r0 + (uint64_t)threadIdx.x * MK_V + v; sp[v] = r < R ? src[r] : MK_DROP; }
    }
#pragma unroll
    for (int v = 0; v < MK_V; v++) mine += sp[v] != MK_DROP;
    uint32_t incl = mine;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
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
#pragma unroll
    for (int v = 0; v < MK_V; v++) if (sp[v] != MK_DROP) srow[pos++] = sp[v];
    __syncthreads();
    const uint32_t nk = total;
    const size_t base = blk_base[blockIdx.x];
    for (uint32_t i = threadIdx.x; i < nk; i += MK_T) okeys[base + i] = keys[srow[i]];
    const uint32_t words = nk * Sw;
    VEC *dst = ocnts + base * Sw;
    for (uint32_t i = threadIdx.x; i < words; i += MK_T) {
        const uint32_t r = Sw == 1 ? i : __umulhi(i, divSw), j = i - r * Sw;     // i < 2^16 * ... exact for i < 2^32 / Sw
        dst[i] = cnts[(size_t)srow[r] * Sw + j];
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
    if (n_mask >= 0xFFFFFFFFull) return dk_fail(ctx, DK_ERR_ARG, "matrix_mask: more than 2^32 mask k-mers; shard by prefix");
    *out = nullptr;
    cudaStream_t st = ctx->stream;
    uint64_t R = m->R;
    unsigned nb = (unsigned)((R + MK_ROWS - 1) / MK_ROWS);
    const uint64_t *mask_dev = nullptr;
    uint64_t *dm = nullptr;
    DK_TRY(mask_to_device(ctx, mask_keys, n_mask, mask_on_device, &mask_dev, &dm));
    uint32_t *src = (uint32_t *)dk_dev_alloc(ctx, 4 * (R + 4));
    uint32_t *blk = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)nb + 1)), *base = (uint32_t *)dk_dev_alloc(ctx, 4 * ((size_t)nb + 2));
    uint32_t *flags = (uint32_t *)dk_dev_alloc(ctx, 64);
    uint32_t *bnd = (uint32_t *)dk_dev_alloc(ctx, 4 * 3 * ((size_t)nb + 1));      // cb | mlo | mup, nb+1 entries each
    if (!src || !blk || !base || !flags || !bnd) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(flags, 0, 64, st));
    uint32_t kept = 0;
    int rc = DK_OK;
    if (R) {
        if (n_mask > 1) { k_mask_sorted<<<(unsigned)((n_mask + 255) / 256), 256, 0, st>>>(mask_dev, n_mask, flags, 1u); DK_LAUNCHED(ctx); }
        k_mask_bounds<<<(nb + 256) / 256, 256, 0, st>>>(m->keys, m->lbase, m->order, (uint32_t)m->n_chunks, R, nb, mask_dev, n_mask,
                                                        bnd, bnd + (nb + 1), bnd + 2 * ((size_t)nb + 1), flags);
        DK_LAUNCHED(ctx);
        k_mask_rows<<<nb, MK_T, 0, st>>>(m->keys, m->lbase, m->order, R, mask_dev, bnd, bnd + (nb + 1), bnd + 2 * ((size_t)nb + 1),
                                         src, blk, flags);
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
            if (m->S % 4 == 0) {
                const uint32_t Sw = (uint32_t)m->S / 4;
                k_mask_copy<int4><<<nb, MK_T, 0, st>>>(m->keys, reinterpret_cast<const int4 *>(m->cnts), R, Sw,
                                                       (uint32_t)((0x100000000ull + Sw - 1) / Sw), src, base, o->keys,
                                                       reinterpret_cast<int4 *>(o->cnts));
            } else {
                const uint32_t Sw = (uint32_t)m->S;
                k_mask_copy<int32_t><<<nb, MK_T, 0, st>>>(m->keys, m->cnts, R, Sw, (uint32_t)((0x100000000ull + Sw - 1) / Sw), src,
                                                          base, o->keys, o->cnts);
            }
            DK_LAUNCHED(ctx);
        }
        unsigned g = (unsigned)((o->n_chunks + 255) / 256);
        k_mask_identity<<<g ? g : 1, 256, 0, st>>>(o->order, o->lbase, o->chunk_rows, (uint32_t)o->n_chunks, o->R); DK_LAUNCHED(ctx);
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { dk_matrix_release(ctx, o); o = nullptr; rc = dk_fail(ctx, DK_ERR_CUDA, "matrix_mask: %s", cudaGetErrorString(e)); }
    }
    dk_dev_release(ctx, bnd); dk_dev_release(ctx, src); dk_dev_release(ctx, blk); dk_dev_release(ctx, base); dk_dev_release(ctx, flags); dk_dev_release(ctx, dm);
    if (rc != DK_OK) return rc;
    *out = o;
    if (n_removed) *n_removed = R - kept;
    return DK_OK;
}
