// dk_format.cu — K8, the text side of joinCounts on the device: rows of the count matrix printed exactly as
// share/joinCounts/joinCounts.c:105-109 does ("%s" k-mer, then "\t%d" per sample, then "\n"), so the CLI only has
// to write() the bytes. SURVEY §8(f) #2: formatting 10^8 rows on one host thread costs more than the whole GPU path.
//
// Two passes over a block of rows: (1) byte length of every row (k + 1 + S tabs + decimal digits), reduced per CTA;
// a scan of the CTA totals places every CTA's text; (2) every thread prints its row into a shared-memory tile at its
// offset inside the CTA's text, and the CTA copies the tile out with coalesced 16-byte stores. Rows whose text does
// not fit the tile (very wide matrices) are written straight to global memory.
#include "dk_common.cuh"

#define FM_T 128                       // rows per CTA
#define FM_TILE (40 * 1024)            // shared-memory text tile per CTA

__device__ __forceinline__ size_t fm_phys_row(const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, uint64_t row)
{
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (lbase[mid] <= row) lo = mid; else hi = mid;
    }
    return (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
}

__device__ __forceinline__ uint32_t fm_digits(uint32_t v)      // decimal digits of v (v = 0 -> 1)
{
    return 1u + (v >= 10u) + (v >= 100u) + (v >= 1000u) + (v >= 10000u) + (v >= 100000u) + (v >= 1000000u) + (v >= 10000000u) +
           (v >= 100000000u) + (v >= 1000000000u);
}
__device__ __forceinline__ uint32_t fm_int_len(int32_t c)      // strlen of "%d"
{
    const uint32_t a = c < 0 ? 0u - (uint32_t)c : (uint32_t)c;
    return fm_digits(a) + (c < 0);
}
__device__ __forceinline__ uint32_t fm_row_len(const int32_t *c, int S, int k)
{
    uint32_t len = (uint32_t)k + (uint32_t)S + 1u;             // k-mer, one tab per sample, newline
    for (int s = 0; s < S; s++) len += fm_int_len(c[s]);
    return len;
}
// prints the row at p, returns the end
__device__ __forceinline__ char *fm_print_row(char *p, uint64_t key, const int32_t *c, int S, int k)
{
    for (int i = k - 1; i >= 0; i--) *p++ = "ACGT"[(key >> (2 * i)) & 3u];
    for (int s = 0; s < S; s++) {
        *p++ = '\t';
        const int32_t v = c[s];
        uint32_t a = v < 0 ? 0u - (uint32_t)v : (uint32_t)v;
        if (v < 0) *p++ = '-';
        const uint32_t nd = fm_digits(a);
        for (uint32_t d = nd; d-- > 0;) { const uint32_t q = a / 10u; p[d] = (char)('0' + (a - q * 10u)); a = q; }
        p += nd;
    }
    *p++ = '\n';
    return p;
}

__global__ void __launch_bounds__(FM_T) k_format_len(const int32_t *cnts, const uint64_t *lbase, const uint32_t *order,
                                                     uint32_t n_chunks, int S, int k, uint64_t row0, uint64_t nrows,
                                                     uint64_t *blk_bytes)
{
    __shared__ uint32_t wsum[FM_T / 32];
    const uint64_t r = (uint64_t)blockIdx.x * FM_T + threadIdx.x;
    uint32_t len = 0;
    if (r < nrows) len = fm_row_len(cnts + fm_phys_row(lbase, order, n_chunks, row0 + r) * (size_t)S, S, k);
    for (int o = 16; o; o >>= 1) len += __shfl_down_sync(0xffffffffu, len, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = len;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t t = 0;
        for (int w = 0; w < FM_T / 32; w++) t += wsum[w];
        blk_bytes[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(FM_T) k_format_rows(const uint64_t *keys, const int32_t *cnts, const uint64_t *lbase,
                                                      const uint32_t *order, uint32_t n_chunks, int S, int k, uint64_t row0,
                                                      uint64_t nrows, const uint64_t *blk_off, char *out)
{
    __shared__ __align__(16) char tile[FM_TILE];
    __shared__ uint32_t wsum[FM_T / 32];
    __shared__ uint32_t total;
    const uint64_t r = (uint64_t)blockIdx.x * FM_T + threadIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    size_t ph = 0;
    uint32_t len = 0;
    if (r < nrows) {
        ph = fm_phys_row(lbase, order, n_chunks, row0 + r);
        len = fm_row_len(cnts + ph * (size_t)S, S, k);
    }
    uint32_t incl = len;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < FM_T / 32; i++) { const uint32_t x = wsum[i]; wsum[i] = t; t += x; }
        total = t;
    }
    __syncthreads();
    const uint32_t off = wsum[w] + incl - len, tot = total;
    char *dst = out + blk_off[blockIdx.x];
    if (tot + 16u <= FM_TILE) {
        // the text sits in the tile at the destination's misalignment, so 16-byte groups line up on both sides
        const uint32_t pad = (uint32_t)((uintptr_t)dst & 15u);
        char *t0 = tile + pad;
        if (r < nrows) fm_print_row(t0 + off, keys[ph], cnts + ph * (size_t)S, S, k);
        __syncthreads();
        const uint32_t mis = (16u - pad) & 15u, head = mis < tot ? mis : tot;
        for (uint32_t i = threadIdx.x; i < head; i += FM_T) dst[i] = t0[i];
        const uint32_t body = (tot - head) >> 4;
        const uint4 *s4 = reinterpret_cast<const uint4 *>(t0 + head);
        uint4 *d4 = reinterpret_cast<uint4 *>(dst + head);
        for (uint32_t i = threadIdx.x; i < body; i += FM_T) d4[i] = s4[i];
        for (uint32_t i = head + 16u * body + threadIdx.x; i < tot; i += FM_T) dst[i] = t0[i];
    } else if (r < nrows) {
        fm_print_row(dst + off, keys[ph], cnts + ph * (size_t)S, S, k);
    }
}

// Text of rows [row0, row0+nrows) exactly as joinCounts prints them; k = k-mer length. The bytes land in the host
// buffer `out` (capacity cap); *len receives their number. When cap is too small nothing is copied, *len holds the
// size needed and DK_ERR_ARG is returned.
extern "C" int dk_matrix_format_rows(dk_ctx *ctx, const dk_matrix *m, uint64_t row0, uint64_t nrows, int k, char *out,
                                     uint64_t cap, uint64_t *len)
{
    if (!ctx || !m || !len || (cap && !out)) return dk_fail(ctx, DK_ERR_ARG, "format_rows: NULL argument");
    if (!m->has_keys) return dk_fail(ctx, DK_ERR_ARG, "format_rows: matrix has no k-mers");
    if (k < 1 || k > 31) return dk_fail(ctx, DK_ERR_ARG, "format_rows: k=%d outside [1,31]", k);
    if (row0 + nrows > m->R) return dk_fail(ctx, DK_ERR_ARG, "format_rows: rows [%llu,+%llu) beyond R=%llu", (unsigned long long)row0,
                                            (unsigned long long)nrows, (unsigned long long)m->R);
    *len = 0;
    if (!nrows) return DK_OK;
    cudaStream_t st = ctx->stream;
    const unsigned nb = (unsigned)((nrows + FM_T - 1) / FM_T);
    uint64_t *bytes = (uint64_t *)dk_dev_alloc(ctx, 8 * ((size_t)nb + 2)), *off = (uint64_t *)dk_dev_alloc(ctx, 8 * ((size_t)nb + 2));
    if (!bytes || !off) return DK_ERR_NOMEM;
    k_format_len<<<nb, FM_T, 0, st>>>(m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks, m->S, k, row0, nrows, bytes); DK_LAUNCHED(ctx);
    DK_TRY(dk_scan_excl_u64(ctx, bytes, off, nb));
    uint64_t total = 0;
    DK_CUDA(ctx, cudaMemcpyAsync(&total, off + nb, 8, cudaMemcpyDeviceToHost, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    *len = total;
    int rc = DK_OK;
    if (total > cap) rc = dk_fail(ctx, DK_ERR_ARG, "format_rows: %llu bytes needed, buffer holds %llu", (unsigned long long)total,
                                  (unsigned long long)cap);
    char *text = rc == DK_OK ? (char *)dk_dev_alloc(ctx, total + 16) : nullptr;
    if (rc == DK_OK && !text) rc = DK_ERR_NOMEM;
    if (rc == DK_OK) {
        k_format_rows<<<nb, FM_T, 0, st>>>(m->keys, m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks, m->S, k, row0, nrows, off, text);
        ctx->launches++;
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpyAsync(out, text, total, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = dk_fail(ctx, DK_ERR_CUDA, "format_rows: %s", cudaGetErrorString(e));
    }
    dk_dev_release(ctx, text); dk_dev_release(ctx, bytes); dk_dev_release(ctx, off);
    return rc;
}
