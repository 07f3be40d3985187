// dk_nf.cu — computeNF on the GPU (reference: share/computeNF/computeNF.c).
//  k_nf_ratios : per row, m = (sum_j log c_j)/S in sample order (computeNF.c:18-25);
//                rows with every c_j > 0 push log c_j - m for every sample (:122-130).
//                Values are stored as order-preserving u64 images, column-major
//                [S][cap], compacted with warp-aggregated atomics (order inside a
//                step is irrelevant to a median).
//  radix select: exact element [size/2] of the ascending order (computeNF.c:38-47),
//                8 rounds of 8 bits: k_nf_hist (shared-memory histograms) + k_nf_pick.
//  DK_NF_REFERENCE replays the step / early-stop logic of computeNF.c:138-168.
#include <algorithm>
#include "dk_common.cuh"

__device__ __forceinline__ uint64_t f64_image(double v)
{
    uint64_t b = (uint64_t)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double image_f64(uint64_t i)
{
    uint64_t b = (i >> 63) ? (i & 0x7FFFFFFFFFFFFFFFull) : ~i;
    return __longlong_as_double((long long)b);
}

// logical row -> physical row (binary search in lbase[0..n_chunks])
__device__ __forceinline__ size_t dk_phys_row(const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, uint64_t row)
{
    uint32_t lo = 0, hi = n_chunks;       // find last oc with lbase[oc] <= row
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (lbase[mid] <= row) lo = mid; else hi = mid;
    }
    return (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
}

#define NF_MAXS_LOCAL 64
__global__ void k_nf_ratios(const int32_t *cnts, const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, int S,
                            uint64_t row0, uint64_t row1, uint64_t *vals, uint64_t cap, unsigned long long *npos,
                            int count_only)
{
    uint64_t row = row0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = row < row1;
    bool pos = false;
    const int32_t *c = nullptr;
    double m = 0;
    if (valid) {
        c = cnts + dk_phys_row(lbase, order, n_chunks, row) * (size_t)S;
        pos = true;
        for (int j = 0; j < S; j++) pos = pos && (c[j] > 0);
        if (pos && !count_only) {
            double sum = 0;
            for (int j = 0; j < S; j++) sum += log((double)c[j]);     // computeNF.c:21-23 (sequential)
            m = sum / (double)S;                                      // :24
        }
    }
    unsigned b = __ballot_sync(0xffffffffu, pos);
    if (b == 0) return;
    int lane = threadIdx.x & 31;
    unsigned long long base = 0;
    if (lane == (__ffs(b) - 1)) base = atomicAdd(npos, (unsigned long long)__popc(b));
    base = __shfl_sync(0xffffffffu, base, __ffs(b) - 1);
    if (pos && !count_only) {
        uint64_t idx = base + __popc(b & ((1u << lane) - 1u));
        if (idx < cap)
            for (int j = 0; j < S; j++) vals[(size_t)j * cap + idx] = f64_image(log((double)c[j]) - m);   // :127
    }
}

// hist[s][256] over entries [0,n) of sample s whose bits above `shift+8` equal prefix[s]
__global__ void k_nf_hist(const uint64_t *vals, uint64_t cap, uint64_t n, int shift, const uint64_t *prefix,
                          unsigned long long *hist)
{
    __shared__ unsigned int sh[256];
    int s = blockIdx.y;
    sh[threadIdx.x] = 0;
    __syncthreads();
    uint64_t pre = prefix[s];
    const uint64_t *v = vals + (size_t)s * cap;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t x = v[i];
        bool match = (shift + 8 >= 64) ? true : ((x >> (shift + 8)) == (pre >> (shift + 8)));
        if (match) atomicAdd(&sh[(x >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(&hist[(size_t)s * 256 + threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}
// per sample: find the digit holding rank k[s]; fold it into prefix[s]; k[s] -= entries below it
// one 256-thread block per sample (launch <<<S, 256>>>): thread d owns bin d. The digit is the first d < 255 whose inclusive
// bin sum exceeds k (else 255) — the sequential walk this replaces took ~25 us per call on one thread per sample.
__global__ void __launch_bounds__(256) k_nf_pick(unsigned long long *hist, int S, int shift, uint64_t *prefix, uint64_t *k)
{
    __shared__ unsigned long long wsum[8];
    const int s = blockIdx.x, d = threadIdx.x, lane = d & 31, w = d >> 5;
    if (s >= S) return;
    unsigned long long *h = hist + (size_t)s * 256;
    const uint64_t kk = k[s];
    const unsigned long long v = h[d];
    h[d] = 0;
    unsigned long long x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    unsigned long long off = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) if (i < w) off += wsum[i];
    const unsigned long long incl = off + x;
    const int dstar = __syncthreads_count(d < 255 && incl <= kk);      // bins walked past by the sequential rule
    if (d == dstar) {
        prefix[s] |= ((uint64_t)dstar) << shift;
        k[s] = kk - (incl - v);
    }
}
__global__ void k_nf_finish(const uint64_t *prefix, int S, double *nf, double *ratio)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    double v = image_f64(prefix[s]);
    ratio[s] = v;
    nf[s] = exp(v);                                                    // computeNF.c:44
}
__global__ void k_nf_setk(uint64_t *prefix, uint64_t *k, int S, uint64_t kval)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < S) { prefix[s] = 0; k[s] = kval; }
}

struct NfWork {
    uint64_t *vals = nullptr; uint64_t cap = 0;
    unsigned long long *npos = nullptr, *hist = nullptr;
    uint64_t *prefix = nullptr, *k = nullptr;
    double *nf = nullptr, *ratio = nullptr;
};

// exact per-sample element [n/2] of the first n stored values -> nf_host[S]
static int nf_select(dk_ctx *ctx, NfWork &w, int S, uint64_t n, double *nf_host)
{
    cudaStream_t st = ctx->stream;
    k_nf_setk<<<(S + 63) / 64, 64, 0, st>>>(w.prefix, w.k, S, n / 2); DK_LAUNCHED(ctx);
    int gx = (int)((n + 256 * 16 - 1) / (256 * 16));
    if (gx < 1) gx = 1;
    if (gx > ctx->sm_count * 8) gx = ctx->sm_count * 8;
    for (int shift = 56; shift >= 0; shift -= 8) {
        k_nf_hist<<<dim3(gx, S), 256, 0, st>>>(w.vals, w.cap, n, shift, w.prefix, w.hist); DK_LAUNCHED(ctx);
        k_nf_pick<<<S, 256, 0, st>>>(w.hist, S, shift, w.prefix, w.k); DK_LAUNCHED(ctx);
    }
    k_nf_finish<<<(S + 63) / 64, 64, 0, st>>>(w.prefix, S, w.nf, w.ratio); DK_LAUNCHED(ctx);
    DK_CUDA(ctx, cudaMemcpyAsync(nf_host, w.nf, sizeof(double) * S, cudaMemcpyDeviceToHost, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    return DK_OK;
}

static int nf_grow(dk_ctx *ctx, NfWork &w, int S, uint64_t need, uint64_t used)
{
    if (need <= w.cap) return DK_OK;
    uint64_t ncap = w.cap ? w.cap : 1;
    while (ncap < need) ncap *= 2;
    uint64_t *nv = (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * ncap * S);
    if (!nv) return DK_ERR_NOMEM;
    if (w.vals && used)
        DK_CUDA(ctx, cudaMemcpy2DAsync(nv, ncap * 8, w.vals, w.cap * 8, used * 8, S, cudaMemcpyDeviceToDevice, ctx->stream));
    dk_dev_release(ctx, w.vals);
    w.vals = nv; w.cap = ncap;
    return DK_OK;
}

// ridx != NULL: m holds ONLY the all-positive rows of a table of R_virtual rows and ridx[i] (ascending, host) is the
// row number of m's row i in that table. Rows that are not all-positive never enter computeNF (their logmean is not
// finite, computeNF.c:122-130), so walking the table in steps of `step` rows and walking these rows by row number is
// the same computation (dk_nf_reference_sparse).
static int compute_nf_impl(dk_ctx *ctx, const dk_matrix *m, int mode, uint64_t step_rows, double *nf_out, uint64_t *info,
                           const uint64_t *ridx, uint64_t R_virtual)
{
    if (!ctx || !m || !nf_out) return dk_fail(ctx, DK_ERR_ARG, "compute_nf: NULL argument");
    const int S = m->S;
    const uint64_t R = ridx ? R_virtual : m->R;
    cudaStream_t st = ctx->stream;
    uint64_t step = step_rows ? step_rows : 1000000;                  // computeNF.c:52
    DK_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
    NfWork w;
    unsigned char *small = (unsigned char *)dk_dev_alloc(ctx, 64 + (size_t)S * (256 * 8 + 8 + 8 + 8 + 8));
    if (!small) return DK_ERR_NOMEM;
    w.npos = (unsigned long long *)small;
    w.hist = (unsigned long long *)(small + 64);
    w.prefix = (uint64_t *)(small + 64 + (size_t)S * 256 * 8);
    w.k = w.prefix + S; w.nf = (double *)(w.k + S); w.ratio = w.nf + S;
    DK_CUDA(ctx, cudaMemsetAsync(small, 0, 64 + (size_t)S * 256 * 8, st));

    std::vector<double> nf(S, 0.0), nw(S, 0.0);
    uint64_t consumed = 0, steps = 0, npos_h = 0;
    int undefined_zone = 0, rc = DK_OK;
    const uint32_t nch = (uint32_t)m->n_chunks;
    auto ratios = [&](uint64_t r0, uint64_t r1) -> int {
        if (ridx) {                                                   // table rows [r0, r1) -> rows of the compact matrix
            r0 = (uint64_t)(std::lower_bound(ridx, ridx + m->R, r0) - ridx);
            r1 = (uint64_t)(std::lower_bound(ridx, ridx + m->R, r1) - ridx);
        }
        if (r1 <= r0) return DK_OK;
        DK_TRY(nf_grow(ctx, w, S, npos_h + (r1 - r0), npos_h));
        k_nf_ratios<<<(unsigned)((r1 - r0 + 127) / 128), 128, 0, st>>>(m->cnts, m->lbase, m->order, nch, S, r0, r1, w.vals,
                                                                   w.cap, w.npos, 0);
        DK_LAUNCHED(ctx);
        DK_CUDA(ctx, cudaMemcpyAsync(&npos_h, w.npos, 8, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
        return DK_OK;
    };
    if (mode == DK_NF_FULL) {
        // size the value store exactly: count all-positive rows first
        if (R) {
            k_nf_ratios<<<(unsigned)((R + 127) / 128), 128, 0, st>>>(m->cnts, m->lbase, m->order, nch, S, 0, R, nullptr, 0, w.npos, 1);
            DK_LAUNCHED(ctx);
        }
        uint64_t cnt = 0;
        DK_CUDA(ctx, cudaMemcpyAsync(&cnt, w.npos, 8, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
        DK_CUDA(ctx, cudaMemsetAsync(w.npos, 0, 8, st));
        if ((rc = nf_grow(ctx, w, S, cnt ? cnt : 1, 0)) != DK_OK) goto done;
        {   // exact capacity: nf_grow rounds up to a power of two; ratios() would re-grow for R rows, so launch directly
            if (R) {
                k_nf_ratios<<<(unsigned)((R + 127) / 128), 128, 0, st>>>(m->cnts, m->lbase, m->order, nch, S, 0, R, w.vals, w.cap, w.npos, 0);
                ctx->launches++;
            }
            npos_h = cnt;
        }
        consumed = R; steps = 1;
        if (npos_h == 0) { rc = dk_fail(ctx, DK_ERR_UNSUPPORTED, "compute_nf: no row has all counts > 0 (reference reads a NULL vector here)"); goto done; }
        rc = nf_select(ctx, w, S, npos_h, nf.data());
    } else {
        // computeNF.c:138-168 replayed: t-th evaluation after t*step rows
        uint64_t nb_step = 0;
        bool have_nf = false, stopped = false;
        while (consumed + step <= R) {
            if ((rc = ratios(consumed, consumed + step)) != DK_OK) goto done;
            consumed += step;
            if (npos_h == 0) { rc = dk_fail(ctx, DK_ERR_UNSUPPORTED, "compute_nf: no all-positive row in the first %llu rows", (unsigned long long)consumed); goto done; }
            if ((rc = nf_select(ctx, w, S, npos_h, nw.data())) != DK_OK) goto done;
            steps++;
            if (nb_step > 0) {
                double sum = 0;
                for (int j = 0; j < S; j++) sum += fabs(nf[j] - nw[j]);   // :145 (abs -> double overload under g++)
                bool stop = nb_step >= 2 && sum < 0.001 * S;              // nb_step==1 compares with uninitialised memory in the reference
                nf = nw; have_nf = true;
                if (stop) { stopped = true; break; }
            } else nf = nw;                                               // reference discards this one
            nb_step++;
        }
        if (nb_step == 0 && !stopped) {                                   // :166-168 fewer than `step` rows
            if ((rc = ratios(consumed, R)) != DK_OK) goto done;
            consumed = R;
            if (npos_h == 0) { rc = dk_fail(ctx, DK_ERR_UNSUPPORTED, "compute_nf: no row has all counts > 0 (reference reads a NULL vector here)"); goto done; }
            rc = nf_select(ctx, w, S, npos_h, nf.data());
            steps++;
        } else if (!have_nf) undefined_zone = 1;                          // step <= R < 2*step
        else if (!stopped) consumed = R;                                  // reference keeps reading to EOF without using the rows
    }
done:
    if (rc == DK_OK) {
        memcpy(nf_out, nf.data(), sizeof(double) * S);
        if (info) { info[0] = consumed; info[1] = steps; info[2] = (uint64_t)undefined_zone; info[3] = npos_h; }
    }
    cudaEventRecord(ctx->ev[1], st);
    cudaStreamSynchronize(st);
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    ctx->stage_ms[1] = ms;
    dk_dev_release(ctx, w.vals);
    dk_dev_release(ctx, small);
    return rc;
}

extern "C" int dk_compute_nf(dk_ctx *ctx, const dk_matrix *m, int mode, uint64_t step_rows, double *nf_out, uint64_t *info)
{
    return compute_nf_impl(ctx, m, mode, step_rows, nf_out, info, nullptr, 0);
}

void dk_matrix_release(dk_ctx *ctx, dk_matrix *m);
extern "C" int dk_nf_reference_sparse(dk_ctx *ctx, int S, uint64_t n_pos, const uint64_t *row_idx, const int32_t *cnts,
                                      uint64_t R_total, uint64_t step_rows, double *nf_out, uint64_t *info)
{
    if (!ctx || !nf_out || (n_pos && (!row_idx || !cnts))) return dk_fail(ctx, DK_ERR_ARG, "nf_reference_sparse: NULL argument");
    if (n_pos == 0) return dk_fail(ctx, DK_ERR_UNSUPPORTED, "compute_nf: no row has all counts > 0 (reference reads a NULL vector here)");
    for (uint64_t i = 0; i < n_pos; i++)
        if (row_idx[i] >= R_total || (i && row_idx[i] <= row_idx[i - 1]))
            return dk_fail(ctx, DK_ERR_ARG, "nf_reference_sparse: row numbers must ascend and lie below R_total");
    dk_matrix *m = nullptr;
    DK_TRY(dk_matrix_from_host(ctx, S, n_pos, nullptr, cnts, &m));
    const int rc = compute_nf_impl(ctx, m, DK_NF_REFERENCE, step_rows, nf_out, info, row_idx, R_total);
    dk_matrix_release(ctx, m);
    return rc;
}

// ---- distributed DK_NF_FULL building blocks (see include/dekupl_b200.h) --------------
struct dk_nf_state {
    NfWork w;
    unsigned char *small = nullptr;
    int S = 0;
    uint64_t n = 0;
    int shift = 56;
};

extern "C" int dk_nf_begin(dk_ctx *ctx, const dk_matrix *m, dk_nf_state **out, uint64_t *n_pos)
{
    if (!ctx || !m || !out) return dk_fail(ctx, DK_ERR_ARG, "nf_begin: NULL argument");
    cudaStream_t st = ctx->stream;
    const int S = m->S;
    dk_nf_state *s = new dk_nf_state();
    s->S = S;
    s->small = (unsigned char *)dk_dev_alloc(ctx, 64 + (size_t)S * (256 * 8 + 32));
    if (!s->small) { delete s; return DK_ERR_NOMEM; }
    NfWork &w = s->w;
    w.npos = (unsigned long long *)s->small;
    w.hist = (unsigned long long *)(s->small + 64);
    w.prefix = (uint64_t *)(s->small + 64 + (size_t)S * 256 * 8);
    w.k = w.prefix + S; w.nf = (double *)(w.k + S); w.ratio = w.nf + S;
    DK_CUDA(ctx, cudaMemsetAsync(s->small, 0, 64 + (size_t)S * (256 * 8 + 32), st));
    const uint32_t nch = (uint32_t)m->n_chunks;
    uint64_t cnt = 0;
    if (m->R) {
        k_nf_ratios<<<(unsigned)((m->R + 127) / 128), 128, 0, st>>>(m->cnts, m->lbase, m->order, nch, S, 0, m->R, nullptr, 0, w.npos, 1);
        DK_LAUNCHED(ctx);
        DK_CUDA(ctx, cudaMemcpyAsync(&cnt, w.npos, 8, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
        DK_CUDA(ctx, cudaMemsetAsync(w.npos, 0, 8, st));
    }
    w.cap = cnt ? cnt : 1;
    w.vals = (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * w.cap * S);
    if (!w.vals) { dk_dev_release(ctx, s->small); delete s; return DK_ERR_NOMEM; }
    if (cnt) {
        k_nf_ratios<<<(unsigned)((m->R + 127) / 128), 128, 0, st>>>(m->cnts, m->lbase, m->order, nch, S, 0, m->R, w.vals, w.cap, w.npos, 0);
        DK_LAUNCHED(ctx);
    }
    s->n = cnt;
    if (n_pos) *n_pos = cnt;
    *out = s;
    return DK_OK;
}
extern "C" int dk_nf_set_rank(dk_ctx *ctx, dk_nf_state *s, uint64_t k)
{
    if (!ctx || !s) return dk_fail(ctx, DK_ERR_ARG, "nf_set_rank: NULL argument");
    k_nf_setk<<<(s->S + 63) / 64, 64, 0, ctx->stream>>>(s->w.prefix, s->w.k, s->S, k); DK_LAUNCHED(ctx);
    s->shift = 56;
    return DK_OK;
}
extern "C" int dk_nf_hist_dev(dk_ctx *ctx, dk_nf_state *s, uint64_t *hist_dev)
{
    if (!ctx || !s || !hist_dev) return dk_fail(ctx, DK_ERR_ARG, "nf_hist: NULL argument");
    if (s->shift < 0) return dk_fail(ctx, DK_ERR_ARG, "nf_hist: all 8 rounds already done");
    cudaStream_t st = ctx->stream;
    DK_CUDA(ctx, cudaMemsetAsync(hist_dev, 0, sizeof(uint64_t) * 256 * s->S, st));
    if (s->n) {
        int gx = (int)((s->n + 4095) / 4096);
        if (gx > ctx->sm_count * 8) gx = ctx->sm_count * 8;
        k_nf_hist<<<dim3(gx, s->S), 256, 0, st>>>(s->w.vals, s->w.cap, s->n, s->shift, s->w.prefix, (unsigned long long *)hist_dev);
        DK_LAUNCHED(ctx);
    }
    return DK_OK;
}
extern "C" int dk_nf_advance_dev(dk_ctx *ctx, dk_nf_state *s, uint64_t *hist_dev)
{
    if (!ctx || !s || !hist_dev) return dk_fail(ctx, DK_ERR_ARG, "nf_advance: NULL argument");
    if (s->shift < 0) return dk_fail(ctx, DK_ERR_ARG, "nf_advance: all 8 rounds already done");
    k_nf_pick<<<s->S, 256, 0, ctx->stream>>>((unsigned long long *)hist_dev, s->S, s->shift, s->w.prefix, s->w.k);
    DK_LAUNCHED(ctx);
    s->shift -= 8;
    return DK_OK;
}
extern "C" int dk_nf_result(dk_ctx *ctx, dk_nf_state *s, double *nf_out)
{
    if (!ctx || !s || !nf_out) return dk_fail(ctx, DK_ERR_ARG, "nf_result: NULL argument");
    if (s->shift >= 0) return dk_fail(ctx, DK_ERR_ARG, "nf_result: radix select not finished");
    k_nf_finish<<<(s->S + 63) / 64, 64, 0, ctx->stream>>>(s->w.prefix, s->S, s->w.nf, s->w.ratio); DK_LAUNCHED(ctx);
    DK_CUDA(ctx, cudaMemcpyAsync(nf_out, s->w.nf, sizeof(double) * s->S, cudaMemcpyDeviceToHost, ctx->stream));
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DK_OK;
}
extern "C" void dk_nf_end(dk_ctx *ctx, dk_nf_state *s)
{
    if (!ctx || !s) return;
    cudaStreamSynchronize(ctx->stream);
    dk_dev_release(ctx, s->w.vals);
    dk_dev_release(ctx, s->small);
    delete s;
}
