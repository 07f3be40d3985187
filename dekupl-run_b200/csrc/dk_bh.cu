// dk_bh.cu — Benjamini-Hochberg + final filter (reference: TtestFilter.c:456-497).
// The reference sorts all R p-values descending and keeps a running minimum of
// R/(R-i)*p. adj(p) = min(1, min_{q>=p} R/#{p'<=q} * q) depends on the VALUE only, and
// any q > P has R/#*q >= q > P, so rows with adj <= P are decided exactly by the
// candidates {p <= P} alone. We therefore sort only the candidates (ascending),
// compute v at the last element of every tie run, take a suffix minimum, and look
// each local candidate up. Multi-GPU: candidates are all-gathered by the caller.
#include <math.h>
#include <stdlib.h>
#include "dk_common.cuh"

void dk_ttest_release(dk_ctx *ctx, dk_ttest *t);

__global__ void k_fill_nan(double *p, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = __longlong_as_double(0x7FF8000000000000ll);
}
// count (cp==nullptr) or append candidates p <= cut. One atomic per 1024-row block (a per-warp atomic on a single
// counter serialises: r01e measured 0.5 TB/s for this pass); order inside the list is irrelevant (it is sorted next)
#define BHC_T 256
#define BHC_V 4
__global__ void __launch_bounds__(BHC_T) k_bh_candidates(const double *p_raw, uint64_t R, double cut, unsigned long long *counter,
                                                         double *cp, uint32_t *crow, uint64_t cap)
{
    __shared__ uint32_t wcnt[BHC_T / 32];
    __shared__ unsigned long long bbase;
    const uint64_t base_row = (uint64_t)blockIdx.x * (BHC_T * BHC_V);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double v[BHC_V];
    uint32_t mine = 0;
#pragma unroll
    for (int j = 0; j < BHC_V; j++) {
        const uint64_t i = base_row + (uint64_t)j * BHC_T + threadIdx.x;
        v[j] = (i < R) ? p_raw[i] : __longlong_as_double(0x7FF8000000000000ll);   // NaN never compares <= cut
        mine += (v[j] <= cut) ? 1u : 0u;
    }
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) wcnt[warp] = incl;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < BHC_T / 32; w++) { uint32_t c = wcnt[w]; if (w < warp) woff += c; tot += c; }
    if (threadIdx.x == 0 && tot) bbase = atomicAdd(counter, (unsigned long long)tot);
    if (!cp) return;
    __syncthreads();
    if (!tot) return;
    uint64_t k = bbase + woff + incl - mine;
#pragma unroll
    for (int j = 0; j < BHC_V; j++) {
        if (v[j] <= cut) {
            if (k < cap) { cp[k] = v[j]; crow[k] = (uint32_t)(base_row + (uint64_t)j * BHC_T + threadIdx.x); }
            k++;
        }
    }
}
__global__ void k_iota_u32(uint32_t *v, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (uint32_t)i;
}
// v[j] at the last element of each tie run (ascending order), +inf elsewhere (TtestFilter.c:467)
__global__ void k_bh_v(const uint64_t *sk, uint64_t n, double Rg, uint64_t n_below, double *v)
{
    uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    bool last = (j + 1 == n) || (sk[j + 1] != sk[j]);
    double p = __longlong_as_double((long long)sk[j]);
    v[j] = last ? (Rg / (double)(n_below + j + 1)) * p : __longlong_as_double(0x7FF0000000000000ll);
}
// suffix minimum in three steps: tile minima, suffix over tiles (one block), apply
#define SM_T 256
#define SM_V 8
__global__ void k_sufmin_tiles(const double *v, uint64_t n, double *tmin)
{
    __shared__ double sh[SM_T / 32];
    uint64_t base = (uint64_t)blockIdx.x * SM_T * SM_V;
    double m = __longlong_as_double(0x7FF0000000000000ll);
    for (int i = 0; i < SM_V; i++) {
        uint64_t k = base + (uint64_t)i * SM_T + threadIdx.x;
        if (k < n) m = fmin(m, v[k]);
    }
    for (int o = 16; o; o >>= 1) m = fmin(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < SM_T / 32; w++) m = fmin(m, sh[w]);
        tmin[blockIdx.x] = m;
    }
}
__global__ void k_sufmin_scan_tiles(double *tmin, uint64_t nt)   // tmin[t] <- min over tiles > t (exclusive suffix)
{
    if (threadIdx.x || blockIdx.x) return;
    double run = __longlong_as_double(0x7FF0000000000000ll);
    for (uint64_t t = nt; t-- > 0;) { double x = tmin[t]; tmin[t] = run; run = fmin(run, x); }
}
__global__ void k_sufmin_apply(double *v, uint64_t n, const double *tmin)
{
    // one thread walks SM_V consecutive items backwards; block-level suffix over thread minima
    __shared__ double sh[SM_T];
    uint64_t base = (uint64_t)blockIdx.x * SM_T * SM_V + (uint64_t)threadIdx.x * SM_V;
    const double INF = __longlong_as_double(0x7FF0000000000000ll);
    double loc[SM_V], m = INF;
    for (int i = 0; i < SM_V; i++) { loc[i] = (base + i < n) ? v[base + i] : INF; m = fmin(m, loc[i]); }
    sh[threadIdx.x] = m;
    __syncthreads();
    for (int o = 1; o < SM_T; o <<= 1) {
        double t = (threadIdx.x + o < SM_T) ? sh[threadIdx.x + o] : INF;
        __syncthreads();
        sh[threadIdx.x] = fmin(sh[threadIdx.x], t);
        __syncthreads();
    }
    double run = fmin(tmin[blockIdx.x], (threadIdx.x + 1 < SM_T) ? sh[threadIdx.x + 1] : INF);
    for (int i = SM_V - 1; i >= 0; i--) {
        run = fmin(run, loc[i]);
        if (base + i < n) v[base + i] = run;
    }
}
// adj for each local candidate: find its p in the sorted global candidate list
__global__ void k_bh_lookup(const double *cp, const uint32_t *crow, uint64_t nc, const uint64_t *sk, const double *F,
                            uint64_t n_all, double *p_adj)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc) return;
    uint64_t key = (uint64_t)__double_as_longlong(cp[i]);
    uint64_t lo = 0, hi = n_all;
    while (lo < hi) { uint64_t mid = (lo + hi) >> 1; if (sk[mid] < key) lo = mid + 1; else hi = mid; }
    double f = (lo < n_all) ? F[lo] : 1.0;
    p_adj[crow[i]] = f > 1.0 ? 1.0 : f;                              // dmin(1,new) :475
}

// n_known: the number of rows with p_raw <= p_cut when the caller has just counted them (dk_ttest_count_le at the same cut),
// else UINT64_MAX and this function counts first
static int ttest_candidates_impl(dk_ctx *ctx, dk_ttest *t, double p_cut, uint64_t n_known, uint64_t *n_cand, const double **cand_p_dev)
{
    if (!ctx || !t) return dk_fail(ctx, DK_ERR_ARG, "ttest_candidates: NULL argument");
    cudaStream_t st = ctx->stream;
    unsigned long long *ctr = (unsigned long long *)dk_dev_alloc(ctx, 64);
    if (!ctr) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(ctr, 0, 64, st));
    unsigned grid = (unsigned)((t->R + BHC_T * BHC_V - 1) / (BHC_T * BHC_V));
    uint64_t nc = 0;
    if (n_known != ~0ull) nc = n_known;
    else if (t->R) {
        k_bh_candidates<<<grid, 256, 0, st>>>(t->p_raw, t->R, p_cut, ctr, nullptr, nullptr, 0); DK_LAUNCHED(ctx);
        DK_CUDA(ctx, cudaMemcpyAsync(&nc, ctr, 8, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
    }
    dk_dev_release(ctx, t->cand_p); dk_dev_release(ctx, t->cand_row);
    // adjusted values belong to the candidate list they were written through (dk_ttest_filter walks that list): a new list voids them
    dk_dev_release(ctx, t->p_adj); t->p_adj = nullptr;
    t->cand_p = (double *)dk_dev_alloc(ctx, sizeof(double) * (nc + 1));
    t->cand_row = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (nc + 1));
    if (!t->cand_p || !t->cand_row) return DK_ERR_NOMEM;
    if (nc) {
        DK_CUDA(ctx, cudaMemsetAsync(ctr, 0, 8, st));
        k_bh_candidates<<<grid, 256, 0, st>>>(t->p_raw, t->R, p_cut, ctr, t->cand_p, t->cand_row, nc); DK_LAUNCHED(ctx);
    }
    t->n_cand = nc; t->p_cut = p_cut; t->cand_cap = nc + 1;
    dk_dev_release(ctx, ctr);
    if (n_cand) *n_cand = nc;
    if (cand_p_dev) *cand_p_dev = t->cand_p;
    return DK_OK;
}

extern "C" int dk_ttest_candidates(dk_ctx *ctx, dk_ttest *t, double p_cut, uint64_t *n_cand, const double **cand_p_dev)
{
    return ttest_candidates_impl(ctx, t, p_cut, ~0ull, n_cand, cand_p_dev);
}

// #{ p_raw <= cut } of this object's rows: the counting pass of the candidate narrowing below (one read of p_raw)
extern "C" int dk_ttest_count_le(dk_ctx *ctx, const dk_ttest *t, double cut, uint64_t *n)
{
    if (!ctx || !t || !n) return dk_fail(ctx, DK_ERR_ARG, "ttest_count_le: NULL argument");
    *n = 0;
    if (!t->R) return DK_OK;
    cudaStream_t st = ctx->stream;
    unsigned long long *ctr = (unsigned long long *)dk_dev_alloc(ctx, 64);
    if (!ctr) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(ctr, 0, 64, st));
    const unsigned grid = (unsigned)((t->R + BHC_T * BHC_V - 1) / (BHC_T * BHC_V));
    k_bh_candidates<<<grid, 256, 0, st>>>(t->p_raw, t->R, cut, ctr, nullptr, nullptr, 0); DK_LAUNCHED(ctx);
    uint64_t nc = 0;
    DK_CUDA(ctx, cudaMemcpyAsync(&nc, ctr, 8, cudaMemcpyDeviceToHost, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    dk_dev_release(ctx, ctr);
    *n = nc;
    return DK_OK;
}
// Candidate narrowing. A row reaches an adjusted p <= P only through some q >= p with R/rank(q) * q <= P, i.e.
// q <= P * rank(q) / R. With n = #{p <= c} over ALL rows (all ranks), every q <= c has rank(q) <= n, hence
// q <= c' = P * n / R: candidates above c' can neither survive nor lower anybody's adjusted value to <= P. Starting from
// c = P the cut falls 0.05 -> P * n0 / R -> ... in a few counting passes (8 M -> a few thousand candidates at cfg 2),
// and sort / rank / lookup shrink with it. The margin keeps rows within the 1e-9 tie window of P in the census.
extern "C" double dk_bh_next_cut(double p_cut, uint64_t n_le_cut_global, uint64_t R_global)
{
    if (!(p_cut < 1.0) || R_global == 0) return p_cut;
    return p_cut * ((double)n_le_cut_global / (double)R_global) * (1.0 + 1e-8) + 1e-300;
}
// dk_ttest_candidates at a NARROWED cut: the list holds p_raw <= cut, and the adjusted values written through it are
// declared exact up to p_cut (the caller derived `cut` from p_cut with dk_bh_next_cut over the global counts)
extern "C" int dk_ttest_candidates_narrowed(dk_ctx *ctx, dk_ttest *t, double cut, double p_cut, uint64_t *n_cand,
                                            const double **cand_p_dev)
{
    if (!(cut <= p_cut)) return dk_fail(ctx, DK_ERR_ARG, "ttest_candidates_narrowed: cut %g above p_cut %g", cut, p_cut);
    DK_TRY(dk_ttest_candidates(ctx, t, cut, n_cand, cand_p_dev));
    t->p_cut = p_cut;
    return DK_OK;
}

// Rank one p-value slice of the (global) candidate list: sorted ascending, then F[j] = min_{i >= j in the slice}
// R/(n_below+i+1) * p_i taken at the last element of every tie run (TtestFilter.c:459-475 restated on ascending order).
// n_below = number of candidates (anywhere) smaller than every value of this slice. A multi-GPU caller owns one
// p-range per rank, then lowers F by the minimum of all higher ranges (the running minimum crosses slice borders).
extern "C" int dk_bh_rank_slice(dk_ctx *ctx, const double *p_dev, uint64_t n, uint64_t n_below, uint64_t R_global,
                                double *sorted_p_dev, double *F_dev)
{
    if (!ctx || (n && (!p_dev || !sorted_p_dev || !F_dev))) return dk_fail(ctx, DK_ERR_ARG, "bh_rank_slice: NULL argument");
    if (n == 0) return DK_OK;
    cudaStream_t st = ctx->stream;
    uint64_t *k0 = (uint64_t *)dk_dev_alloc(ctx, 8 * n), *k1 = (uint64_t *)dk_dev_alloc(ctx, 8 * n);
    uint32_t *v0 = (uint32_t *)dk_dev_alloc(ctx, 4 * n), *v1 = (uint32_t *)dk_dev_alloc(ctx, 4 * n);
    const uint64_t nt = (n + SM_T * SM_V - 1) / (SM_T * SM_V);
    double *tmin = (double *)dk_dev_alloc(ctx, 8 * nt);
    if (!k0 || !k1 || !v0 || !v1 || !tmin) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemcpyAsync(k0, p_dev, 8 * n, cudaMemcpyDeviceToDevice, st));
    k_iota_u32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(v0, n); DK_LAUNCHED(ctx);
    uint64_t *sk; uint32_t *sv;
    DK_TRY(dk_sort_pairs_u64(ctx, k0, v0, k1, v1, n, &sk, &sv));
    k_bh_v<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(sk, n, (double)R_global, n_below, F_dev); DK_LAUNCHED(ctx);
    k_sufmin_tiles<<<(unsigned)nt, SM_T, 0, st>>>(F_dev, n, tmin); DK_LAUNCHED(ctx);
    k_sufmin_scan_tiles<<<1, 32, 0, st>>>(tmin, nt); DK_LAUNCHED(ctx);
    k_sufmin_apply<<<(unsigned)nt, SM_T, 0, st>>>(F_dev, n, tmin); DK_LAUNCHED(ctx);
    DK_CUDA(ctx, cudaMemcpyAsync(sorted_p_dev, sk, 8 * n, cudaMemcpyDeviceToDevice, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    dk_dev_release(ctx, k0); dk_dev_release(ctx, k1); dk_dev_release(ctx, v0); dk_dev_release(ctx, v1); dk_dev_release(ctx, tmin);
    return DK_OK;
}

// ---- dk_bh_rank_slice without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops). The scratch the
// allocating entry point takes from the context's cache (two key buffers, the sort's tile counters, the tile minima)
// lives in ONE file-scope device array of fixed size: every GPU that loads the library holds these ~4.3 MB until exit.
// In host code `g_bh_scratch` is only a host-side stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
#define BHN_MAX DK_BH_NOALLOC_MAX
#define BHN_TILES ((BHN_MAX + 4095) / 4096)                    /* tiles of the radix sort (RS_TILE = 4096) */
#define BHN_OFF_K1 ((size_t)8 * BHN_MAX)
#define BHN_OFF_CNT ((size_t)16 * BHN_MAX)
#define BHN_OFF_TMIN (BHN_OFF_CNT + (size_t)4 * 256 * BHN_TILES)
#define BHN_BYTES (BHN_OFF_TMIN + (size_t)8 * ((BHN_MAX + SM_T * SM_V - 1) / (SM_T * SM_V)) + 256)
static __device__ __align__(256) unsigned char g_bh_scratch[BHN_BYTES];
#define BHN_MAX_DEVICES 64
static unsigned char *g_bh_scratch_addr[BHN_MAX_DEVICES];       // device address of the array, per device ordinal

extern "C" int dk_bh_rank_slice_noalloc(dk_ctx *ctx, const double *p_dev, uint64_t n, uint64_t n_below, uint64_t R_global,
                                        double *sorted_p_dev, double *F_dev)
{
    if (!ctx || (n && (!p_dev || !sorted_p_dev || !F_dev))) return dk_fail(ctx, DK_ERR_ARG, "bh_rank_slice_noalloc: NULL argument");
    if (n > (uint64_t)BHN_MAX)
        return dk_fail(ctx, DK_ERR_ARG, "bh_rank_slice_noalloc: %llu p-values exceed the fixed scratch (DK_BH_NOALLOC_MAX = %d); "
                       "use dk_bh_rank_slice", (unsigned long long)n, (int)BHN_MAX);
    if (n == 0) return DK_OK;
    int cur = -1;
    DK_CUDA(ctx, cudaGetDevice(&cur));
    if (cur != ctx->device || cur < 0 || cur >= BHN_MAX_DEVICES)
        return dk_fail(ctx, DK_ERR_ARG, "bh_rank_slice_noalloc: current device %d is not the context's device %d", cur, ctx->device);
    if (!g_bh_scratch_addr[cur]) {                              // set-up, once per device, that device current
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_bh_scratch);
        if (e != cudaSuccess || !ptr)
            return dk_fail(ctx, DK_ERR_CUDA, "bh_rank_slice_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_bh_scratch_addr[cur] = (unsigned char *)ptr;
    }
    unsigned char *sc = g_bh_scratch_addr[cur];
    uint64_t *k0 = (uint64_t *)sc, *k1 = (uint64_t *)(sc + BHN_OFF_K1);
    uint32_t *cnt = (uint32_t *)(sc + BHN_OFF_CNT);
    double *tmin = (double *)(sc + BHN_OFF_TMIN);
    cudaStream_t st = ctx->stream;
    // every scratch region is fully written before it is read (keys copied in, counters rewritten by each sort pass,
    // tile minima written by k_sufmin_tiles): nothing of the previous call's contents is ever consumed
    const uint64_t nt = (n + SM_T * SM_V - 1) / (SM_T * SM_V);
    DK_CUDA(ctx, cudaMemcpyAsync(k0, p_dev, 8 * n, cudaMemcpyDeviceToDevice, st));
    DK_TRY(dk_sort_keys_u64_scratch(ctx, k0, k1, n, cnt));      // positive doubles order like their bit patterns
    k_bh_v<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(k0, n, (double)R_global, n_below, F_dev); DK_LAUNCHED(ctx);
    k_sufmin_tiles<<<(unsigned)nt, SM_T, 0, st>>>(F_dev, n, tmin); DK_LAUNCHED(ctx);
    k_sufmin_scan_tiles<<<1, 32, 0, st>>>(tmin, nt); DK_LAUNCHED(ctx);
    k_sufmin_apply<<<(unsigned)nt, SM_T, 0, st>>>(F_dev, n, tmin); DK_LAUNCHED(ctx);
    DK_CUDA(ctx, cudaMemcpyAsync(sorted_p_dev, k0, 8 * n, cudaMemcpyDeviceToDevice, st));
    return DK_OK;                                               // asynchronous on the context's stream: dk_sync before reading
}

// adjusted p of this shard's candidates from the global ascending candidate list and its F values
extern "C" int dk_ttest_apply_adjusted(dk_ctx *ctx, dk_ttest *t, const double *sorted_p_dev, const double *F_dev, uint64_t n_all)
{
    if (!ctx || !t) return dk_fail(ctx, DK_ERR_ARG, "ttest_apply_adjusted: NULL argument");
    cudaStream_t st = ctx->stream;
    dk_dev_release(ctx, t->p_adj);
    t->p_adj = (double *)dk_dev_alloc(ctx, sizeof(double) * (t->R + 1));
    if (!t->p_adj) return DK_ERR_NOMEM;
    if (t->R) { k_fill_nan<<<(unsigned)((t->R + 255) / 256), 256, 0, st>>>(t->p_adj, t->R); DK_LAUNCHED(ctx); }
    if (n_all == 0 || t->n_cand == 0) return DK_OK;
    k_bh_lookup<<<(unsigned)((t->n_cand + 255) / 256), 256, 0, st>>>(t->cand_p, t->cand_row, t->n_cand, (const uint64_t *)sorted_p_dev,
                                                                   F_dev, n_all, t->p_adj);
    DK_LAUNCHED(ctx);
    return DK_OK;
}

extern "C" int dk_ttest_adjust_global(dk_ctx *ctx, dk_ttest *t, const double *all_cand_p_dev, uint64_t n_all, uint64_t R_global)
{
    if (!ctx || !t) return dk_fail(ctx, DK_ERR_ARG, "ttest_adjust: NULL argument");
    if (n_all == 0 || t->n_cand == 0) return dk_ttest_apply_adjusted(ctx, t, nullptr, nullptr, 0);
    double *sp = (double *)dk_dev_alloc(ctx, 8 * n_all), *F = (double *)dk_dev_alloc(ctx, 8 * n_all);
    if (!sp || !F) return DK_ERR_NOMEM;
    int rc = dk_bh_rank_slice(ctx, all_cand_p_dev, n_all, 0, R_global, sp, F);
    if (rc == DK_OK) rc = dk_ttest_apply_adjusted(ctx, t, sp, F, n_all);
    cudaStreamSynchronize(ctx->stream);
    dk_dev_release(ctx, sp); dk_dev_release(ctx, F);
    return rc;
}

// internal (dk_comm.cu): the candidate list at a narrowed cut when the caller already knows this rank's count at that cut
int dk_ttest_candidates_known(dk_ctx *ctx, dk_ttest *t, double cut, double p_cut, uint64_t n_known, uint64_t *n_cand,
                              const double **cand_p_dev)
{
    if (!(cut <= p_cut)) return dk_fail(ctx, DK_ERR_ARG, "ttest_candidates: cut %g above p_cut %g", cut, p_cut);
    DK_TRY(ttest_candidates_impl(ctx, t, cut, n_known, n_cand, cand_p_dev));
    t->p_cut = p_cut;
    return DK_OK;
}

extern "C" int dk_ttest_adjust(dk_ctx *ctx, dk_ttest *t, double p_cut)
{
    if (!ctx || !t) return dk_fail(ctx, DK_ERR_ARG, "ttest_adjust: NULL argument");
    uint64_t nc = 0, n_known = ~0ull;
    const double *cp = nullptr;
    double cut = p_cut;
    if (p_cut < 1.0 && t->R && !getenv("DK_BH_NO_NARROW")) {
        for (int it = 0; it < 4; it++) {                          // each pass: one read of p_raw (dk_bh_next_cut)
            uint64_t n = 0;
            DK_TRY(dk_ttest_count_le(ctx, t, cut, &n));
            n_known = n;                                          // the compaction below need not count again at this cut
            const double next = dk_bh_next_cut(p_cut, n, t->R);
            if (n <= DK_BH_SMALL_LIST || !(next < cut * 0.75)) break;   // a list this short sorts for free: no more passes
            cut = next; n_known = ~0ull;
        }
    }
    if (!(cut <= p_cut)) return dk_fail(ctx, DK_ERR_ARG, "ttest_adjust: cut %g above p_cut %g", cut, p_cut);
    DK_TRY(ttest_candidates_impl(ctx, t, cut, n_known, &nc, &cp));
    t->p_cut = p_cut;                                             // adjusted values are exact up to the REQUESTED cut
    return dk_ttest_adjust_global(ctx, t, cp, nc, t->R);
}
