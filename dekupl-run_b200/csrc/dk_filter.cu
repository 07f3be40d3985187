// dk_filter.cu — survivor selection (TtestFilter.c:486) and the gathers that feed the
// formatter (TtestFilter.c:487-489): adjusted p, meanA, meanB, log2FC, normalised counts.
// Compiled with -fmad=false (means are recomputed with the reference's operation order).
#include <math.h>
#include "dk_common.cuh"

__device__ __forceinline__ size_t phys_row(const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, uint64_t row)
{
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (lbase[mid] <= row) lo = mid; else hi = mid;
    }
    return (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
}

__global__ void k_filter(const double *p_adj, const double *l2fc, uint64_t R, double pmax, double fmin_,
                         unsigned long long *counter, uint64_t *surv /*may be null: count only*/, uint64_t cap,
                         unsigned long long *ties)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool keep = false;
    if (i < R) {
        double p = p_adj[i], f = fabs(l2fc[i]);
        keep = (p <= pmax) && (f >= fmin_);                              // TtestFilter.c:486
        if (!surv) {   // tie census on the counting pass only
            if (fabs(p - pmax) <= 1e-9 * fabs(pmax) && (f >= fmin_)) atomicAdd(&ties[0], 1ull);
            if (fmin_ != 0 && fabs(f - fmin_) <= 1e-9 * fabs(fmin_) && (p <= pmax)) atomicAdd(&ties[1], 1ull);
        }
    }
    unsigned b = __ballot_sync(0xffffffffu, keep);
    if (!b) return;
    int lane = threadIdx.x & 31, leader = __ffs(b) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(b));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (keep && surv) {
        uint64_t k = base + __popc(b & ((1u << lane) - 1u));
        if (k < cap) surv[k] = i;
    }
}
// the same test over the BH candidates only: p_adj is NaN (never <= pmax, never a tie) on every other row, because
// dk_ttest_apply_adjusted writes adjusted values at the candidate rows alone — 20x fewer rows to look at than R at P = 0.05
__global__ void k_filter_cand(const uint32_t *crow, uint64_t nc, const double *p_adj, const double *l2fc, double pmax, double fmin_,
                              unsigned long long *counter, uint64_t *surv /*may be null: count only*/, uint64_t cap,
                              unsigned long long *ties)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool keep = false;
    uint64_t row = 0;
    if (i < nc) {
        row = crow[i];
        double p = p_adj[row], f = fabs(l2fc[row]);
        keep = (p <= pmax) && (f >= fmin_);                              // TtestFilter.c:486
        if (!surv) {   // tie census on the counting pass only
            if (fabs(p - pmax) <= 1e-9 * fabs(pmax) && (f >= fmin_)) atomicAdd(&ties[0], 1ull);
            if (fmin_ != 0 && fabs(f - fmin_) <= 1e-9 * fabs(fmin_) && (p <= pmax)) atomicAdd(&ties[1], 1ull);
        }
    }
    unsigned b = __ballot_sync(0xffffffffu, keep);
    if (!b) return;
    int lane = threadIdx.x & 31, leader = __ffs(b) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(b));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (keep && surv) {
        uint64_t k = base + __popc(b & ((1u << lane) - 1u));
        if (k < cap) surv[k] = row;
    }
}
__global__ void k_u64_to_u32(const uint64_t *in, uint32_t *out, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (uint32_t)in[i];
}

extern "C" int dk_ttest_filter(dk_ctx *ctx, dk_ttest *t, double pmax, double fmin_, uint64_t *n_keep, uint64_t *ties)
{
    if (!ctx || !t || !t->p_adj) return dk_fail(ctx, DK_ERR_ARG, "ttest_filter: call dk_ttest_adjust first");
    // adjusted values exist at the candidate rows (p_raw <= p_cut) only and are exact only up to p_cut
    if (t->p_cut < 1.0 && pmax > t->p_cut)
        return dk_fail(ctx, DK_ERR_ARG, "ttest_filter: pmax %g exceeds the cut %g the adjusted p-values were computed for", pmax, t->p_cut);
    cudaStream_t st = ctx->stream;
    // log2FC is read at the candidate rows (or at every row without a candidate list)
    DK_TRY(dk_ttest_need_log2fc(ctx, t, t->cand_row, t->cand_row ? t->n_cand : 0));
    unsigned long long *ctr = (unsigned long long *)dk_dev_alloc(ctx, 64);
    if (!ctr) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(ctr, 0, 64, st));
    DK_CUDA(ctx, cudaMemsetAsync(t->tie_counts, 0, 16, st));          // keep [2] (sd ties from the row pass)
    unsigned grid = (unsigned)((t->R + 255) / 256);
    const bool by_cand = t->cand_row != nullptr;                      // the list the adjusted values were written through
    const unsigned cgrid = (unsigned)((t->n_cand + 255) / 256);
    uint64_t D = 0;
    if (t->R && !(by_cand && t->n_cand == 0)) {
        if (by_cand) k_filter_cand<<<cgrid, 256, 0, st>>>(t->cand_row, t->n_cand, t->p_adj, t->log2fc, pmax, fmin_, ctr, nullptr, 0,
                                                         (unsigned long long *)t->tie_counts);
        else k_filter<<<grid, 256, 0, st>>>(t->p_adj, t->log2fc, t->R, pmax, fmin_, ctr, nullptr, 0, (unsigned long long *)t->tie_counts);
        DK_LAUNCHED(ctx);
        DK_CUDA(ctx, cudaMemcpyAsync(&D, ctr, 8, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
    }
    dk_dev_release(ctx, t->surv_row);
    t->surv_row = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (D + 1));
    if (!t->surv_row) return DK_ERR_NOMEM;
    if (D) {
        uint64_t *k0 = (uint64_t *)dk_dev_alloc(ctx, 8 * D), *k1 = (uint64_t *)dk_dev_alloc(ctx, 8 * D);
        uint32_t *v0 = (uint32_t *)dk_dev_alloc(ctx, 4 * D), *v1 = (uint32_t *)dk_dev_alloc(ctx, 4 * D);
        if (!k0 || !k1 || !v0 || !v1) return DK_ERR_NOMEM;
        DK_CUDA(ctx, cudaMemsetAsync(ctr, 0, 8, st));
        DK_CUDA(ctx, cudaMemsetAsync(v0, 0, 4 * D, st));
        if (by_cand) k_filter_cand<<<cgrid, 256, 0, st>>>(t->cand_row, t->n_cand, t->p_adj, t->log2fc, pmax, fmin_, ctr, k0, D, nullptr);
        else k_filter<<<grid, 256, 0, st>>>(t->p_adj, t->log2fc, t->R, pmax, fmin_, ctr, k0, D, nullptr);
        DK_LAUNCHED(ctx);
        uint64_t *sk; uint32_t *sv;
        DK_TRY(dk_sort_pairs_u64(ctx, k0, v0, k1, v1, D, &sk, &sv));  // ascending row index == reference order
        k_u64_to_u32<<<(unsigned)((D + 255) / 256), 256, 0, st>>>(sk, t->surv_row, D); DK_LAUNCHED(ctx);
        dk_dev_release(ctx, k0); dk_dev_release(ctx, k1); dk_dev_release(ctx, v0); dk_dev_release(ctx, v1);
    }
    t->D = D;
    if (n_keep) *n_keep = D;
    if (ties) {
        DK_CUDA(ctx, cudaMemcpyAsync(ties, t->tie_counts, 24, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
    }
    dk_dev_release(ctx, ctr);
    return DK_OK;
}

// one thread per survivor: key, adj p, means (recomputed as TtestFilter.c:392-405), log2FC, counts/NF
__global__ void k_gather_survivors(const uint32_t *surv, uint64_t D, const uint64_t *keys, const int32_t *cnts,
                                   const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, int S,
                                   const double *nf, const int *idx, int nA, int nB, const double *p_adj,
                                   const double *l2fc, uint64_t *o_row, uint64_t *o_key, double *o_p, double *o_mA,
                                   double *o_mB, double *o_fc, double *o_norm)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D) return;
    uint64_t row = surv[i];
    size_t ph = phys_row(lbase, order, n_chunks, row);
    const int32_t *c = cnts + ph * (size_t)S;
    if (o_row) o_row[i] = row;
    if (o_key) o_key[i] = keys ? keys[ph] : 0;
    if (o_p) o_p[i] = p_adj[row];
    if (o_fc) o_fc[i] = l2fc[row];
    if (o_mA) { double s = 0; for (int k = 0; k < nA; k++) s += (double)c[idx[k]] / nf[idx[k]] + 1; o_mA[i] = s / nA; }
    if (o_mB) { double s = 0; for (int k = 0; k < nB; k++) s += (double)c[idx[nA + k]] / nf[idx[nA + k]] + 1; o_mB[i] = s / nB; }
    if (o_norm) for (int j = 0; j < S; j++) o_norm[i * (size_t)S + j] = (double)c[j] / nf[j];
}

extern "C" int dk_ttest_export_survivors(dk_ctx *ctx, const dk_ttest *t, uint64_t *row_idx, uint64_t *keys, double *p_adj,
                                         double *meanA, double *meanB, double *log2fc, double *norm_cnts)
{
    if (!ctx || !t) return dk_fail(ctx, DK_ERR_ARG, "export_survivors: NULL argument");
    uint64_t D = t->D;
    if (!D) return DK_OK;
    cudaStream_t st = ctx->stream;
    const dk_matrix *m = t->m;
    size_t per = 8 * 6 + 8 * (size_t)t->S;
    unsigned char *buf = (unsigned char *)dk_dev_alloc(ctx, per * D);
    if (!buf) return DK_ERR_NOMEM;
    uint64_t *d_row = (uint64_t *)buf, *d_key = d_row + D;
    double *d_p = (double *)(d_key + D), *d_mA = d_p + D, *d_mB = d_mA + D, *d_fc = d_mB + D, *d_norm = d_fc + D;
    k_gather_survivors<<<(unsigned)((D + 127) / 128), 128, 0, st>>>(
        t->surv_row, D, m->has_keys ? m->keys : nullptr, m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks, t->S, t->nf_dev,
        t->idx_dev, t->nA, t->nB, t->p_adj, t->log2fc, d_row, d_key, d_p, d_mA, d_mB, d_fc, norm_cnts ? d_norm : nullptr);
    DK_LAUNCHED(ctx);
    if (row_idx) DK_CUDA(ctx, cudaMemcpyAsync(row_idx, d_row, 8 * D, cudaMemcpyDeviceToHost, st));
    if (keys) DK_CUDA(ctx, cudaMemcpyAsync(keys, d_key, 8 * D, cudaMemcpyDeviceToHost, st));
    if (p_adj) DK_CUDA(ctx, cudaMemcpyAsync(p_adj, d_p, 8 * D, cudaMemcpyDeviceToHost, st));
    if (meanA) DK_CUDA(ctx, cudaMemcpyAsync(meanA, d_mA, 8 * D, cudaMemcpyDeviceToHost, st));
    if (meanB) DK_CUDA(ctx, cudaMemcpyAsync(meanB, d_mB, 8 * D, cudaMemcpyDeviceToHost, st));
    if (log2fc) DK_CUDA(ctx, cudaMemcpyAsync(log2fc, d_fc, 8 * D, cudaMemcpyDeviceToHost, st));
    if (norm_cnts) DK_CUDA(ctx, cudaMemcpyAsync(norm_cnts, d_norm, 8 * D * t->S, cudaMemcpyDeviceToHost, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    dk_dev_release(ctx, buf);
    return DK_OK;
}

extern "C" int dk_ttest_export_rows(dk_ctx *ctx, const dk_ttest *t, double *p_raw, double *log2fc, double *p_adj)
{
    if (!ctx || !t) return dk_fail(ctx, DK_ERR_ARG, "export_rows: NULL argument");
    cudaStream_t st = ctx->stream;
    if (log2fc) DK_TRY(dk_ttest_need_log2fc(ctx, t, nullptr, 0));
    if (p_raw) DK_CUDA(ctx, cudaMemcpyAsync(p_raw, t->p_raw, 8 * t->R, cudaMemcpyDeviceToHost, st));
    if (log2fc) DK_CUDA(ctx, cudaMemcpyAsync(log2fc, t->log2fc, 8 * t->R, cudaMemcpyDeviceToHost, st));
    if (p_adj) {
        if (!t->p_adj) return dk_fail(ctx, DK_ERR_ARG, "export_rows: p_adj requested before dk_ttest_adjust");
        DK_CUDA(ctx, cudaMemcpyAsync(p_adj, t->p_adj, 8 * t->R, cudaMemcpyDeviceToHost, st));
    }
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    return DK_OK;
}
