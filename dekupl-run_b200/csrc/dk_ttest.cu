// dk_ttest.cu — TtestFilter pass 2 on the GPU (reference: share/TtestFilter/TtestFilter.c:373-440):
// one streaming pass over the count matrix, one thread per row: normalise by NF, +1,
// log, group means / sds of the logs, Welch t and df, two-sided Student-t tail, log2FC.
// Arithmetic mirrors the reference operation by operation (sequential sums in group
// order, no FMA contraction: this file is compiled with -fmad=false).
#include "dk_common.cuh"

// ---- Student-t tail: 2*Q(|t|;v) = I_{v/(v+t^2)}(v/2, 1/2)  (replaces the Boost call at :146-147)
__device__ double dk_beta_cf(double a, double b, double x)
{
    const double TINY = 1e-300, EPS = 1e-16;
    double qab = a + b, qap = a + 1.0, qam = a - 1.0;
    double c = 1.0, d = 1.0 - qab * x / qap;
    if (fabs(d) < TINY) d = TINY;
    d = 1.0 / d;
    double h = d;
    for (int m = 1; m <= 10000; m++) {
        double m2 = 2.0 * m;
        double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
        d = 1.0 + aa * d; if (fabs(d) < TINY) d = TINY;
        c = 1.0 + aa / c; if (fabs(c) < TINY) c = TINY;
        d = 1.0 / d; h *= d * c;
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
        d = 1.0 + aa * d; if (fabs(d) < TINY) d = TINY;
        c = 1.0 + aa / c; if (fabs(c) < TINY) c = TINY;
        d = 1.0 / d;
        double del = d * c;
        h *= del;
        if (fabs(del - 1.0) < EPS) break;
    }
    return h;
}
// log( Gamma(v+1/2) / Gamma(v) ): direct for small v, asymptotic series for v >= 40
// (lgamma(v+.5)-lgamma(v) would lose ~|lgamma| * 1e-16 to cancellation at large v)
__device__ double dk_lgamma_ratio_half(double v)
{
    if (v < 40.0) return lgamma(v + 0.5) - lgamma(v);
    double r = 1.0 / v, r2 = r * r;
    return 0.5 * log(v) + r * (-1.0 / 8.0 + r2 * (1.0 / 192.0 + r2 * (-1.0 / 640.0 + r2 * (17.0 / 14336.0))));
}

// I_x(v, 1/2) with y = 1-x supplied by the caller (no cancellation)
__device__ double dk_ibeta_half(double v, double x, double y)
{
    if (x <= 0.0) return 0.0;
    if (y <= 0.0) return 1.0;
    const double LOG_SQRT_PI = 0.57236494292470008707;
    double lx = (y < 0.5) ? log1p(-y) : log(x);
    double ly = (x < 0.5) ? log1p(-x) : log(y);
    double lbt = dk_lgamma_ratio_half(v) - LOG_SQRT_PI + v * lx + 0.5 * ly;
    if (x < (v + 1.0) / (v + 2.5))
        return exp(lbt) * dk_beta_cf(v, 0.5, x) / v;
    return 1.0 - exp(lbt) * dk_beta_cf(0.5, v, y) / 0.5;
}

__device__ double dk_student_t_2tail(double t, double df)
{
    double t2 = t * t;
    if (isinf(df)) return erfc(fabs(t) / sqrt(2.0));
    double x = df / (df + t2), y = t2 / (df + t2);
    return dk_ibeta_half(0.5 * df, x, y);
}

// TtestFilter.c:120-150, evaluation order as written there
__device__ double dk_welch_p(double m1, double m2, int n1, int n2, double sd1, double sd2)
{
    if (sd1 == 0 && sd2 == 0) return 0.5;
    double df = sd1 * sd1 / n1 + sd2 * sd2 / n2;
    df *= df;
    double t1 = sd1 * sd1 / n1; t1 *= t1; t1 /= (n1 - 1);
    double t2 = sd2 * sd2 / n2; t2 *= t2; t2 /= (n2 - 1);
    df /= (t1 + t2);
    double t = (m1 - m2) / sqrt(sd1 * sd1 / n1 + sd2 * sd2 / n2);
    return dk_student_t_2tail(fabs(t), df);
}

#define TT_THREADS 128
// group moments of one row; la scratch is [n][TT_THREADS] in shared memory (conflict-free columns)
__device__ __forceinline__ void dk_group_moments(const int32_t *c, const double *nf, const int *idx, int n, double *la,
                                                 double &meanv, double &mlog, double &sd)
{
    double sa = 0, sl = 0;
    for (int i = 0; i < n; i++) {
        int k = idx[i];
        double x = (double)c[k] / nf[k];                 // TtestFilter.c:380-382
        double a = x + 1;                                // :394
        double l = log(x + 1);                           // :395
        sa += a; sl += l;                                // sum() :77-84
        la[i * TT_THREADS] = l;
    }
    meanv = sa / n;                                      // mean() :87-89
    mlog = sl / n;
    double s = 0;
    for (int i = 0; i < n; i++) { double d = la[i * TT_THREADS] - mlog; s += d * d; }   // sd() :91-98, pow(d,2)==d*d
    sd = sqrt(s / (n - 1));
}

__global__ void __launch_bounds__(TT_THREADS)
k_ttest_rows(const int32_t *cnts, const uint32_t *order, const uint64_t *lbase, uint32_t n_chunks, int S,
             const double *nf_g, const int *idx_g, int nA, int nB, double *p_raw, double *log2fc,
             unsigned long long *tie_sd)
{
    extern __shared__ __align__(16) unsigned char sm[];
    int nmax = nA > nB ? nA : nB;
    double *la = (double *)sm;                                   // [nmax][TT_THREADS]
    double *nf = la + (size_t)nmax * TT_THREADS;                 // [S]
    int *idx = (int *)(nf + S);                                  // [nA+nB]
    for (int i = threadIdx.x; i < S; i += TT_THREADS) nf[i] = nf_g[i];
    for (int i = threadIdx.x; i < nA + nB; i += TT_THREADS) idx[i] = idx_g[i];
    __syncthreads();
    uint32_t oc = blockIdx.x * (TT_THREADS / DK_CH) + threadIdx.x / DK_CH;
    uint32_t j = threadIdx.x % DK_CH;
    if (oc >= n_chunks) return;
    uint64_t b0 = lbase[oc], b1 = lbase[oc + 1];
    if (j >= b1 - b0) return;
    const int32_t *c = cnts + ((size_t)order[oc] * DK_CH + j) * (size_t)S;
    double *myla = la + threadIdx.x;
    double mA, mB, m1, m2, sd1, sd2;
    dk_group_moments(c, nf, idx, nA, myla, mA, m1, sd1);
    dk_group_moments(c, nf, idx + nA, nB, myla, mB, m2, sd2);
    double p = dk_welch_p(m1, m2, nA, nB, sd1, sd2);
    uint64_t row = b0 + j;
    p_raw[row] = p;
    log2fc[row] = log(mB / mA) / log(2.0);                       // TtestFilter.c:72-75,433
    // libm-sensitive zero test: both sds ~0 but the logs are not all exactly 0
    const double eps = 1e-12;
    if (sd1 <= eps * (1 + fabs(m1)) && sd2 <= eps * (1 + fabs(m2)) && (m1 != 0 || m2 != 0))
        atomicAdd(tie_sd, 1ull);
}

void dk_ttest_release(dk_ctx *ctx, dk_ttest *t)
{
    if (!t) return;
    dk_dev_release(ctx, t->nf_dev); dk_dev_release(ctx, t->idx_dev); dk_dev_release(ctx, t->p_raw);
    dk_dev_release(ctx, t->log2fc); dk_dev_release(ctx, t->p_adj); dk_dev_release(ctx, t->cand_p);
    dk_dev_release(ctx, t->cand_row); dk_dev_release(ctx, t->surv_row); dk_dev_release(ctx, t->tie_counts);
    delete t;
}

extern "C" int dk_ttest_rows(dk_ctx *ctx, const dk_matrix *m, const double *nf, const int *idxA, int nA, const int *idxB,
                             int nB, dk_ttest **out)
{
    if (!ctx || !m || !nf || !idxA || !idxB || !out) return dk_fail(ctx, DK_ERR_ARG, "ttest_rows: NULL argument");
    *out = nullptr;
    const int S = m->S;
    if (nA < 2 || nB < 2)
        return dk_fail(ctx, DK_ERR_UNSUPPORTED,
                       "ttest: groups of %d and %d samples; the reference aborts inside Boost for a group < 2", nA, nB);
    for (int i = 0; i < nA; i++) if (idxA[i] < 0 || idxA[i] >= S) return dk_fail(ctx, DK_ERR_ARG, "ttest: idxA[%d] out of range", i);
    for (int i = 0; i < nB; i++) if (idxB[i] < 0 || idxB[i] >= S) return dk_fail(ctx, DK_ERR_ARG, "ttest: idxB[%d] out of range", i);
    if (m->R >= 0xFFFFFFFFull) return dk_fail(ctx, DK_ERR_ARG, "ttest: more than 2^32 rows on one GPU; shard by prefix");
    cudaStream_t st = ctx->stream;
    DK_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
    dk_ttest *t = new dk_ttest();
    t->m = m; t->R = m->R; t->S = S; t->nA = nA; t->nB = nB;
    t->nf_dev = (double *)dk_dev_alloc(ctx, sizeof(double) * S);
    t->idx_dev = (int *)dk_dev_alloc(ctx, sizeof(int) * (nA + nB));
    t->p_raw = (double *)dk_dev_alloc(ctx, sizeof(double) * (t->R + 1));
    t->log2fc = (double *)dk_dev_alloc(ctx, sizeof(double) * (t->R + 1));
    t->tie_counts = (uint64_t *)dk_dev_alloc(ctx, 64);
    if (!t->nf_dev || !t->idx_dev || !t->p_raw || !t->log2fc || !t->tie_counts) { dk_ttest_release(ctx, t); return DK_ERR_NOMEM; }
    std::vector<int> idx(nA + nB);
    memcpy(idx.data(), idxA, sizeof(int) * nA);
    memcpy(idx.data() + nA, idxB, sizeof(int) * nB);
    DK_CUDA(ctx, cudaMemcpyAsync(t->nf_dev, nf, sizeof(double) * S, cudaMemcpyHostToDevice, st));
    DK_CUDA(ctx, cudaMemcpyAsync(t->idx_dev, idx.data(), sizeof(int) * (nA + nB), cudaMemcpyHostToDevice, st));
    DK_CUDA(ctx, cudaMemsetAsync(t->tie_counts, 0, 64, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));          // idx is a local staging vector
    int nmax = nA > nB ? nA : nB;
    size_t smem = sizeof(double) * ((size_t)nmax * TT_THREADS + S) + sizeof(int) * (nA + nB);
    DK_CUDA(ctx, cudaFuncSetAttribute(k_ttest_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint32_t nch = (uint32_t)m->n_chunks;
    if (nch) {
        unsigned grid = (nch + (TT_THREADS / DK_CH) - 1) / (TT_THREADS / DK_CH);
        k_ttest_rows<<<grid, TT_THREADS, smem, st>>>(m->cnts, m->order, m->lbase, nch, S, t->nf_dev, t->idx_dev, nA, nB,
                                                     t->p_raw, t->log2fc, (unsigned long long *)(t->tie_counts + 2));
        DK_LAUNCHED(ctx);
    }
    DK_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    ctx->stage_ms[2] = ms;
    *out = t;
    return DK_OK;
}
extern "C" uint64_t dk_ttest_rows_count(const dk_ttest *t) { return t ? t->R : 0; }
extern "C" void dk_ttest_free(dk_ctx *ctx, dk_ttest *t) { if (ctx) dk_ttest_release(ctx, t); }
