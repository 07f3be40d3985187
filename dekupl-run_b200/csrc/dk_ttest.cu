// dk_ttest.cu — TtestFilter pass 2 on the GPU (reference: share/TtestFilter/TtestFilter.c:373-440):
// one streaming pass over the count matrix, one thread per row: normalise by NF, +1,
// log, group means / sds of the logs, Welch t and df, two-sided Student-t tail, log2FC.
// Arithmetic mirrors the reference operation by operation (sequential sums in group
// order, no FMA contraction: this file is compiled with -fmad=false).
#include "dk_common.cuh"

// ---- Student-t tail: 2*Q(|t|;v) = I_{v/(v+t^2)}(v/2, 1/2)  (replaces the Boost call at :146-147)
// The per-row cost of this pass is fp64 issue, not HBM (r01 ncu), so the tail is written to avoid
// fp64 divisions: the continued fraction of the incomplete beta is evaluated by the Wallis forward
// recurrence on (numerator, denominator) pairs with the partial quotients d_i = N_i/D_i cleared of
// their denominators (A_{i+1} = D_i A_i + D_{i-1} N_i A_{i-1}), convergence is tested by cross
// multiplication, and Gamma(v+1/2)/Gamma(v) comes from an upward shift to v >= 20 plus a 5-term
// asymptotic series (no lgamma). Checked against scipy.special.stdtr to 4e-13 relative over
// df in [1,2000], |t| up to 1e4 (tests/test_gpu_parity.py).
__device__ double dk_beta_cf(double a, double b, double x)
{
    double Ap = 0.0, A = 1.0, Bp = 1.0, B = 1.0, Dp = 1.0;
    for (int m = 0; m < 2000; m++) {
        const double fm = (double)m, a2 = a + 2.0 * fm;
        double N = -(a + fm) * (a + b + fm) * x, D = a2 * (a2 + 1.0);              // d_{2m+1}
        double t = Dp * N;
        double An = __fma_rn(D, A, t * Ap), Bn = __fma_rn(D, B, t * Bp);     // explicit fma: this file is built -fmad=false
        Ap = A; A = An; Bp = B; B = Bn; Dp = D;
        const double m1 = fm + 1.0, a3 = a2 + 1.0;
        N = m1 * (b - m1) * x; D = a3 * (a3 + 1.0);                                // d_{2m+2}
        t = Dp * N;
        An = __fma_rn(D, A, t * Ap); Bn = __fma_rn(D, B, t * Bp);
        const bool conv = fabs(__fma_rn(An, B, -(A * Bn))) <= 1e-15 * fabs(Bn * B);
        Ap = A; A = An; Bp = B; B = Bn; Dp = D;
        if (fabs(B) > 1e100) { Ap *= 1e-100; A *= 1e-100; Bp *= 1e-100; B *= 1e-100; }
        if (conv) break;
    }
    return A / B;
}

// I_x(v, 1/2) with y = 1-x supplied by the caller (no cancellation)
__device__ double dk_ibeta_half(double v, double x, double y)
{
    if (x <= 0.0) return 0.0;
    if (y <= 0.0) return 1.0;
    const double INV_SQRT_PI = 0.56418958354775628695;
    const double lx = (y < 0.5) ? log1p(-y) : log(x);
    // Gamma(v+1/2)/Gamma(v) = sqrt(w) * exp(ser(w)) * prod_{i<n} (v+i)/(v+i+1/2),  w = v+n >= 20
    double num = 1.0, den = 1.0, w = v;
    while (w < 20.0) { num *= w; den *= (w + 0.5); w += 1.0; }
    const double r = 1.0 / w, r2 = r * r;
    const double ser = r * (-1.0 / 8.0 + r2 * (1.0 / 192.0 + r2 * (-1.0 / 640.0 + r2 * (17.0 / 14336.0 + r2 * (-31.0 / 18432.0)))));
    // x^v (1-x)^(1/2) / B(v,1/2)
    const double front = sqrt(w * y) * INV_SQRT_PI * (num / den) * exp(ser + v * lx);
    if (x < (v + 1.0) / (v + 2.5)) return front * dk_beta_cf(v, 0.5, x) / v;
    return 1.0 - front * dk_beta_cf(0.5, v, y) / 0.5;
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

#ifndef TT_THREADS
#define TT_THREADS 1024          // launch bound; the host picks the block size (a multiple of DK_CH) per job
#endif
#ifndef TT_MINB
#define TT_MINB 1
#endif
#define TT_PO_THREADS 128        // Poisson rows kernel
#ifndef TT_TAB_BYTES
#define TT_TAB_BYTES 65536       // (a, l) tables per CTA
#endif
// group moments of one row; la scratch is [n][blockDim.x] in shared memory (conflict-free columns)
// c/nf as the reference computes it (TtestFilter.c:380-382), without a per-element fp64 division:
// with r = RN(1/nf), two Newton corrections through exact fma residuals give the correctly rounded
// quotient (Markstein); the result is a deterministic function of (c, nf), which is all the
// sd1==0 && sd2==0 test needs. Used only when every nf is a normal number in [1e-30, 1e30].
__device__ __forceinline__ double dk_div_by_nf(double c, double nf, double r)
{
    double q = c * r;
    double e = __fma_rn(-nf, q, c);
    q = __fma_rn(e, r, q);
    e = __fma_rn(-nf, q, c);
    return __fma_rn(e, r, q);
}

// Small counts dominate real and synthetic dumps, and (count, sample) determines x = count/NF exactly, so every
// persistent block tabulates a = x+1 and l = log(x+1) for count < TN per sample in shared memory, with the SAME
// device code as the direct path (bit-identical results); larger counts take the direct path.
template <bool FASTDIV>
__device__ __forceinline__ void dk_norm_log(int32_t cc, double nf, double rnf, double &a, double &l)
{
    double x = FASTDIV ? dk_div_by_nf((double)cc, nf, rnf) : (double)cc / nf;   // TtestFilter.c:380-382
    a = x + 1;                                       // :394
    l = log(x + 1);                                  // :395
}

template <bool FASTDIV>
__device__ __noinline__ void dk_group_moments(const int32_t *c, const double *nf, const double *rnf, const int *idx, int n,
                                              const double2 *tab, uint32_t TN, double *la, double &meanv, double &mlog, double &sd)
{
    double sa = 0, sl = 0;
    const uint32_t T = blockDim.x;
    for (int i = 0; i < n; i++) {
        const int k = idx[i];
        const int32_t cc = c[k];
        double a, l;
        if ((uint32_t)cc < TN) { const double2 t = tab[(uint32_t)k * TN + (uint32_t)cc]; a = t.x; l = t.y; }
        else dk_norm_log<FASTDIV>(cc, nf[k], rnf[k], a, l);
        sa += a; sl += l;                                // sum() :77-84
        la[i * T] = l;
    }
    meanv = sa / n;                                      // mean() :87-89
    mlog = sl / n;
    double s = 0;
    for (int i = 0; i < n; i++) { double d = la[i * T] - mlog; s += d * d; }   // sd() :91-98, pow(d,2)==d*d
    sd = sqrt(s / (n - 1));
}

template <bool FASTDIV>
__global__ void __launch_bounds__(TT_THREADS, TT_MINB)
k_ttest_rows(const int32_t *cnts, const uint32_t *order, const uint64_t *lbase, uint32_t n_chunks, int S,
             const double *nf_g, const int *idx_g, int nA, int nB, uint32_t TN, double *p_raw, double *log2fc,
             unsigned long long *tie_sd)
{
    extern __shared__ __align__(16) unsigned char sm[];
    int nmax = nA > nB ? nA : nB;
    const uint32_t T = blockDim.x, cpb = T / DK_CH;              // block size chosen by the host; chunks per block
    double2 *tab = (double2 *)sm;                                // [S][TN] (a, l)
    double *la = (double *)(tab + (size_t)S * TN);               // [nmax][T]
    double *nf = la + (size_t)nmax * T;                          // [S]
    double *rnf = nf + S;                                        // [S] RN(1/nf)
    int *idx = (int *)(rnf + S);                                 // [nA+nB]
    for (int i = threadIdx.x; i < S; i += T) { nf[i] = nf_g[i]; rnf[i] = 1.0 / nf_g[i]; }
    for (int i = threadIdx.x; i < nA + nB; i += T) idx[i] = idx_g[i];
    __syncthreads();
    for (uint32_t e = threadIdx.x; e < (uint32_t)S * TN; e += T) {
        const uint32_t k = e / TN, cc = e - k * TN;
        double a, l;
        dk_norm_log<FASTDIV>((int32_t)cc, nf[k], rnf[k], a, l);
        tab[e] = make_double2(a, l);
    }
    __syncthreads();
    double *myla = la + threadIdx.x;
    const uint32_t nblk = (n_chunks + cpb - 1) / cpb;
    unsigned long long ties = 0;
    for (uint32_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {      // persistent: the tables are built once per CTA
        const uint32_t oc = blk * cpb + threadIdx.x / DK_CH;
        const uint32_t j = threadIdx.x % DK_CH;
        if (oc >= n_chunks) continue;
        const uint64_t b0 = lbase[oc], b1 = lbase[oc + 1];
        if (j >= b1 - b0) continue;
        const int32_t *c = cnts + ((size_t)order[oc] * DK_CH + j) * (size_t)S;
        double mA, mB, m1, m2, sd1, sd2;
        dk_group_moments<FASTDIV>(c, nf, rnf, idx, nA, tab, TN, myla, mA, m1, sd1);
        dk_group_moments<FASTDIV>(c, nf, rnf, idx + nA, nB, tab, TN, myla, mB, m2, sd2);
        const double p = dk_welch_p(m1, m2, nA, nB, sd1, sd2);
        const uint64_t row = b0 + j;
        p_raw[row] = p;
        log2fc[row] = log(mB / mA) / log(2.0);                   // TtestFilter.c:72-75,433
        // libm-sensitive zero test: both sds ~0 but the logs are not all exactly 0
        const double eps = 1e-12;
        if (sd1 <= eps * (1 + fabs(m1)) && sd2 <= eps * (1 + fabs(m2)) && (m1 != 0 || m2 != 0)) ties++;
    }
    if (ties) atomicAdd(tie_sd, ties);
}

// ---- fast row kernel (groups in column order, <= TT8_N samples per group, S % 4 == 0, S <= 2 * TT8_N) ----------
// r2e ncu of k_ttest_rows at cfg 2 (profiles/r02_ttest_before.txt): 2937 warp instructions per 32 rows at 19.5 active
// lanes. 26 % of them are a SECOND copy of the continued-fraction loop (the x < (v+1)/(v+2.5) branch of the incomplete
// beta, taken by 12 % of the rows, running with 3.8 lanes), 5 % the direct log path of counts beyond the shared-memory
// table (1 lane), and 60 % straight-line fp64 code: 25 divisions per row (18 of them by the group sizes, which are
// constants), 4 square roots, 2 logs, an exp. This kernel computes the same values with
//   * ONE continued-fraction loop: the branch only selects its parameters (a, b, x), so all lanes iterate together;
//   * a second-level table in global memory (count < TT8_TN2) behind the shared-memory one: a miss is one 16-byte load;
//   * reciprocals of the group-size constants, sd^2 taken straight from the sum of squares (no sqrt followed by a
//     square), t^2 instead of t, x and y from one reciprocal, the Gamma ratio's denominator and the fraction's B folded
//     into the final division: 7 divisions, 1 square root per row;
//   * the logs of a group kept in registers (no shared-memory scratch), so two 448-thread CTAs share an SM;
//   * (r02final ncu: 70 % of the instructions sat before the sort, three quarters of them control and moves) both table
//     levels read without a branch, one log for the fraction's front factor, log2FC left to dk_ttest_need_log2fc.
// Arithmetic that decides exact comparisons is untouched: the logs and their sequential sums, mean = sum / n by a true
// division, the sum of squared deviations (sd == 0 <=> that sum == 0; TtestFilter.c:128). Everything else moves by
// a few ulp at most, far inside the 1e-9 contract (tests/test_gpu_parity.py).
#define TT8_N 8
#ifndef TT8_THREADS
#define TT8_THREADS 448
#endif
#ifndef TT8_TN
#define TT8_TN 256
#endif
#define TT8_TN2 4096
// the small-group kernel's second-level table is larger (L2-resident: S * tn2 * 16 bytes <= 16 MB) so that a row with a
// count beyond it — the only case that leaves the branch-free look-ups — is rare even for a heavy-tailed spectrum
#define TT8_TN2F 65536
#ifndef TT8_FAST
#define TT8_FAST 1        // 0: branchy three-level look-up; 1: the two table levels branch-free (a single generic load
                          // for both levels was measured too: 15.8 against 15.5 ms)
#endif
#ifndef TT8_ONELOG
#define TT8_ONELOG 1
#endif
#ifndef TT8_GW
#define TT8_GW 8.0
#endif
// log2FC is read at the rows that pass the p-value cut only (TtestFilter.c:486-489 prints it for the survivors): the
// small-group kernel leaves it out — no second log, no division, no sums of the normalised counts, 8 bytes per row less
// to write — and dk_ttest_need_log2fc computes it where and when somebody reads it, from the same look-up arithmetic
#ifndef TT8_LAZYFC
#define TT8_LAZYFC 1
#endif
// (a shared-memory table of the logs alone, 8-byte entries for twice as many counts, was measured slower: 14.6 against
// 14.2 ms at cfg 2)

// continued fraction of I_x(a,b) (Wallis form, see dk_beta_cf): returns A and B unreduced
__device__ __forceinline__ void dk_beta_cf2(double a, double b, double x, double &Ao, double &Bo)
{
    double Ap = 0.0, A = 1.0, Bp = 1.0, B = 1.0, Dp = 1.0;
    const double ab = a + b;
    for (int m = 0; m < 2000; m++) {
        const double fm = (double)m, a2 = a + 2.0 * fm;
        double N = -(a + fm) * (ab + fm) * x, D = a2 * (a2 + 1.0);               // d_{2m+1}
        double t = Dp * N;
        double An = __fma_rn(D, A, t * Ap), Bn = __fma_rn(D, B, t * Bp);
        Ap = A; A = An; Bp = B; B = Bn; Dp = D;
        const double m1 = fm + 1.0, a3 = a2 + 1.0;
        N = m1 * (b - m1) * x; D = a3 * (a3 + 1.0);                              // d_{2m+2}
        t = Dp * N;
        An = __fma_rn(D, A, t * Ap); Bn = __fma_rn(D, B, t * Bp);
        const bool conv = fabs(__fma_rn(An, B, -(A * Bn))) <= 1e-15 * fabs(Bn * B);
        Ap = A; A = An; Bp = B; B = Bn; Dp = D;
        if (fabs(B) > 1e100) { Ap *= 1e-100; A *= 1e-100; Bp *= 1e-100; B *= 1e-100; }
        if (conv) break;
    }
    Ao = A; Bo = B;
}

struct TT8Consts { double rnA, rnB, rnA1, rnB1; int p2A, p2B; };   // 1/nA, 1/nB, 1/(nA-1), 1/(nB-1); group size is a power of two

__global__ void k_tt8_table(const double *nf_g, int S, double2 *tab2, uint32_t tn2)
{
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (uint32_t)S * tn2) return;
    const uint32_t k = e / tn2, cc = e - k * tn2;
    const double nf = nf_g[k];
    double a, l;
    dk_norm_log<true>((int32_t)cc, nf, 1.0 / nf, a, l);
    tab2[e] = make_double2(a, l);
}

__device__ __forceinline__ void tt8_lookup(int32_t cc, int k, const double2 *tab, const double2 *tab2, const double *nf,
                                           const double *rnf, double &a, double &l, const uint32_t tn = TT8_TN,
                                           const uint32_t tn2 = TT8_TN2)
{
    if ((uint32_t)cc < tn) { const double2 t = tab[(uint32_t)k * tn + (uint32_t)cc]; a = t.x; l = t.y; }
    else if ((uint32_t)cc < tn2) { const double2 t = __ldg(tab2 + (size_t)k * tn2 + (uint32_t)cc); a = t.x; l = t.y; }
    else dk_norm_log<true>(cc, nf[k], rnf[k], a, l);
}
// the small-group kernel's form: the two table levels without a branch (a clamped shared-memory load, a predicated
// second-level load), so the 16 look-ups of a row are independent loads the scheduler keeps in flight together (r02final
// ncu: 16 % of the first phase's stall samples were loads, and the three-way branch cost ~8 control instructions per
// look-up); only a count beyond the second-level table (or a negative one) branches to the direct computation
#ifndef TT8_DIRECTCALL
#define TT8_DIRECTCALL 1
#endif
__device__ __noinline__ void tt8_direct(int32_t cc, double nf, double rnf, double *o)
{
    double a, l;
    dk_norm_log<true>(cc, nf, rnf, a, l);
    o[0] = a; o[1] = l;
}
__device__ __forceinline__ void tt8_lookup_fast(int32_t cc, int k, const double2 *tab, const double2 *tab2, const uint32_t tn2,
                                                const double *nf, const double *rnf, double &a, double &l)
{
    const uint32_t u = (uint32_t)cc;
    double2 t = tab[(uint32_t)k * TT8_TN + min(u, (uint32_t)TT8_TN - 1u)];
    if (u >= (uint32_t)TT8_TN) t = __ldg(tab2 + (size_t)k * tn2 + min(u, tn2 - 1u));
    a = t.x; l = t.y;
    if (u >= tn2) {
#if TT8_DIRECTCALL
        double o[2];
        tt8_direct(cc, nf[k], rnf[k], o);
        a = o[0]; l = o[1];
#else
        dk_norm_log<true>(cc, nf[k], rnf[k], a, l);
#endif
    }
}

// group moments in two passes over the row's counts with the general look-up (large groups)
__device__ __forceinline__ void tt8_moments_2pass(const int32_t *c, int nA, int nB, const double2 *tab, uint32_t tn, const double2 *tab2,
                                                  uint32_t tn2, const double *nf, const double *rnf, const TT8Consts &K, double &sa,
                                                  double &sb, double &m1, double &m2, double &s1, double &s2)
{
    double sl = 0;
    for (int i = 0; i < nA; i++) { double a, l; tt8_lookup(__ldg(c + i), i, tab, tab2, nf, rnf, a, l, tn, tn2); sa += a; sl += l; }
    m1 = K.p2A ? sl * K.rnA : sl / nA;
    for (int i = 0; i < nA; i++) { double a, l; tt8_lookup(__ldg(c + i), i, tab, tab2, nf, rnf, a, l, tn, tn2); const double d = l - m1; s1 += d * d; }
    sl = 0;
    for (int i = 0; i < nB; i++) { double a, l; tt8_lookup(__ldg(c + nA + i), nA + i, tab, tab2, nf, rnf, a, l, tn, tn2); sb += a; sl += l; }
    m2 = K.p2B ? sl * K.rnB : sl / nB;
    for (int i = 0; i < nB; i++) { double a, l; tt8_lookup(__ldg(c + nA + i), nA + i, tab, tab2, nf, rnf, a, l, tn, tn2); const double d = l - m2; s2 += d * d; }
}

#ifndef TT8_MINB
#define TT8_MINB 2
#endif
#define TT8_BINS 64
// The continued fraction needs 1 to ~20 steps depending on where x sits between 0 and the switch point, so rows that
// share a warp by row order keep 12 of 32 lanes busy in its loop (r2g ncu: 44 % of the kernel's instructions). Each
// tile of rows is therefore processed in two phases: (1) one thread per row does everything up to the fraction's
// arguments and queues (x, v, front, denominator, row); the queue is ordered by a 64-bin counting sort on x (steps
// needed grow with x, per branch); (2) thread i evaluates the fraction of queue entry i, so a warp holds rows of
// similar x. The values computed per row do not depend on the order.
// VEC (S % 4 == 0, nA % 4 == 0): the tile's rows are fetched by cp.async one tile ahead into a per-thread 64-byte
// slot (r2i ncu: after the re-binning the kernel waited on its 16 dependent 4-byte count loads, long scoreboard
// 3.2 per issue), and read back with 16-byte shared loads (slot segments swizzled: conflict-free)
// BIG (groups of more than TT8_N samples, cfg 3 / cfg 4): the logs of a group do not fit registers, so the moments take
// two passes over the row's counts (means, then squared deviations: the table look-ups are cheap next to the fp64 work)
// and the shared-memory table holds tn <= TT8_TN counts per sample, sized by the host.
template <bool VEC, bool BIG>
__global__ void __launch_bounds__(TT8_THREADS, TT8_MINB)
k_ttest_rows8(const int32_t *cnts, const uint32_t *order, const uint64_t *lbase, uint32_t n_chunks, int S, const double *nf_g,
              const double2 *tab2, int nA, int nB, TT8Consts K, double *p_raw, double *log2fc, unsigned long long *tie_sd, uint32_t tn_arg, uint32_t tn2)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const uint32_t TNv = BIG ? tn_arg : (uint32_t)TT8_TN;
    double2 *tab = (double2 *)sm;                                // [S][TNv] (a, l)
    double *nf = (double *)(tab + (size_t)S * TNv);              // [S]
    double *rnf = nf + S;                                        // [S]
    const uint32_t T = blockDim.x, cpb = T / DK_CH;
    double *q_x = rnf + S, *q_v = q_x + T, *q_f = q_v + T, *q_d = q_f + T;   // [T] each: queue of fraction arguments
    unsigned long long *q_row = (unsigned long long *)(q_d + T);             // [T] row | lower << 63
    uint32_t *hist = (uint32_t *)(q_row + T);                                // [BINS] + [BINS] bases + total
    uint32_t *base = hist + TT8_BINS;
    int32_t *slot = (int32_t *)(hist + 2 * TT8_BINS + 4) + (size_t)threadIdx.x * 16;   // VEC: my row of the tile in flight
    const uint32_t swz = (threadIdx.x >> 1) & 3u;
    // VEC: start the copy of my row of tile `b` (nothing for a thread without a row there)
    auto fetch = [&](uint32_t b) {
        const uint32_t oc = b * cpb + threadIdx.x / DK_CH, j = threadIdx.x % DK_CH;
        if (b < (n_chunks + cpb - 1) / cpb && oc < n_chunks && j < lbase[oc + 1] - lbase[oc]) {
            const int32_t *src = cnts + ((size_t)order[oc] * DK_CH + j) * (size_t)S;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(slot);
            for (int q = 0; q < S / 4; q++)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + ((q ^ swz) << 4)), "l"(src + 4 * q) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int i = threadIdx.x; i < S; i += T) { nf[i] = nf_g[i]; rnf[i] = 1.0 / nf_g[i]; }
    for (uint32_t e = threadIdx.x; e < (uint32_t)S * TNv; e += T) {
        const uint32_t k = e / TNv, cc = e - k * TNv;
        tab[e] = tab2[(size_t)k * tn2 + cc];
    }
    __syncthreads();
    const uint32_t nblk = (n_chunks + cpb - 1) / cpb;
    unsigned long long ties = 0;
    if (VEC) fetch(blockIdx.x);
    for (uint32_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
        if (threadIdx.x < TT8_BINS) hist[threadIdx.x] = 0;
        const uint32_t oc = blk * cpb + threadIdx.x / DK_CH;
        const uint32_t j = threadIdx.x % DK_CH;
        bool act = oc < n_chunks;
        uint64_t b0 = 0;
        if (act) { b0 = lbase[oc]; act = j < lbase[oc + 1] - b0; }
        bool needcf = false;
        uint32_t bin = 0;
        double c_x = 0, c_v = 0, c_f = 0, c_d = 1;
        bool lower = false;
        if (act) {
            const int32_t *c = cnts + ((size_t)order[oc] * DK_CH + j) * (size_t)S;
            double sa = 0, sb = 0, m1 = 0, m2 = 0, s1 = 0, s2 = 0;
            if constexpr (!BIG) {
                int32_t cA[TT8_N], cB[TT8_N];
                if (VEC) {
                    asm volatile("cp.async.wait_group 0;" ::: "memory");           // my own copy: no barrier needed
                    const int4 *s4 = reinterpret_cast<const int4 *>(slot);
                    const uint32_t qb = (uint32_t)nA >> 2;
                    const int4 a0 = s4[0 ^ swz], a1 = s4[1 ^ swz], b0v = s4[(qb & 3u) ^ swz], b1v = s4[((qb + 1u) & 3u) ^ swz];
                    cA[0] = a0.x; cA[1] = a0.y; cA[2] = a0.z; cA[3] = a0.w; cA[4] = a1.x; cA[5] = a1.y; cA[6] = a1.z; cA[7] = a1.w;
                    cB[0] = b0v.x; cB[1] = b0v.y; cB[2] = b0v.z; cB[3] = b0v.w; cB[4] = b1v.x; cB[5] = b1v.y; cB[6] = b1v.z; cB[7] = b1v.w;
                } else {
    #pragma unroll
                    for (int i = 0; i < TT8_N; i++) { cA[i] = i < nA ? __ldg(c + i) : 0; cB[i] = i < nB ? __ldg(c + nA + i) : 0; }
                }
                {
#if TT8_FAST
#define TT8_LK(cc, k, a, l) tt8_lookup_fast(cc, k, tab, tab2, tn2, nf, rnf, a, l)
#else
#define TT8_LK(cc, k, a, l) tt8_lookup(cc, k, tab, tab2, nf, rnf, a, l, TT8_TN, tn2)
#endif
                // ---- group moments (TtestFilter.c:379-411; sum/mean/sd :77-98), sums sequential in column order
                double la[TT8_N], sl = 0;
    #pragma unroll
                for (int i = 0; i < TT8_N; i++)
                    if (i < nA) { double a, l; TT8_LK(cA[i], i, a, l); sa += a; sl += l; la[i] = l; }
                m1 = K.p2A ? sl * K.rnA : sl / nA;    // mean() of the logs decides sd == 0: a true division, or for a group of 2^k
                                                      // samples the exact multiply by 2^-k (same bits)
    #pragma unroll
                for (int i = 0; i < TT8_N; i++) if (i < nA) { const double d = la[i] - m1; s1 += d * d; }
                sl = 0;
    #pragma unroll
                for (int i = 0; i < TT8_N; i++)
                    if (i < nB) { double a, l; TT8_LK(cB[i], nA + i, a, l); sb += a; sl += l; la[i] = l; }
                m2 = K.p2B ? sl * K.rnB : sl / nB;
    #pragma unroll
                for (int i = 0; i < TT8_N; i++) if (i < nB) { const double d = la[i] - m2; s2 += d * d; }
                }
#undef TT8_LK
            }
            if constexpr (BIG)      // two passes (TtestFilter.c:379-411; sum/mean/sd :77-98): the same sums in the same order
                tt8_moments_2pass(c, nA, nB, tab, TNv, tab2, tn2, nf, rnf, K, sa, sb, m1, m2, s1, s2);
            // ---- Welch (TtestFilter.c:120-150) on sd^2 = s/(n-1)
            double p = 0.5;                                      // sd1 == 0 && sd2 == 0 (:128)
            if (!(s1 == 0 && s2 == 0)) {
                const double q1 = s1 * K.rnA1, q2 = s2 * K.rnB1; // sd1^2, sd2^2
                const double v1 = q1 * K.rnA, v2 = q2 * K.rnB, vv = v1 + v2;
                const double den = v1 * v1 * K.rnA1 + v2 * v2 * K.rnB1;
                const double df = vv * vv / den;                 // :132-138
                const double dm = m1 - m2, dm2 = dm * dm;        // t^2 = dm2 / vv (:139-143)
                if (isinf(df) || !(den > 0)) p = dk_student_t_2tail(fabs(dm / sqrt(vv)), df);  // degenerate variances: the old route
                else {
                    // x = df/(df+t^2), y = t^2/(df+t^2) from one reciprocal
                    const double dv = df * vv, r2 = 1.0 / (dv + dm2);
                    const double x = dv * r2, y = dm2 * r2, v = 0.5 * df;
                    if (!(x > 0.0)) p = 0.0;
                    else if (!(y > 0.0)) p = 1.0;
                    else {
                        const double INV_SQRT_PI = 0.56418958354775628695;
#if TT8_ONELOG
                        // one log for both halves: for x > 1/2 its absolute error (an ulp of x, ~1e-16) enters exp(v * lx)
                        // as a relative 1e-16 * v, the same order as everything else here; the log1p/log pair ran BOTH
                        // branches in nearly every warp
                        const double lx = log(x);
#else
                        const double lx = (y < 0.5) ? log1p(-y) : log(x);
#endif
                        // Gamma(v+1/2)/Gamma(v): shift up to w >= 12, then the asymptotic series with 6 terms (the next one,
                        // 5461/(65536*7) / w^13, is below 1e-16 there; stopping at 12 instead of 20 saves 8 trips of the loop)
                        double num = 1.0, dg = 1.0, w = v;
                        while (w < TT8_GW) { num *= w; dg *= (w + 0.5); w += 1.0; }
                        const double r = 1.0 / w, rr = r * r;
                        const double ser = r * (-1.0 / 8.0 + rr * (1.0 / 192.0 + rr * (-1.0 / 640.0 + rr * (17.0 / 14336.0 +
                                           rr * (-31.0 / 18432.0 + rr * (691.0 / 180224.0))))));
                        lower = x * (v + 2.5) < (v + 1.0);       // x < (v+1)/(v+2.5)
                        c_x = lower ? x : y;
                        c_v = v;
                        c_f = sqrt(w * y) * INV_SQRT_PI * num * exp(ser + v * lx);   // (its 1/dg joins the last division)
                        c_d = dg * (lower ? v : 0.5);
                        needcf = true;
                        const int bq = lower ? (int)(c_x * 38.0) : (int)(c_x * 80.0);
                        bin = (lower ? 32u : 0u) + (uint32_t)(bq > 31 ? 31 : bq);
                    }
                }
            }
            const uint64_t row = b0 + j;
            if (!needcf) p_raw[row] = p;
            if (BIG || !TT8_LAZYFC)
                log2fc[row] = log((sb * K.rnB) / (sa * K.rnA)) * 1.4426950408889634074;   // log(meanB/meanA)/log(2) (:72-75,433)
            const double eps2 = 1e-24;
            const double e1 = 1 + fabs(m1), e2 = 1 + fabs(m2);
            if (s1 * K.rnA1 <= eps2 * e1 * e1 && s2 * K.rnB1 <= eps2 * e2 * e2 && (m1 != 0 || m2 != 0)) ties++;
        }
        if (VEC) fetch(blk + gridDim.x);                         // my slot is free: the next tile's row starts flying
        __syncthreads();                                         // hist is zero
        uint32_t inbin = 0;
        if (needcf) inbin = atomicAdd(&hist[bin], 1u);
        __syncthreads();
        if (threadIdx.x < 32) {                                  // exclusive prefix of the 64 bins
            const uint32_t h0 = hist[2 * threadIdx.x], h1 = hist[2 * threadIdx.x + 1];
            uint32_t x = h0 + h1;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)threadIdx.x >= o) x += y; }
            base[2 * threadIdx.x] = x - h0 - h1;
            base[2 * threadIdx.x + 1] = x - h1;
            if (threadIdx.x == 31) base[TT8_BINS] = x;
        }
        __syncthreads();
        if (needcf) {
            const uint32_t pos = base[bin] + inbin;
            q_x[pos] = c_x; q_v[pos] = c_v; q_f[pos] = c_f; q_d[pos] = c_d;
            q_row[pos] = (b0 + j) | (lower ? 1ull << 63 : 0ull);
        }
        __syncthreads();
        if (threadIdx.x < base[TT8_BINS]) {
            const uint32_t i = threadIdx.x;
            const unsigned long long rw = q_row[i];
            const bool lw = (rw >> 63) != 0;
            const double v = q_v[i];
            double A, B;
            dk_beta_cf2(lw ? v : 0.5, lw ? 0.5 : v, q_x[i], A, B);
            const double res = q_f[i] * A / (q_d[i] * B);
            p_raw[rw & ~(1ull << 63)] = lw ? res : 1.0 - res;
        }
        // (the next tile's first barrier orders these reads before the queue is rewritten)
    }
    if (ties) atomicAdd(tie_sd, ties);
}

// log2FC of the listed rows (rows != NULL: n logical rows) or of every row (n = 64 * n_chunks slots) of a table whose row
// kernel left it out: the same sums of normalised counts in column order and the same expression as k_ttest_rows8
__global__ void k_tt8_log2fc(const int32_t *cnts, const uint32_t *order, const uint64_t *lbase, uint32_t n_chunks, int S,
                             const double *nf_g, int nA, int nB, double rnA, double rnB, const uint32_t *rows, uint64_t n,
                             double *log2fc)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t row;
    size_t ph;
    if (rows) {
        row = rows[i];
        uint32_t lo = 0, hi = n_chunks;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (lbase[mid] <= row) lo = mid; else hi = mid;
        }
        ph = (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
    } else {
        const uint32_t oc = (uint32_t)(i / DK_CH), j = (uint32_t)(i % DK_CH);
        const uint64_t b0 = lbase[oc];
        if (j >= lbase[oc + 1] - b0) return;
        row = b0 + j;
        ph = (size_t)order[oc] * DK_CH + j;
    }
    const int32_t *c = cnts + ph * (size_t)S;
    double sa = 0, sb = 0;
    for (int k = 0; k < nA; k++) { const double nf = nf_g[k]; double a, l; dk_norm_log<true>(c[k], nf, 1.0 / nf, a, l); sa += a; }
    for (int k = 0; k < nB; k++) { const double nf = nf_g[nA + k]; double a, l; dk_norm_log<true>(c[nA + k], nf, 1.0 / nf, a, l); sb += a; }
    log2fc[row] = log((sb * rnB) / (sa * rnA)) * 1.4426950408889634074;
}

int dk_ttest_need_log2fc(dk_ctx *ctx, const dk_ttest *t, const uint32_t *rows, uint64_t n)
{
    if (!t->fc_lazy || !t->R) return DK_OK;
    const dk_matrix *m = t->m;
    cudaStream_t st = ctx->stream;
    if (!rows) n = (uint64_t)m->n_chunks * DK_CH;
    if (n) {
        k_tt8_log2fc<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m->cnts, m->order, m->lbase, (uint32_t)m->n_chunks, t->S, t->nf_dev,
                                                                 t->nA, t->nB, 1.0 / t->nA, 1.0 / t->nB, rows, n, t->log2fc);
        DK_LAUNCHED(ctx);
    }
    if (!rows) t->fc_lazy = false;                               // every row holds its value now
    return DK_OK;
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
    // fast kernel: groups in column order (what a conditions file listing the samples in matrix order gives), small groups
    bool ident = true;
    for (int i = 0; i < nA; i++) if (idxA[i] != i) ident = false;
    for (int i = 0; i < nB; i++) if (idxB[i] != nA + i) ident = false;
    bool nf_ok = true;
    for (int s = 0; s < S; s++) if (!(nf[s] >= 1e-30 && nf[s] <= 1e30)) nf_ok = false;
    const char *force_old = getenv("DK_TTEST_OLD");
    const bool small_groups = nA <= TT8_N && nB <= TT8_N;
    uint32_t tn_big = TT8_TN;                                        // large groups: the table shrinks until two CTAs fit an SM
    while (tn_big > 16 && (size_t)S * tn_big * 16 + 16 * S + (size_t)TT8_THREADS * 40 + 2048 > (size_t)(227 * 1024) / TT8_MINB - 1024) tn_big >>= 1;
    const bool fits = small_groups ? (size_t)S * TT8_TN * 16 + 16 * S + (size_t)TT8_THREADS * 104 + 1024 <= (size_t)(227 * 1024) / TT8_MINB - 1024
                                   : (size_t)S * tn_big * 16 + 16 * S + (size_t)TT8_THREADS * 40 + 2048 <= (size_t)(227 * 1024) / TT8_MINB - 1024;
    if (ident && nf_ok && nA + nB <= S && !(force_old && atoi(force_old)) && fits) {
        // second-level table: 4096 counts per sample; for a large matrix of small groups as many as 16 MB hold (L2-resident)
        uint32_t tn2 = TT8_TN2;
        if (TT8_FAST && small_groups && t->R >= (1u << 20))
            while (tn2 < (uint32_t)TT8_TN2F && (size_t)S * (tn2 * 2) * sizeof(double2) <= ((size_t)16 << 20)) tn2 *= 2;
        double2 *tab2 = (double2 *)dk_dev_alloc(ctx, sizeof(double2) * (size_t)S * tn2);
        if (!tab2) { dk_ttest_release(ctx, t); return DK_ERR_NOMEM; }
        k_tt8_table<<<(unsigned)(((size_t)S * tn2 + 255) / 256), 256, 0, st>>>(t->nf_dev, S, tab2, tn2); DK_LAUNCHED(ctx);
        const bool vec = small_groups && (S % 4 == 0) && (nA % 4 == 0) && S <= 16;
        auto k8 = !small_groups ? k_ttest_rows8<false, true> : (vec ? k_ttest_rows8<true, false> : k_ttest_rows8<false, false>);
        const uint32_t tn8 = small_groups ? (uint32_t)TT8_TN : tn_big;
        const size_t smem8 = (size_t)S * tn8 * 16 + 16 * S + (size_t)TT8_THREADS * 40 + (2 * TT8_BINS + 4) * 4 + (vec ? (size_t)TT8_THREADS * 64 : 0);
        DK_CUDA(ctx, cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem8));
        const uint32_t nch8 = (uint32_t)m->n_chunks;
        if (nch8) {
            const unsigned cpb = TT8_THREADS / DK_CH;
            unsigned grid = (nch8 + cpb - 1) / cpb;
            int per_sm = 1;
            DK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k8, TT8_THREADS, smem8));
            if (per_sm < 1) per_sm = 1;
            if (grid > (unsigned)(ctx->sm_count * per_sm)) grid = (unsigned)(ctx->sm_count * per_sm);
            TT8Consts K;
            K.rnA = 1.0 / nA; K.rnB = 1.0 / nB; K.rnA1 = 1.0 / (nA - 1); K.rnB1 = 1.0 / (nB - 1);
            K.p2A = (nA & (nA - 1)) == 0; K.p2B = (nB & (nB - 1)) == 0;
            k8<<<grid, TT8_THREADS, smem8, st>>>(m->cnts, m->order, m->lbase, nch8, S, t->nf_dev, tab2, nA, nB, K,
                                                            t->p_raw, t->log2fc, (unsigned long long *)(t->tie_counts + 2), tn8, tn2);
            DK_LAUNCHED(ctx);
            t->fc_lazy = TT8_LAZYFC && small_groups;             // log2FC on demand (dk_ttest_need_log2fc)
        }
        DK_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
        dk_dev_release(ctx, tab2);
        float ms8 = 0;
        cudaEventElapsedTime(&ms8, ctx->ev[0], ctx->ev[1]);
        ctx->stage_ms[2] = ms8;
        *out = t;
        return DK_OK;
    }
    int nmax = nA > nB ? nA : nB;
    // Block size and (a, l) table size. The kernel is latency-bound (dependent fp64 chains, scattered count loads:
    // r01j ncu), so resident warps matter most: take the largest block whose per-thread scratch (nmax doubles) still
    // leaves room for a useful table (>= 32 counts per sample, up to TT_TAB_BYTES).
    uint32_t TN = 0, T = 0;
    size_t smem = 0;
    size_t budget = ctx->smem_optin ? ctx->smem_optin : 200 * 1024;
    if (TT_MINB > 1) budget = (budget + 1024) / TT_MINB - 1024;      // TT_MINB blocks share the SM (1 KB reserved per block)
    for (uint32_t want = 32; want >= 4 && !T; want >>= 3)            // second pass: any table at all (very wide matrices)
        for (uint32_t tt = TT_THREADS; tt >= 64 && !T; tt = tt > 1024 ? 1024 : (tt & (tt - 1)) ? 512 : tt >> 1) {
            const size_t fixed = sizeof(double) * ((size_t)nmax * tt + 2 * S) + sizeof(int) * (nA + nB) + 1024;
            if (fixed >= budget) continue;
            size_t tabmax = budget - fixed < TT_TAB_BYTES ? budget - fixed : TT_TAB_BYTES;
            uint32_t tn = 0;
            for (uint32_t t = 4; t <= 1024 && (size_t)16 * S * t <= tabmax; t <<= 1) tn = t;
            if (tn >= want) { T = tt; TN = tn; smem = fixed - 1024 + (size_t)16 * S * tn; }
        }
    if (!T) { dk_ttest_release(ctx, t); return dk_fail(ctx, DK_ERR_UNSUPPORTED, "ttest: %d samples per group exceed shared memory", nmax); }
    bool fastdiv = true;                                // reciprocal path needs normal, moderate factors
    for (int s = 0; s < S; s++) if (!(nf[s] >= 1e-30 && nf[s] <= 1e30)) fastdiv = false;
    auto kfn = fastdiv ? k_ttest_rows<true> : k_ttest_rows<false>;
    DK_CUDA(ctx, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint32_t nch = (uint32_t)m->n_chunks;
    if (nch) {
        unsigned grid = (nch + (T / DK_CH) - 1) / (T / DK_CH);
        int per_sm = 1;
        DK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, (int)T, smem));
        if (per_sm < 1) per_sm = 1;
        if (grid > (unsigned)(ctx->sm_count * per_sm)) grid = (unsigned)(ctx->sm_count * per_sm);
        kfn<<<grid, T, smem, st>>>(m->cnts, m->order, m->lbase, nch, S, t->nf_dev, t->idx_dev, nA, nB, TN,
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
// ---- Poisson likelihood-ratio mode (TtestFilter -l; TtestFilter.c:152-185,340-362,419-424) ---------------
// pass 1 (:340-362): totals of the raw counts over the two groups, all rows
__global__ void k_group_totals(const int32_t *cnts, const uint32_t *order, const uint64_t *lbase, uint32_t n_chunks, int S,
                               const int *idx_g, int nA, int nB, unsigned long long *tot)
{
    unsigned long long a = 0, b = 0;
    for (uint32_t oc = blockIdx.x; oc < n_chunks; oc += gridDim.x) {
        const uint32_t rows = (uint32_t)(lbase[oc + 1] - lbase[oc]);
        const int32_t *base = cnts + (size_t)order[oc] * DK_CH * S;
        for (uint32_t r = threadIdx.x; r < rows; r += blockDim.x) {
            const int32_t *c = base + (size_t)r * S;
            for (int i = 0; i < nA; i++) a += (unsigned long long)(long long)c[idx_g[i]];
            for (int i = 0; i < nB; i++) b += (unsigned long long)(long long)c[idx_g[nA + i]];
        }
    }
    for (int o = 16; o; o >>= 1) { a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&tot[0], a); atomicAdd(&tot[1], b); }
}
// logpoisson (:152-164): log of lambda^k e^-lambda / Gamma(k+1) through the density itself, like the reference
// (so an underflowing density gives the reference's constant 0.001)
__device__ double dk_logpoisson(double lambda, double k)
{
    const double t = exp(-lambda + k * log(lambda) - lgamma(k + 1.0));
    if (t <= 0) return 0.001;
    return log(t);
}
__global__ void __launch_bounds__(TT_PO_THREADS)
k_poisson_rows(const int32_t *cnts, const uint32_t *order, const uint64_t *lbase, uint32_t n_chunks, int S, const int *idx_g,
               int nA, int nB, const unsigned long long *tot, double *p_raw, double *log2fc, unsigned long long *neg_stat)
{
    const uint32_t oc = blockIdx.x * (TT_PO_THREADS / DK_CH) + threadIdx.x / DK_CH, j = threadIdx.x % DK_CH;
    if (oc >= n_chunks) return;
    const uint64_t b0 = lbase[oc];
    if (j >= lbase[oc + 1] - b0) return;
    const int32_t *c = cnts + ((size_t)order[oc] * DK_CH + j) * (size_t)S;
    double sa = 0, sb = 0;
    for (int i = 0; i < nA; i++) sa += (double)c[idx_g[i]] + 1;          // a[i] = count + 1 (:394), counts NOT normalised (:380-382)
    for (int i = 0; i < nB; i++) sb += (double)c[idx_g[nA + i]] + 1;
    const double K1 = sa / nA, K2 = sb / nB;                             // "sumA/sumB" are means (:421-422)
    const long long N1 = (long long)tot[0], N2 = (long long)tot[1];
    const double ll1 = dk_logpoisson(K1, K1) + dk_logpoisson(K2, K2);    // :175
    const double theta = (K1 + K2) / (double)(N1 + N2);                  // :176
    const double ll2 = dk_logpoisson(theta * (double)N1, K1) + dk_logpoisson(theta * (double)N2, K2);
    const double t_stat = 2 * (ll1 - ll2);
    double p;
    if (!(t_stat >= 0)) { p = 1.0; atomicAdd(neg_stat, 1ull); }          // Boost throws here (the reference aborts)
    else p = 1.0 - erf(sqrt(t_stat / 2.0));                              // 1 - cdf(chi_squared(1), t_stat) (:182-183)
    p_raw[b0 + j] = p;
    log2fc[b0 + j] = log(K2 / K1) / log(2.0);                            // :433
}

extern "C" int dk_poisson_rows(dk_ctx *ctx, const dk_matrix *m, const int *idxA, int nA, const int *idxB, int nB, dk_ttest **out,
                               uint64_t *info)
{
    if (!ctx || !m || !idxA || !idxB || !out) return dk_fail(ctx, DK_ERR_ARG, "poisson_rows: NULL argument");
    *out = nullptr;
    const int S = m->S;
    if (nA < 1 || nB < 1) return dk_fail(ctx, DK_ERR_UNSUPPORTED, "poisson: empty group (%d, %d samples)", nA, nB);
    for (int i = 0; i < nA; i++) if (idxA[i] < 0 || idxA[i] >= S) return dk_fail(ctx, DK_ERR_ARG, "poisson: idxA[%d] out of range", i);
    for (int i = 0; i < nB; i++) if (idxB[i] < 0 || idxB[i] >= S) return dk_fail(ctx, DK_ERR_ARG, "poisson: idxB[%d] out of range", i);
    if (m->R >= 0xFFFFFFFFull) return dk_fail(ctx, DK_ERR_ARG, "poisson: more than 2^32 rows on one GPU; shard by prefix");
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
    std::vector<double> ones(S, 1.0);                      // survivors print the RAW counts in this mode (count / 1.0)
    DK_CUDA(ctx, cudaMemcpyAsync(t->nf_dev, ones.data(), sizeof(double) * S, cudaMemcpyHostToDevice, st));
    DK_CUDA(ctx, cudaMemcpyAsync(t->idx_dev, idx.data(), sizeof(int) * (nA + nB), cudaMemcpyHostToDevice, st));
    DK_CUDA(ctx, cudaMemsetAsync(t->tie_counts, 0, 64, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    unsigned long long *tot = (unsigned long long *)(t->tie_counts + 4);      // [4]=N1 [5]=N2 [6]=negative statistics
    const uint32_t nch = (uint32_t)m->n_chunks;
    if (nch) {
        unsigned g1 = nch < (unsigned)ctx->sm_count * 8 ? nch : (unsigned)ctx->sm_count * 8;
        k_group_totals<<<g1, 256, 0, st>>>(m->cnts, m->order, m->lbase, nch, S, t->idx_dev, nA, nB, tot); DK_LAUNCHED(ctx);
        unsigned grid = (nch + (TT_PO_THREADS / DK_CH) - 1) / (TT_PO_THREADS / DK_CH);
        k_poisson_rows<<<grid, TT_PO_THREADS, 0, st>>>(m->cnts, m->order, m->lbase, nch, S, t->idx_dev, nA, nB, tot, t->p_raw, t->log2fc,
                                                   tot + 2);
        DK_LAUNCHED(ctx);
    }
    uint64_t h[3] = {0, 0, 0};
    DK_CUDA(ctx, cudaMemcpyAsync(h, tot, sizeof h, cudaMemcpyDeviceToHost, st));
    DK_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    ctx->stage_ms[2] = ms;
    if (info) { info[0] = h[0]; info[1] = h[1]; info[2] = h[2]; }
    *out = t;
    return DK_OK;
}

extern "C" uint64_t dk_ttest_rows_count(const dk_ttest *t) { return t ? t->R : 0; }
extern "C" void dk_ttest_free(dk_ctx *ctx, dk_ttest *t) { if (ctx) dk_ttest_release(ctx, t); }
