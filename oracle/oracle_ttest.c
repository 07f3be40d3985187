/* CPU ORACLE — test infrastructure only (see dekupl_oracle.h).
 * PARITY UNPINNED for the p-value: the reference calls Boost.Math students_t
 * (TtestFilter.c:10,146-147; Boost absent here, version unpinned). We restate the
 * published identity  2*Q(|t|;v) = I_{v/(v+t^2)}(v/2, 1/2)  with a local fp64
 * regularised incomplete beta (modified-Lentz continued fraction), and pin it
 * against scipy (tests/golden/ttest_scipy.json). */
#include "dekupl_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* continued fraction for I_x(a,b), valid/fast for x < (a+1)/(a+b+2) */
static double beta_cf(double a, double b, double x)
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

/* log( Gamma(v+1/2) / Gamma(v) ): direct for small v, asymptotic series for v >= 40
 * (lgamma(v+.5)-lgamma(v) would lose ~|lgamma| * 1e-16 to cancellation at large v) */
static double lgamma_ratio_half(double v)
{
    if (v < 40.0) return lgamma(v + 0.5) - lgamma(v);
    double r = 1.0 / v, r2 = r * r;
    return 0.5 * log(v) + r * (-1.0 / 8.0 + r2 * (1.0 / 192.0 + r2 * (-1.0 / 640.0 + r2 * (17.0 / 14336.0))));
}

/* I_x(v, 1/2) with y = 1-x supplied by the caller (no cancellation) */
static double ibeta_half(double v, double x, double y)
{
    if (x <= 0.0) return 0.0;
    if (y <= 0.0) return 1.0;
    const double LOG_SQRT_PI = 0.57236494292470008707;
    double lx = (y < 0.5) ? log1p(-y) : log(x);
    double ly = (x < 0.5) ? log1p(-x) : log(y);
    double lbt = lgamma_ratio_half(v) - LOG_SQRT_PI + v * lx + 0.5 * ly;
    if (x < (v + 1.0) / (v + 2.5))
        return exp(lbt) * beta_cf(v, 0.5, x) / v;
    return 1.0 - exp(lbt) * beta_cf(0.5, v, y) / 0.5;
}

double orc_student_t_2tail(double t, double df)
{
    double t2 = t * t;
    if (isinf(df)) return erfc(fabs(t) / sqrt(2.0));
    double x = df / (df + t2), y = t2 / (df + t2);
    return ibeta_half(0.5 * df, x, y);
}

/* ---- second evaluation path: the branch structure of Boost.Math's students_t cdf complement ----------
 * cdf(complement(students_t(v), |t|)) with t2 = t*t:
 *     v > 2 t2 :  z = t2 / (v + t2),  P = ibetac(1/2, v/2, z) / 2
 *     else     :  z = v / (v + t2),   P = ibeta(v/2, 1/2, z) / 2
 * and TtestFilter.c:146-147 doubles it. Each routine below receives ONLY its z (as Boost's does) and forms 1 - z itself;
 * the branch keeps z <= 2/3, which is what makes that subtraction harmless. orc_student_t_2tail above always evaluates
 * I_{v/(v+t2)}(v/2, 1/2) with both x and 1-x supplied exactly: a different route to the same number. The two must agree
 * to ~1e-13 (tests/test_oracle.py); on a box with Boost, `make -C oracle ref` + the same grid closes the pin. */
static double lbeta_half(double v) { return 0.57236494292470008707 - lgamma_ratio_half(v); }   /* log B(v, 1/2) */
static double ibeta_v_half(double v, double z)          /* I_z(v, 1/2) from z alone */
{
    if (z <= 0.0) return 0.0;
    if (z >= 1.0) return 1.0;
    const double y = 1.0 - z;
    const double bt = exp(v * log(z) + 0.5 * log1p(-z) - lbeta_half(v));
    if (z < (v + 1.0) / (v + 2.5)) return bt * beta_cf(v, 0.5, z) / v;
    return 1.0 - bt * beta_cf(0.5, v, y) / 0.5;
}
static double ibetac_half_v(double v, double z)         /* 1 - I_z(1/2, v) from z alone */
{
    if (z <= 0.0) return 1.0;
    if (z >= 1.0) return 0.0;
    const double y = 1.0 - z;
    const double bt = exp(0.5 * log(z) + v * log1p(-z) - lbeta_half(v));   /* log1p: v * log(1 - z) must not lose the low bits of a tiny z */
    if (z < 1.5 / (v + 2.5)) return 1.0 - bt * beta_cf(0.5, v, z) / 0.5;
    return bt * beta_cf(v, 0.5, y) / v;
}
double orc_student_t_2tail_boost_branches(double t, double df)
{
    const double t2 = t * t;
    if (isinf(df)) return erfc(fabs(t) / sqrt(2.0));
    double P;
    if (df > 2.0 * t2) P = ibetac_half_v(0.5 * df, t2 / (df + t2)) / 2.0;
    else P = ibeta_v_half(0.5 * df, df / (df + t2)) / 2.0;
    return 2.0 * P;
}

/* TtestFilter.c:77-98 */
static double o_sum(const double *v, int n) { double s = 0; for (int i = 0; i < n; i++) s += v[i]; return s; }
static double o_mean(const double *v, int n) { return o_sum(v, n) / n; }
static double o_sd(const double *v, int n, double m) {
    double s = 0;
    for (int i = 0; i < n; i++) { double d = v[i] - m; s += d * d; }   /* pow(d,2) == d*d under g++ -O2 */
    return sqrt(s / (n - 1));
}

/* TtestFilter.c:120-150 */
static double o_ttest(double m1, double m2, int n1, int n2, double sd1, double sd2)
{
    if (sd1 == 0 && sd2 == 0) return 0.5;
    double df = sd1 * sd1 / n1 + sd2 * sd2 / n2;
    df *= df;
    double t1 = sd1 * sd1 / n1; t1 *= t1; t1 /= (n1 - 1);
    double t2 = sd2 * sd2 / n2; t2 *= t2; t2 /= (n2 - 1);
    df /= (t1 + t2);
    double t = (m1 - m2) / sqrt(sd1 * sd1 / n1 + sd2 * sd2 / n2);
    return orc_student_t_2tail(fabs(t), df);
}

void orc_ttest_rows(const int32_t *cnts, size_t R, int S, const double *nf,
                    const int *idxA, int nA, const int *idxB, int nB,
                    double *p_raw, double *meanA, double *meanB, double *log2fc)
{
    double *a = (double *)malloc(sizeof(double) * (nA + 1)), *la = (double *)malloc(sizeof(double) * (nA + 1));
    double *b = (double *)malloc(sizeof(double) * (nB + 1)), *lb = (double *)malloc(sizeof(double) * (nB + 1));
    for (size_t r = 0; r < R; r++) {
        const int32_t *c = cnts + r * S;
        for (int i = 0; i < nA; i++) {                            /* TtestFilter.c:379-396 */
            double x = (double)c[idxA[i]] / nf[idxA[i]];
            a[i] = x + 1; la[i] = log(x + 1);
        }
        for (int i = 0; i < nB; i++) {
            double x = (double)c[idxB[i]] / nf[idxB[i]];
            b[i] = x + 1; lb[i] = log(x + 1);
        }
        double mA = o_mean(a, nA), mB = o_mean(b, nB);            /* :404-405 */
        double m1 = o_mean(la, nA), m2 = o_mean(lb, nB);          /* :408 */
        double sd1 = o_sd(la, nA, m1), sd2 = o_sd(lb, nB, m2);    /* :411 */
        p_raw[r] = o_ttest(m1, m2, nA, nB, sd1, sd2);
        meanA[r] = mA; meanB[r] = mB;
        log2fc[r] = log(mB / mA) / log(2);                        /* :72-75,433 */
    }
    free(a); free(la); free(b); free(lb);
}

/* TtestFilter.c:100-118,459-476 */
static int cmp_rev_ptr(const void *a, const void *b) {
    double aa = **(double *const *)a, bb = **(double *const *)b;
    return aa > bb ? -1 : (bb > aa ? 1 : 0);
}
void orc_bh(const double *p_raw, size_t R, double *p_adj)
{
    if (R == 0) return;
    memcpy(p_adj, p_raw, sizeof(double) * R);
    double **ptr = (double **)malloc(sizeof(double *) * R);
    for (size_t i = 0; i < R; i++) ptr[i] = &p_adj[i];
    qsort(ptr, R, sizeof(double *), cmp_rev_ptr);
    double mn = 1;
    for (size_t i = 0; i < R; i++) {
        double v = (double)R / ((double)R - (double)i) * (*ptr[i]);
        if (v > mn) v = mn; else mn = v;
        *ptr[i] = v > 1 ? 1 : v;                                  /* dmin(1,v) */
    }
    free(ptr);
}

size_t orc_filter(const double *p_adj, const double *log2fc, size_t R,
                  double pmax, double fmin, uint8_t *keep)
{
    size_t D = 0;
    for (size_t i = 0; i < R; i++) {
        keep[i] = (p_adj[i] <= pmax && fabs(log2fc[i]) >= fmin) ? 1 : 0;
        D += keep[i];
    }
    return D;
}

/* ---- Poisson likelihood-ratio mode (-l), TtestFilter.c:152-185,340-362,419-424 -------------------------
 * logpoisson (:152-164): log of the Poisson density at a NON-integer k. Boost evaluates
 * pdf(poisson(lambda), k) = lambda^k e^-lambda / Gamma(k+1) (gamma_p_derivative(k+1, lambda)); the reference then
 * takes log1p(t-1) == log(t), or returns the constant 0.001 when the density underflows to 0 (:157-161). */
static double o_logpoisson(double lambda, double k)
{
    double t = exp(-lambda + k * log(lambda) - lgamma(k + 1.0));
    if (t <= 0) return 0.001;
    return log(t);
}
/* :173-185. cdf(chi_squared(1), x) = P(1/2, x/2) = erf(sqrt(x/2)); the reference forms 1.0 - cdf (cancellation
 * included). Boost throws for x < 0 (the reference aborts); the oracle and the product return p = 1 there. */
double orc_poisson_lr(double K1, double K2, int64_t N1, int64_t N2)
{
    double ll1 = o_logpoisson(K1, K1) + o_logpoisson(K2, K2);
    double theta = (K1 + K2) / (N1 + N2);
    double ll2 = o_logpoisson(theta * N1, K1) + o_logpoisson(theta * N2, K2);
    double llr = ll1 - ll2;
    double t_stat = 2 * llr;
    if (!(t_stat >= 0)) return 1.0;
    return 1.0 - erf(sqrt(t_stat / 2.0));
}
/* pass 1 (:340-362) totals over all rows, then per row (:392-433) K = mean(count+1) per group (counts are NOT
 * normalised in this mode, :380-382) */
void orc_poisson_rows(const int32_t *cnts, size_t R, int S, const int *idxA, int nA, const int *idxB, int nB,
                      double *p_raw, double *meanA, double *meanB, double *log2fc, int64_t *N_out)
{
    uint64_t N1 = 0, N2 = 0;
    for (size_t r = 0; r < R; r++) {
        const int32_t *c = cnts + r * S;
        for (int i = 0; i < nA; i++) N1 += (double)c[idxA[i]];
        for (int i = 0; i < nB; i++) N2 += (double)c[idxB[i]];
    }
    if (N_out) { N_out[0] = (int64_t)N1; N_out[1] = (int64_t)N2; }
    double *a = (double *)malloc(sizeof(double) * (nA + 1)), *b = (double *)malloc(sizeof(double) * (nB + 1));
    for (size_t r = 0; r < R; r++) {
        const int32_t *c = cnts + r * S;
        for (int i = 0; i < nA; i++) a[i] = (double)c[idxA[i]] + 1;
        for (int i = 0; i < nB; i++) b[i] = (double)c[idxB[i]] + 1;
        double mA = o_mean(a, nA), mB = o_mean(b, nB);
        p_raw[r] = orc_poisson_lr(mA, mB, (int64_t)N1, (int64_t)N2);
        meanA[r] = mA; meanB[r] = mB;
        log2fc[r] = log(mB / mA) / log(2);
    }
    free(a); free(b);
}
