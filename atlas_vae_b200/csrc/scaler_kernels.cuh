// Non-affine scaler transforms on the device: apply_scaling (OE-VAE/utils.py:406-417) for
//   QuantileTransformer(output_distribution='normal', n_quantiles=10000, random_state=0)   utils.py:394 (the scaler of vae.sh:9-10)
//   PowerTransformer()  (yeo-johnson, standardize)                                          utils.py:396
// The arithmetic is scikit-learn's (1.9.0: QuantileTransformer._transform_col, PowerTransformer.transform, StandardScaler.transform),
// NumPy's interp (compiled_base.c arr_interp) and SciPy's ndtri (Cephes ndtri.c), restated element by element exactly as
// oracle/scalers.py does (which is pinned bit-for-bit against the libraries on the CPU).  Explicit round-to-nearest intrinsics keep
// ptxas from contracting a*b+c into FMAs the host libraries do not use.
// HBM/L2-bound table look-ups: the per-column quantile tables (D x 10000 doubles = 32 MB at D = 400) live in L2.
#pragma once
#include "common.cuh"

namespace oevae {
namespace scal {

// np.interp for one point: xp(j), fp(j) are accessors of a sorted table of n points
template <typename XP, typename FP>
__device__ __forceinline__ double interp1(double x, int n, XP xp, FP fp) {
    if (isnan(x)) return x;
    if (x > xp(n - 1)) return fp(n - 1);
    if (x < xp(0)) return fp(0);
    int lo = 0, hi = n;                                      // j = last index with xp[j] <= x
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (xp(mid) <= x) lo = mid + 1; else hi = mid; }
    const int j = lo - 1;
    const double xj = xp(j), fj = fp(j);
    if (j == n - 1 || xj == x) return fj;
    const double xj1 = xp(j + 1), fj1 = fp(j + 1);
    const double slope = __ddiv_rn(__dsub_rn(fj1, fj), __dsub_rn(xj1, xj));
    double v = __dadd_rn(__dmul_rn(slope, __dsub_rn(x, xj)), fj);
    if (isnan(v)) {                                          // "if we get nan in one direction, try the other"
        v = __dadd_rn(__dmul_rn(slope, __dsub_rn(x, xj1)), fj1);
        if (isnan(v) && fj == fj1) v = fj;
    }
    return v;
}

__device__ __forceinline__ double polevl(double x, const double* c, int n) {   // c[0] x^n + ... + c[n]
    double a = c[0];
    for (int k = 1; k <= n; ++k) a = __dadd_rn(__dmul_rn(a, x), c[k]);
    return a;
}
__device__ __forceinline__ double p1evl(double x, const double* c, int n) {    // x^n + c[0] x^(n-1) + ... + c[n-1]
    double a = __dadd_rn(x, c[0]);
    for (int k = 1; k < n; ++k) a = __dadd_rn(__dmul_rn(a, x), c[k]);
    return a;
}
__constant__ double kP0[5] = {-5.99633501014107895267E1, 9.80010754185999661536E1, -5.66762857469070293439E1, 1.39312609387279679503E1, -1.23916583867381258016E0};
__constant__ double kQ0[8] = {1.95448858338141759834E0, 4.67627912898881538453E0, 8.63602421390890590575E1, -2.25462687854119370527E2, 2.00260212380060660359E2,
                              -8.20372256168333339912E1, 1.59056225126211695515E1, -1.18331621121330003142E0};
__constant__ double kP1[9] = {4.05544892305962419923E0, 3.15251094599893866154E1, 5.71628192246421288162E1, 4.40805073893200834700E1, 1.46849561928858024014E1,
                              2.18663306850790267539E0, -1.40256079171354495875E-1, -3.50424626827848203418E-2, -8.57456785154685413611E-4};
__constant__ double kQ1[8] = {1.57799883256466749731E1, 4.53907635128879210584E1, 4.13172038254672030440E1, 1.50425385692907503408E1, 2.50464946208309415979E0,
                              -1.42182922854787788574E-1, -3.80806407691578277194E-2, -9.33259480895457427372E-4};
__constant__ double kP2[9] = {3.23774891776946035970E0, 6.91522889068984211695E0, 3.93881025292474443415E0, 1.33303460815807542389E0, 2.01485389549179081538E-1,
                              1.23716634817820021358E-2, 3.01581553508235416007E-4, 2.65806974686737550832E-6, 6.23974539184983293730E-9};
__constant__ double kQ2[8] = {6.02427039364742014255E0, 3.67983563856160859403E0, 1.37702099489081330271E0, 2.16236993594496635890E-1, 1.34204006088543189037E-2,
                              3.28014464682127739104E-4, 2.89247864745380683936E-6, 6.79019408009981274425E-9};

// Cephes ndtri (scipy.special.ndtri / stats.norm.ppf)
__device__ __forceinline__ double ndtri(double y0) {
    if (y0 == 0.0) return -INFINITY;
    if (y0 == 1.0) return INFINITY;
    if (!(y0 > 0.0 && y0 < 1.0)) return NAN;
    const double expm2 = 0.13533528323661269189;
    bool code = true;
    double y = y0;
    if (y > __dsub_rn(1.0, expm2)) { y = __dsub_rn(1.0, y); code = false; }
    if (y > expm2) {
        y = __dsub_rn(y, 0.5);
        const double y2 = __dmul_rn(y, y);
        double x = __dadd_rn(y, __dmul_rn(y, __ddiv_rn(__dmul_rn(y2, polevl(y2, kP0, 4)), p1evl(y2, kQ0, 8))));
        return __dmul_rn(x, 2.50662827463100050242E0);
    }
    double x = sqrt(__dmul_rn(-2.0, log(y)));
    const double x0 = __dsub_rn(x, __ddiv_rn(log(x), x));
    const double z = __ddiv_rn(1.0, x);
    const double x1 = x < 8.0 ? __ddiv_rn(__dmul_rn(z, polevl(z, kP1, 8)), p1evl(z, kQ1, 8)) : __ddiv_rn(__dmul_rn(z, polevl(z, kP2, 8)), p1evl(z, kQ2, 8));
    x = __dsub_rn(x0, x1);
    return code ? -x : x;
}

// qt: [D][nq] float64 (column f's quantiles contiguous), ref: [nq] float64.  One thread per element; consecutive threads walk a row.
__global__ void quantile_normal_kernel(const float* __restrict__ x, float* __restrict__ y, long n_elems, int D, const double* __restrict__ qt,
                                       const double* __restrict__ ref, int nq, double clip_min, double clip_max) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_elems) return;
    const int f = (int)(i % D);
    const double* q = qt + (long)f * nq;
    const float xf = x[i];
    const bool lower = (double)__fsub_rn(xf, 1e-7f) < q[0];          // X_col - BOUNDS_THRESHOLD < quantiles[0]  (float32 arithmetic, float64 compare)
    const bool upper = (double)__fadd_rn(xf, 1e-7f) > q[nq - 1];
    const double xd = (double)xf;
    const double a = interp1(xd, nq, [&](int j) { return q[j]; }, [&](int j) { return ref[j]; });
    const double b = interp1(-xd, nq, [&](int j) { return -q[nq - 1 - j]; }, [&](int j) { return -ref[nq - 1 - j]; });
    float t = (float)__dmul_rn(0.5, __dsub_rn(a, b));                 // stored into the float32 column
    if (upper) t = 1.f;
    if (lower) t = 0.f;
    const float v = (float)ndtri((double)t);                          // SciPy's float32 loop: float32 result
    double c = (double)v;
    c = c < clip_min ? clip_min : (c > clip_max ? clip_max : c);      // np.clip keeps NaN
    y[i] = (float)c;
}

// yeo-johnson in float32 (a transformer fitted on float32 data holds float32 lambdas_) + StandardScaler with float32 mean / scale
__global__ void power_yj_kernel(const float* __restrict__ x, float* __restrict__ y, long n_elems, int D, const float* __restrict__ lambdas,
                                const float* __restrict__ mean, const float* __restrict__ scale) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_elems) return;
    const int f = (int)(i % D);
    const float xf = x[i], lm = lambdas[f];
    const float eps = 1.1920929e-07f;
    float v;
    if (xf >= 0.f) v = fabsf(lm) < eps ? log1pf(xf) : __fdiv_rn(expm1f(__fmul_rn(lm, log1pf(xf))), lm);
    else if (xf < 0.f) { const float l2 = __fsub_rn(2.f, lm); v = fabsf(__fsub_rn(lm, 2.f)) > eps ? -__fdiv_rn(expm1f(__fmul_rn(l2, log1pf(-xf))), l2) : -log1pf(-xf); }
    else v = xf;                                                      // NaN: `pos = x >= 0` is False, log1p(-NaN) = NaN
    if (mean) v = __fsub_rn(v, mean[f]);
    if (scale) v = __fdiv_rn(v, scale[f]);
    y[i] = v;
}

}  // namespace scal
}  // namespace oevae
