// Armijo line-search helpers of Ceres 1.14 (ArmijoLineSearch + polynomial.cc MinimizeInterpolatingPolynomial), restated
// for the device.  Only the rare path of the bounded single-frame problems uses them (the step size 1 normally already
// satisfies the sufficient-decrease test), and then one warp runs them.
#pragma once
#include <cfloat>

namespace b3d {

constexpr double kSufficientDecrease = 1e-4, kMinLineSearchStepSize = 1e-9;
constexpr int kMaxLineSearchIters = 20;
constexpr double kLsContractLo = 1e-3, kLsContractHi = 0.6;   // max_/min_line_search_step_contraction (Solver::Options defaults)

// ---- polynomial helpers for the Armijo interpolation ---------------------------------------------------------------
struct LsSample { double x, value, gradient; bool valid; };

__device__ double poly_eval(const double* p, int nc, double x) { double v = 0; for (int i = 0; i < nc; ++i) v = v * x + p[i]; return v; }

// Real roots of a polynomial of degree <= 4 (coefficients highest first, nc of them) inside [lo, hi], ascending.  Ceres hands the
// derivative of the interpolating polynomial to a companion-matrix eigenvalue solver and tries the real part of EVERY root as a step
// size; but the minimum of the interpolant over [x_min, x_max] sits at an end point or at a real root of its derivative inside the
// interval — the other candidates (real parts of complex pairs, roots outside) can never give a smaller value — so only those are
// needed.  They are isolated exactly by the derivative chain: between two consecutive roots of p' the polynomial p is monotone, so
// every such piece holds at most one root, found by safeguarded Newton (bisection fallback) to the last bit; the pieces of a level
// are refined by different lanes of the calling warp.  (The simultaneous complex iteration this replaces was 14 % of fit_single_kernel's cycles.)
__device__ unsigned long long g_poly_cyc[8];    // debug: cycles in the interpolation, the end-point values, the root search; calls; refinements, their iterations
template <int NC>
__device__ __forceinline__ double poly_eval_fixed(const double (&p)[NC], double x) {
    double v = p[0];
#pragma unroll
    for (int i = 1; i < NC; ++i) v = fma(v, x, p[i]);
    return v;
}
// Root of p inside (a, b), p(a) and p(b) of opposite signs (fa = p(a) != 0), by the WHOLE warp: every round the 32 lanes evaluate p
// at 32 points that cut the bracket into 33 parts and the bracket shrinks to the part with the sign change (6 rounds: 7.7e-10 of the
// first bracket), then one secant step inside the last bracket.  No divisions in the loop, no dependence on how well Newton's method
// likes the polynomial (safeguarded Newton needed 18 iterations of ~450 cycles on these interpolants).  All lanes call it with the same arguments and get the same result.
template <int NC>
__device__ __noinline__ double poly_refine_root(const double (&p)[NC], double a, double b, double fa) {
    const int lane = threadIdx.x & 31;
    const bool neg_a = fa < 0.0;
    double f_lo = fa, f_hi = poly_eval_fixed<NC>(p, b);       // values at the bracket ends (kept for the final secant step)
    for (int round = 0; round < 6; ++round) {                  // 33^-6 = 7.7e-10 of the first bracket; the secant step below squares that
        const double h = (b - a) * (1.0 / 33.0);
        const double f = poly_eval_fixed<NC>(p, a + h * (double)(lane + 1));
        // lanes whose point is already on b's side of the root (or on it); every lane then knows between which two points the sign
        // changes and evaluates the new end points itself (no shuffles: the bracket stays identical in all lanes by construction)
        const unsigned m = __ballot_sync(0xffffffffu, (f < 0.0) != neg_a || f == 0.0);
        const int first = m ? __ffs(m) - 1 : 32;              // between point first - 1 and point first (point -1 = a, point 32 = b)
        const double a0 = a;
        if (first > 0) { a = a0 + h * (double)first; f_lo = poly_eval_fixed<NC>(p, a); }
        if (first < 32) { b = a0 + h * (double)(first + 1); f_hi = poly_eval_fixed<NC>(p, b); }
        if (f_hi == 0.0) return b;
        if (f_lo == 0.0) return a;
        if (!(fabs(b - a) > 4.5e-16 * fmax(fabs(a), fabs(b)))) break;
    }
    // secant step inside the last bracket (p is monotone on it)
    const double t = f_lo / (f_lo - f_hi);
    const double x = a + (b - a) * t;
    return (x >= fmin(a, b) && x <= fmax(a, b)) ? x : 0.5 * (a + b);
}
// real roots of a x^2 + b x + c inside the open interval (lo, hi), ascending (a = 0: the linear root)
__device__ __forceinline__ int quad_roots_in(double a, double b, double c, double lo, double hi, double* out) {
    double r0, r1;
    int n = 0;
    if (a == 0.0) {
        if (b == 0.0) return 0;
        r0 = -c / b; n = 1;
    } else {
        const double D = b * b - 4.0 * a * c;
        if (!(D >= 0.0)) return 0;
        const double sD = sqrt(D);
        if (b >= 0) { r0 = (-b - sD) / (2.0 * a); r1 = (2.0 * c) / (-b - sD); }
        else { r0 = (2.0 * c) / (-b + sD); r1 = (-b + sD) / (2.0 * a); }
        n = 2;
        if (r1 < r0) { const double t = r0; r0 = r1; r1 = t; }
    }
    int m = 0;
    if (r0 > lo && r0 < hi) out[m++] = r0;
    if (n == 2 && r1 > lo && r1 < hi && !(m == 1 && r1 == out[0])) out[m++] = r1;
    return m;
}
// one level of the chain: the roots of q (NC coefficients) in [lo, hi], given the nb roots of q' in (lo, hi) (ascending) as break
// points: q is monotone on every piece between them.  All lanes of the warp call it with the same arguments.
template <int NC>
__device__ __forceinline__ int poly_level_roots(const double (&q)[NC], double lo, double hi, const double* brk_in, int nb, double* brk_out) {
    int n = 0;
    double a = lo, fa = poly_eval_fixed<NC>(q, lo);
    for (int j = 0; j <= nb; ++j) {
        const double b = j < nb ? brk_in[j] : hi;
        const double fb = poly_eval_fixed<NC>(q, b);
        if (fa == 0.0) { if (n == 0 || brk_out[n - 1] != a) brk_out[n++] = a; }
        else if ((fa < 0.0) != (fb < 0.0) && fb != 0.0) brk_out[n++] = poly_refine_root<NC>(q, a, b, fa);
        // (a root exactly at b is the next piece's a; at b = hi it is the interval's end point, which the caller tries anyway)
        a = b; fa = fb;
    }
    return n;
}
// Real roots in [lo, hi] of the derivative of the interpolant: a quadratic (two samples; pin has 3 coefficients) or a quartic (three
// samples; 5 coefficients).  Called by ALL lanes of one warp with the same arguments.  __noinline__ on purpose, here and on
// poly_refine_root: inlined into a caller whose control flow the compiler cannot prove warp-uniform, the __ballot_sync of the
// refinement was observed to see a partial warp (the standalone test hook returned no roots); the call boundary keeps it convergent.
// tests/test_fit_single_gpu.py::test_line_search_polynomial_minimiser_vs_numpy pins this function against numpy.
__device__ __noinline__ int poly_real_roots_in(const double* pin, int nc, double lo, double hi, double* out) {
    if (!(lo <= hi)) return 0;
    if (nc == 3) return quad_roots_in(pin[0], pin[1], pin[2], lo, hi, out);
    if (nc != 5) return 0;
    const double q4[5] = {pin[0], pin[1], pin[2], pin[3], pin[4]};
    const double q3[4] = {4.0 * pin[0], 3.0 * pin[1], 2.0 * pin[2], pin[3]};
    double b2[2], b3[4];
    const int n2 = quad_roots_in(12.0 * pin[0], 6.0 * pin[1], 2.0 * pin[2], lo, hi, b2);     // where the cubic q4' turns
    const int n3 = poly_level_roots<4>(q3, lo, hi, b2, n2, b3);                              // where the quartic turns
    return poly_level_roots<5>(q4, lo, hi, b3, n3, out);
}

// Coefficients (highest degree first) of the polynomial of degree 2 NS - 1 that interpolates value and gradient at NS distinct
// points: ceres::FindInterpolatingPolynomial.  Ceres solves the 2 NS x 2 NS Vandermonde-type system with Eigen's FullPivLU; the same
// polynomial is formed here through Hermite divided differences (nodes x_0, x_0, x_1, x_1, ...) expanded into the monomial
// basis — every index is a compile-time constant, so the whole table lives in registers (the LU on a local-memory matrix was the
// longest one-warp stretch of a line search: profiles/r02 fit_single notes).  Returns false if two sample points coincide.
template <int NS>
__device__ __forceinline__ bool hermite_coefficients(const LsSample* s, double* coef /* 2 NS */) {
    constexpr int N = 2 * NS;
    double z[N], d[N];
#pragma unroll
    for (int q = 0; q < NS; ++q) { z[2 * q] = s[q].x; z[2 * q + 1] = s[q].x; d[2 * q] = s[q].value; d[2 * q + 1] = s[q].value; }
    // first-order differences: the gradient inside a doubled node, a secant between different nodes
    bool ok = true;
#pragma unroll
    for (int i = N - 1; i >= 1; --i) {
        if (i & 1) d[i] = s[i >> 1].gradient;
        else {
            const double h = z[i] - z[i - 1];
            ok = ok && h != 0.0;
            d[i] = (d[i] - d[i - 1]) / h;
        }
    }
    // higher orders, in place from the bottom of each column
#pragma unroll
    for (int k = 2; k < N; ++k)
#pragma unroll
        for (int i = N - 1; i >= k; --i) {
            const double h = z[i] - z[i - k];
            ok = ok && h != 0.0;
            d[i] = (d[i] - d[i - 1]) / h;
        }
    // Newton form sum_k d_k prod_{m < k} (x - z_m)  ->  monomial coefficients, lowest degree first in c[]
    double c[N];
#pragma unroll
    for (int i = 0; i < N; ++i) c[i] = 0.0;
    c[0] = d[N - 1];
#pragma unroll
    for (int k = N - 2; k >= 0; --k) {
        // c(x) <- c(x) (x - z_k) + d_k ; the polynomial held so far has degree N - 2 - k
#pragma unroll
        for (int i = N - 1; i >= 1; --i) c[i] = c[i - 1] - z[k] * c[i];
        c[0] = d[k] - z[k] * c[0];
    }
#pragma unroll
    for (int i = 0; i < N; ++i) coef[i] = c[N - 1 - i];
    return ok;
}

// ceres MinimizeInterpolatingPolynomial: every sample carries value and gradient.  Called by all lanes of one warp (uniform).
__device__ double minimize_interpolating_polynomial(const LsSample* s, int ns, double x_min, double x_max) {
    const int nc = 2 * ns, degree = nc - 1;
    double coef[6];
    double best_x = (x_min + x_max) / 2.0;
    const long long t0 = clock64();
    const bool ok = ns == 3 ? hermite_coefficients<3>(s, coef) : hermite_coefficients<2>(s, coef);
    const long long t1 = clock64();
    if (!ok) return best_x;
    double best_v = poly_eval(coef, nc, best_x);
    const double vmin = poly_eval(coef, nc, x_min);
    if (vmin < best_v) { best_v = vmin; best_x = x_min; }
    const double vmax = poly_eval(coef, nc, x_max);
    if (vmax < best_v) { best_v = vmax; best_x = x_max; }
    double dcoef[5], roots[5];
    for (int i = 0; i < degree; ++i) dcoef[i] = (degree - i) * coef[i];
    const long long t2 = clock64();
    const int nr = poly_real_roots_in(dcoef, degree, x_min, x_max, roots);
    if ((threadIdx.x & 31) == 0) {
        const long long t3 = clock64();
        atomicAdd(&g_poly_cyc[0], (unsigned long long)(t1 - t0)); atomicAdd(&g_poly_cyc[1], (unsigned long long)(t2 - t1));
        atomicAdd(&g_poly_cyc[2], (unsigned long long)(t3 - t2)); atomicAdd(&g_poly_cyc[3], 1ull);
    }
    for (int i = 0; i < nr; ++i) {
        const double v = poly_eval(coef, nc, roots[i]);
        if (v < best_v) { best_v = v; best_x = roots[i]; }
    }
    return best_x;
}


}  // namespace b3d
