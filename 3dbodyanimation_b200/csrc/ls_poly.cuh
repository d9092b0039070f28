// Armijo line-search helpers of Ceres 1.14 (ArmijoLineSearch + polynomial.cc MinimizeInterpolatingPolynomial), restated
// for the device.  Only the rare path of the bounded single-frame problems uses them (the step size 1 normally already
// satisfies the sufficient-decrease test), and then one warp runs them.
#pragma once
#include <cfloat>

namespace b3d {

constexpr double kSufficientDecrease = 1e-4, kMinLineSearchStepSize = 1e-9;
constexpr int kMaxLineSearchIters = 20;
constexpr double kLsContractLo = 1e-3, kLsContractHi = 0.6;   // max_/min_line_search_step_contraction (Solver::Options defaults)

// ---- polynomial helpers for the (rare) Armijo interpolation: warp 0, all lanes --------------------------------------
struct LsSample { double x, value, gradient; bool valid; };

__device__ double poly_eval(const double* p, int nc, double x) { double v = 0; for (int i = 0; i < nc; ++i) v = v * x + p[i]; return v; }

// real parts of the roots of a polynomial of degree <= 4 (coefficients highest first).  Called by ALL lanes of one warp:
// degrees 1 and 2 in closed form; degrees 3 and 4 by the Aberth-Ehrlich simultaneous iteration with one lane per root (the
// other lanes idle along), which is what made the single-thread version the longest serial stretch of a line search.
__device__ int poly_root_real_parts(const double* pin, int nc, double* out) {
    int lead = 0;
    while (lead + 1 < nc && pin[lead] == 0.0) ++lead;
    const double* p = pin + lead;
    const int deg = nc - lead - 1;
    if (deg <= 0) return 0;
    if (deg == 1) { out[0] = -p[1] / p[0]; return 1; }
    if (deg == 2) {
        const double a = p[0], b = p[1], c = p[2];
        const double D = b * b - 4 * a * c, sD = sqrt(fabs(D));
        if (D >= 0) {
            if (b >= 0) { out[0] = (-b - sD) / (2.0 * a); out[1] = (2.0 * c) / (-b - sD); }
            else { out[0] = (2.0 * c) / (-b + sD); out[1] = (-b + sD) / (2.0 * a); }
        } else { out[0] = out[1] = -b / (2.0 * a); }
        return 2;
    }
    const int lane = threadIdx.x & 31;
    double ar[6];
    const double ilead = 1.0 / p[0];
    for (int i = 0; i <= deg; ++i) ar[i] = p[i] * ilead;
    double rad = 0;
    for (int i = 1; i <= deg; ++i) rad = fmax(rad, fabs(ar[i]));
    rad = 1.0 + rad;
    double zr, zi;
    {
        double sn, cs;
        sincos(2 * 3.14159265358979323846 * (lane % deg) / deg + 0.4, &sn, &cs);
        zr = rad * 0.5 * cs; zi = rad * 0.5 * sn;
    }
    for (int it = 0; it < 64; ++it) {
        double fr = ar[0], fi = 0, dr = 0, di = 0;
        for (int k = 1; k <= deg; ++k) {
            double t = dr * zr - di * zi + fr; di = dr * zi + di * zr + fi; dr = t;
            t = fr * zr - fi * zi + ar[k]; fi = fr * zi + fi * zr; fr = t;
        }
        double ozr[4], ozi[4];                      // every root, gathered outside any divergent code
#pragma unroll
        for (int j = 0; j < 4; ++j) { ozr[j] = __shfl_sync(0xffffffffu, zr, j); ozi[j] = __shfl_sync(0xffffffffu, zi, j); }
        double wr = 0.0, wi = 0.0;
        if (!(fr == 0.0 && fi == 0.0)) {
            const double idd = 1.0 / (dr * dr + di * di);
            const double rr = (fr * dr + fi * di) * idd, ri = (fi * dr - fr * di) * idd;   // ratio = f / f'
            double sr = 0, si = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j < deg && j != lane) { const double ur = zr - ozr[j], ui = zi - ozi[j], iu = 1.0 / (ur * ur + ui * ui); sr += ur * iu; si -= ui * iu; }
            const double er = 1.0 - (rr * sr - ri * si), ei = -(rr * si + ri * sr);
            const double iee = 1.0 / (er * er + ei * ei);
            wr = (rr * er + ri * ei) * iee; wi = (ri * er - rr * ei) * iee;
        }
        zr -= wr; zi -= wi;
        double change = lane < deg ? sqrt(wr * wr + wi * wi) : 0.0;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) change = fmax(change, __shfl_xor_sync(0xffffffffu, change, o));
        change = __shfl_sync(0xffffffffu, change, 0);
        if (change < 1e-14 * rad) break;   // warp-uniform
    }
    for (int i = 0; i < deg; ++i) out[i] = __shfl_sync(0xffffffffu, zr, i);
    return deg;
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
    const bool ok = ns == 3 ? hermite_coefficients<3>(s, coef) : hermite_coefficients<2>(s, coef);
    if (!ok) return best_x;
    double best_v = poly_eval(coef, nc, best_x);
    const double vmin = poly_eval(coef, nc, x_min);
    if (vmin < best_v) { best_v = vmin; best_x = x_min; }
    const double vmax = poly_eval(coef, nc, x_max);
    if (vmax < best_v) { best_v = vmax; best_x = x_max; }
    double dcoef[5], roots[5];
    for (int i = 0; i < degree; ++i) dcoef[i] = (degree - i) * coef[i];
    const int nr = poly_root_real_parts(dcoef, degree, roots);
    for (int i = 0; i < nr; ++i) {
        if (roots[i] < x_min || roots[i] > x_max) continue;
        const double v = poly_eval(coef, nc, roots[i]);
        if (v < best_v) { best_v = v; best_x = roots[i]; }
    }
    return best_x;
}


}  // namespace b3d
