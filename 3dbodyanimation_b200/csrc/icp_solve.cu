// SURVEY 8(f) rank 4, last piece: the ceres::Solve of ark::AvatarOptimizer::optimize (external/avatar/AvatarOptimizer.cpp:1486 with
// the options of :1313-1339): minimizer_type = LINE_SEARCH, line_search_direction_type = BFGS, line_search_interpolation_type =
// CUBIC, line_search_type = WOLFE (Ceres default), function_tolerance = 1e-4, max_num_iterations = maxItersPerICP — for a BATCH of
// bodies at once.  Ceres 1.14 (absent from the reference tree) is restated from its published algorithm, as in oracle/icp_solve.cpp:
// LineSearchMinimizer::Minimize, BFGS::NextDirection (dense inverse Hessian, secant guard 1e-14), WolfeLineSearch (bracketing + zoom,
// Nocedal & Wright 3.5 / 3.6 with Ceres' guards), MinimizeInterpolatingPolynomial for two samples with value and gradient.
//
// B200 shape of the loop: the objective (b3d_icp_eval_dev + b3d_icp_add_priors_dev: fused shape blend + skinning + residual +
// gradient over all vertices of all bodies) is the only heavy step and is evaluated for the whole batch per launch; every body is a
// small state machine (phase = which evaluation it is waiting for: first point, bracketing sample, zoom sample, accepted point) that
// icp_bfgs_advance_kernel steps once per evaluation — CTA = body, the 85 x 85 inverse Hessian of each body stays in HBM (58 KB per
// body, read and written once per accepted step).  Bodies that are done keep their trial point (their evaluation is wasted work until
// the slowest body finishes; the host looks at the done counter every few rounds, without draining the stream in between).
#include <algorithm>
#include <vector>

#include "internal.hpp"

namespace b3d {

namespace {

constexpr int NT = 3 + 3 * NJ + NS;          // 85 tangent coordinates: p | 24 x rotation tangent | w
constexpr int SOLVE_THREADS = 128;
constexpr int TRACE_W = 7;
constexpr double kMinStep = 1e-9, kSuffDecrease = 1e-4, kSuffCurvature = 0.9, kMaxExpansion = 10.0, kSecant = 1e-14;
constexpr int kMaxLsIters = 20, kMaxRestarts = 5;

enum { PH_INIT = 0, PH_BRACKET = 1, PH_ZOOM = 2, PH_FINAL = 3, PH_DONE = 4 };
enum { TERM_NO_CONVERGENCE = 0, TERM_GRADIENT_TOL = 1, TERM_PARAMETER_TOL = 2, TERM_FUNCTION_TOL = 3, TERM_FAILURE = 4 };

struct Sample { double x, value, gradient; int valid, pad; };

struct BodyState {
    double cost, prev_cost, prev_step, gmax, dirderiv, dmax, optimal, xnorm, initial_cost, trial_step;
    Sample init, previous, current, low, high;
    int phase, iteration, restarts, ls_iters, ls_evals, evals, term, nrec, last_dir_ok, pad;
};

struct SolveArrays {
    BodyState* st;
    double *xw, *xp, *xq;        // accepted point
    double *tw, *tp, *tq;        // trial point (what the next evaluation sees)
    double *cost_t, *grad_t;     // its evaluation
    double *g, *g_prev, *dir, *H;
    double* trace;
    int32_t* done_count;
    int trace_cap, shape, max_iterations;
    double function_tolerance, gradient_tolerance, parameter_tolerance;
};

// polynomial.cc MinimizeInterpolatingPolynomial for two samples with value and gradient (cubic), as in oracle/icp_solve.cpp
__device__ double minimize_cubic(const Sample& a, const Sample& b, double x_min, double x_max) {
    const double h = b.x - a.x;
    double best_x = (x_min + x_max) / 2.0;
    if (h == 0.0) return best_x;
    const double d1 = a.gradient, d2 = (b.value - a.value) / h, d3 = b.gradient;
    const double e1 = (d2 - d1) / h, e2 = (d3 - d2) / h;
    const double f1 = (e2 - e1) / h;
    const double A = a.x, Bx = b.x;
    const double c3 = f1;
    const double c2 = e1 - f1 * (2.0 * A + Bx);
    const double c1 = d1 - 2.0 * e1 * A + f1 * (A * A + 2.0 * A * Bx);
    const double c0 = a.value - d1 * A + e1 * A * A - f1 * A * A * Bx;
    auto eval = [&](double x) { return ((c3 * x + c2) * x + c1) * x + c0; };
    double best_v = eval(best_x);
    const double vmin = eval(x_min);
    if (vmin < best_v) { best_v = vmin; best_x = x_min; }
    const double vmax = eval(x_max);
    if (vmax < best_v) { best_v = vmax; best_x = x_max; }
    const double qa = 3.0 * c3, qb = 2.0 * c2, qc = c1;
    double roots[2];
    int nr = 0;
    if (qa == 0.0) {
        if (qb != 0.0) roots[nr++] = -qc / qb;
    } else {
        const double D = qb * qb - 4.0 * qa * qc, sD = sqrt(fabs(D));
        if (D >= 0) {
            if (qb >= 0) { roots[nr++] = (-qb - sD) / (2.0 * qa); roots[nr++] = (2.0 * qc) / (-qb - sD); }
            else { roots[nr++] = (2.0 * qc) / (-qb + sD); roots[nr++] = (-qb + sD) / (2.0 * qa); }
        } else { roots[nr++] = -qb / (2.0 * qa); roots[nr++] = -qb / (2.0 * qa); }
    }
    for (int i = 0; i < nr; ++i) {
        const double rt = roots[i];
        if (!(rt >= x_min && rt <= x_max)) continue;
        const double v = eval(rt);
        if (v < best_v) { best_v = v; best_x = rt; }
    }
    return best_x;
}

__device__ double interpolating_step(const Sample& lowerbound, const Sample& current, double min_step, double max_step) {
    if (!current.valid) return fmin(fmax(current.x * 0.5, min_step), max_step);
    return minimize_cubic(lowerbound, current, min_step, max_step);
}

// FakeQuaternionParameterization::Plus (AvatarOptimizer.cpp:122-141), coefficients (x, y, z, w)
__device__ void quat_plus(const double* x, const double* d, double* out) {
    const double n = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (n > 0.0) {
        const double s = sin(n) / n;
        const double a0 = s * d[0], a1 = s * d[1], a2 = s * d[2], a3 = cos(n);
        out[3] = a3 * x[3] - a0 * x[0] - a1 * x[1] - a2 * x[2];
        out[0] = a3 * x[0] + a0 * x[3] + a1 * x[2] - a2 * x[1];
        out[1] = a3 * x[1] + a1 * x[3] + a2 * x[0] - a0 * x[2];
        out[2] = a3 * x[2] + a2 * x[3] + a0 * x[1] - a1 * x[0];
    } else {
        for (int k = 0; k < 4; ++k) out[k] = x[k];
    }
}

__device__ double block_sum128(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    const double s = (red[0] + red[1]) + (red[2] + red[3]);
    __syncthreads();
    return s;
}
__device__ double block_max128(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    const double s = fmax(fmax(red[0], red[1]), fmax(red[2], red[3]));
    __syncthreads();
    return s;
}

__global__ void __launch_bounds__(SOLVE_THREADS) icp_bfgs_init_kernel(SolveArrays A, int B, const double* __restrict__ w, const double* __restrict__ p,
                                                                      const double* __restrict__ q) {
    const int b = blockIdx.x, tid = threadIdx.x;
    if (b >= B) return;
    for (int k = tid; k < NS; k += SOLVE_THREADS) A.xw[(size_t)b * NS + k] = A.tw[(size_t)b * NS + k] = w[(size_t)b * NS + k];
    for (int k = tid; k < 3; k += SOLVE_THREADS) A.xp[(size_t)b * 3 + k] = A.tp[(size_t)b * 3 + k] = p[(size_t)b * 3 + k];
    for (int k = tid; k < 4 * NJ; k += SOLVE_THREADS) A.xq[(size_t)b * 4 * NJ + k] = A.tq[(size_t)b * 4 * NJ + k] = q[(size_t)b * 4 * NJ + k];
    double* H = A.H + (size_t)b * NT * NT;
    for (int k = tid; k < NT * NT; k += SOLVE_THREADS) H[k] = (k / NT == k % NT) ? 1.0 : 0.0;
    if (tid == 0) {
        BodyState s;
        memset(&s, 0, sizeof(s));
        s.phase = PH_INIT;
        A.st[b] = s;
    }
}

// One step of every body's state machine: consumes the evaluation (cost_t, grad_t) of the body's trial point.
__global__ void __launch_bounds__(SOLVE_THREADS) icp_bfgs_advance_kernel(SolveArrays A, int B) {
    const int b = blockIdx.x, tid = threadIdx.x;
    __shared__ BodyState S;
    __shared__ double gt[NT], g[NT], dir[NT], sv[NT], yv[NT], Hy[NT], xq[4 * NJ], xp[3], xw[NS], red[8];
    __shared__ int action;      // what the vector part has to do after thread 0's decision
    enum { ACT_NONE = 0, ACT_TRIAL = 1, ACT_BEGIN = 2, ACT_ACCEPT = 3 };
    if (tid == 0) S = A.st[b];
    __syncthreads();
    if (S.phase == PH_DONE) return;
    const int n = A.shape ? NT : NT - NS;
    // the evaluation that just finished
    double bad = 0.0;
    for (int k = tid; k < NT; k += SOLVE_THREADS) {
        double v = A.grad_t[(size_t)b * NT + k];
        if (k >= n) v = 0.0;
        gt[k] = v;
        bad += isfinite(v) ? 0.0 : 1.0;
        g[k] = A.g[(size_t)b * NT + k];
        dir[k] = A.dir[(size_t)b * NT + k];
    }
    for (int k = tid; k < 4 * NJ; k += SOLVE_THREADS) xq[k] = A.xq[(size_t)b * 4 * NJ + k];
    if (tid < 3) xp[tid] = A.xp[(size_t)b * 3 + tid];
    if (tid < NS) xw[tid] = A.xw[(size_t)b * NS + tid];
    const double cost_t = A.cost_t[b];
    const bool valid = block_sum128(bad, red) == 0.0 && isfinite(cost_t);
    double dd = 0.0;
    for (int k = tid; k < n; k += SOLVE_THREADS) dd += gt[k] * dir[k];
    const double slope_t = block_sum128(dd, red);      // directional derivative at the trial point (LineSearchFunction::Evaluate)

    // x - Plus(x, -grad) in the ambient space, max norm (EvaluateGradientNorms) — of the TRIAL point's gradient at the TRIAL point;
    // only used when the trial point becomes the accepted one (PH_INIT, PH_FINAL)
    auto grad_max_norm_trial = [&]() -> double {
        double m = 0.0;
        if (tid < NJ) {
            double neg[3] = {-gt[3 + 3 * tid], -gt[4 + 3 * tid], -gt[5 + 3 * tid]}, q2[4];
            const double* qq = A.tq + (size_t)b * 4 * NJ + 4 * tid;
            const double q0[4] = {qq[0], qq[1], qq[2], qq[3]};
            quat_plus(q0, neg, q2);
            for (int c = 0; c < 4; ++c) m = fmax(m, fabs(q0[c] - q2[c]));
        } else if (tid >= 32 && tid < 35) {
            const double pp = A.tp[(size_t)b * 3 + tid - 32];
            m = fabs(pp - (pp - gt[tid - 32]));
        } else if (tid >= 64 && tid < 64 + NS && A.shape) {
            const double ww = A.tw[(size_t)b * NS + tid - 64];
            m = fabs(ww - (ww - gt[3 + 3 * NJ + tid - 64]));
        }
        return block_max128(m, red);
    };
    auto push_record = [&](double cost, double cost_change, double gmax, double step_norm, double step_size) {
        if (A.trace && S.nrec < A.trace_cap) {
            double* t = A.trace + ((size_t)b * A.trace_cap + S.nrec) * TRACE_W;
            t[0] = cost; t[1] = cost_change; t[2] = gmax; t[3] = step_norm; t[4] = step_size; t[5] = S.ls_evals; t[6] = S.ls_iters;
        }
        ++S.nrec;
    };
    auto finish = [&](int term) { S.phase = PH_DONE; S.term = term; atomicAdd(A.done_count, 1); };
    // top of the zoom loop: either asks for the next sample (ACT_TRIAL) or ends the search with S.optimal
    auto zoom_next = [&]() -> int {   // thread 0
        if (S.ls_iters >= kMaxLsIters || fabs(S.high.x - S.low.x) * S.dmax < kMinStep) { S.optimal = S.low.x; S.phase = PH_FINAL; S.trial_step = S.optimal; return ACT_TRIAL; }
        ++S.ls_iters;
        const Sample& lo_step = S.low.x < S.high.x ? S.low : S.high;
        const Sample& hi_step = S.low.x < S.high.x ? S.high : S.low;
        S.trial_step = interpolating_step(lo_step, hi_step, lo_step.x, hi_step.x);
        S.phase = PH_ZOOM;
        return ACT_TRIAL;
    };

    if (tid == 0) { action = ACT_NONE; ++S.evals; }
    __syncthreads();
    if (S.phase == PH_INIT) {
        const double gmax = grad_max_norm_trial();
        if (tid == 0) {
            if (!valid) finish(TERM_FAILURE);
            else {
                S.cost = S.initial_cost = cost_t; S.gmax = gmax;
                push_record(cost_t, 0.0, gmax, 0.0, 0.0);
                if (gmax <= A.gradient_tolerance) finish(TERM_GRADIENT_TOL);
                else action = ACT_BEGIN;
            }
        }
        for (int k = tid; k < NT; k += SOLVE_THREADS) { g[k] = gt[k]; A.g[(size_t)b * NT + k] = gt[k]; }
    } else if (S.phase == PH_BRACKET) {
        if (tid == 0) {
            ++S.ls_evals;
            Sample& cur = S.current;
            cur.value = cost_t; cur.gradient = slope_t; cur.valid = valid ? 1 : 0;
            bool ended = false, failed = false, do_zoom = false;
            ++S.ls_iters;
            if (cur.valid && (cur.value > S.init.value + kSuffDecrease * S.init.gradient * cur.x || (S.previous.valid && cur.value > S.previous.value))) {
                do_zoom = true; S.low = S.previous; S.high = cur; ended = true;
            } else if (cur.valid && fabs(cur.gradient) <= -kSuffCurvature * S.init.gradient) {
                S.low = cur; S.high = cur; ended = true;
            } else if (cur.valid && cur.gradient >= 0) {
                do_zoom = true; S.low = cur; S.high = S.previous; ended = true;
            } else if (cur.valid && fabs(cur.x - S.previous.x) * S.dmax < kMinStep) {
                S.low = cur; ended = true;
            } else if (S.ls_iters >= kMaxLsIters) {
                if (cur.valid && cur.value < S.low.value) S.low = cur;
                ended = true;
            }
            if (!ended) {
                const double max_step = cur.valid ? cur.x * kMaxExpansion : cur.x;
                const double step = interpolating_step(S.previous, cur, S.previous.x, max_step);
                if (step * S.dmax < kMinStep) failed = true;
                else {
                    if (cur.valid) S.previous = cur;
                    S.current.x = step; S.trial_step = step; action = ACT_TRIAL;
                }
            } else {
                if (do_zoom && fabs(S.high.x - S.low.x) * S.dmax < kMinStep) do_zoom = false;
                if (!do_zoom) { S.optimal = S.low.x; S.phase = PH_FINAL; S.trial_step = S.optimal; action = ACT_TRIAL; }
                else if (S.low.gradient * (S.high.x - S.low.x) >= 0) failed = true;     // ZoomPhase: invalid bracket
                else action = zoom_next();
            }
            if (failed) finish(TERM_FAILURE);
        }
    } else if (S.phase == PH_ZOOM) {
        if (tid == 0) {
            ++S.ls_evals;
            Sample sol;
            sol.x = S.trial_step; sol.value = cost_t; sol.gradient = slope_t; sol.valid = valid ? 1 : 0; sol.pad = 0;
            if (!sol.valid) finish(TERM_FAILURE);
            else if (sol.value > S.init.value + kSuffDecrease * S.init.gradient * sol.x || sol.value >= S.low.value) {
                S.high = sol;
                action = zoom_next();
            } else if (fabs(sol.gradient) <= -kSuffCurvature * S.init.gradient) {
                S.optimal = sol.x; S.phase = PH_FINAL; S.trial_step = S.optimal; action = ACT_TRIAL;
            } else {
                if (sol.gradient * (S.high.x - S.low.x) >= 0) S.high = S.low;
                S.low = sol;
                action = zoom_next();
            }
        }
    } else if (S.phase == PH_FINAL) {
        // the accepted point was evaluated again (LineSearchMinimizer evaluates x_plus_delta itself): it becomes x
        const double gmax = grad_max_norm_trial();
        double sn = 0.0;
        for (int k = tid; k < n; k += SOLVE_THREADS) { const double d = S.optimal * dir[k]; sn += d * d; }
        const double step_norm = sqrt(block_sum128(sn, red));
        if (!valid) {
            if (tid == 0) finish(TERM_FAILURE);
        } else {
            for (int k = tid; k < NT; k += SOLVE_THREADS) {
                A.g_prev[(size_t)b * NT + k] = g[k];
                A.g[(size_t)b * NT + k] = gt[k];
                sv[k] = g[k];          // the old gradient, kept for the BFGS update below
                g[k] = gt[k];
            }
            for (int k = tid; k < NS; k += SOLVE_THREADS) A.xw[(size_t)b * NS + k] = xw[k] = A.tw[(size_t)b * NS + k];
            for (int k = tid; k < 3; k += SOLVE_THREADS) A.xp[(size_t)b * 3 + k] = xp[k] = A.tp[(size_t)b * 3 + k];
            for (int k = tid; k < 4 * NJ; k += SOLVE_THREADS) A.xq[(size_t)b * 4 * NJ + k] = xq[k] = A.tq[(size_t)b * 4 * NJ + k];
            __syncthreads();
            if (tid == 0) {
                S.prev_cost = S.cost; S.cost = cost_t; S.prev_step = S.optimal; S.gmax = gmax;
                const double cost_change = S.prev_cost - S.cost;
                push_record(S.cost, cost_change, gmax, step_norm, S.optimal);
                if (gmax <= A.gradient_tolerance) finish(TERM_GRADIENT_TOL);
                else if (step_norm <= A.parameter_tolerance * (S.xnorm + A.parameter_tolerance)) finish(TERM_PARAMETER_TOL);
                else if (fabs(cost_change) <= A.function_tolerance * S.prev_cost) finish(TERM_FUNCTION_TOL);
                else action = ACT_BEGIN;
            }
        }
    }
    __syncthreads();

    if (action == ACT_BEGIN) {
        // ---- next iteration: direction, initial step size, first bracketing sample ----
        double xn = 0.0;
        for (int k = tid; k < 4 * NJ; k += SOLVE_THREADS) xn += xq[k] * xq[k];
        if (tid < 3) xn += xp[tid] * xp[tid];
        if (tid < NS && A.shape) xn += xw[tid] * xw[tid];
        const double xnorm = sqrt(block_sum128(xn, red));
        bool stop = false;
        if (tid == 0) {
            S.xnorm = xnorm;
            if (S.iteration >= A.max_iterations) { finish(TERM_NO_CONVERGENCE); stop = true; }
            else ++S.iteration;
            action = stop ? ACT_NONE : ACT_BEGIN;
        }
        __syncthreads();
        if (action == ACT_BEGIN) {
            double* H = A.H + (size_t)b * NT * NT;
            bool dir_ok = true;
            if (S.iteration == 1) {
                for (int k = tid; k < NT; k += SOLVE_THREADS) dir[k] = -g[k];
                __syncthreads();
            } else {
                // BFGS::NextDirection: s = previous direction * previous step size, y = gradient difference (old gradient in sv)
                double sy = 0.0;
                for (int k = tid; k < n; k += SOLVE_THREADS) {
                    const double y = g[k] - sv[k];
                    yv[k] = y;
                    const double s = dir[k] * S.prev_step;
                    sy += s * y;
                }
                __syncthreads();
                for (int k = tid; k < n; k += SOLVE_THREADS) sv[k] = dir[k] * S.prev_step;
                sy = block_sum128(sy, red);
                if (sy > kSecant) {
                    double yHy = 0.0;
                    for (int i = tid; i < n; i += SOLVE_THREADS) {
                        double a = 0.0;
                        for (int j = 0; j < n; ++j) a += H[(size_t)i * NT + j] * yv[j];
                        Hy[i] = a;
                        yHy += yv[i] * a;
                    }
                    yHy = block_sum128(yHy, red);
                    const double rho = 1.0 / sy, c1 = (1.0 + rho * yHy) * rho;
                    for (int idx = tid; idx < n * n; idx += SOLVE_THREADS) {
                        const int i = idx / n, j = idx % n;
                        H[(size_t)i * NT + j] += c1 * sv[i] * sv[j] - rho * (Hy[i] * sv[j] + sv[i] * Hy[j]);
                    }
                    __syncthreads();
                }
                double dg = 0.0;
                for (int i = tid; i < n; i += SOLVE_THREADS) {
                    double a = 0.0;
                    for (int j = 0; j < n; ++j) a += H[(size_t)i * NT + j] * g[j];
                    dir[i] = -a;
                    dg += -a * g[i];
                }
                dg = block_sum128(dg, red);
                if (dg >= 0.0) dir_ok = false;
            }
            bool give_up = false;
            if (!dir_ok) {     // block-uniform
                if (S.restarts >= kMaxRestarts) give_up = true;
                else {
                    for (int k = tid; k < NT * NT; k += SOLVE_THREADS) H[k] = (k / NT == k % NT) ? 1.0 : 0.0;
                    for (int k = tid; k < NT; k += SOLVE_THREADS) dir[k] = -g[k];
                }
                __syncthreads();
            }
            for (int k = n + tid; k < NT; k += SOLVE_THREADS) dir[k] = 0.0;
            __syncthreads();
            double dd2 = 0.0, dm = 0.0;
            for (int k = tid; k < n; k += SOLVE_THREADS) { dd2 += g[k] * dir[k]; dm = fmax(dm, fabs(dir[k])); }
            const double dirderiv = block_sum128(dd2, red);
            const double dmax = block_max128(dm, red);
            for (int k = tid; k < NT; k += SOLVE_THREADS) A.dir[(size_t)b * NT + k] = dir[k];
            if (tid == 0) {
                if (give_up) { finish(TERM_FAILURE); action = ACT_NONE; }
                else {
                    if (!dir_ok) ++S.restarts;
                    S.dirderiv = dirderiv; S.dmax = dmax;
                    const double step0 = (S.iteration == 1 || !dir_ok) ? fmin(1.0, 1.0 / S.gmax) : fmin(1.0, 2.0 * (S.cost - S.prev_cost) / dirderiv);
                    if (step0 < 0.0) { finish(TERM_FAILURE); action = ACT_NONE; }
                    else {
                        S.init.x = 0.0; S.init.value = S.cost; S.init.gradient = dirderiv; S.init.valid = 1; S.init.pad = 0;
                        S.low = S.high = S.previous = S.init;
                        S.current = S.init; S.current.x = step0; S.current.valid = 0;
                        S.ls_iters = 0; S.ls_evals = 0;
                        S.trial_step = step0;
                        S.phase = PH_BRACKET;
                        action = ACT_TRIAL;
                    }
                }
            }
            __syncthreads();
        }
    }
    if (action == ACT_TRIAL) {
        // trial point = Plus(x, step * direction)
        const double step = S.trial_step;
        if (tid < NJ) {
            const double d[3] = {step * dir[3 + 3 * tid], step * dir[4 + 3 * tid], step * dir[5 + 3 * tid]};
            double q2[4];
            quat_plus(xq + 4 * tid, d, q2);
            for (int c = 0; c < 4; ++c) A.tq[(size_t)b * 4 * NJ + 4 * tid + c] = q2[c];
        } else if (tid >= 32 && tid < 35) {
            A.tp[(size_t)b * 3 + tid - 32] = xp[tid - 32] + step * dir[tid - 32];
        } else if (tid >= 64 && tid < 64 + NS) {
            A.tw[(size_t)b * NS + tid - 64] = xw[tid - 64] + step * dir[3 + 3 * NJ + tid - 64];
        }
    }
    __syncthreads();
    if (tid == 0) A.st[b] = S;
}

__global__ void icp_solve_summary_kernel(SolveArrays A, int B, double* __restrict__ w, double* __restrict__ p, double* __restrict__ q,
                                         int32_t* __restrict__ nrec, double* __restrict__ summary) {
    const int b = blockIdx.x, tid = threadIdx.x;
    if (b >= B) return;
    for (int k = tid; k < NS; k += blockDim.x) w[(size_t)b * NS + k] = A.xw[(size_t)b * NS + k];
    for (int k = tid; k < 3; k += blockDim.x) p[(size_t)b * 3 + k] = A.xp[(size_t)b * 3 + k];
    for (int k = tid; k < 4 * NJ; k += blockDim.x) q[(size_t)b * 4 * NJ + k] = A.xq[(size_t)b * 4 * NJ + k];
    if (tid == 0) {
        const BodyState& s = A.st[b];
        if (nrec) nrec[b] = s.nrec;
        if (summary) {
            double* o = summary + (size_t)b * 6;
            o[0] = s.initial_cost; o[1] = s.cost; o[2] = s.iteration; o[3] = s.evals; o[4] = s.restarts;
            o[5] = s.phase == PH_DONE ? s.term : TERM_NO_CONVERGENCE;
        }
    }
}

}  // namespace

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_icp_solve_dev(b3d_icp* h, const b3d_gmm* gmm, const b3d_icp_solve_opts* o, double* w, double* p, double* q, double* trace,
                      int32_t trace_cap, int32_t* nrec, double* summary, int32_t* evaluations_out, void* stream) {
    if (!h || !o || !w || !p || !q) return set_error(B3D_EINVAL, "b3d_icp_solve: null argument");
    if (o->max_iterations < 0) return set_error(B3D_EINVAL, "b3d_icp_solve: max_iterations < 0");
    if (o->beta_pose > 0.0 && !gmm) return set_error(B3D_EINVAL, "b3d_icp_solve: beta_pose > 0 needs the pose prior");
    if (trace && trace_cap <= 0) return set_error(B3D_EINVAL, "b3d_icp_solve: trace without capacity");
    cudaStream_t st = (cudaStream_t)stream;
    const int B = b3d_icp_num_bodies(h);
    if (B <= 0) return B3D_OK;
    const size_t nb = (size_t)B;
    // one allocation: state machines, points, gradients, inverse Hessians
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t s_st = al(sizeof(BodyState) * nb), s_w = al(8 * nb * NS), s_p = al(8 * nb * 3), s_q = al(8 * nb * 4 * NJ), s_c = al(8 * nb),
                 s_g = al(8 * nb * NT), s_H = al(8 * nb * NT * NT), s_d = 256;
    const size_t total = s_st + 2 * (s_w + s_p + s_q) + s_c + 4 * s_g + s_H + s_d;
    char* base = nullptr;
    B3D_CUDA(cudaMalloc(&base, total));
    SolveArrays A;
    char* c = base;
    A.st = (BodyState*)c; c += s_st;
    A.xw = (double*)c; c += s_w; A.xp = (double*)c; c += s_p; A.xq = (double*)c; c += s_q;
    A.tw = (double*)c; c += s_w; A.tp = (double*)c; c += s_p; A.tq = (double*)c; c += s_q;
    A.cost_t = (double*)c; c += s_c;
    A.grad_t = (double*)c; c += s_g; A.g = (double*)c; c += s_g; A.g_prev = (double*)c; c += s_g; A.dir = (double*)c; c += s_g;
    A.H = (double*)c; c += s_H;
    A.done_count = (int32_t*)c;
    A.trace = trace; A.trace_cap = trace_cap;
    A.shape = (o->flags & B3D_ICP_POSE_ONLY) ? 0 : 1;
    A.max_iterations = o->max_iterations;
    A.function_tolerance = o->function_tolerance; A.gradient_tolerance = o->gradient_tolerance; A.parameter_tolerance = o->parameter_tolerance;
    int rc = B3D_OK;
    cudaError_t e = cudaMemsetAsync(base + s_st + 2 * (s_w + s_p + s_q) + s_c, 0, 4 * s_g, st);      // gradients / direction start at zero
    if (e == cudaSuccess) e = cudaMemsetAsync(A.done_count, 0, s_d, st);
    int evals = 0;
    if (e == cudaSuccess) {
        icp_bfgs_init_kernel<<<B, SOLVE_THREADS, 0, st>>>(A, B, w, p, q);
        count_launch();
        // every body needs at most 1 + max_iterations * (line-search evaluations + the accepted point) evaluations
        const long long cap = 1 + (long long)o->max_iterations * (kMaxLsIters + 2);
        int32_t done = 0;
        for (long long it = 0; it < cap && !rc; ++it) {
            rc = b3d_icp_eval_dev(h, A.tw, A.tp, A.tq, o->flags & B3D_ICP_POSE_ONLY, A.cost_t, A.grad_t, st);
            if (!rc) rc = b3d_icp_add_priors_dev(h, gmm, o->beta_pose, o->beta_shape, A.tw, A.tq, A.cost_t, A.grad_t, st);
            if (rc) break;
            icp_bfgs_advance_kernel<<<B, SOLVE_THREADS, 0, st>>>(A, B);
            count_launch();
            ++evals;
            if ((it & 3) == 3 || it + 1 == cap) {     // look at the done counter every fourth round
                e = cudaMemcpyAsync(&done, A.done_count, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaStreamSynchronize(st);
                if (e != cudaSuccess || done >= B) break;
            }
        }
        if (!rc && e == cudaSuccess) {
            icp_solve_summary_kernel<<<B, 128, 0, st>>>(A, B, w, p, q, nrec, summary);
            count_launch();
            e = cudaGetLastError();
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
    }
    cudaFree(base);
    if (evaluations_out) *evaluations_out = evals;
    if (rc) return rc;
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_icp_solve: %s", cudaGetErrorString(e));
    return B3D_OK;
}

int b3d_icp_solve(b3d_icp* h, const b3d_gmm* gmm, const b3d_icp_solve_opts* o, double* w, double* p, double* q, double* trace,
                  int32_t trace_cap, int32_t* nrec, double* summary, int32_t* evaluations_out) {
    if (!h || !o || !w || !p || !q) return set_error(B3D_EINVAL, "b3d_icp_solve: null argument");
    const size_t B = (size_t)b3d_icp_num_bodies(h);
    if (B == 0) return B3D_OK;
    const size_t n_w = B * NS, n_p = B * 3, n_q = B * 4 * NJ, n_t = trace ? B * (size_t)std::max(trace_cap, 0) * TRACE_W : 0, n_s = B * 6;
    double* d = nullptr;
    int32_t* dn = nullptr;
    B3D_CUDA(cudaMalloc(&d, sizeof(double) * (n_w + n_p + n_q + n_t + n_s)));
    cudaError_t e = cudaMalloc(&dn, sizeof(int32_t) * B);
    double *dw = d, *dp = dw + n_w, *dq = dp + n_p, *dt = dq + n_q, *ds = dt + n_t;
    if (e == cudaSuccess) e = cudaMemcpy(dw, w, sizeof(double) * n_w, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dp, p, sizeof(double) * n_p, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dq, q, sizeof(double) * n_q, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n_t) e = cudaMemset(dt, 0, sizeof(double) * n_t);
    int rc = B3D_OK;
    if (e == cudaSuccess) {
        rc = b3d_icp_solve_dev(h, gmm, o, dw, dp, dq, trace ? dt : nullptr, trace_cap, dn, ds, evaluations_out, nullptr);
        if (!rc) {
            e = cudaMemcpy(w, dw, sizeof(double) * n_w, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(p, dp, sizeof(double) * n_p, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(q, dq, sizeof(double) * n_q, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && n_t) e = cudaMemcpy(trace, dt, sizeof(double) * n_t, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && nrec) e = cudaMemcpy(nrec, dn, sizeof(int32_t) * B, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && summary) e = cudaMemcpy(summary, ds, sizeof(double) * n_s, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(d);
    cudaFree(dn);
    if (rc) return rc;
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_icp_solve: %s", cudaGetErrorString(e));
    return B3D_OK;
}

}  // extern "C"
