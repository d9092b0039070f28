// CPU oracle (test infrastructure only): the minimiser ark::AvatarOptimizer::optimize hands its ICP objective to
// (external/avatar/AvatarOptimizer.cpp:1313-1339 options, :1486 ceres::Solve):
//     minimizer_type = LINE_SEARCH, line_search_direction_type = BFGS, line_search_interpolation_type = CUBIC,
//     line_search_type = WOLFE (Ceres default), max_num_iterations = maxItersPerICP, function_tolerance = 1e-4,
//     every other Solver::Options field at its Ceres 1.14 default.
// Ceres Solver 1.14 is a third-party dependency that is absent from /root/reference (README.md:26, CMakeLists.txt:12), so what is
// restated here is its PUBLISHED algorithm: LineSearchMinimizer::Minimize (line_search_minimizer.cc), BFGS::NextDirection
// (line_search_direction.cc: dense inverse-Hessian update, secant-condition guard 1e-14, no eigenvalue scaling),
// WolfeLineSearch::DoSearch / BracketingPhase / ZoomPhase (line_search.cc, Nocedal & Wright alg. 3.5 / 3.6 with Ceres' guards) and
// MinimizeInterpolatingPolynomial (polynomial.cc) for two samples with value and gradient.  PARITY UNPINNED: nothing in the
// reference exercises this loop in a test and Ceres cannot be built here; the restatement is pinned only by its own properties
// (tests/test_icp_cpu.py: monotone cost, strong-Wolfe step sizes, convergence on a quadratic to the closed-form minimum).
#include <algorithm>
#include <cmath>
#include <vector>

#include "oracle.hpp"

namespace orc {

namespace {

struct LsSample {
    double x = 0, value = 0, gradient = 0;
    bool valid = false;
};

// polynomial.cc MinimizeInterpolatingPolynomial for two samples carrying value and gradient: the cubic through them, minimised over
// [x_min, x_max] among the interval's midpoint, its ends and the real parts of the roots of the derivative
double minimize_cubic(const LsSample& a, const LsSample& b, double x_min, double x_max) {
    // Hermite divided differences on nodes a, a, b, b
    const double h = b.x - a.x;
    double best_x = (x_min + x_max) / 2.0;
    if (h == 0.0) return best_x;
    const double d1 = a.gradient, d2 = (b.value - a.value) / h, d3 = b.gradient;
    const double e1 = (d2 - d1) / h, e2 = (d3 - d2) / h;
    const double f1 = (e2 - e1) / h;
    // p(x) = a.value + d1 (x - a) + e1 (x - a)^2 + f1 (x - a)^2 (x - b)  ->  monomial coefficients c3 x^3 + c2 x^2 + c1 x + c0
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
    // derivative 3 c3 x^2 + 2 c2 x + c1 (leading zeros removed like RemoveLeadingZeros)
    double qa = 3.0 * c3, qb = 2.0 * c2, qc = c1;
    double roots[2];
    int nr = 0;
    if (qa == 0.0) {
        if (qb != 0.0) roots[nr++] = -qc / qb;
    } else {
        const double D = qb * qb - 4.0 * qa * qc, sD = std::sqrt(std::fabs(D));
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

// LineSearch::InterpolatingPolynomialMinimizingStepSize with interpolation_type = CUBIC and no third sample
double interpolating_step(const LsSample& lowerbound, const LsSample& current, double min_step, double max_step) {
    if (!current.valid) return std::min(std::max(current.x * 0.5, min_step), max_step);
    return minimize_cubic(lowerbound, current, min_step, max_step);
}

struct Objective {
    const Model& M;
    const GMM* gmm;
    double beta_pose, beta_shape;
    int ncorr;
    const int *mpt, *dpt;
    const double* data;
    int shape;
    double total;   // number of ICP residual blocks (matched pairs)
    int evals = 0;
    // x = w (S) | p (3) | q (4 J); grad in the tangent layout of icp_eval (3 + 3 J + S)
    bool eval(const double* w, const double* p, const double* q, double* cost, double* grad) {
        ++evals;
        icp_eval(M, w, p, q, ncorr, mpt, dpt, data, shape, nullptr, nullptr, cost, grad);
        icp_priors(gmm, beta_pose, beta_shape, total, M.S, w, q, cost, grad);
        if (!shape) for (int s = 0; s < M.S; ++s) grad[3 + 3 * M.J + s] = 0.0;
        bool ok = std::isfinite(*cost);
        for (int k = 0; k < 3 + 3 * M.J + M.S; ++k) ok = ok && std::isfinite(grad[k]);
        return ok;
    }
};

}  // namespace

// Plus of the whole parameter vector (tangent layout: p | 24 x rotation tangent | w)
static void plus_all(int J, int S, const double* w, const double* p, const double* q, const double* d, double* w2, double* p2, double* q2) {
    for (int c = 0; c < 3; ++c) p2[c] = p[c] + d[c];
    for (int j = 0; j < J; ++j) quat_plus(q + 4 * j, d + 3 + 3 * j, q2 + 4 * j);
    for (int s = 0; s < S; ++s) w2[s] = w[s] + d[3 + 3 * J + s];
}

int icp_minimize(const Model& M, const GMM* gmm, double beta_pose, double beta_shape, int ncorr, const int* mpt, const int* dpt,
                 const double* data, int shape_enabled, const IcpSolveOptions& o, double* w, double* p, double* q,
                 std::vector<IcpSolveRecord>* trace, IcpSolveSummary* summary) {
    const int J = M.J, S = M.S, NC = 3 + 3 * J + S, n = shape_enabled ? NC : NC - S;
    double total = 0;
    for (int e = 0; e < ncorr; ++e) total += (mpt[e] >= 0) ? 1.0 : 0.0;
    Objective F{M, gmm, beta_pose, beta_shape, ncorr, mpt, dpt, data, shape_enabled, total};
    // Solver::Options defaults of Ceres 1.14
    const double min_step_size = 1e-9, sufficient_decrease = 1e-4, sufficient_curvature_decrease = 0.9, max_step_expansion = 10.0;
    const int max_ls_iterations = 20, max_restarts = 5;

    std::vector<double> g(NC), g_prev(NC), dir(NC, 0.0), delta(NC), H((size_t)NC * NC, 0.0), Hy(NC), yv(NC), sv(NC), gt(NC);
    std::vector<double> tw(S), tq(4 * J);
    double tp[3];
    auto reset_H = [&]() { std::fill(H.begin(), H.end(), 0.0); for (int i = 0; i < NC; ++i) H[(size_t)i * NC + i] = 1.0; };
    reset_H();
    // gradient norms of Ceres' EvaluateGradientNorms: x - Plus(x, -g) in the ambient space
    auto grad_max_norm = [&](const double* ww, const double* pp, const double* qq, const double* gg) {
        std::vector<double> neg(NC), w2(S), q2(4 * J);
        double p2[3];
        for (int k = 0; k < NC; ++k) neg[k] = -gg[k];
        plus_all(J, S, ww, pp, qq, neg.data(), w2.data(), p2, q2.data());
        double m = 0.0;
        for (int c = 0; c < 3; ++c) m = std::max(m, std::fabs(pp[c] - p2[c]));
        for (int k = 0; k < 4 * J; ++k) m = std::max(m, std::fabs(qq[k] - q2[k]));
        if (shape_enabled) for (int s = 0; s < S; ++s) m = std::max(m, std::fabs(ww[s] - w2[s]));
        return m;
    };
    auto x_norm = [&]() {
        double a = 0.0;
        for (int c = 0; c < 3; ++c) a += p[c] * p[c];
        for (int k = 0; k < 4 * J; ++k) a += q[k] * q[k];
        if (shape_enabled) for (int s = 0; s < S; ++s) a += w[s] * w[s];
        return std::sqrt(a);
    };

    IcpSolveSummary sm;
    double cost = 0.0, prev_cost = 0.0, prev_step = 0.0;
    if (!F.eval(w, p, q, &cost, g.data())) { sm.termination = ICP_TERM_FAILURE; if (summary) *summary = sm; return sm.termination; }
    double gmax = grad_max_norm(w, p, q, g.data());
    sm.initial_cost = cost;
    IcpSolveRecord rec{};
    rec.cost = cost; rec.gradient_max_norm = gmax;
    if (trace) trace->push_back(rec);
    int iteration = 0, restarts = 0;
    sm.termination = ICP_TERM_NO_CONVERGENCE;
    if (gmax <= o.gradient_tolerance) sm.termination = ICP_TERM_GRADIENT_TOL;
    else
        for (;;) {
            if (iteration >= o.max_iterations) { sm.termination = ICP_TERM_NO_CONVERGENCE; break; }
            ++iteration;
            bool dir_ok = true;
            if (iteration == 1) {
                for (int k = 0; k < NC; ++k) dir[k] = -g[k];
            } else {
                // BFGS::NextDirection: s = previous direction * previous step size, y = gradient difference
                double sy = 0.0;
                for (int k = 0; k < n; ++k) { sv[k] = dir[k] * prev_step; yv[k] = g[k] - g_prev[k]; sy += sv[k] * yv[k]; }
                if (sy > 1e-14) {
                    double yHy = 0.0;
                    for (int i = 0; i < n; ++i) {
                        double a = 0.0;
                        for (int j = 0; j < n; ++j) a += H[(size_t)i * NC + j] * yv[j];
                        Hy[i] = a;
                    }
                    for (int i = 0; i < n; ++i) yHy += yv[i] * Hy[i];
                    const double rho = 1.0 / sy, c1 = (1.0 + rho * yHy) * rho;
                    for (int i = 0; i < n; ++i)
                        for (int j = 0; j < n; ++j) H[(size_t)i * NC + j] += c1 * sv[i] * sv[j] - rho * (Hy[i] * sv[j] + sv[i] * Hy[j]);
                }
                double dg = 0.0;
                for (int i = 0; i < n; ++i) {
                    double a = 0.0;
                    for (int j = 0; j < n; ++j) a += H[(size_t)i * NC + j] * g[j];
                    dir[i] = -a;
                    dg += dir[i] * g[i];
                }
                if (dg >= 0.0) dir_ok = false;   // not a descent direction
            }
            if (!dir_ok) {
                if (restarts >= max_restarts) { sm.termination = ICP_TERM_FAILURE; break; }
                ++restarts;
                reset_H();
                for (int k = 0; k < NC; ++k) dir[k] = -g[k];
            }
            for (int k = n; k < NC; ++k) dir[k] = 0.0;
            double dirderiv = 0.0, dmax = 0.0;
            for (int k = 0; k < n; ++k) { dirderiv += g[k] * dir[k]; dmax = std::max(dmax, std::fabs(dir[k])); }
            const double initial_step = (iteration == 1 || !dir_ok) ? std::min(1.0, 1.0 / gmax)
                                                                    : std::min(1.0, 2.0 * (cost - prev_cost) / dirderiv);
            if (initial_step < 0.0) { sm.termination = ICP_TERM_FAILURE; break; }

            // ---- WolfeLineSearch::DoSearch ----
            int ls_iters = 0, ls_evals = 0;
            auto sample_at = [&](double step) {
                LsSample s;
                s.x = step;
                for (int k = 0; k < NC; ++k) delta[k] = step * dir[k];
                plus_all(J, S, w, p, q, delta.data(), tw.data(), tp, tq.data());
                double c = 0.0;
                s.valid = F.eval(tw.data(), tp, tq.data(), &c, gt.data());
                ++ls_evals;
                s.value = c;
                double d = 0.0;
                for (int k = 0; k < n; ++k) d += gt[k] * dir[k];
                s.gradient = d;
                return s;
            };
            LsSample init;
            init.x = 0.0; init.value = cost; init.gradient = dirderiv; init.valid = true;
            LsSample low = init, high = init, previous = init, current = sample_at(initial_step);
            bool do_zoom = false, ls_ok = true;
            for (;;) {   // BracketingPhase
                ++ls_iters;
                if (current.valid && (current.value > init.value + sufficient_decrease * init.gradient * current.x ||
                                      (previous.valid && current.value > previous.value))) {
                    do_zoom = true; low = previous; high = current; break;
                }
                if (current.valid && std::fabs(current.gradient) <= -sufficient_curvature_decrease * init.gradient) {
                    low = current; high = current; break;
                } else if (current.valid && current.gradient >= 0) {
                    do_zoom = true; low = current; high = previous; break;
                } else if (current.valid && std::fabs(current.x - previous.x) * dmax < min_step_size) {
                    low = current; break;
                } else if (ls_iters >= max_ls_iterations) {
                    if (current.valid && current.value < low.value) low = current;
                    break;
                }
                const double max_step = current.valid ? current.x * max_step_expansion : current.x;
                const double step = interpolating_step(previous, current, previous.x, max_step);
                if (step * dmax < min_step_size) { ls_ok = false; break; }
                if (current.valid) previous = current;
                current = sample_at(step);
            }
            if (ls_ok && do_zoom && std::fabs(high.x - low.x) * dmax < min_step_size) do_zoom = false;
            double optimal = 0.0;
            if (ls_ok && !do_zoom) {
                optimal = low.x;
            } else if (ls_ok) {   // ZoomPhase
                LsSample solution;
                bool zoom_ok = true;
                if (low.gradient * (high.x - low.x) >= 0) { zoom_ok = false; solution.valid = false; }
                while (zoom_ok) {
                    solution = low;
                    if (ls_iters >= max_ls_iterations) { zoom_ok = false; break; }
                    if (std::fabs(high.x - low.x) * dmax < min_step_size) { zoom_ok = false; break; }
                    ++ls_iters;
                    const LsSample& lo_step = low.x < high.x ? low : high;
                    const LsSample& hi_step = low.x < high.x ? high : low;
                    const double step = interpolating_step(lo_step, hi_step, lo_step.x, hi_step.x);
                    solution = sample_at(step);
                    if (!solution.valid) { zoom_ok = false; break; }
                    if (solution.value > init.value + sufficient_decrease * init.gradient * solution.x || solution.value >= low.value) {
                        high = solution;
                        continue;
                    }
                    if (std::fabs(solution.gradient) <= -sufficient_curvature_decrease * init.gradient) break;
                    else if (solution.gradient * (high.x - low.x) >= 0) high = low;
                    low = solution;
                }
                if (!zoom_ok && !solution.valid) ls_ok = false;
                else {
                    if (!solution.valid || solution.value > low.value) solution = low;
                    optimal = solution.x;
                }
            }
            if (!ls_ok) { sm.termination = ICP_TERM_FAILURE; break; }

            // ---- take the step: x <- Plus(x, optimal * direction), evaluated again like LineSearchMinimizer does ----
            for (int k = 0; k < NC; ++k) delta[k] = optimal * dir[k];
            const double xn = x_norm();
            plus_all(J, S, w, p, q, delta.data(), tw.data(), tp, tq.data());
            double new_cost = 0.0;
            g_prev = g;
            if (!F.eval(tw.data(), tp, tq.data(), &new_cost, g.data())) { sm.termination = ICP_TERM_FAILURE; break; }
            for (int s = 0; s < S; ++s) w[s] = tw[s];
            for (int c = 0; c < 3; ++c) p[c] = tp[c];
            for (int k = 0; k < 4 * J; ++k) q[k] = tq[k];
            prev_cost = cost; cost = new_cost; prev_step = optimal;
            gmax = grad_max_norm(w, p, q, g.data());
            double sn = 0.0;
            for (int k = 0; k < n; ++k) sn += delta[k] * delta[k];
            rec = IcpSolveRecord{};
            rec.cost = cost; rec.cost_change = prev_cost - cost; rec.gradient_max_norm = gmax; rec.step_norm = std::sqrt(sn);
            rec.step_size = optimal; rec.line_search_evals = ls_evals; rec.line_search_iters = ls_iters;
            if (trace) trace->push_back(rec);
            if (gmax <= o.gradient_tolerance) { sm.termination = ICP_TERM_GRADIENT_TOL; break; }
            if (rec.step_norm <= o.parameter_tolerance * (xn + o.parameter_tolerance)) { sm.termination = ICP_TERM_PARAMETER_TOL; break; }
            if (std::fabs(rec.cost_change) <= o.function_tolerance * prev_cost) { sm.termination = ICP_TERM_FUNCTION_TOL; break; }
        }
    sm.final_cost = cost;
    sm.iterations = iteration;
    sm.evaluations = F.evals;
    sm.restarts = restarts;
    if (summary) *summary = sm;
    return sm.termination;
}

}  // namespace orc
