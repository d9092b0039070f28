// CPU ORACLE (test infrastructure only) — Levenberg-Marquardt trust-region loop as configured by the
// reference's three ceres::Solve call sites (Sim3BA.h:473-479, 641-647; MultiFrameBA.h:145-151):
// Ceres Solver 1.14 defaults + DENSE_QR.  Ceres is an external dependency that is NOT in /root/reference
// (README.md:26 pins 1.14); this file restates its published algorithm:
//   TrustRegionMinimizer::{Minimize, IterationZero, EvaluateGradientAndJacobian, ComputeTrustRegionStep,
//   HandleInvalidStep, DoLineSearch, ComputeCandidatePointAndEvaluateCost, ParameterToleranceReached,
//   FunctionToleranceReached, IsStepSuccessful, HandleSuccessfulStep, HandleUnsuccessfulStep,
//   FinalizeIterationAndCheckIfMinimizerCanContinue}, LevenbergMarquardtStrategy::{ComputeStep, StepAccepted,
//   StepRejected}, DenseQRSolver, ArmijoLineSearch, polynomial.cc (MinimizeInterpolatingPolynomial).
#include <complex>

#include "oracle.hpp"

namespace orc {

namespace {

constexpr double kInitialRadius = 1e4, kMaxRadius = 1e16, kMinRadius = 1e-32;
constexpr double kMinDiagonal = 1e-6, kMaxDiagonal = 1e32;
constexpr double kMinRelativeDecrease = 1e-3;
constexpr double kFunctionTolerance = 1e-6, kGradientTolerance = 1e-10, kParameterTolerance = 1e-8;
constexpr int kMaxConsecutiveInvalidSteps = 5;
constexpr double kSufficientDecrease = 1e-4, kMinLineSearchStepSize = 1e-9;
constexpr int kMaxLineSearchIters = 20;

double norm2(const std::vector<double>& v) { double s = 0; for (double x : v) s += x * x; return std::sqrt(s); }

bool all_finite(const std::vector<double>& v) { for (double x : v) if (!std::isfinite(x)) return false; return true; }

// ---- polynomial helpers (coefficients highest degree first, like ceres/internal/polynomial.cc) ----
double poly_eval(const std::vector<double>& p, double x) { double v = 0; for (double c : p) v = v * x + c; return v; }

std::vector<double> poly_diff(const std::vector<double>& p) {
    const int deg = (int)p.size() - 1;
    if (deg == 0) return {0.0};
    std::vector<double> d(deg);
    for (int i = 0; i < deg; ++i) d[i] = (deg - i) * p[i];
    return d;
}

// real parts of all (complex) roots — ceres checks every real part ("a bit of an overkill")
bool poly_root_real_parts(std::vector<double> p, std::vector<double>* out) {
    out->clear();
    size_t lead = 0;
    while (lead + 1 < p.size() && p[lead] == 0.0) ++lead;   // RemoveLeadingZeros
    p.erase(p.begin(), p.begin() + lead);
    const int deg = (int)p.size() - 1;
    if (deg <= 0) return deg == 0 ? true : false;
    if (deg == 1) { out->push_back(-p[1] / p[0]); return true; }
    if (deg == 2) {
        const double a = p[0], b = p[1], c = p[2];
        const double D = b * b - 4 * a * c, sD = std::sqrt(std::fabs(D));
        if (D >= 0) {
            if (b >= 0) { out->push_back((-b - sD) / (2.0 * a)); out->push_back((2.0 * c) / (-b - sD)); }
            else { out->push_back((2.0 * c) / (-b + sD)); out->push_back((-b + sD) / (2.0 * a)); }
        } else { out->push_back(-b / (2.0 * a)); out->push_back(-b / (2.0 * a)); }
        return true;
    }
    // higher degree: ceres uses companion-matrix eigenvalues; Aberth-Ehrlich iteration gives the same roots
    using C = std::complex<double>;
    std::vector<C> a(deg + 1);
    for (int i = 0; i <= deg; ++i) a[i] = p[i] / p[0];
    double rad = 0;
    for (int i = 1; i <= deg; ++i) rad = std::max(rad, std::abs(a[i]));
    rad = 1.0 + rad;
    std::vector<C> z(deg);
    for (int i = 0; i < deg; ++i) z[i] = std::polar(rad * 0.5, 2 * M_PI * i / deg + 0.4);
    for (int it = 0; it < 500; ++it) {
        double change = 0;
        for (int i = 0; i < deg; ++i) {
            C f = a[0], df = 0;
            for (int k = 1; k <= deg; ++k) { df = df * z[i] + f; f = f * z[i] + a[k]; }
            if (std::abs(f) == 0) continue;
            C ratio = f / df, sum = 0;
            for (int j = 0; j < deg; ++j) if (j != i) sum += C(1.0) / (z[i] - z[j]);
            C dz = ratio / (C(1.0) - ratio * sum);
            z[i] -= dz;
            change = std::max(change, std::abs(dz));
        }
        if (change < 1e-15 * rad) break;
    }
    for (int i = 0; i < deg; ++i) out->push_back(z[i].real());
    return true;
}

struct Sample { double x = 0, value = 0, gradient = 0; bool value_valid = false, gradient_valid = false; };

// MinimizeInterpolatingPolynomial(samples, x_min, x_max)
double minimize_interpolating_polynomial(const std::vector<Sample>& samples, double x_min, double x_max) {
    int nc = 0;
    for (auto& s : samples) nc += (s.value_valid ? 1 : 0) + (s.gradient_valid ? 1 : 0);
    const int degree = nc - 1;
    std::vector<double> lhs((size_t)nc * nc, 0.0), rhs(nc, 0.0), coef(nc, 0.0);
    int row = 0;
    for (auto& s : samples) {
        if (s.value_valid) {
            for (int j = 0; j <= degree; ++j) lhs[(size_t)row * nc + j] = std::pow(s.x, degree - j);
            rhs[row++] = s.value;
        }
        if (s.gradient_valid) {
            for (int j = 0; j < degree; ++j) lhs[(size_t)row * nc + j] = (degree - j) * std::pow(s.x, degree - j - 1);
            rhs[row++] = s.gradient;
        }
    }
    double best_x = (x_min + x_max) / 2.0;
    if (!lu_fullpiv_solve(lhs.data(), nc, rhs.data(), coef.data())) return best_x;
    double best_v = poly_eval(coef, best_x);
    const double vmin = poly_eval(coef, x_min);
    if (vmin < best_v) { best_v = vmin; best_x = x_min; }
    const double vmax = poly_eval(coef, x_max);
    if (vmax < best_v) { best_v = vmax; best_x = x_max; }
    if (coef.size() <= 2) return best_x;
    std::vector<double> roots;
    if (!poly_root_real_parts(poly_diff(coef), &roots)) return best_x;
    for (double rt : roots) {
        if (rt < x_min || rt > x_max) continue;
        const double v = poly_eval(coef, rt);
        if (v < best_v) { best_v = v; best_x = rt; }
    }
    return best_x;
}

// ---- linear solvers: y = argmin |J y - r|^2 + |D y|^2 (J already column scaled) ----
bool solve_dense_qr(const SparseRows& J, const std::vector<double>& r, const std::vector<double>& D, std::vector<double>& y) {
    const int m = J.rows(), n = J.n;
    std::vector<double> A((size_t)(m + n) * n, 0.0), b(m + n, 0.0);
    for (int i = 0; i < m; ++i) {
        for (int e = J.rowptr[i]; e < J.rowptr[i + 1]; ++e) A[(size_t)i * n + J.col[e]] += J.val[e];
        b[i] = r[i];
    }
    for (int k = 0; k < n; ++k) A[(size_t)(m + k) * n + k] = D[k];
    return householder_qr_solve(A.data(), m + n, n, b.data(), y.data());
}

bool solve_dense_chol(const SparseRows& J, const std::vector<double>& r, const std::vector<double>& D, std::vector<double>& y) {
    const int m = J.rows(), n = J.n;
    std::vector<double> H((size_t)n * n, 0.0);
    std::fill(y.begin(), y.end(), 0.0);
    for (int i = 0; i < m; ++i)
        for (int e = J.rowptr[i]; e < J.rowptr[i + 1]; ++e) {
            y[J.col[e]] += J.val[e] * r[i];
            for (int f = J.rowptr[i]; f < J.rowptr[i + 1]; ++f) H[(size_t)J.col[e] * n + J.col[f]] += J.val[e] * J.val[f];
        }
    for (int k = 0; k < n; ++k) H[(size_t)k * n + k] += D[k] * D[k];
    if (!cholesky_lower(H.data(), n)) return false;
    chol_solve(H.data(), n, y.data());
    return true;
}

// block tridiagonal (F blocks of nb) + arrowhead (na) normal equations, block Cholesky + Schur complement
bool solve_block_tridiag(const SparseRows& J, const std::vector<double>& r, const std::vector<double>& D,
                         const BlockStructure& bs, std::vector<double>& y) {
    const int F = bs.F, nb = bs.nb, na = bs.na, n = J.n, m = J.rows();
    if (F * nb + na != n) return false;
    const size_t bb = (size_t)nb * nb;
    std::vector<double> Hd(F * bb, 0.0), Ho((size_t)std::max(F - 1, 0) * bb, 0.0);   // Ho[f] = H(f+1, f) block
    std::vector<double> Ha((size_t)F * nb * na, 0.0), Hw((size_t)na * na, 0.0), g(n, 0.0);
    for (int i = 0; i < m; ++i)
        for (int e = J.rowptr[i]; e < J.rowptr[i + 1]; ++e) {
            const int ce = J.col[e];
            g[ce] += J.val[e] * r[i];
            for (int f2 = J.rowptr[i]; f2 < J.rowptr[i + 1]; ++f2) {
                const int cf = J.col[f2];
                const double v = J.val[e] * J.val[f2];
                const bool ea = ce >= F * nb, fa = cf >= F * nb;
                if (ea && fa) Hw[(size_t)(ce - F * nb) * na + (cf - F * nb)] += v;
                else if (!ea && fa) Ha[(size_t)ce * na + (cf - F * nb)] += v;
                else if (!ea && !fa) {
                    const int be = ce / nb, bf = cf / nb;
                    if (be == bf) Hd[be * bb + (size_t)(ce % nb) * nb + (cf % nb)] += v;
                    else if (be == bf + 1) Ho[bf * bb + (size_t)(ce % nb) * nb + (cf % nb)] += v;
                    else if (be + 1 != bf) return false;   // structure violated
                }
            }
        }
    for (int k = 0; k < F * nb; ++k) Hd[(k / nb) * bb + (size_t)(k % nb) * nb + (k % nb)] += D[k] * D[k];
    for (int k = 0; k < na; ++k) Hw[(size_t)k * na + k] += D[F * nb + k] * D[F * nb + k];
    // forward block Cholesky: L_ff L_ff^T = Hd_f - L_{f,f-1} L_{f,f-1}^T ; L_{f+1,f} = Ho_f L_ff^-T
    // carry along the arrow rows: W_f = (Ha_f^T - W_{f-1} L_{f,f-1}^T) L_ff^-T  (na x nb), and rhs.
    std::vector<double> W((size_t)F * na * nb, 0.0);
    for (int f = 0; f < F; ++f) {
        double* Lf = &Hd[f * bb];
        if (f > 0) {
            const double* Lo = &Ho[(f - 1) * bb];   // L_{f,f-1}  (nb x nb)
            for (int a = 0; a < nb; ++a)
                for (int b = 0; b <= a; ++b) {
                    double s = 0;
                    for (int k = 0; k < nb; ++k) s += Lo[(size_t)a * nb + k] * Lo[(size_t)b * nb + k];
                    Lf[(size_t)a * nb + b] -= s;
                }
        }
        if (!cholesky_lower(Lf, nb)) return false;
        // arrow rows
        double* Wf = &W[(size_t)f * na * nb];
        for (int a = 0; a < na; ++a) {
            for (int c = 0; c < nb; ++c) {
                double s = Ha[((size_t)f * nb + c) * na + a];
                if (f > 0) {
                    const double* Wp = &W[(size_t)(f - 1) * na * nb];
                    const double* Lo = &Ho[(f - 1) * bb];
                    for (int k = 0; k < nb; ++k) s -= Wp[(size_t)a * nb + k] * Lo[(size_t)c * nb + k];
                }
                Wf[(size_t)a * nb + c] = s;
            }
            // solve row * L^T = rhs  => forward substitution over columns
            for (int c = 0; c < nb; ++c) {
                double s = Wf[(size_t)a * nb + c];
                for (int k = 0; k < c; ++k) s -= Wf[(size_t)a * nb + k] * Lf[(size_t)c * nb + k];
                Wf[(size_t)a * nb + c] = s / Lf[(size_t)c * nb + c];
            }
        }
        if (f + 1 < F) {
            double* Lo = &Ho[f * bb];   // rows of block f+1, cols of block f: Lo <- Ho L_ff^-T
            for (int a = 0; a < nb; ++a)
                for (int c = 0; c < nb; ++c) {
                    double s = Lo[(size_t)a * nb + c];
                    for (int k = 0; k < c; ++k) s -= Lo[(size_t)a * nb + k] * Lf[(size_t)c * nb + k];
                    Lo[(size_t)a * nb + c] = s / Lf[(size_t)c * nb + c];
                }
        }
    }
    // Schur complement for the arrow block
    for (int f = 0; f < F; ++f) {
        const double* Wf = &W[(size_t)f * na * nb];
        for (int a = 0; a < na; ++a)
            for (int b = 0; b <= a; ++b) {
                double s = 0;
                for (int k = 0; k < nb; ++k) s += Wf[(size_t)a * nb + k] * Wf[(size_t)b * nb + k];
                Hw[(size_t)a * na + b] -= s;
            }
    }
    if (na > 0 && !cholesky_lower(Hw.data(), na)) return false;
    // forward solve L z = g
    std::vector<double> z(g);
    for (int f = 0; f < F; ++f) {
        const double* Lf = &Hd[f * bb];
        double* zf = &z[(size_t)f * nb];
        if (f > 0) {
            const double* Lo = &Ho[(f - 1) * bb];
            const double* zp = &z[(size_t)(f - 1) * nb];
            for (int a = 0; a < nb; ++a) { double s = 0; for (int k = 0; k < nb; ++k) s += Lo[(size_t)a * nb + k] * zp[k]; zf[a] -= s; }
        }
        for (int a = 0; a < nb; ++a) { double s = zf[a]; for (int k = 0; k < a; ++k) s -= Lf[(size_t)a * nb + k] * zf[k]; zf[a] = s / Lf[(size_t)a * nb + a]; }
    }
    double* za = &z[(size_t)F * nb];
    for (int f = 0; f < F; ++f) {
        const double* Wf = &W[(size_t)f * na * nb];
        for (int a = 0; a < na; ++a) { double s = 0; for (int k = 0; k < nb; ++k) s += Wf[(size_t)a * nb + k] * z[(size_t)f * nb + k]; za[a] -= s; }
    }
    for (int a = 0; a < na; ++a) { double s = za[a]; for (int k = 0; k < a; ++k) s -= Hw[(size_t)a * na + k] * za[k]; za[a] = s / Hw[(size_t)a * na + a]; }
    // backward solve L^T y = z
    for (int a = na - 1; a >= 0; --a) { double s = za[a]; for (int k = a + 1; k < na; ++k) s -= Hw[(size_t)k * na + a] * za[k]; za[a] = s / Hw[(size_t)a * na + a]; }
    for (int f = F - 1; f >= 0; --f) {
        const double* Lf = &Hd[f * bb];
        const double* Wf = &W[(size_t)f * na * nb];
        double* zf = &z[(size_t)f * nb];
        for (int a = 0; a < nb; ++a) {
            double s = 0;
            for (int k = 0; k < na; ++k) s += Wf[(size_t)k * nb + a] * za[k];
            if (f + 1 < F) {
                const double* Lo = &Ho[f * bb];
                const double* zn = &z[(size_t)(f + 1) * nb];
                for (int k = 0; k < nb; ++k) s += Lo[(size_t)k * nb + a] * zn[k];
            }
            zf[a] -= s;
        }
        for (int a = nb - 1; a >= 0; --a) { double s = zf[a]; for (int k = a + 1; k < nb; ++k) s -= Lf[(size_t)k * nb + a] * zf[k]; zf[a] = s / Lf[(size_t)a * nb + a]; }
    }
    y = z;
    return true;
}

}  // namespace

LMSummary lm_minimize(LMProblem& P, const LMOptions& opt, double* x_inout, std::vector<IterRecord>* trace) {
    const int n = P.n;
    LMSummary sum;
    std::vector<double> x(x_inout, x_inout + n), cand(n), delta(n, 0.0), step(n), g(n), gc(n), scale(n, 1.0), diag(n), lmd(n), tmp(n), y(n);
    std::vector<double> r, rc;
    SparseRows J;
    double x_cost = 0, cand_cost = 0, radius = kInitialRadius, decrease_factor = 2.0, model_cost_change = 0;
    bool reuse_diagonal = false;
    int num_invalid = 0, iteration = 0;
    double x_norm = norm2(x);
    IterRecord it{};

    auto eval_grad_jac = [&](bool first) -> bool {
        if (!P.evaluate(x.data(), &x_cost, &r, &J, &g)) return false;
        if (!std::isfinite(x_cost)) return false;
        it.cost = x_cost;
        if (first) {
            std::vector<double> cs(n, 0.0);
            for (size_t e = 0; e < J.col.size(); ++e) cs[J.col[e]] += J.val[e] * J.val[e];
            for (int k = 0; k < n; ++k) scale[k] = 1.0 / (1.0 + std::sqrt(cs[k]));
        }
        for (size_t e = 0; e < J.col.size(); ++e) J.val[e] *= scale[J.col[e]];
        for (int k = 0; k < n; ++k) tmp[k] = -g[k];
        P.plus(x.data(), tmp.data(), cand.data());
        double mx = 0;
        for (int k = 0; k < n; ++k) mx = std::max(mx, std::fabs(x[k] - cand[k]));
        it.gradient_max_norm = mx;
        return true;
    };
    auto push = [&]() {
        it.radius = radius;
        it.aux = P.aux_state();
        if (trace) trace->push_back(it);
        ++sum.num_iterations;
    };

    // ---- IterationZero ----
    it = IterRecord{};
    if (P.constrained) {
        std::fill(delta.begin(), delta.end(), 0.0);
        P.plus(x.data(), delta.data(), cand.data());
        x = cand;
        x_norm = norm2(x);
    }
    if (!eval_grad_jac(true)) {
        sum.termination = T_FAILURE_EVAL; sum.usable = false;
        return sum;
    }
    sum.initial_cost = x_cost;
    it.flags = 3;
    bool step_successful = true;

    for (;;) {
        // ---- FinalizeIterationAndCheckIfMinimizerCanContinue ----
        push();
        if (iteration >= opt.max_iters) { sum.termination = T_NO_CONVERGENCE_ITERS; break; }
        if (step_successful && it.gradient_max_norm <= kGradientTolerance) { sum.termination = T_CONVERGENCE_GRAD; break; }
        if (!(radius > kMinRadius)) { sum.termination = T_NO_CONVERGENCE_RADIUS; break; }

        const double prev_gmax = it.gradient_max_norm;
        it = IterRecord{};
        ++iteration;
        step_successful = false;

        // ---- ComputeTrustRegionStep (LevenbergMarquardtStrategy::ComputeStep) ----
        if (!reuse_diagonal) {
            std::fill(diag.begin(), diag.end(), 0.0);
            for (size_t e = 0; e < J.col.size(); ++e) diag[J.col[e]] += J.val[e] * J.val[e];
            for (int k = 0; k < n; ++k) diag[k] = std::min(std::max(diag[k], kMinDiagonal), kMaxDiagonal);
        }
        for (int k = 0; k < n; ++k) lmd[k] = std::sqrt(diag[k] / radius);
        bool solved = false;
        switch (opt.linear_solver) {
            case LS_DENSE_QR: solved = solve_dense_qr(J, r, lmd, y); break;
            case LS_DENSE_CHOLESKY: solved = solve_dense_chol(J, r, lmd, y); break;
            default: solved = solve_block_tridiag(J, r, lmd, opt.blocks, y); break;
        }
        reuse_diagonal = true;
        bool step_valid = false;
        if (solved && all_finite(y)) {
            for (int k = 0; k < n; ++k) step[k] = -y[k];
            // model_cost_change = -(J step)^T (r + J step / 2)
            double mcc = 0;
            for (int i = 0; i < J.rows(); ++i) {
                double js = 0;
                for (int e = J.rowptr[i]; e < J.rowptr[i + 1]; ++e) js += J.val[e] * step[J.col[e]];
                mcc -= js * (r[i] + js / 2.0);
            }
            model_cost_change = mcc;
            step_valid = mcc > 0.0;
        }
        if (!step_valid) {
            // ---- HandleInvalidStep ----
            if (++num_invalid >= kMaxConsecutiveInvalidSteps) { sum.termination = T_FAILURE_INVALID_STEPS; break; }
            radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;   // StepIsInvalid -> StepRejected(0)
            it.cost = x_cost; it.cost_change = 0; it.gradient_max_norm = prev_gmax; it.step_norm = 0; it.relative_decrease = 0;
            it.flags = 0;
            continue;
        }
        num_invalid = 0;
        for (int k = 0; k < n; ++k) delta[k] = step[k] * scale[k];
        it.flags = 1;

        // ---- DoLineSearch (bounded problems only) ----
        if (P.constrained) {
            double gd = 0, dmax = 0;
            for (int k = 0; k < n; ++k) { gd += g[k] * delta[k]; dmax = std::max(dmax, std::fabs(delta[k])); }
            Sample init; init.x = 0; init.value = x_cost; init.gradient = gd; init.value_valid = init.gradient_valid = true;
            Sample prev, cur;
            auto ls_eval = [&](double a, Sample& s) {
                s = Sample{}; s.x = a;
                for (int k = 0; k < n; ++k) tmp[k] = a * delta[k];
                P.plus(x.data(), tmp.data(), cand.data());
                double c = 0;
                if (!P.evaluate(cand.data(), &c, nullptr, nullptr, &gc) || !std::isfinite(c)) return;
                double gg = 0;
                for (int k = 0; k < n; ++k) gg += delta[k] * gc[k];
                if (!std::isfinite(gg)) return;
                s.value = c; s.gradient = gg; s.value_valid = s.gradient_valid = true;
            };
            ls_eval(1.0, cur);
            bool success = true;
            int ls_iters = 0;
            while (!cur.value_valid || cur.value > x_cost + kSufficientDecrease * gd * cur.x) {
                ++ls_iters;
                if (ls_iters >= kMaxLineSearchIters) { success = false; break; }
                double step_size;
                const double lo = opt.max_line_search_step_contraction * cur.x, hi = opt.min_line_search_step_contraction * cur.x;
                if (!cur.value_valid) step_size = std::min(std::max(cur.x * 0.5, lo), hi);
                else {
                    std::vector<Sample> ss{init, cur};
                    if (prev.value_valid) ss.push_back(prev);
                    step_size = minimize_interpolating_polynomial(ss, lo, hi);
                }
                if (step_size * dmax < kMinLineSearchStepSize) { success = false; break; }
                prev = cur;
                ls_eval(step_size, cur);
            }
            sum.line_search_steps += ls_iters;
            if (success && cur.x != 1.0) {
                for (int k = 0; k < n; ++k) delta[k] *= cur.x;
                it.flags |= 4;
            }
        }

        // ---- ComputeCandidatePointAndEvaluateCost ----
        P.plus(x.data(), delta.data(), cand.data());
        if (!P.evaluate(cand.data(), &cand_cost, nullptr, nullptr, nullptr) || !std::isfinite(cand_cost))
            cand_cost = std::numeric_limits<double>::max();

        // ---- ParameterToleranceReached ----
        double sn = 0;
        for (int k = 0; k < n; ++k) sn += (x[k] - cand[k]) * (x[k] - cand[k]);
        it.step_norm = std::sqrt(sn);
        if (it.step_norm <= kParameterTolerance * (x_norm + kParameterTolerance)) { sum.termination = T_CONVERGENCE_PARAM; break; }
        // ---- FunctionToleranceReached ----
        it.cost_change = x_cost - cand_cost;
        if (std::fabs(it.cost_change) <= kFunctionTolerance * x_cost) { sum.termination = T_CONVERGENCE_FUNC; break; }
        // ---- IsStepSuccessful ----
        it.relative_decrease = (x_cost - cand_cost) / model_cost_change;
        if (it.relative_decrease > kMinRelativeDecrease) {
            // ---- HandleSuccessfulStep ----
            x = cand;
            x_norm = norm2(x);
            if (!eval_grad_jac(false)) { sum.termination = T_FAILURE_EVAL; break; }
            step_successful = true;
            it.flags |= 2;
            const double q = it.relative_decrease;
            radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * q - 1.0, 3));
            radius = std::min(kMaxRadius, radius);
            decrease_factor = 2.0;
            reuse_diagonal = false;
        } else {
            // ---- HandleUnsuccessfulStep ----
            radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
            it.cost = cand_cost;   // gradient_max_norm stays 0 in the pushed summary, as in Ceres
        }
    }
    sum.final_cost = x_cost;
    sum.usable = (sum.termination >= T_CONVERGENCE_PARAM && sum.termination <= T_NO_CONVERGENCE_RADIUS);
    std::memcpy(x_inout, x.data(), sizeof(double) * n);
    return sum;
}

}  // namespace orc
