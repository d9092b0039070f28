// CPU ORACLE (test infrastructure only) — the reference's three problem builders, restated.
//   SingleProblem  : OptimizePoseReprojection (Sim3BA.h:515-683) and OptimizePoseShapeReprojection (Sim3BA.h:348-511)
//   MultiProblem   : OptimizeMultiFrame (MultiFrameBA.h:33-177)
//   priors         : PosePriorAAAnalytic (Sim3BA.h:245-320), ShapePriorL2Analytic (Sim3BA.h:323-346)
//   temporal       : Vec3DiffCost (MultiFrameBA.h:20-28)
//   robust loss    : ceres::HuberLoss(3.0) + Corrector (Ceres 1.14)
#include "oracle.hpp"
#include "resid.hpp"

namespace orc {

namespace {

using JetF = Jet<NFULL>;

void make_skeleton(const Model& M, bool with_shape, Skeleton& sk) {
    sk.J = M.J; sk.S = M.S;
    for (int j = 0; j < M.J; ++j) sk.parent[j] = M.parent[j];
    rest_offsets(M, sk.offset);
    sk.JSR = with_shape ? M.JSR.data() : nullptr;
}

// one reprojection block: robustified residual rows + (optional) Jacobian rows in the 86-layout
struct ReprojRows { double r[2]; double J[2 * NFULL]; double cost; };

void eval_reproj(const Skeleton& sk, const KP& kp, const double cam[4], const double R0[9], const double* full, double huber,
                 bool want_jac, bool use_jets, ReprojRows& out) {
    if (want_jac && use_jets) {
        // ceres::DynamicAutoDiffCostFunction: Jets seeded on every parameter
        static thread_local std::vector<JetF> xj(NFULL);
        for (int k = 0; k < NFULL; ++k) xj[k] = JetF(full[k], k);
        JetF res[2];
        reproj_residual<JetF>(sk, kp.jid, kp.u, kp.v, cam, R0, xj.data(), res);
        for (int i = 0; i < 2; ++i) { out.r[i] = res[i].a; for (int k = 0; k < NFULL; ++k) out.J[i * NFULL + k] = res[i].v[k]; }
    } else if (want_jac) {
        reproj_analytic(sk, kp.jid, kp.u, kp.v, cam, R0, full, out.r, out.J);
    } else {
        reproj_residual<double>(sk, kp.jid, kp.u, kp.v, cam, R0, full, out.r);
    }
    // HuberLoss(a) + Corrector (rho'' <= 0 branch): r <- sqrt(rho') r, J <- sqrt(rho') J, cost = rho/2
    const double s = out.r[0] * out.r[0] + out.r[1] * out.r[1];
    const double b = huber * huber;
    double rho0, rho1;
    if (s > b) { const double rr = std::sqrt(s); rho0 = 2.0 * huber * rr - b; rho1 = std::max(std::numeric_limits<double>::min(), huber / rr); }
    else { rho0 = s; rho1 = 1.0; }
    out.cost = 0.5 * rho0;
    const double sc = std::sqrt(rho1);
    out.r[0] *= sc; out.r[1] *= sc;
    if (want_jac) for (int k = 0; k < 2 * NFULL; ++k) out.J[k] *= sc;
}

struct SingleProblem : LMProblem {
    const Model& M; const GMM* gmm; FitOptions o; Skeleton sk;
    std::vector<KP> kps; double R0[9]; double cam[4];
    double full[NFULL];
    std::vector<int> free_idx, col_of;   // free -> full, full -> free (-1 fixed)
    bool use_gmm = false, has_w = false;
    int comp = -1;
    SingleProblem(const Model& M_, const GMM* g, const FitOptions& o_) : M(M_), gmm(g), o(o_) {}
    int aux_state() const override { return comp; }

    bool evaluate(const double* x, double* cost, std::vector<double>* r, SparseRows* J, std::vector<double>* g) override {
        for (int k = 0; k < n; ++k) full[free_idx[k]] = x[k];
        const bool want_jac = (J != nullptr) || (g != nullptr);
        SparseRows Jl;
        SparseRows* Jp = J ? J : &Jl;
        std::vector<double> rl;
        std::vector<double>* rp = r ? r : &rl;
        if (want_jac) Jp->clear(n);
        rp->clear();
        double c = 0.0;
        ReprojRows rr;
        for (const KP& kp : kps) {
            eval_reproj(sk, kp, cam, R0, full, o.huber, want_jac, o.use_jets != 0, rr);
            c += rr.cost;
            for (int i = 0; i < 2; ++i) {
                rp->push_back(rr.r[i]);
                if (want_jac) {
                    for (int k = 0; k < NFULL; ++k) { const double v = rr.J[i * NFULL + k]; if (v != 0.0 && col_of[k] >= 0) Jp->push(col_of[k], v); }
                    Jp->end_row();
                }
            }
        }
        if (o.beta_pose > 0.0) {
            const double* xp = full + IDX_JAA;
            const int nd = 3 * (M.J - 1);
            if (use_gmm) {
                std::vector<double> res(nd + 1);
                comp = gmm->residual(xp, res.data());
                const double* L = gmm->prec_cho[comp].data();
                for (int k = 0; k <= nd; ++k) {
                    const double v = res[k] * o.beta_pose;
                    rp->push_back(v);
                    c += 0.5 * v * v;
                    if (want_jac) {
                        // J = beta * L^T (top nd rows), last row zero; NOTE: no sqrt(.5) (reference quirk, Sim3BA.h:298-299)
                        if (k < nd)
                            for (int d = k; d < nd; ++d) { const int col = col_of[IDX_JAA + d]; if (col >= 0) { const double lv = L[(size_t)d * nd + k] * o.beta_pose; if (lv != 0.0) Jp->push(col, lv); } }
                        Jp->end_row();
                    }
                }
            } else {
                for (int k = 0; k < nd; ++k) {
                    const double v = xp[k] * o.beta_pose;
                    rp->push_back(v);
                    c += 0.5 * v * v;
                    if (want_jac) { const int col = col_of[IDX_JAA + k]; if (col >= 0) Jp->push(col, o.beta_pose); Jp->end_row(); }
                }
            }
        }
        if (has_w) {
            for (int s = 0; s < M.S; ++s) {
                const double v = full[IDX_W + s] * o.beta_shape;
                rp->push_back(v);
                c += 0.5 * v * v;
                if (want_jac) { Jp->push(col_of[IDX_W + s], o.beta_shape); Jp->end_row(); }
            }
        }
        *cost = c;
        if (g) {
            g->assign(n, 0.0);
            for (int i = 0; i < Jp->rows(); ++i)
                for (int e = Jp->rowptr[i]; e < Jp->rowptr[i + 1]; ++e) (*g)[Jp->col[e]] += Jp->val[e] * (*rp)[i];
        }
        return std::isfinite(c);
    }
};

struct MultiProblem : LMProblem {
    const Model& M; FitOptions o; Skeleton sk; int F;
    const KP* kps; const int* nkp; const double* R0; double cam[4];
    bool shape_prior;
    MultiProblem(const Model& M_, const FitOptions& o_, int F_) : M(M_), o(o_), F(F_) {}

    bool evaluate(const double* x, double* cost, std::vector<double>* r, SparseRows* J, std::vector<double>* g) override {
        const bool want_jac = (J != nullptr) || (g != nullptr);
        SparseRows Jl; SparseRows* Jp = J ? J : &Jl;
        std::vector<double> rl; std::vector<double>* rp = r ? r : &rl;
        if (want_jac) Jp->clear(n);
        rp->clear();
        double c = 0.0;
        const double* w = x + (size_t)F * NPF;
        const int wcol = F * NPF;
        double full[NFULL];
        ReprojRows rr;
        for (int f = 0; f < F; ++f) {
            const double* xf = x + (size_t)f * NPF;
            for (int k = 0; k < NPF; ++k) full[k] = xf[k];
            for (int s = 0; s < NS; ++s) full[IDX_W + s] = w[s];
            for (int q = 0; q < nkp[f]; ++q) {
                eval_reproj(sk, kps[(size_t)f * MAXKP + q], cam, R0 + 9 * (size_t)f, full, o.huber, want_jac, o.use_jets != 0, rr);
                c += rr.cost;
                for (int i = 0; i < 2; ++i) {
                    rp->push_back(rr.r[i]);
                    if (want_jac) {
                        for (int k = 0; k < NFULL; ++k) { const double v = rr.J[i * NFULL + k]; if (v != 0.0) Jp->push(k < NPF ? f * NPF + k : wcol + (k - NPF), v); }
                        Jp->end_row();
                    }
                }
            }
            if (o.beta_pose > 0.0 && M.J > 1) {   // L2 prior only: gmm == nullptr always (MultiFrameBA.h:109)
                for (int k = 0; k < 3 * (M.J - 1); ++k) {
                    const double v = xf[IDX_JAA + k] * o.beta_pose;
                    rp->push_back(v); c += 0.5 * v * v;
                    if (want_jac) { Jp->push(f * NPF + IDX_JAA + k, o.beta_pose); Jp->end_row(); }
                }
            }
        }
        if (shape_prior) {
            for (int s = 0; s < NS; ++s) {
                const double v = w[s] * o.beta_shape;
                rp->push_back(v); c += 0.5 * v * v;
                if (want_jac) { Jp->push(wcol + s, o.beta_shape); Jp->end_row(); }
            }
        }
        if (o.lambda_temp > 0.0 && F > 1) {
            auto diff3 = [&](int f, int off) {
                for (int i = 0; i < 3; ++i) {
                    const double v = (x[(size_t)f * NPF + off + i] - x[(size_t)(f + 1) * NPF + off + i]) * o.lambda_temp;
                    rp->push_back(v); c += 0.5 * v * v;
                    if (want_jac) { Jp->push(f * NPF + off + i, o.lambda_temp); Jp->push((f + 1) * NPF + off + i, -o.lambda_temp); Jp->end_row(); }
                }
            };
            for (int f = 0; f < F - 1; ++f) {
                diff3(f, IDX_T);
                diff3(f, IDX_RAA);
                for (int j = 1; j < M.J; ++j) diff3(f, idx_jaa(j));
            }
        }
        *cost = c;
        if (g) {
            g->assign(n, 0.0);
            for (int i = 0; i < Jp->rows(); ++i)
                for (int e = Jp->rowptr[i]; e < Jp->rowptr[i + 1]; ++e) (*g)[Jp->col[e]] += Jp->val[e] * (*rp)[i];
        }
        return std::isfinite(c);
    }
};

}  // namespace

bool fit_single(const Model& M, const GMM* gmm, const FitOptions& o, const KP* kps, int nkp, const double* R0, double* sim3,
                double* w, double* jointAA, std::vector<IterRecord>* trace, LMSummary* summary) {
    SingleProblem P(M, gmm, o);
    // mode 1 passes jointShapeReg iff betaShape > 0 && useJointShapeRegressor (Sim3BA.h:417); mode 0 never does
    make_skeleton(M, o.mode == 1 && o.beta_shape > 0.0, P.sk);
    P.kps.assign(kps, kps + nkp);
    for (int k = 0; k < 9; ++k) P.R0[k] = R0[k];
    P.cam[0] = o.fx; P.cam[1] = o.fy; P.cam[2] = o.cx; P.cam[3] = o.cy;
    P.use_gmm = gmm && gmm->nComps > 0;                    // silent L2 degrade otherwise (Sim3BA.h:266)
    P.has_w = o.beta_shape > 0.0 && M.S > 0;               // shape prior block (Sim3BA.h:466-470 / 630-638)
    for (int k = 0; k < 7; ++k) P.full[k] = sim3[k];
    for (int k = 0; k < 3 * (M.J - 1); ++k) P.full[IDX_JAA + k] = 0.0;   // jointAA start at zero (Sim3BA.h:401-404)
    for (int s = 0; s < M.S; ++s) P.full[IDX_W + s] = w[s];
    P.col_of.assign(NFULL, -1);
    auto add = [&](int fi) { P.col_of[fi] = (int)P.free_idx.size(); P.free_idx.push_back(fi); };
    for (int k = 0; k < 7; ++k) add(k);
    for (int j = 1; j < M.J; ++j) {
        // pose-only mode holds joints 10, 11, 22, 23 constant (Sim3BA.h:608-611)
        if (o.mode == 0 && (j == 10 || j == 11 || j == 22 || j == 23)) continue;
        for (int c = 0; c < 3; ++c) add(idx_jaa(j) + c);
    }
    if (P.has_w) for (int s = 0; s < M.S; ++s) add(IDX_W + s);
    P.n = (int)P.free_idx.size();
    P.constrained = true;                                   // scale in [0.3, 3.0] (Sim3BA.h:450-451 / 613-614)
    P.lower.assign(P.n, -std::numeric_limits<double>::infinity());
    P.upper.assign(P.n, std::numeric_limits<double>::infinity());
    P.lower[0] = o.scale_lo; P.upper[0] = o.scale_hi;
    std::vector<double> x(P.n);
    for (int k = 0; k < P.n; ++k) x[k] = P.full[P.free_idx[k]];
    LMOptions lo;
    lo.max_iters = o.max_iters;
    lo.linear_solver = (o.linear_solver == LS_BLOCK_TRIDIAG) ? LS_DENSE_CHOLESKY : o.linear_solver;
    LMSummary s = lm_minimize(P, lo, x.data(), trace);
    for (int k = 0; k < P.n; ++k) P.full[P.free_idx[k]] = x[k];
    for (int k = 0; k < 7; ++k) sim3[k] = P.full[k];
    for (int c = 0; c < 3; ++c) jointAA[c] = 0.0;
    for (int k = 0; k < 3 * (M.J - 1); ++k) jointAA[3 + k] = P.full[IDX_JAA + k];
    if (P.has_w) for (int sidx = 0; sidx < M.S; ++sidx) w[sidx] = P.full[IDX_W + sidx];
    if (summary) *summary = s;
    return s.usable;
}

bool fit_window(const Model& M, const FitOptions& o, int F, const KP* kps, const int* nkp, const double* R0, double* poses,
                double* w, std::vector<IterRecord>* trace, LMSummary* summary) {
    MultiProblem P(M, o, F);
    make_skeleton(M, o.beta_shape > 0.0, P.sk);              // MultiFrameBA.h:87
    P.kps = kps; P.nkp = nkp; P.R0 = R0;
    P.cam[0] = o.fx; P.cam[1] = o.fy; P.cam[2] = o.cx; P.cam[3] = o.cy;
    P.shape_prior = M.S > 0 && o.beta_shape > 0.0;           // MultiFrameBA.h:115-118
    P.n = F * NPF + NS;                                      // w block always present (MultiFrameBA.h:67-68)
    P.constrained = false;                                   // no bounds in multi
    std::vector<double> x(P.n);
    std::memcpy(x.data(), poses, sizeof(double) * (size_t)F * NPF);
    for (int s = 0; s < NS; ++s) x[(size_t)F * NPF + s] = w[s];
    LMOptions lo;
    lo.max_iters = o.max_iters;
    lo.linear_solver = o.linear_solver;
    lo.blocks.F = F; lo.blocks.nb = NPF; lo.blocks.na = NS;
    LMSummary s = lm_minimize(P, lo, x.data(), trace);
    std::memcpy(poses, x.data(), sizeof(double) * (size_t)F * NPF);
    for (int k = 0; k < NS; ++k) w[k] = x[(size_t)F * NPF + k];
    if (summary) *summary = s;
    return s.usable;
}

}  // namespace orc
