// CPU ORACLE (test infrastructure only) — vertex-level ICP residuals of ark::AvatarOptimizer with their analytic Jacobians
// (SURVEY 8(f) rank 4).  PARITY UNPINNED like the rest of the oracle (no reference tests / golden vectors for this path; the
// reference needs Eigen + Ceres + nanoflann); pinned by finite differences through the reference's own Plus() (tests/).
//
// Literal restatement, table for table, of
//   external/avatar/AvatarOptimizer.cpp:161-456  AvatarEvaluationCommonData  (ancestor lists, CalcShape, R(j,i) / t(j,i) tables,
//                                                localJacobian, accumulated shape deltas H)
//   external/avatar/AvatarOptimizer.cpp:459-600  AvatarCostFunctorCache::updateData / getICPJacobians
//   external/avatar/AvatarOptimizer.cpp:602-641  AvatarICPCostFunctor::Evaluate (residual = model point - data point)
//   external/avatar/AvatarOptimizer.cpp:118-153  FakeQuaternionParameterization::Plus
// The Jacobian is returned dense in the tangent layout Ceres would assemble after the (fake) local parameterisation:
//   column 0..2 root position p | 3 + 3 j + c rotation tangent of joint j | 75 + s shape key s      (85 columns)
#include <algorithm>

#include "oracle.hpp"

namespace orc {

namespace {

struct Anc {   // AvatarEvaluationCommonData::Ancestor (:386-424)
    int jid;
    std::vector<int> assign;
    std::vector<double> weight;
};

// Eigen::Quaternion::toRotationMatrix, coefficients (x, y, z, w); row-major 3x3 out
void quat_to_matrix(const double* q, double* R) {
    const double tx = 2 * q[0], ty = 2 * q[1], tz = 2 * q[2];
    const double twx = tx * q[3], twy = ty * q[3], twz = tz * q[3];
    const double txx = tx * q[0], txy = ty * q[0], txz = tz * q[0];
    const double tyy = ty * q[1], tyz = tz * q[1], tzz = tz * q[2];
    R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
    R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
}
inline void mm3(const double* A, const double* B, double* C) {   // row-major
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
inline void mv3(const double* A, const double* x, double* y) {
    for (int i = 0; i < 3; ++i) y[i] = A[3 * i] * x[0] + A[3 * i + 1] * x[1] + A[3 * i + 2] * x[2];
}

}  // namespace

// One evaluation of all ICP residual blocks at (w, p, q).  corr: ncorr pairs (model point, data point).
void icp_eval(const Model& M, const double* w, const double* p, const double* q, int ncorr, const int* mpt, const int* dpt,
              const double* data, int shape_enabled, double* resid_out, double* jac_out, double* cost_out, double* grad_out) {
    const int J = M.J, S = M.S, V = M.V, NSP = J + 1, NC = 3 + 3 * J + S;
    // assignedJoints (AvatarModel.cpp:74-94): weights > 1e-12, heaviest first
    std::vector<std::vector<std::pair<double, int>>> assigned(V);
    for (int v = 0; v < V; ++v) {
        for (int e = M.wptr[v]; e < M.wptr[v + 1]; ++e)
            if (M.wval[e] > 1e-12) assigned[v].push_back({M.wval[e], M.wjoint[e]});
        std::sort(assigned[v].begin(), assigned[v].end(), std::greater<std::pair<double, int>>());
    }
    // CalcShape (:258-293)
    std::vector<double> shaped(3 * (size_t)V), jpos(3 * (size_t)J), jvec(3 * (size_t)J);
    for (size_t i = 0; i < 3 * (size_t)V; ++i) {
        double acc = 0.0;
        for (int s = 0; s < S; ++s) acc += M.key[i * S + s] * w[s];
        shaped[i] = acc + M.base[i];
    }
    for (int i = 0; i < 3 * J; ++i) {
        double acc = 0.0;
        for (int s = 0; s < S; ++s) acc += M.JSR[(size_t)i * S + s] * w[s];
        jpos[i] = M.J0[i] + acc;
    }
    const double off[3] = {jpos[0], jpos[1], jpos[2]};
    for (int v = 0; v < V; ++v) for (int c = 0; c < 3; ++c) shaped[3 * (size_t)v + c] -= off[c];
    for (int j = 0; j < J; ++j) for (int c = 0; c < 3; ++c) jpos[3 * j + c] -= off[c];
    jvec = jpos;
    for (int i = J - 1; i >= 1; --i) for (int c = 0; c < 3; ++c) jvec[3 * i + c] -= jvec[3 * M.parent[i] + c];
    // PrepareForEvaluation (:295-362)
    for (int c = 0; c < 3; ++c) jvec[c] = p[c];
    std::vector<double> LJ(12 * (size_t)J);   // localJacobian, 4x3 row-major
    for (int i = 0; i < J; ++i) {
        const double* x = q + 4 * i;
        const double l[12] = {x[3], x[2], -x[1], -x[2], x[3], x[0], x[1], -x[0], x[3], -x[0], -x[1], -x[2]};
        std::copy(l, l + 12, &LJ[12 * (size_t)i]);
    }
    std::vector<double> Rt(9 * (size_t)NSP * NSP, 0.0), tt(3 * (size_t)NSP * NSP, 0.0);
    auto R = [&](int ja, int j) { return &Rt[9 * ((size_t)NSP * (ja + 1) + j + 1)]; };
    auto t = [&](int ja, int j) { return &tt[3 * ((size_t)NSP * (ja + 1) + j + 1)]; };
    { double* I = R(-1, -1); I[0] = I[4] = I[8] = 1.0; }
    for (int i = 0; i < J; ++i) {
        double* I = R(i, i); I[0] = I[4] = I[8] = 1.0;
        double rot[9];
        quat_to_matrix(q + 4 * i, rot);
        const int pa = M.parent[i];
        for (int j = pa;; j = M.parent[j]) {
            mm3(R(j, pa), rot, R(j, i));
            double tmp[3];
            mv3(R(j, pa), &jvec[3 * i], tmp);
            for (int c = 0; c < 3; ++c) t(j, i)[c] = tmp[c] + t(j, pa)[c];
            if (j == -1) break;
        }
    }
    // shape deltas (:193-252, :347-353): S_j = jointShapeReg rows, Sp_j = S_j - S_parent, H_j = R(-1, parent) Sp_j + H_parent
    std::vector<double> Hs(3 * (size_t)S * J, 0.0);
    if (shape_enabled)
        for (int j = 1; j < J; ++j) {
            const int pa = M.parent[j];
            const double* Rg = R(-1, pa);
            for (int s = 0; s < S; ++s) {
                double sp[3], o[3];
                for (int c = 0; c < 3; ++c) sp[c] = M.JSR[(size_t)(3 * j + c) * S + s] - M.JSR[(size_t)(3 * pa + c) * S + s];
                mv3(Rg, sp, o);
                for (int c = 0; c < 3; ++c) Hs[((size_t)j * 3 + c) * S + s] = o[c] + Hs[((size_t)pa * 3 + c) * S + s];
            }
        }
    double cost = 0.0;
    std::vector<double> grad(NC, 0.0), Jrow(3 * (size_t)NC);
    for (int e = 0; e < ncorr; ++e) {
        const int pt = mpt[e];
        const double* x0 = &shaped[3 * (size_t)pt];
        // ancestors: deduplicated, sorted by joint id, each with the assigned joints below it (:181-204)
        std::vector<Anc> anc;
        for (auto& wj : assigned[pt]) {
            for (int j = wj.second; j != -1; j = M.parent[j]) {
                auto it = std::find_if(anc.begin(), anc.end(), [j](const Anc& a) { return a.jid == j; });
                if (it == anc.end()) { anc.push_back({j, {}, {}}); it = anc.end() - 1; }
                it->assign.push_back(wj.second);
                it->weight.push_back(wj.first);
            }
        }
        std::sort(anc.begin(), anc.end(), [](const Anc& a, const Anc& b) { return a.jid < b.jid; });
        // updateData (:503-577)
        double res[3] = {0, 0, 0};
        for (auto& wj : assigned[pt]) {
            const int k = wj.second;
            double d[3] = {x0[0] - jpos[3 * k], x0[1] - jpos[3 * k + 1], x0[2] - jpos[3 * k + 2]}, o[3];
            mv3(R(-1, k), d, o);
            for (int c = 0; c < 3; ++c) res[c] += wj.first * (o[c] + t(-1, k)[c]);
        }
        std::fill(Jrow.begin(), Jrow.end(), 0.0);
        for (int c = 0; c < 3; ++c) Jrow[(size_t)c * NC + c] = 1.0;   // root position block = identity (:476-480)
        for (auto& a : anc) {
            const int j = a.jid;
            double v[3] = {0, 0, 0};
            for (size_t s = 0; s < a.assign.size(); ++s) {
                const int k = a.assign[s];
                double d[3] = {x0[0] - jpos[3 * k], x0[1] - jpos[3 * k + 1], x0[2] - jpos[3 * k + 2]}, o[3];
                mv3(R(j, k), d, o);
                for (int c = 0; c < 3; ++c) v[c] += a.weight[s] * (o[c] + t(j, k)[c]);
            }
            const double* qq = q + 4 * j;
            const double u[3] = {2 * qq[0], 2 * qq[1], 2 * qq[2]}, ww = 2 * qq[3];
            const double dR[12] = {   // quaternion-vector rotation pseudo-Jacobian, 3x4 (:533-546)
                u[1] * v[1] + v[2] * u[2], ww * v[2] + u[0] * v[1] - 2 * u[1] * v[0], -ww * v[1] - 2 * v[0] * u[2] + u[0] * v[2], u[1] * v[2] - v[1] * u[2],
                -ww * v[2] - 2 * u[0] * v[1] + v[0] * u[1], v[2] * u[2] + u[0] * v[0], ww * v[0] + u[1] * v[2] - 2 * v[1] * u[2], v[0] * u[2] - u[0] * v[2],
                ww * v[1] + v[0] * u[2] - 2 * u[0] * v[2], -ww * v[0] - 2 * u[1] * v[2] + v[1] * u[2], u[0] * v[0] + v[1] * u[1], u[0] * v[1] - v[0] * u[1]};
            // icpJacobian = R(-1, parent[j]) * dRot * localJacobian[j]  (:547-553)
            double dl[9];
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) {
                    double acc = 0.0;
                    for (int k = 0; k < 4; ++k) acc += dR[4 * r + k] * LJ[12 * (size_t)j + 3 * k + c];
                    dl[3 * r + c] = acc;
                }
            double Jb[9];
            mm3(R(-1, M.parent[j]), dl, Jb);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) Jrow[(size_t)r * NC + 3 + 3 * j + c] = Jb[3 * r + c];
        }
        if (shape_enabled)   // (:556-567)
            for (auto& wj : assigned[pt]) {
                const int j = wj.second;
                for (int s = 0; s < S; ++s) {
                    double d[3], o[3];
                    for (int c = 0; c < 3; ++c) d[c] = M.key[(size_t)(3 * pt + c) * S + s] - M.JSR[(size_t)(3 * j + c) * S + s];
                    mv3(R(-1, j), d, o);
                    for (int c = 0; c < 3; ++c) Jrow[(size_t)c * NC + 3 + 3 * J + s] += (o[c] + Hs[((size_t)j * 3 + c) * S + s]) * wj.first;
                }
            }
        // AvatarICPCostFunctor::Evaluate (:630-636)
        for (int c = 0; c < 3; ++c) res[c] -= data[3 * (size_t)dpt[e] + c];
        for (int c = 0; c < 3; ++c) {
            cost += 0.5 * res[c] * res[c];
            for (int k = 0; k < NC; ++k) grad[k] += Jrow[(size_t)c * NC + k] * res[c];
        }
        if (resid_out) for (int c = 0; c < 3; ++c) resid_out[3 * (size_t)e + c] = res[c];
        if (jac_out) std::copy(Jrow.begin(), Jrow.end(), jac_out + (size_t)e * 3 * NC);
    }
    if (cost_out) *cost_out = cost;
    if (grad_out) std::copy(grad.begin(), grad.end(), grad_out);
}

// FakeQuaternionParameterization::Plus (:122-141): x_plus_delta = [sin|d|/|d| d, cos|d|] * x, coefficients (x, y, z, w)
void quat_plus(const double* x, const double* d, double* out) {
    const double n = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (n > 0.0) {
        const double s = std::sin(n) / n;
        const double a[4] = {s * d[0], s * d[1], s * d[2], std::cos(n)};
        // Eigen quaternion product a * x
        out[3] = a[3] * x[3] - a[0] * x[0] - a[1] * x[1] - a[2] * x[2];
        out[0] = a[3] * x[0] + a[0] * x[3] + a[1] * x[2] - a[2] * x[1];
        out[1] = a[3] * x[1] + a[1] * x[3] + a[2] * x[0] - a[0] * x[2];
        out[2] = a[3] * x[2] + a[2] * x[3] + a[0] * x[1] - a[1] * x[0];
    } else {
        for (int k = 0; k < 4; ++k) out[k] = x[k];
    }
}


// findNN, invert = true (AvatarOptimizer.cpp:841-907): every data point is matched to the nearest VISIBLE model point that carries
// its part label; -1 when the part has none.  The reference asks a nanoflann kd-tree per part for the exact nearest neighbour
// (metric_L2_Simple, eps = 0); exhaustive search returns the same point (ties, measure zero on real data, go to the first model
// point in index order here).  correspondences[model point] then lists its data points in data order.
void icp_nearest(int V, const double* cloud, const int* model_part, const uint8_t* visible, int nd, const double* data,
                 const int* data_part, int* out) {
    for (int i = 0; i < nd; ++i) {
        const double* d = data + 3 * (size_t)i;
        double best = std::numeric_limits<double>::infinity();
        int arg = -1;
        for (int v = 0; v < V; ++v) {
            if (model_part[v] != data_part[i] || (visible && !visible[v])) continue;
            const double dx = d[0] - cloud[3 * (size_t)v], dy = d[1] - cloud[3 * (size_t)v + 1], dz = d[2] - cloud[3 * (size_t)v + 2];
            const double r = (dx * dx + dy * dy) + dz * dz;
            if (r < best) { best = r; arg = v; }
        }
        out[i] = arg;
    }
}


// Prior blocks of AvatarOptimizer::optimize: AvatarPosePriorCostFunctor (:644-700) and AvatarShapePriorCostFunctor (:703-726) with
// the weights scaled by the number of ICP residual blocks (:1412-1414).  Adds their cost and their gradient (tangent layout of
// icp_eval) to cost / grad.  The pose block's Jacobian is the reference's: L.middleRows<3>(3 i)^T * 0.707106781186548 * beta for
// joint i + 1, zero for the constant last residual.
void icp_priors(const GMM* gmm, double beta_pose, double beta_shape, double total_residuals, int S, const double* w, const double* q,
                double* cost, double* grad) {
    const double sbp = beta_pose * std::sqrt(total_residuals) / 15.0, sbs = beta_shape * std::sqrt(total_residuals) / 15.0;
    if (beta_pose > 0.0 && gmm && gmm->nComps > 0) {
        const int nd = gmm->nDims, nj = nd / 3;
        std::vector<double> x(nd), r(nd + 1);
        for (int i = 0; i < nj; ++i) {   // Eigen::AngleAxisd(q): axis * angle
            const double* qq = q + 4 * (i + 1);
            double n = std::sqrt(qq[0] * qq[0] + qq[1] * qq[1] + qq[2] * qq[2]);
            if (n != 0.0) {
                const double angle = 2.0 * std::atan2(n, std::fabs(qq[3]));
                if (qq[3] < 0.0) n = -n;
                for (int c = 0; c < 3; ++c) x[3 * i + c] = qq[c] / n * angle;
            } else {
                for (int c = 0; c < 3; ++c) x[3 * i + c] = 0.0;
            }
        }
        const int comp = gmm->residual(x.data(), r.data());
        for (int k = 0; k <= nd; ++k) { r[k] *= sbp; *cost += 0.5 * r[k] * r[k]; }
        const double* L = gmm->prec_cho[comp].data();
        for (int i = 0; i < nj; ++i)
            for (int a = 0; a < 3; ++a) {
                double acc = 0.0;
                for (int k = 0; k < nd; ++k) acc += L[(size_t)(3 * i + a) * nd + k] * 0.707106781186548 * sbp * r[k];
                grad[3 + 3 * (i + 1) + a] += acc;
            }
    }
    if (beta_shape > 0.0)
        for (int s = 0; s < S; ++s) {
            const double rs = w[s] * sbs;
            *cost += 0.5 * rs * rs;
            grad[3 + 3 * 24 + s] += sbs * rs;
        }
}

// pointVisible (AvatarOptimizer.cpp:1337-1352)
void icp_visible(int V, int F, const uint32_t* faces, const double* cloud, uint8_t* visible) {
    std::fill(visible, visible + V, (uint8_t)0);
    for (int f = 0; f < F; ++f) {
        const uint32_t i1 = faces[3 * f], i2 = faces[3 * f + 1], i3 = faces[3 * f + 2];
        const double* p1 = cloud + 3 * (size_t)i1; const double* p2 = cloud + 3 * (size_t)i2; const double* p3 = cloud + 3 * (size_t)i3;
        const double ax = p2[0] - p1[0], ay = p2[1] - p1[1], bx = p1[0] - p3[0], by = p1[1] - p3[1];
        if (ax * by - ay * bx > 1e-4) visible[i1] = visible[i2] = visible[i3] = 1;
    }
}

}  // namespace orc
