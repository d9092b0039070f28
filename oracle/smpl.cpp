// CPU ORACLE (test infrastructure only) — SMPL model derivation and forward pass.
// Follows external/avatar/AvatarModel.cpp:23-127 and external/avatar/Avatar.cpp:22-75.
#include "oracle.hpp"

namespace orc {

// AvatarModel::AvatarModel, npz branch (AvatarModel.cpp:23-127).
void Model::derive(int V_, int J_, int S_, const double* v_template, const double* shapedirs,
                   const double* J_regressor, const double* weights, const uint32_t* kintree_row0,
                   const double* posedirs_or_null, int P_) {
    V = V_; J = J_; S = S_; P = posedirs_or_null ? P_ : 0;
    // kintree row 0, uint32 -> int (0xFFFFFFFF -> -1), AvatarModel.cpp:35-41
    parent.resize(J);
    for (int j = 0; j < J; ++j) parent[j] = (int)kintree_row0[j];
    // baseCloud = v_template flattened (:44-48)
    base.assign(v_template, v_template + 3 * (size_t)V);
    // keyClouds (3V x S) from shapedirs (V,3,S) (:93-102)
    key.assign(shapedirs, shapedirs + 3 * (size_t)V * S);
    // LBS weights, sparse J x V; sparseView drops exact zeros only (:66-71)
    wptr.assign(1, 0); wjoint.clear(); wval.clear();
    for (int v = 0; v < V; ++v) {
        for (int j = 0; j < J; ++j) {
            const double x = weights[(size_t)v * J + j];
            if (x != 0.0) { wjoint.push_back(j); wval.push_back(x); }
        }
        wptr.push_back((int)wjoint.size());
    }
    // initialJointPos = template(3 x V) * jointRegressor(V x J) (:113-117)
    J0.assign(3 * (size_t)J, 0.0);
    for (int j = 0; j < J; ++j)
        for (int v = 0; v < V; ++v) {
            const double g = J_regressor[(size_t)j * V + v];
            if (g == 0.0) continue;
            for (int c = 0; c < 3; ++c) J0[3 * j + c] += base[3 * (size_t)v + c] * g;
        }
    // jointShapeReg column s = (key cloud s as 3 x V) * jointRegressor (:121-127)
    JSR.assign(3 * (size_t)J * S, 0.0);
    for (int j = 0; j < J; ++j)
        for (int v = 0; v < V; ++v) {
            const double g = J_regressor[(size_t)j * V + v];
            if (g == 0.0) continue;
            for (int c = 0; c < 3; ++c)
                for (int s = 0; s < S; ++s) JSR[(size_t)(3 * j + c) * S + s] += key[(size_t)(3 * v + c) * S + s] * g;
        }
    posedirs.clear();
    if (P > 0) posedirs.assign(posedirs_or_null, posedirs_or_null + 3 * (size_t)V * P);
}

// Avatar::update (Avatar.cpp:22-75).  r = J col-major 3x3 rotation matrices.
void Model::forward(const double* w, const double* p, const double* r, double* cloud, double* jointPos,
                    double* jointTrans) const {
    std::vector<double> shaped(3 * (size_t)V);
    // shapedCloudVec = keyClouds * w + baseCloud (:26)
    for (size_t i = 0; i < 3 * (size_t)V; ++i) {
        double acc = 0.0;
        for (int s = 0; s < S; ++s) acc += key[i * S + s] * w[s];
        shaped[i] = acc + base[i];
    }
    if (P > 0) {
        // standard SMPL pose blendshapes (absent from the reference: AvatarModel.cpp:104-109 commented out).
        // pose feature = vec_rowmajor(R_j - I), j = 1..J-1
        std::vector<double> pf(P);
        for (int j = 1; j < J; ++j)
            for (int a = 0; a < 3; ++a)
                for (int b = 0; b < 3; ++b)
                    pf[(j - 1) * 9 + a * 3 + b] = r[9 * j + b * 3 + a] - (a == b ? 1.0 : 0.0);
        for (size_t i = 0; i < 3 * (size_t)V; ++i) {
            double acc = 0.0;
            for (int k = 0; k < P; ++k) acc += posedirs[i * P + k] * pf[k];
            shaped[i] += acc;
        }
    }
    // jointPos = initialJointPos + jointShapeReg * w (:32-36)
    std::vector<double> jp(3 * (size_t)J), jt(12 * (size_t)J);
    for (int i = 0; i < 3 * J; ++i) {
        double acc = 0.0;
        for (int s = 0; s < S; ++s) acc += JSR[(size_t)i * S + s] * w[s];
        jp[i] = J0[i] + acc;
    }
    // root: [r0 | p] (:47-49)
    for (int k = 0; k < 9; ++k) jt[k] = r[k];
    for (int c = 0; c < 3; ++c) jt[9 + c] = p[c];
    // children: local = [r_i | jp_i - jp_parent]; global = parent o local (mulAffine, Util.h:191-199) (:50-57)
    for (int i = 1; i < J; ++i) {
        const int pa = parent[i];
        const double* A = &jt[12 * (size_t)pa];
        const double* Rl = &r[9 * (size_t)i];
        double tl[3] = {jp[3 * i] - jp[3 * pa], jp[3 * i + 1] - jp[3 * pa + 1], jp[3 * i + 2] - jp[3 * pa + 2]};
        double* B = &jt[12 * (size_t)i];
        for (int col = 0; col < 3; ++col)
            for (int row = 0; row < 3; ++row)
                B[3 * col + row] = A[row] * Rl[3 * col] + A[3 + row] * Rl[3 * col + 1] + A[6 + row] * Rl[3 * col + 2];
        for (int row = 0; row < 3; ++row) B[9 + row] = A[9 + row] + (A[row] * tl[0] + A[3 + row] * tl[1] + A[6 + row] * tl[2]);
    }
    // jointPos <- translation; translation <- t - R * Jinit (:59-64)
    for (int i = 0; i < J; ++i) {
        double* B = &jt[12 * (size_t)i];
        const double ji[3] = {jp[3 * i], jp[3 * i + 1], jp[3 * i + 2]};
        for (int c = 0; c < 3; ++c) jp[3 * i + c] = B[9 + c];
        for (int row = 0; row < 3; ++row) B[9 + row] -= B[row] * ji[0] + B[3 + row] * ji[1] + B[6 + row] * ji[2];
    }
    if (jointPos) std::memcpy(jointPos, jp.data(), sizeof(double) * 3 * J);
    if (jointTrans) std::memcpy(jointTrans, jt.data(), sizeof(double) * 12 * J);
    if (!cloud) return;
    // pointTrans = jointTrans * weights ; cloud_i = pointTrans_i * [shaped_i; 1] (:69-73)
    for (int v = 0; v < V; ++v) {
        double pt[12];
        for (int k = 0; k < 12; ++k) pt[k] = 0.0;
        for (int e = wptr[v]; e < wptr[v + 1]; ++e) {
            const double* B = &jt[12 * (size_t)wjoint[e]];
            const double wt = wval[e];
            for (int k = 0; k < 12; ++k) pt[k] += wt * B[k];
        }
        const double* sv = &shaped[3 * (size_t)v];
        for (int row = 0; row < 3; ++row)
            cloud[3 * (size_t)v + row] = pt[row] * sv[0] + pt[3 + row] * sv[1] + pt[6 + row] * sv[2] + pt[9 + row];
    }
}

// Rest-pose bone offsets (Sim3BA.h:533-555 / 367-392, MultiFrameBA.h:53-60): base avatar w=0, r=I, p=0,
// root translated to origin, offset[j] = J_j - J_parent.
void rest_offsets(const Model& M, double* offset) {
    std::vector<double> w(M.S, 0.0), r(9 * (size_t)M.J, 0.0), jp(3 * (size_t)M.J);
    for (int j = 0; j < M.J; ++j) r[9 * j] = r[9 * j + 4] = r[9 * j + 8] = 1.0;
    const double p[3] = {0, 0, 0};
    // forward() with P>0 would add posedirs*(I-I)=0, so offsets are identical either way
    M.forward(w.data(), p, r.data(), nullptr, jp.data(), nullptr);
    const double root[3] = {jp[0], jp[1], jp[2]};
    for (int j = 0; j < M.J; ++j)
        for (int c = 0; c < 3; ++c) jp[3 * j + c] -= root[c];
    for (int c = 0; c < 3; ++c) offset[c] = 0.0;
    for (int j = 1; j < M.J; ++j) {
        const int pa = M.parent[j];
        for (int c = 0; c < 3; ++c) offset[3 * j + c] = pa >= 0 ? jp[3 * j + c] - jp[3 * pa + c] : jp[3 * j + c];
    }
}

}  // namespace orc
