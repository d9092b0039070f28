// b3d_shim.hpp — the reference's own C++ call signatures for the hot path, implemented on top of the C-ABI (b3d.h).
//
// The reference has no FFI layer: its boundary is three inline functions and the public fields of ark::Avatar /
// ark::AvatarModel.  A maintainer of jonH34400/3DBodyAnimation switches the two drivers
// (src/main_single_frame.cpp, src/main_multi_frame.cpp) to the B200 path by including this header instead of
// Sim3BA.h / MultiFrameBA.h / Avatar.h and linking lib3dba_b200.so — see INTEGRATION.md.  Header-only, C++14, no Eigen /
// Ceres: matrices are plain column-major arrays (the memory layout of Eigen::Matrix3d and of the 3 x V `cloud`).
//
//   b3d::shim::AvatarModel                    <- ark::AvatarModel            external/avatar/include/Avatar.h:64-147
//   b3d::shim::Avatar (+ update())            <- ark::Avatar                 external/avatar/include/Avatar.h:155-220, Avatar.cpp:22-75
//   b3d::shim::GaussianMixture                <- ark::GaussianMixture        external/avatar/include/GaussianMixture.h:9-49
//   PixelKP / Sim3Params / FramePoseParams    <- include/Sim3BA.h:9-19, include/MultiFrameBA.h:9-14
//   OptimizePoseReprojection                  <- include/Sim3BA.h:515-683
//   OptimizePoseShapeReprojection             <- include/Sim3BA.h:348-511
//   OptimizeMultiFrame                        <- include/MultiFrameBA.h:33-177
//   mean_pixel_error                          <- include/Utils.h:102-115
//
// Error behaviour follows the reference: the Optimize* functions return {IsSolutionUsable(), report} and never throw on
// non-convergence; hard failures (bad model, CUDA error) throw std::runtime_error where the reference calls std::exit(1).
#ifndef B3D_SHIM_HPP_
#define B3D_SHIM_HPP_

#include <array>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "b3d.h"

namespace b3d {
namespace shim {

inline void check(int rc, const char* what) {
    if (rc != B3D_OK) throw std::runtime_error(std::string(what) + ": " + b3d_last_error());
}

struct PixelKP { int jid; double u, v; };                                   // include/Sim3BA.h:9

struct Sim3Params {                                                          // include/Sim3BA.h:11-19
    double data[7] = {1.0, 0, 0, 0, 0, 0, 0};                                // [s, aa(3), t(3)]
    double& s() { return data[0]; }
    double* aa() { return data + 1; }
    double* t() { return data + 4; }
};

struct FramePoseParams {                                                     // include/MultiFrameBA.h:9-14
    double scale = 1.0;
    double rootAA[3] = {0, 0, 0};
    double rootT[3] = {0, 0, 0};
    std::vector<std::array<double, 3>> jointAA;                              // size J, [0] unused
};

class GaussianMixture {                                                      // ark::GaussianMixture
public:
    int nComps = -1, nDims = 0;                                              // nComps = -1: prior file missing (GaussianMixture.cpp:15-19)
    b3d_gmm* handle = nullptr;
    GaussianMixture() = default;
    GaussianMixture(const GaussianMixture&) = delete;
    GaussianMixture& operator=(const GaussianMixture&) = delete;
    ~GaussianMixture() { b3d_gmm_destroy(handle); }
    void load(const std::string& path) {
        b3d_gmm_destroy(handle);
        handle = nullptr; nComps = -1; nDims = 0;
        const int rc = b3d_gmm_load(path.c_str(), &handle);
        if (rc == B3D_ENOFILE) { std::fprintf(stderr, "WARNING: pose prior %s not found\n", path.c_str()); return; }
        check(rc, "GaussianMixture::load");
        int32_t nc = 0, nd = 0;
        check(b3d_gmm_dims(handle, &nc, &nd), "b3d_gmm_dims");
        nComps = nc; nDims = nd;
    }
};

// ark::AvatarModel built from the raw arrays of SMPL's model.npz (reading the .npz container itself is the caller's job:
// the reference does it with its vendored cnpy, external/avatar/AvatarModel.cpp:23-34).
class AvatarModel {
public:
    b3d_model* handle = nullptr;
    int numPoints = 0, numJoints = B3D_NUM_JOINTS, numShapeKeys = B3D_NUM_SHAPE;
    bool useJointShapeRegressor = true;                                      // AvatarModel.cpp:111
    std::array<int32_t, B3D_NUM_JOINTS> parent{};
    GaussianMixture posePrior;                                               // AvatarModel.cpp:296
    explicit AvatarModel(const b3d_model_desc& desc, const std::string& pose_prior_path = "") {
        check(b3d_model_create(&desc, &handle), "AvatarModel");
        numPoints = b3d_model_num_verts(handle);
        check(b3d_model_get_parents(handle, parent.data()), "b3d_model_get_parents");
        if (!pose_prior_path.empty()) posePrior.load(pose_prior_path);
    }
    AvatarModel(const AvatarModel&) = delete;
    AvatarModel& operator=(const AvatarModel&) = delete;
    ~AvatarModel() { b3d_model_destroy(handle); }
};

// ark::Avatar: public w, p, r in; cloud, jointPos, jointTrans out of update().
class Avatar {
public:
    const AvatarModel& model;
    std::vector<double> w;                         // S
    double p[3] = {0, 0, 0};
    std::vector<std::array<double, 9>> r;          // J col-major 3x3
    std::vector<double> cloud;                     // 3 x V col-major
    std::vector<double> jointPos;                  // 3 x J
    std::vector<double> jointTrans;                // 12 x J
    explicit Avatar(const AvatarModel& m)
        : model(m), w(B3D_NUM_SHAPE, 0.0), r(B3D_NUM_JOINTS, {1, 0, 0, 0, 1, 0, 0, 0, 1}), cloud(3 * (size_t)m.numPoints, 0.0),
          jointPos(3 * B3D_NUM_JOINTS, 0.0), jointTrans(12 * B3D_NUM_JOINTS, 0.0) {}
    void update() {                                // external/avatar/Avatar.cpp:22-75
        check(b3d_smpl_forward(model.handle, 1, B3D_F64, w.data(), p, r[0].data(), cloud.data(), jointPos.data(), jointTrans.data(),
                               B3D_FWD_DEFAULT), "Avatar::update");
    }
};

namespace detail {
inline void pack_kps(const std::vector<PixelKP>& kps, const std::vector<int>& valid, std::vector<b3d_kp>& out, int32_t& n) {
    out.assign(B3D_MAX_KP, b3d_kp{0, 0, 0.0, 0.0});
    n = 0;
    for (const PixelKP& k : kps) {
        bool keep = valid.empty();                 // Sim3BA.h:410-413: an empty list does not filter
        for (int id : valid) keep = keep || id == k.jid;
        if (!keep) continue;
        if (n >= B3D_MAX_KP) throw std::runtime_error("more than 17 keypoints in a frame");
        out[n++] = b3d_kp{k.jid, 0, k.u, k.v};
    }
}
inline b3d_fit_opts opts(int mode, int max_iters, double bp, double bs, double lam, double fx, double fy, double cx, double cy) {
    b3d_fit_opts o;
    o.mode = mode; o.max_iters = max_iters; o.beta_pose = bp; o.beta_shape = bs; o.lambda_temp = lam;
    o.fx = fx; o.fy = fy; o.cx = cx; o.cy = cy; o.huber = 3.0; o.scale_lo = 0.3; o.scale_hi = 3.0;
    return o;
}
inline std::pair<bool, std::string> single(int mode, const AvatarModel& model, Avatar& avatar, const std::vector<PixelKP>& kps, double fx,
                                           double fy, double cx, double cy, const std::vector<int>& valid, Sim3Params& init, int max_iters,
                                           double bp, double bs, const GaussianMixture* gmm) {
    std::vector<b3d_kp> k;
    int32_t n = 0;
    pack_kps(kps, valid, k, n);
    const b3d_fit_opts o = opts(mode, max_iters, bp, bs, 0.0, fx, fy, cx, cy);
    double jaa[B3D_NUM_JOINTS * 3];
    uint8_t usable = 0;
    int32_t iters = 0, term = 0;
    const b3d_gmm* g = (gmm && gmm->nComps > 0) ? gmm->handle : nullptr;   // silent L2 degrade (Sim3BA.h:266)
    double R0[9];
    for (int i = 0; i < 9; ++i) R0[i] = avatar.r[0][i];
    check(b3d_fit_single(model.handle, g, &o, 1, k.data(), &n, R0, init.data, avatar.w.data(), jaa, &usable, nullptr, 0, &iters, &term),
          "b3d_fit_single");
    // write-back (Sim3BA.h:481-507 / 649-679)
    double pose[B3D_POSE_DIM];
    for (int i = 0; i < 7; ++i) pose[i] = init.data[i];
    for (int i = 0; i < 69; ++i) pose[7 + i] = jaa[3 + i];
    std::vector<double> r(B3D_NUM_JOINTS * 9);
    check(b3d_writeback(1, pose, R0, r.data(), avatar.p), "b3d_writeback");
    for (int j = 0; j < B3D_NUM_JOINTS; ++j)
        for (int i = 0; i < 9; ++i) avatar.r[j][i] = r[9 * j + i];
    char buf[96];
    std::snprintf(buf, sizeof(buf), "iterations %d, termination %d", iters, term);
    return {usable != 0, buf};
}
}  // namespace detail

inline std::pair<bool, std::string> OptimizePoseReprojection(const AvatarModel& model, Avatar& avatar, const std::vector<PixelKP>& kps,
                                                             double fx, double fy, double cx, double cy,
                                                             const std::vector<int>& valid_joint_ids, Sim3Params& initSim3,
                                                             int max_iters = 100, double betaPose = 0.0, double betaShape = 0.0,
                                                             const GaussianMixture* gmm = nullptr) {
    return detail::single(B3D_FIT_POSE, model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters, betaPose, betaShape, gmm);
}

inline std::pair<bool, std::string> OptimizePoseShapeReprojection(const AvatarModel& model, Avatar& avatar, const std::vector<PixelKP>& kps,
                                                                  double fx, double fy, double cx, double cy,
                                                                  const std::vector<int>& valid_joint_ids, Sim3Params& initSim3,
                                                                  int max_iters = 100, double betaPose = 0.0, double betaShape = 0.0,
                                                                  const GaussianMixture* gmm = nullptr) {
    return detail::single(B3D_FIT_POSE_SHAPE, model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters, betaPose, betaShape, gmm);
}

// include/MultiFrameBA.h:33-43.  Mutates avatars[f]->{r, p, w} and calls update() on each, as :154-174 does.
inline std::pair<bool, std::string> OptimizeMultiFrame(const AvatarModel& model, const std::vector<Avatar*>& avatars,
                                                       const std::vector<std::vector<PixelKP>>& kps_per_frame, double fx, double fy,
                                                       double cx, double cy, const std::vector<int>& valid_ids,
                                                       std::vector<FramePoseParams>& poses, double betaPose, double betaShape,
                                                       double lambdaTemp, int max_iters = 100) {
    const int F = (int)avatars.size();
    if (F == 0 || (int)poses.size() != F || (int)kps_per_frame.size() != F) return {false, "bad sizes"};
    std::vector<b3d_kp> kps((size_t)F * B3D_MAX_KP), one;
    std::vector<int32_t> nkp(F);
    std::vector<double> R0((size_t)F * 9), P((size_t)F * B3D_POSE_DIM), w(avatars.front()->w);   // shared block: the front avatar's w (:67)
    for (int f = 0; f < F; ++f) {
        detail::pack_kps(kps_per_frame[f], valid_ids, one, nkp[f]);
        for (int i = 0; i < B3D_MAX_KP; ++i) kps[(size_t)f * B3D_MAX_KP + i] = one[i];
        for (int i = 0; i < 9; ++i) R0[9 * (size_t)f + i] = avatars[f]->r[0][i];
        double* q = &P[(size_t)f * B3D_POSE_DIM];
        q[0] = poses[f].scale;
        for (int i = 0; i < 3; ++i) { q[1 + i] = poses[f].rootAA[i]; q[4 + i] = poses[f].rootT[i]; }
        for (int j = 1; j < B3D_NUM_JOINTS; ++j)
            for (int i = 0; i < 3; ++i) q[7 + 3 * (j - 1) + i] = poses[f].jointAA[j][i];
    }
    const b3d_fit_opts o = detail::opts(B3D_FIT_POSE_SHAPE, max_iters, betaPose, betaShape, lambdaTemp, fx, fy, cx, cy);
    uint8_t usable = 0;
    int32_t iters = 0, term = 0;
    check(b3d_fit_window(model.handle, &o, F, kps.data(), nkp.data(), R0.data(), P.data(), w.data(), &usable, nullptr, 0, &iters, &term),
          "b3d_fit_window");
    std::vector<double> r((size_t)F * B3D_NUM_JOINTS * 9), p((size_t)F * 3);
    check(b3d_writeback(F, P.data(), R0.data(), r.data(), p.data()), "b3d_writeback");
    for (int f = 0; f < F; ++f) {
        const double* q = &P[(size_t)f * B3D_POSE_DIM];
        poses[f].scale = q[0];
        for (int i = 0; i < 3; ++i) { poses[f].rootAA[i] = q[1 + i]; poses[f].rootT[i] = q[4 + i]; }
        for (int j = 1; j < B3D_NUM_JOINTS; ++j)
            for (int i = 0; i < 3; ++i) poses[f].jointAA[j][i] = q[7 + 3 * (j - 1) + i];
        Avatar& a = *avatars[f];
        for (int j = 0; j < B3D_NUM_JOINTS; ++j)
            for (int i = 0; i < 9; ++i) a.r[j][i] = r[((size_t)f * B3D_NUM_JOINTS + j) * 9 + i];
        for (int i = 0; i < 3; ++i) a.p[i] = p[3 * (size_t)f + i];
    }
    avatars.front()->w = w;                         // Ceres optimised the front avatar's block in place
    for (Avatar* a : avatars) a->update();          // :173
    char buf[96];
    std::snprintf(buf, sizeof(buf), "iterations %d, termination %d", iters, term);
    return {usable != 0, buf};
}

inline double mean_pixel_error(const std::vector<PixelKP>& kps, const Avatar& avatar, double fx, double fy, double cx, double cy) {
    std::vector<b3d_kp> k;
    int32_t n = 0;
    detail::pack_kps(kps, {}, k, n);
    double err = 0.0;
    check(b3d_mean_pixel_error(1, k.data(), &n, avatar.jointPos.data(), fx, fy, cx, cy, &err), "b3d_mean_pixel_error");
    return err;
}

}  // namespace shim
}  // namespace b3d

#endif  // B3D_SHIM_HPP_
