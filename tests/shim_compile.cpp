// Compile-and-link check of include/b3d_shim.hpp (tests/test_capi_cpu.py): the reference-signature shim must build with a
// plain host compiler against lib3dba_b200.so.  Run without a GPU it must fail loudly (B3D_ECUDA -> exception), never fall back.
#include <cstdio>
#include <cstring>

#include "b3d_shim.hpp"

using namespace b3d::shim;

int main() {
    std::vector<float> vt(3 * 8, 0.f), sd(3 * 8 * 10, 0.f), jr(24 * 8, 0.f), wt(8 * 24, 0.f);
    // SMPL's kinematic tree (the fit kernels size their ancestor chains for it); a small blob of vertices per joint
    const int parent[24] = {-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21};
    std::vector<uint32_t> kt(24, 0);
    for (int j = 0; j < 24; ++j) kt[j] = (uint32_t)parent[j];
    for (int v = 0; v < 8; ++v) {
        vt[3 * v] = 0.1f * (float)(v % 3); vt[3 * v + 1] = 0.2f * (float)v - 0.7f; vt[3 * v + 2] = 0.05f * (float)(v % 2);
        wt[v * 24 + (v * 3) % 24] = 1.f;
    }
    for (int j = 0; j < 24; ++j) jr[j * 8 + j % 8] = 1.f;
    b3d_model_desc d;
    std::memset(&d, 0, sizeof(d));
    d.num_verts = 8; d.num_joints = 24; d.num_shape = 10; d.dtype = B3D_F32;
    d.v_template = vt.data(); d.shapedirs = sd.data(); d.J_regressor = jr.data(); d.weights = wt.data(); d.kintree_row0 = kt.data();
    try {
        AvatarModel model(d);
        Avatar av(model);
        av.update();
        std::vector<PixelKP> kps = {{1, 10.0, 20.0}};
        Sim3Params s3;
        s3.t()[2] = 3.0;
        auto r1 = OptimizePoseReprojection(model, av, kps, 432, 432, 240, 135, {1}, s3, 5, 20.0, 30.0, nullptr);
        auto r2 = OptimizePoseShapeReprojection(model, av, kps, 432, 432, 240, 135, {1}, s3, 5, 20.0, 30.0, &model.posePrior);
        std::vector<Avatar*> avs = {&av};
        std::vector<FramePoseParams> poses(1);
        poses[0].jointAA.assign(24, {0, 0, 0});
        poses[0].rootT[2] = 3.0;
        auto r3 = OptimizeMultiFrame(model, avs, {kps}, 432, 432, 240, 135, {1}, poses, 5.0, 25.0, 3.0, 5);
        std::printf("ran on a GPU: %d %d %d err %.3f\n", (int)r1.first, (int)r2.first, (int)r3.first, mean_pixel_error(kps, av, 432, 432, 240, 135));
    } catch (const std::exception& e) {
        std::printf("no usable device: %s\n", e.what());
        return 3;
    }
    return 0;
}
