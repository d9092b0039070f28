// GPU parity driver for include/b3d_shim.hpp (tests/test_shim_gpu.py): runs the three reference-signature calls on inputs the test
// wrote to a raw binary file and prints every result in full precision, so that the test can compare them with the ctypes path.
//   file: int32 V, nframes ; float32 v_template[V*3], shapedirs[V*3*10], J_regressor[24*V], weights[V*24] ; uint32 kintree[24] ;
//         per frame: int32 nkp, then 17 x { int32 jid, int32 pad, double u, double v } ; double R0[9] (col-major) ; pose_prior path (c-string)
#include <cstdio>
#include <cstring>
#include <fstream>

#include "b3d_shim.hpp"

using namespace b3d::shim;

template <typename T>
static void rd(std::ifstream& f, T* p, size_t n) { f.read(reinterpret_cast<char*>(p), (std::streamsize)(sizeof(T) * n)); }

static void print_vec(const char* name, const double* v, int n) {
    std::printf("%s", name);
    for (int i = 0; i < n; ++i) std::printf(" %.17g", v[i]);
    std::printf("\n");
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    std::ifstream f(argv[1], std::ios::binary);
    if (!f) return 2;
    int32_t V = 0, nframes = 0;
    rd(f, &V, 1); rd(f, &nframes, 1);
    std::vector<float> vt(3 * (size_t)V), sd(30 * (size_t)V), jr(24 * (size_t)V), wt(24 * (size_t)V);
    std::vector<uint32_t> kt(24);
    rd(f, vt.data(), vt.size()); rd(f, sd.data(), sd.size()); rd(f, jr.data(), jr.size()); rd(f, wt.data(), wt.size()); rd(f, kt.data(), 24);
    std::vector<std::vector<PixelKP>> kps(nframes);
    for (int i = 0; i < nframes; ++i) {
        int32_t n = 0;
        rd(f, &n, 1);
        b3d_kp raw[B3D_MAX_KP];
        rd(f, raw, B3D_MAX_KP);
        for (int k = 0; k < n; ++k) kps[i].push_back(PixelKP{raw[k].jid, raw[k].u, raw[k].v});
    }
    double R0[9];
    rd(f, R0, 9);
    std::string prior;
    std::getline(f, prior, '\0');
    if (!f) return 2;
    b3d_model_desc d;
    std::memset(&d, 0, sizeof(d));
    d.num_verts = V; d.num_joints = 24; d.num_shape = 10; d.dtype = B3D_F32;
    d.v_template = vt.data(); d.shapedirs = sd.data(); d.J_regressor = jr.data(); d.weights = wt.data(); d.kintree_row0 = kt.data();
    try {
        AvatarModel model(d, prior);
        // ---- OptimizePoseReprojection and OptimizePoseShapeReprojection on frame 0 (src/main_single_frame.cpp:205-255) ----
        for (int mode = 0; mode < 2; ++mode) {
            Avatar av(model);
            for (int i = 0; i < 9; ++i) av.r[0][i] = R0[i];
            av.p[2] = 3.0;
            av.update();
            Sim3Params s3;
            s3.t()[2] = 3.0;
            std::vector<int> valid;
            for (const PixelKP& k : kps[0]) valid.push_back(k.jid);
            auto res = mode == 0 ? OptimizePoseReprojection(model, av, kps[0], 432, 432, 240, 135, valid, s3, 100, 20.0, 30.0, nullptr)
                                 : OptimizePoseShapeReprojection(model, av, kps[0], 432, 432, 240, 135, valid, s3, 100, 20.0, 30.0, &model.posePrior);
            av.update();
            std::printf("single%d usable %d %s\n", mode, (int)res.first, res.second.c_str());
            print_vec(mode ? "single1 sim3" : "single0 sim3", s3.data, 7);
            print_vec(mode ? "single1 w" : "single0 w", av.w.data(), 10);
            print_vec(mode ? "single1 r" : "single0 r", av.r[0].data(), 9 * 24);
            const double e = mean_pixel_error(kps[0], av, 432, 432, 240, 135);
            print_vec(mode ? "single1 px_err" : "single0 px_err", &e, 1);
        }
        // ---- OptimizeMultiFrame over all frames (include/MultiFrameBA.h:33-177) ----
        std::vector<Avatar> store;
        store.reserve(nframes);
        std::vector<Avatar*> avs;
        std::vector<FramePoseParams> poses(nframes);
        for (int i = 0; i < nframes; ++i) {
            store.emplace_back(model);
            for (int k = 0; k < 9; ++k) store.back().r[0][k] = R0[k];
            avs.push_back(&store.back());
            poses[i].jointAA.assign(24, {0, 0, 0});
            poses[i].rootT[2] = 3.0;
        }
        auto r3 = OptimizeMultiFrame(model, avs, kps, 432, 432, 240, 135, {}, poses, 20.0, 30.0, 3.0, 40);
        std::printf("multi usable %d %s\n", (int)r3.first, r3.second.c_str());
        print_vec("multi w", avs[0]->w.data(), 10);
        for (int i = 0; i < nframes; ++i) {
            double q[76];
            q[0] = poses[i].scale;
            for (int k = 0; k < 3; ++k) { q[1 + k] = poses[i].rootAA[k]; q[4 + k] = poses[i].rootT[k]; }
            for (int j = 1; j < 24; ++j)
                for (int k = 0; k < 3; ++k) q[7 + 3 * (j - 1) + k] = poses[i].jointAA[j][k];
            print_vec("multi pose", q, 76);
            print_vec("multi jointPos", avs[i]->jointPos.data(), 72);
        }
    } catch (const std::exception& e) {
        std::printf("exception: %s\n", e.what());
        return 3;
    }
    return 0;
}
