// b3d_fit_frames: the per-frame body of the reference's single-frame driver (src/main_single_frame.cpp:205-255), batched over
// frames and kept on the device between stages: H2D inputs -> LM fit -> write-back -> Avatar::update -> mean pixel error -> D2H.
#include "internal.hpp"

struct b3d_workspace {
    cudaStream_t stream = nullptr;
    char* buf = nullptr;
    size_t cap = 0;
    void* cloud = nullptr;      // inside buf
    int cloud_dtype = B3D_F64;
};

using namespace b3d;

namespace {
inline size_t al256(size_t x) { return (x + 255) / 256 * 256; }
}

extern "C" {

int b3d_workspace_create(b3d_workspace** out) {
    if (!out) return set_error(B3D_EINVAL, "b3d_workspace_create: null out");
    *out = nullptr;
    if (int rc = require_device()) return rc;
    b3d_workspace* ws = new b3d_workspace();
    cudaError_t e = cudaStreamCreateWithFlags(&ws->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete ws; return set_error(B3D_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
    *out = ws;
    return B3D_OK;
}

void b3d_workspace_destroy(b3d_workspace* ws) {
    if (!ws) return;
    cudaFree(ws->buf);
    if (ws->stream) cudaStreamDestroy(ws->stream);
    delete ws;
}

const void* b3d_workspace_cloud(const b3d_workspace* ws, int32_t* dtype) {
    if (!ws) return nullptr;
    if (dtype) *dtype = ws->cloud_dtype;
    return ws->cloud;
}

int b3d_fit_frames(b3d_workspace* ws, const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* opts, int32_t B, const b3d_kp* kps,
                   const int32_t* nkp, const double* R0, double* sim3, double* w, double* jointAA, uint8_t* usable, int32_t* niters,
                   int32_t* term, double* r, double* p, double* jointPos, double* px_err, void* cloud, int32_t cloud_dtype) {
    if (!ws || !m || !opts || !kps || !nkp || !R0 || !sim3 || !w || !jointAA) return set_error(B3D_EINVAL, "b3d_fit_frames: null argument");
    if (cloud_dtype != B3D_F32 && cloud_dtype != B3D_F64) return set_error(B3D_EINVAL, "b3d_fit_frames: bad cloud dtype");
    if (B <= 0) return B < 0 ? set_error(B3D_EINVAL, "b3d_fit_frames: B < 0") : B3D_OK;
    if (int rc = validate_keypoints(kps, nkp, (size_t)B, "b3d_fit_frames")) return rc;
    const size_t nb = (size_t)B, es = cloud_dtype == B3D_F32 ? 4 : 8;
    const size_t s_kp = al256(sizeof(b3d_kp) * nb * MAXKP), s_nk = al256(4 * nb), s_r0 = al256(8 * nb * 9), s_s3 = al256(8 * nb * 7),
                 s_w = al256(8 * nb * NS), s_ja = al256(8 * nb * NJ * 3), s_us = al256(nb), s_ni = al256(4 * nb), s_po = al256(8 * nb * NPF),
                 s_r = al256(8 * nb * NJ * 9), s_p = al256(8 * nb * 3), s_jp = al256(8 * nb * NJ * 3), s_jt = al256(8 * nb * NJ * 12),
                 s_pe = al256(8 * nb), s_cl = al256(es * nb * 3 * m->V), s_cv = al256(es * nb * (NS + 3 + NJ * 9 + NJ * 3 + NJ * 12));
    const size_t total = s_kp + s_nk + s_r0 + s_s3 + s_w + s_ja + s_us + 2 * s_ni + s_po + s_r + s_p + s_jp + s_jt + s_pe + s_cl + s_cv;
    if (total > ws->cap) {
        cudaFree(ws->buf);
        ws->buf = nullptr; ws->cap = 0; ws->cloud = nullptr;
        B3D_CUDA(cudaMalloc(&ws->buf, total));
        ws->cap = total;
    }
    char* q = ws->buf;
    auto take = [&](size_t s) { char* t = q; q += s; return t; };
    b3d_kp* d_kp = (b3d_kp*)take(s_kp); int32_t* d_nk = (int32_t*)take(s_nk); double* d_r0 = (double*)take(s_r0);
    double* d_s3 = (double*)take(s_s3); double* d_w = (double*)take(s_w); double* d_ja = (double*)take(s_ja);
    uint8_t* d_us = (uint8_t*)take(s_us); int32_t* d_ni = (int32_t*)take(s_ni); int32_t* d_te = (int32_t*)take(s_ni);
    double* d_po = (double*)take(s_po); double* d_r = (double*)take(s_r); double* d_p = (double*)take(s_p);
    double* d_jp = (double*)take(s_jp); double* d_jt = (double*)take(s_jt); double* d_pe = (double*)take(s_pe);
    char* d_cl = take(s_cl); char* d_cv = take(s_cv);
    cudaStream_t st = ws->stream;
    B3D_CUDA(cudaMemcpyAsync(d_kp, kps, sizeof(b3d_kp) * nb * MAXKP, cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(d_nk, nkp, 4 * nb, cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(d_r0, R0, 8 * nb * 9, cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(d_s3, sim3, 8 * nb * 7, cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(d_w, w, 8 * nb * NS, cudaMemcpyHostToDevice, st));
    if (int rc = b3d_fit_single_dev(m, gmm, opts, B, d_kp, d_nk, d_r0, d_s3, d_w, d_ja, d_us, nullptr, 0, d_ni, d_te, st)) return rc;
    // poses = [sim3, jointAA[1..23]] ; write-back ; update ; pixel error
    B3D_CUDA(cudaMemcpy2DAsync(d_po, 8 * NPF, d_s3, 8 * 7, 8 * 7, nb, cudaMemcpyDeviceToDevice, st));
    B3D_CUDA(cudaMemcpy2DAsync(d_po + 7, 8 * NPF, d_ja + 3, 8 * NJ * 3, 8 * (NJ - 1) * 3, nb, cudaMemcpyDeviceToDevice, st));
    if (int rc = b3d_writeback_dev(B, d_po, d_r0, d_r, d_p, st)) return rc;
    if (cloud_dtype == B3D_F64) {
        if (int rc = b3d_smpl_forward_dev(m, B, B3D_F64, d_w, d_p, d_r, d_cl, d_jp, d_jt, B3D_FWD_DEFAULT, st)) return rc;
    } else {
        // float32 cloud: joints in float64 for the pixel error, then the float32 evaluator for the 6890-vertex cloud
        if (int rc = b3d_smpl_forward_dev(m, B, B3D_F64, d_w, d_p, d_r, nullptr, d_jp, d_jt, B3D_FWD_DEFAULT, st)) return rc;
        float* cw = (float*)d_cv; float* cp = cw + nb * NS; float* cr = cp + nb * 3; float* cjp = cr + nb * NJ * 9; float* cjt = cjp + nb * NJ * 3;
        if (int rc = b3d::b3d_convert_f64_f32_dev(d_w, cw, nb * NS, st)) return rc;
        if (int rc = b3d::b3d_convert_f64_f32_dev(d_p, cp, nb * 3, st)) return rc;
        if (int rc = b3d::b3d_convert_f64_f32_dev(d_r, cr, nb * NJ * 9, st)) return rc;
        if (int rc = b3d_smpl_forward_dev(m, B, B3D_F32, cw, cp, cr, d_cl, cjp, cjt, B3D_FWD_DEFAULT, st)) return rc;
    }
    ws->cloud = d_cl; ws->cloud_dtype = cloud_dtype;
    if (int rc = b3d_mean_pixel_error_dev(B, d_kp, d_nk, d_jp, opts->fx, opts->fy, opts->cx, opts->cy, d_pe, st)) return rc;
    B3D_CUDA(cudaMemcpyAsync(sim3, d_s3, 8 * nb * 7, cudaMemcpyDeviceToHost, st));
    B3D_CUDA(cudaMemcpyAsync(w, d_w, 8 * nb * NS, cudaMemcpyDeviceToHost, st));
    B3D_CUDA(cudaMemcpyAsync(jointAA, d_ja, 8 * nb * NJ * 3, cudaMemcpyDeviceToHost, st));
    if (usable) B3D_CUDA(cudaMemcpyAsync(usable, d_us, nb, cudaMemcpyDeviceToHost, st));
    if (niters) B3D_CUDA(cudaMemcpyAsync(niters, d_ni, 4 * nb, cudaMemcpyDeviceToHost, st));
    if (term) B3D_CUDA(cudaMemcpyAsync(term, d_te, 4 * nb, cudaMemcpyDeviceToHost, st));
    if (r) B3D_CUDA(cudaMemcpyAsync(r, d_r, 8 * nb * NJ * 9, cudaMemcpyDeviceToHost, st));
    if (p) B3D_CUDA(cudaMemcpyAsync(p, d_p, 8 * nb * 3, cudaMemcpyDeviceToHost, st));
    if (jointPos) B3D_CUDA(cudaMemcpyAsync(jointPos, d_jp, 8 * nb * NJ * 3, cudaMemcpyDeviceToHost, st));
    if (px_err) B3D_CUDA(cudaMemcpyAsync(px_err, d_pe, 8 * nb, cudaMemcpyDeviceToHost, st));
    if (cloud) B3D_CUDA(cudaMemcpyAsync(cloud, d_cl, es * nb * 3 * m->V, cudaMemcpyDeviceToHost, st));
    B3D_CUDA(cudaStreamSynchronize(st));
    return B3D_OK;
}

}  // extern "C"
