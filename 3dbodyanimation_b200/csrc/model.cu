// Model handle: one-time derivation (host, float64 — as ark::AvatarModel does at load, external/avatar/AvatarModel.cpp:23-127)
// and upload of the per-vertex tables in the structure-of-arrays layout the forward kernels read.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

#include "internal.hpp"

namespace b3d {

static thread_local std::string g_last_error;
static std::atomic<uint64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int require_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "no CUDA device: %s (this library has no CPU fallback)", cudaGetErrorString(e));
    int major = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
    if (major != 10) return set_error(B3D_ECUDA, "device %d has compute capability %d.x; the kernels are built for sm_100a only", dev, major);
    // The per-call scratch buffers come from the stream-ordered pool (cudaMallocAsync).  Its default release threshold is 0: at
    // every synchronisation the pool hands its free memory back to the driver, so a caller that synchronises after each call
    // (b3d_fit_frames, every host-pointer entry point) would pay a real allocation per call — measured as 14 -> 18-58 ms per
    // 1,000-frame step.  Keep the memory in the pool.
    static bool pool_set[64] = {};
    if (dev >= 0 && dev < 64 && !pool_set[dev]) {
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            uint64_t keep = UINT64_MAX;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        pool_set[dev] = true;
    }
    return B3D_OK;
}

template <typename T>
static inline double load_as_double(const void* p, size_t i, int dtype) {
    return dtype == B3D_F64 ? ((const double*)p)[i] : (double)((const float*)p)[i];
}

template <typename T>
static int upload_tables(const b3d_model& m, const std::vector<double>& vert, const std::vector<double>& wval,
                         const std::vector<int32_t>& wjoint, VertexTables<T>& t) {
    std::vector<T> tmp(vert.size());
    for (size_t i = 0; i < vert.size(); ++i) tmp[i] = (T)vert[i];
    B3D_CUDA(cudaMalloc(&t.vert, sizeof(T) * tmp.size()));
    B3D_CUDA(cudaMemcpy(t.vert, tmp.data(), sizeof(T) * tmp.size(), cudaMemcpyHostToDevice));
    tmp.resize(wval.size());
    for (size_t i = 0; i < wval.size(); ++i) tmp[i] = (T)wval[i];
    B3D_CUDA(cudaMalloc(&t.wval, sizeof(T) * tmp.size()));
    B3D_CUDA(cudaMemcpy(t.wval, tmp.data(), sizeof(T) * tmp.size(), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMalloc(&t.wjoint, sizeof(int32_t) * wjoint.size()));
    B3D_CUDA(cudaMemcpy(t.wjoint, wjoint.data(), sizeof(int32_t) * wjoint.size(), cudaMemcpyHostToDevice));
    return B3D_OK;
}

// Tile-sorted float tables (SortedTables, internal.hpp): slots of a 256-vertex tile ordered by joint tuple, padding last.
static int upload_sorted_tables(b3d_model& m, const std::vector<double>& vert, const std::vector<double>& wval,
                                const std::vector<int32_t>& wjoint) {
    constexpr int TILE = 256;
    const size_t Vp = (size_t)m.Vpad;
    const int rows = 3 + 3 * NS;
    std::vector<float> sv(rows * Vp, 0.f), sw(4 * Vp, 0.f);
    std::vector<int32_t> sj(4 * Vp, 0), loc(Vp, 0);
    std::vector<int> order(TILE);
    for (size_t base = 0; base < Vp; base += TILE) {
        for (int i = 0; i < TILE; ++i) order[i] = i;
        auto key = [&](int i) -> uint64_t {
            const size_t v = base + i;
            if ((int)v >= m.V) return ~0ull;
            uint64_t k = 0;
            for (int q = 0; q < 4; ++q) k = k * 32 + (uint64_t)(wval[q * Vp + v] != 0.0 ? wjoint[q * Vp + v] : 31);
            return k;
        };
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return key(a) < key(b); });
        for (int t = 0; t < TILE; ++t) {
            const size_t src = base + order[t], dst = base + t;
            loc[dst] = order[t];
            for (int r = 0; r < rows; ++r) sv[r * Vp + dst] = (float)vert[r * Vp + src];
            for (int q = 0; q < 4; ++q) {
                sw[q * Vp + dst] = (float)wval[q * Vp + src];
                // a zero-weight slot re-reads the joint of slot 0, so that it does not add a distinct shared-memory address
                sj[q * Vp + dst] = 12 * (wval[q * Vp + src] != 0.0 ? wjoint[q * Vp + src] : wjoint[src]);
            }
        }
    }
    // records of the body-lane kernel
    std::vector<float> rec(40 * Vp, 0.f);
    std::vector<int32_t> j4(4 * Vp, 0);
    for (size_t slot = 0; slot < Vp; ++slot) {
        float* r = &rec[40 * slot];
        for (int c = 0; c < 3; ++c) {
            r[c] = sv[c * Vp + slot];
            for (int s2 = 0; s2 < NS; ++s2) r[4 + NS * c + s2] = sv[(3 + NS * c + s2) * Vp + slot];
        }
        const int32_t l3 = 3 * loc[slot];
        std::memcpy(&r[3], &l3, 4);
        bool changed = (slot % 32) == 0;
        for (int q = 0; q < 4; ++q) {
            r[34 + q] = sw[q * Vp + slot];
            j4[4 * slot + q] = sj[q * Vp + slot];
            if (slot > 0 && sj[q * Vp + slot] != sj[q * Vp + slot - 1]) changed = true;
        }
        const int32_t flag = changed ? 1 : 0;
        std::memcpy(&r[38], &flag, 4);
    }
    B3D_CUDA(cudaMalloc(&m.ts.rec, sizeof(float) * rec.size()));
    B3D_CUDA(cudaMemcpy(m.ts.rec, rec.data(), sizeof(float) * rec.size(), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMalloc(&m.ts.joff4, sizeof(int32_t) * j4.size()));
    B3D_CUDA(cudaMemcpy(m.ts.joff4, j4.data(), sizeof(int32_t) * j4.size(), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMalloc(&m.ts.vert, sizeof(float) * sv.size()));
    B3D_CUDA(cudaMemcpy(m.ts.vert, sv.data(), sizeof(float) * sv.size(), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMalloc(&m.ts.wval, sizeof(float) * sw.size()));
    B3D_CUDA(cudaMemcpy(m.ts.wval, sw.data(), sizeof(float) * sw.size(), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMalloc(&m.ts.wjoff, sizeof(int32_t) * sj.size()));
    B3D_CUDA(cudaMemcpy(m.ts.wjoff, sj.data(), sizeof(int32_t) * sj.size(), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMalloc(&m.ts.loc, sizeof(int32_t) * loc.size()));
    B3D_CUDA(cudaMemcpy(m.ts.loc, loc.data(), sizeof(int32_t) * loc.size(), cudaMemcpyHostToDevice));
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

const char* b3d_last_error(void) { return g_last_error.c_str(); }
const char* b3d_version(void) { return "3dba-b200 0.1 (sm_100a)"; }
uint64_t b3d_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int b3d_model_create(const b3d_model_desc* d, b3d_model** out) {
    if (!d || !out) return set_error(B3D_EINVAL, "b3d_model_create: null argument");
    *out = nullptr;
    if (d->num_joints != NJ || d->num_shape != NS)
        return set_error(B3D_EINVAL, "b3d_model_create: need %d joints and %d shape keys (got %d, %d)", NJ, NS, d->num_joints, d->num_shape);
    if (d->num_verts <= 0 || !d->v_template || !d->shapedirs || !d->J_regressor || !d->weights || !d->kintree_row0)
        return set_error(B3D_EINVAL, "b3d_model_create: missing array");
    if (d->dtype != B3D_F32 && d->dtype != B3D_F64) return set_error(B3D_EINVAL, "b3d_model_create: bad dtype");
    if (int rc = require_device()) return rc;

    b3d_model* m = new b3d_model();
    cudaGetDevice(&m->device);
    const int V = d->num_verts, dt = d->dtype;
    m->V = V;
    m->Vpad = (V + 255) / 256 * 256;   // multiple of the LBS tile (256) and of the posedirs tile (128)
    m->F = d->faces ? d->num_faces : 0;
    // kintree row 0: uint32 -> int, root 0xFFFFFFFF -> -1 (AvatarModel.cpp:35-41)
    for (int j = 0; j < NJ; ++j) m->sk.parent[j] = (int32_t)d->kintree_row0[j];
    if (m->sk.parent[0] != -1) { delete m; return set_error(B3D_EINVAL, "kintree root parent must be -1 (_ARK_ASSERT_EQ, AvatarModel.cpp:41)"); }
    m->sk.depth[0] = 0;
    m->sk.max_depth = 0;
    for (int j = 1; j < NJ; ++j) {
        const int pa = m->sk.parent[j];
        if (pa < 0 || pa >= j) { delete m; return set_error(B3D_EINVAL, "joint %d: parent %d must precede it (Avatar.cpp:50-57 walks joints in order)", j, pa); }
        m->sk.depth[j] = m->sk.depth[pa] + 1;
        if (m->sk.depth[j] > m->sk.max_depth) m->sk.max_depth = m->sk.depth[j];
    }
    for (int l = 0, pos = 0; l <= NJ + 1; ++l) {
        m->sk.lvstart[l] = pos;
        for (int j = 0; j < NJ; ++j)
            if (m->sk.depth[j] == l) m->sk.order[pos++] = j;
    }
    // initialJointPos = template * jointRegressor; jointShapeReg column s = shapedir s * jointRegressor (AvatarModel.cpp:113-127)
    for (int i = 0; i < 3 * NJ; ++i) m->sk.J0[i] = 0.0;
    for (int i = 0; i < 3 * NJ * NS; ++i) m->sk.JSR[i] = 0.0;
    for (int j = 0; j < NJ; ++j)
        for (int v = 0; v < V; ++v) {
            const double g = load_as_double<double>(d->J_regressor, (size_t)j * V + v, dt);
            if (g == 0.0) continue;
            for (int c = 0; c < 3; ++c) {
                m->sk.J0[3 * j + c] += load_as_double<double>(d->v_template, 3 * (size_t)v + c, dt) * g;
                for (int s = 0; s < NS; ++s)
                    m->sk.JSR[(3 * j + c) * NS + s] += load_as_double<double>(d->shapedirs, (3 * (size_t)v + c) * NS + s, dt) * g;
            }
        }
    // rest bone vectors: base avatar (w = 0, r = I, p = 0), root moved to the origin, offset[j] = J_j - J_parent (Sim3BA.h:533-555)
    {
        double posed[3 * NJ];   // posed joints of the base avatar: root at p = 0, child = parent + (J0_child - J0_parent)
        for (int c = 0; c < 3; ++c) { posed[c] = 0.0; m->sk.offset[c] = 0.0; }
        for (int j = 1; j < NJ; ++j) {
            const int pa = m->sk.parent[j];
            for (int c = 0; c < 3; ++c) posed[3 * j + c] = posed[3 * pa + c] + (m->sk.J0[3 * j + c] - m->sk.J0[3 * pa + c]);
        }
        for (int j = 1; j < NJ; ++j)
            for (int c = 0; c < 3; ++c)
                m->sk.offset[3 * j + c] = (posed[3 * j + c] - posed[c]) - (posed[3 * m->sk.parent[j] + c] - posed[c]);
    }

    // per-vertex tables
    const size_t Vp = (size_t)m->Vpad;
    int maxnz = 0;
    for (int v = 0; v < V; ++v) {
        int nz = 0;
        for (int j = 0; j < NJ; ++j) nz += load_as_double<double>(d->weights, (size_t)v * NJ + j, dt) != 0.0;
        if (nz > maxnz) maxnz = nz;
    }
    m->nw = maxnz <= 4 ? 4 : (maxnz <= 8 ? 8 : NJ);
    std::vector<double> vert((3 + 3 * NS) * Vp, 0.0), wval((size_t)m->nw * Vp, 0.0);
    std::vector<int32_t> wjoint((size_t)m->nw * Vp, 0);
    for (int v = 0; v < V; ++v) {
        for (int c = 0; c < 3; ++c) {
            vert[c * Vp + v] = load_as_double<double>(d->v_template, 3 * (size_t)v + c, dt);
            for (int s = 0; s < NS; ++s)
                vert[(3 + NS * c + s) * Vp + v] = load_as_double<double>(d->shapedirs, (3 * (size_t)v + c) * NS + s, dt);
        }
        int k = 0;
        for (int j = 0; j < NJ; ++j) {
            const double x = load_as_double<double>(d->weights, (size_t)v * NJ + j, dt);
            if (x != 0.0) { wval[k * Vp + v] = x; wjoint[k * Vp + v] = j; ++k; }   // sparseView drops exact zeros only
        }
    }
    int rc = upload_tables<float>(*m, vert, wval, wjoint, m->t32);
    if (!rc) rc = upload_tables<double>(*m, vert, wval, wjoint, m->t64);
    if (!rc && m->nw == 4) rc = upload_sorted_tables(*m, vert, wval, wjoint);
    if (!rc && m->nw == 4) rc = build_tc_tables(*m, vert, wval, wjoint);
    if (!rc && d->posedirs) {
        const int P = B3D_POSEDIRS_DIM;
        std::vector<float> p32((size_t)P * 3 * Vp, 0.f);
        std::vector<double> p64((size_t)P * 3 * Vp, 0.0);
        for (int v = 0; v < V; ++v)
            for (int c = 0; c < 3; ++c)
                for (int k = 0; k < P; ++k) {
                    const double x = load_as_double<double>(d->posedirs, (3 * (size_t)v + c) * P + k, dt);
                    p64[((size_t)k * 3 + c) * Vp + v] = x;
                    p32[((size_t)k * 3 + c) * Vp + v] = (float)x;
                }
        cudaError_t e = cudaMalloc(&m->posedirs32, sizeof(float) * p32.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->posedirs32, p32.data(), sizeof(float) * p32.size(), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMalloc(&m->posedirs64, sizeof(double) * p64.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->posedirs64, p64.data(), sizeof(double) * p64.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) rc = set_error(B3D_ECUDA, "posedirs upload: %s", cudaGetErrorString(e));
        if (!rc && m->nw == 4) rc = build_tc_pose_tables(*m, p32);
        m->has_posedirs = (rc == 0);
    }
    if (!rc) {
        cudaError_t e = cudaMalloc(&m->d_sk, sizeof(SkeletonConst));
        if (e == cudaSuccess) e = cudaMemcpy(m->d_sk, &m->sk, sizeof(SkeletonConst), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) rc = set_error(B3D_ECUDA, "skeleton upload: %s", cudaGetErrorString(e));
    }
    if (!rc && m->F > 0) {
        m->faces.assign(d->faces, d->faces + 3 * (size_t)m->F);
        for (uint32_t v : m->faces)
            if (v >= (uint32_t)m->V) { rc = set_error(B3D_EFORMAT, "face index %u out of range (V = %d)", v, m->V); break; }
        if (!rc) {
            cudaError_t e = cudaMalloc(&m->d_faces, sizeof(int32_t) * m->faces.size());
            if (e == cudaSuccess) e = cudaMemcpy(m->d_faces, m->faces.data(), sizeof(int32_t) * m->faces.size(), cudaMemcpyHostToDevice);
            if (e != cudaSuccess) rc = set_error(B3D_ECUDA, "faces upload: %s", cudaGetErrorString(e));
        }
    }
    if (rc) { b3d_model_destroy(m); return rc; }
    *out = m;
    return B3D_OK;
}

void b3d_model_destroy(b3d_model* m) {
    if (!m) return;
    cudaFree(m->t32.vert); cudaFree(m->t32.wval); cudaFree(m->t32.wjoint);
    cudaFree(m->t64.vert); cudaFree(m->t64.wval); cudaFree(m->t64.wjoint);
    cudaFree(m->ts.vert); cudaFree(m->ts.wval); cudaFree(m->ts.wjoff); cudaFree(m->ts.loc);
    cudaFree(m->ts.rec); cudaFree(m->ts.joff4); cudaFree(m->tc.blob); cudaFree(m->tc.pose_b);
    cudaFree(m->posedirs32); cudaFree(m->posedirs64); cudaFree(m->d_sk); cudaFree(m->d_faces);
    delete m;
}

int b3d_model_num_verts(const b3d_model* m) { return m ? m->V : -1; }

int b3d_model_get_parents(const b3d_model* m, int32_t* parent) {
    if (!m || !parent) return set_error(B3D_EINVAL, "b3d_model_get_parents: null argument");
    for (int j = 0; j < NJ; ++j) parent[j] = m->sk.parent[j];
    return B3D_OK;
}

int b3d_model_get_constants(const b3d_model* m, double* J0, double* JSR, double* offset) {
    if (!m) return set_error(B3D_EINVAL, "b3d_model_get_constants: null model");
    if (J0) std::memcpy(J0, m->sk.J0, sizeof(m->sk.J0));
    if (JSR) std::memcpy(JSR, m->sk.JSR, sizeof(m->sk.JSR));
    if (offset) std::memcpy(offset, m->sk.offset, sizeof(m->sk.offset));
    return B3D_OK;
}

}  // extern "C"
