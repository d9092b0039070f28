// Correspondence search of ark::AvatarOptimizer's ICP loop: findNN with invert = true
// (external/avatar/AvatarOptimizer.cpp:841-907, called at :1372-1375) — every data point is matched to the nearest VISIBLE model
// point that carries its part label.  The reference builds a nanoflann kd-tree per part and asks it for the exact nearest
// neighbour (metric_L2_Simple, eps = 0); here the search is exhaustive, which returns the same point: a body's posed cloud is
// staged once per CTA in shared memory, grouped by part (invisible points as +inf so that they never win), threads are data
// points and walk their part's slice — lanes of one part read the same address (broadcast).  The squared distance is formed as
// (dx*dx + dy*dy) + dz*dz with separately rounded operations, so that the float64 instantiation reproduces the reference's
// comparisons bit for bit; ties go to the lowest model index.  Output: matched model point per data point, -1 for none —
// exactly the pair list b3d_icp_create takes (dpt = NULL).
#include <algorithm>
#include <limits>
#include <vector>

#include "internal.hpp"

namespace b3d {

constexpr int NN_THREADS = 512, NN_PER_THREAD = 4;   // a work item: up to 32 * NN_PER_THREAD points of one part
constexpr int NN_LOAD = 4;                           // 2048 points per CTA chunk (the local index fits the 12 low bits of `order`):
                                                     // larger chunks give fuller per-part buckets

template <typename T> __device__ __forceinline__ T inf_();
template <> __device__ __forceinline__ float inf_<float>() { return __int_as_float(0x7f800000); }
template <> __device__ __forceinline__ double inf_<double>() { return __longlong_as_double(0x7ff0000000000000LL); }
__device__ __forceinline__ float mul_rn_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn_(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double mul_rn_(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn_(double a, double b) { return __dadd_rn(a, b); }

// float64: the reference's own operation order, every operation rounded separately (indices bit-identical to the CPU search);
// float32 (the fast path, no bit-exactness claimed against a float64 reference): fused
__device__ __forceinline__ double dist2_(double dx, double dy, double dz) { return add_rn_(add_rn_(mul_rn_(dx, dx), mul_rn_(dy, dy)), mul_rn_(dz, dz)); }
__device__ __forceinline__ float dist2_(float dx, float dy, float dz) { return fmaf(dz, dz, fmaf(dy, dy, dx * dx)); }

template <typename T>
__device__ __forceinline__ void nn_test(T px, T py, T pz, T mx, T my, T mz, int k, T& best, int& arg) {
    const T dx = px - mx, dy = py - my, dz = pz - mz;
    const T r = dist2_(dx, dy, dz);
    if (r < best) { best = r; arg = k; }
}

// One work item: c <= 32 U points of one part (bucket-order positions pos0 ...), lane takes positions lane, lane + 32, ...
template <typename T, int U>
__device__ __forceinline__ void nn_item(const T* sx, const T* sy, const T* sz, const T* qx, const T* qy, const T* qz, const int32_t* order,
                                        int k0, int k1, int pos0, int c, int lane, const int32_t* __restrict__ part_idx, int32_t* __restrict__ out) {
    int li[U], arg[U];
    T px[U], py[U], pz[U], best[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const bool ok = lane + 32 * u < c;
        li[u] = ok ? (order[pos0 + lane + 32 * u] & 4095) : -1;
        const int q = ok ? li[u] : 0;
        px[u] = qx[q]; py[u] = qy[q]; pz[u] = qz[q];
        best[u] = inf_<T>(); arg[u] = -1;
    }
    for (int k = k0; k < k1; ++k) {
        const T mx = sx[k], my = sy[k], mz = sz[k];
#pragma unroll
        for (int u = 0; u < U; ++u) nn_test(px[u], py[u], pz[u], mx, my, mz, k, best[u], arg[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
        if (li[u] >= 0) out[li[u]] = arg[u] >= 0 ? part_idx[arg[u]] : -1;
}

// CTA = chunks of up to NN_THREADS * NN_LOAD data points of one body.  The points are bucketed by part in shared memory first, so that
// a thread's NN_PER_THREAD points (adjacent in bucket order) and the lanes of a warp nearly always share a part: one broadcast
// read of a model point then serves NN_PER_THREAD * 32 distance tests.
template <typename T>
__global__ void __launch_bounds__(NN_THREADS) icp_nearest_kernel(int V, int nparts, const int32_t* __restrict__ part_ptr,
                                                                 const int32_t* __restrict__ part_idx, const T* __restrict__ cloud,
                                                                 const uint8_t* __restrict__ visible, const int64_t* __restrict__ dptr,
                                                                 const T* __restrict__ data, const int32_t* __restrict__ data_part,
                                                                 int32_t* __restrict__ out) {
    constexpr int NPTS = NN_THREADS * NN_LOAD;
    extern __shared__ __align__(16) unsigned char nn_smem[];
    T* sx = reinterpret_cast<T*>(nn_smem);
    T* sy = sx + V;
    T* sz = sy + V;
    T* qx = sz + V;                                       // the CTA's data points
    T* qy = qx + NPTS;
    T* qz = qy + NPTS;
    int32_t* sp = reinterpret_cast<int32_t*>(qz + NPTS);  // [nparts + 1] slices of the part-ordered cloud
    int32_t* cnt = sp + nparts + 1;                       // [nparts + 2] bucket cursors (bucket nparts = no valid part)
    int32_t* order = cnt + nparts + 2;                    // [NPTS] (bucket << 12) | local index, in bucket order
    const int b = blockIdx.y;
    const int64_t d0 = dptr[b], nd = dptr[b + 1] - d0;
    if ((int64_t)blockIdx.x * NPTS >= nd) return;
    const T* cb = cloud + 3 * (size_t)V * b;
    const T inf = inf_<T>();
    for (int k = threadIdx.x; k < V; k += NN_THREADS) {
        const int v = part_idx[k];
        const bool vis = !visible || visible[(size_t)V * b + v];
        sx[k] = vis ? cb[3 * (size_t)v] : inf;
        sy[k] = vis ? cb[3 * (size_t)v + 1] : inf;
        sz[k] = vis ? cb[3 * (size_t)v + 2] : inf;
    }
    for (int k = threadIdx.x; k <= nparts; k += NN_THREADS) sp[k] = part_ptr[k];
    // the staged cloud serves every chunk of data points this CTA takes
    for (int64_t first = (int64_t)blockIdx.x * NPTS; first < nd; first += (int64_t)gridDim.x * NPTS) {
    const int n_here = (int)min((int64_t)NPTS, nd - first);
    __syncthreads();   // the previous chunk's readers of cnt / order / q are done; the cloud is staged
    for (int k = threadIdx.x; k < nparts + 2; k += NN_THREADS) cnt[k] = 0;
    __syncthreads();
    int bucket[NN_LOAD];
#pragma unroll
    for (int u = 0; u < NN_LOAD; ++u) {
        const int i = u * NN_THREADS + threadIdx.x;
        bucket[u] = -1;
        if (i < n_here) {
            const int64_t g = d0 + first + i;
            qx[i] = data[3 * g]; qy[i] = data[3 * g + 1]; qz[i] = data[3 * g + 2];
            const int part = data_part[g];
            bucket[u] = (part >= 0 && part < nparts) ? part : nparts;
            atomicAdd(&cnt[bucket[u] + 1], 1);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int k = 1; k < nparts + 2; ++k) cnt[k] += cnt[k - 1];   // cnt[k] = first position of bucket k
    __syncthreads();
#pragma unroll
    for (int u = 0; u < NN_LOAD; ++u)
        if (bucket[u] >= 0) order[atomicAdd(&cnt[bucket[u]], 1)] = (bucket[u] << 12) | (u * NN_THREADS + threadIdx.x);
    __syncthreads();
    // After the scatter cnt[k] is the END of bucket k.  Work items = (bucket, block of up to 128 of its points), dealt to the warps
    // round robin; the lanes of a warp take the item's points strided by 32, so a warp always works on ONE part: a model point
    // is one broadcast read for up to 128 distance tests, and short buckets still fill all 32 lanes.
    {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        int item = 0;
        for (int k = 0; k < nparts; ++k) {
            const int start = k ? cnt[k - 1] : 0, count = cnt[k] - start;
            for (int o = 0; o < count; o += 32 * NN_PER_THREAD, ++item) {
                if (item % (NN_THREADS / 32) != warp) continue;
                const int c = min(32 * NN_PER_THREAD, count - o);
                if (c > 64) nn_item<T, 4>(sx, sy, sz, qx, qy, qz, order, sp[k], sp[k + 1], start + o, c, lane, part_idx, out + d0 + first);
                else if (c > 32) nn_item<T, 2>(sx, sy, sz, qx, qy, qz, order, sp[k], sp[k + 1], start + o, c, lane, part_idx, out + d0 + first);
                else nn_item<T, 1>(sx, sy, sz, qx, qy, qz, order, sp[k], sp[k + 1], start + o, c, lane, part_idx, out + d0 + first);
            }
        }
        // data points without a valid part label
        for (int pos = (nparts ? cnt[nparts - 1] : 0) + threadIdx.x; pos < n_here; pos += NN_THREADS) out[d0 + first + (order[pos] & 4095)] = -1;
    }
    }
}

// pointVisible of the ICP loop (AvatarOptimizer.cpp:1337-1352): a vertex is visible if it belongs to a face whose normal has
// ((p2 - p1) x (p1 - p3)).z > 1e-4 in the posed cloud.  thread = (body, face); the three byte stores race benignly (all write 1).
template <typename T>
__global__ void __launch_bounds__(256) icp_visible_kernel(int B, int V, int F, const int32_t* __restrict__ faces, const T* __restrict__ cloud,
                                                          uint8_t* __restrict__ visible) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= (size_t)B * F) return;
    const int b = (int)(i / F), f = (int)(i % F);
    const int i1 = faces[3 * f], i2 = faces[3 * f + 1], i3 = faces[3 * f + 2];
    const T* c = cloud + 3 * (size_t)V * b;
    const T ax = c[3 * i2] - c[3 * i1], ay = c[3 * i2 + 1] - c[3 * i1 + 1];
    const T bx = c[3 * i1] - c[3 * i3], by = c[3 * i1 + 1] - c[3 * i3 + 1];
    const T z = add_rn_(mul_rn_(ax, by), -mul_rn_(ay, bx));   // Eigen's cross(): a.x * b.y - a.y * b.x, no contraction
    if (z > (T)1e-4) {
        uint8_t* vb = visible + (size_t)V * b;
        vb[i1] = 1; vb[i2] = 1; vb[i3] = 1;
    }
}

}  // namespace b3d

using namespace b3d;

extern "C" int b3d_icp_visible(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, int32_t flags, uint8_t* visible) {
    if (!m || !cloud || !visible) return set_error(B3D_EINVAL, "b3d_icp_visible: null argument");
    if (B <= 0) return set_error(B3D_EINVAL, "b3d_icp_visible: B <= 0");
    if (dtype != B3D_F32 && dtype != B3D_F64) return set_error(B3D_EINVAL, "b3d_icp_visible: bad dtype %d", dtype);
    if (!m->d_faces || m->F <= 0) return set_error(B3D_EINVAL, "b3d_icp_visible: the model was created without faces");
    if (int rc = require_device()) return rc;
    const bool dev = (flags & B3D_ICP_DEVICE_INPUT) != 0;
    const size_t es = dtype == B3D_F32 ? 4 : 8, nc = es * 3 * (size_t)m->V * B, nv = (size_t)m->V * B;
    char* buf = nullptr;
    const void* pc = cloud;
    uint8_t* pv = visible;
    cudaError_t e = cudaSuccess;
    if (!dev) {
        e = cudaMalloc(&buf, nc + nv);
        if (e == cudaSuccess) e = cudaMemcpy(buf, cloud, nc, cudaMemcpyHostToDevice);
        pc = buf; pv = (uint8_t*)(buf + nc);
    }
    if (e == cudaSuccess) e = cudaMemsetAsync(pv, 0, nv, nullptr);
    if (e == cudaSuccess) {
        const unsigned grid = (unsigned)(((size_t)B * m->F + 255) / 256);
        if (dtype == B3D_F32) icp_visible_kernel<float><<<grid, 256>>>(B, m->V, m->F, m->d_faces, (const float*)pc, pv);
        else icp_visible_kernel<double><<<grid, 256>>>(B, m->V, m->F, m->d_faces, (const double*)pc, pv);
        count_launch();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(nullptr);
    if (e == cudaSuccess && !dev) e = cudaMemcpy(visible, pv, nv, cudaMemcpyDeviceToHost);
    cudaFree(buf);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_icp_visible: %s", cudaGetErrorString(e));
    return B3D_OK;
}


extern "C" int b3d_icp_nearest(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, int32_t nparts, const int32_t* model_part,
                               const uint8_t* visible, const int64_t* dptr, const void* data, const int32_t* data_part, int32_t flags,
                               int32_t* nn_out) {
    if (!m || !cloud || !model_part || !dptr || !data || !data_part || !nn_out) return set_error(B3D_EINVAL, "b3d_icp_nearest: null argument");
    if (B <= 0 || nparts <= 0 || nparts > 4096) return set_error(B3D_EINVAL, "b3d_icp_nearest: bad B / nparts");
    if (dtype != B3D_F32 && dtype != B3D_F64) return set_error(B3D_EINVAL, "b3d_icp_nearest: bad dtype %d", dtype);
    if (int rc = require_device()) return rc;
    const int V = m->V;
    const size_t es = dtype == B3D_F32 ? 4 : 8;
    // model points grouped by part, ascending index inside a part (= the order findNN fills its part clouds in, :865-874)
    std::vector<int32_t> pptr(nparts + 1, 0), pidx(V);
    int labelled = 0;
    for (int v = 0; v < V; ++v)
        if (model_part[v] >= 0 && model_part[v] < nparts) { ++pptr[model_part[v] + 1]; ++labelled; }
    for (int k = 0; k < nparts; ++k) pptr[k + 1] += pptr[k];
    {
        std::vector<int32_t> fill(pptr.begin(), pptr.end() - 1);
        for (int v = 0; v < V; ++v)
            if (model_part[v] >= 0 && model_part[v] < nparts) pidx[fill[model_part[v]]++] = v;
    }
    for (int k = labelled; k < V; ++k) pidx[k] = 0;   // unlabelled model points are in no part's slice
    int64_t maxnd = 0;
    for (int b = 0; b < B; ++b) {
        if (dptr[b + 1] < dptr[b]) return set_error(B3D_EINVAL, "b3d_icp_nearest: offsets of body %d decrease", b);
        maxnd = std::max(maxnd, dptr[b + 1] - dptr[b]);
    }
    const int64_t ndt = dptr[B];
    if (maxnd == 0) return B3D_OK;
    const bool dev = (flags & B3D_ICP_DEVICE_INPUT) != 0;
    int32_t* d_meta = nullptr;   // part_ptr | part_idx
    int64_t* d_dptr = nullptr;
    char* d_in = nullptr;
    cudaError_t e = cudaMalloc(&d_meta, sizeof(int32_t) * (nparts + 1 + V));
    if (e == cudaSuccess) e = cudaMalloc(&d_dptr, sizeof(int64_t) * (B + 1));
    if (e == cudaSuccess) e = cudaMemcpy(d_meta, pptr.data(), sizeof(int32_t) * (nparts + 1), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_meta + nparts + 1, pidx.data(), sizeof(int32_t) * V, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_dptr, dptr, sizeof(int64_t) * (B + 1), cudaMemcpyHostToDevice);
    const void *p_cloud = cloud, *p_data = data;
    const uint8_t* p_vis = visible;
    const int32_t* p_dpart = data_part;
    int32_t* p_out = nn_out;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    if (e == cudaSuccess && !dev) {
        const size_t s_cloud = al(es * 3 * (size_t)V * B), s_data = al(es * 3 * (size_t)ndt), s_part = al(4 * (size_t)ndt), s_vis = visible ? al((size_t)V * B) : 0;
        e = cudaMalloc(&d_in, s_cloud + s_data + 2 * s_part + s_vis);
        if (e == cudaSuccess) {
            char *c = d_in, *d = c + s_cloud, *dp = d + s_data, *o = dp + s_part, *vs = o + s_part;
            e = cudaMemcpy(c, cloud, es * 3 * (size_t)V * B, cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(d, data, es * 3 * (size_t)ndt, cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(dp, data_part, 4 * (size_t)ndt, cudaMemcpyHostToDevice);
            if (e == cudaSuccess && visible) e = cudaMemcpy(vs, visible, (size_t)V * B, cudaMemcpyHostToDevice);
            p_cloud = c; p_data = d; p_dpart = (const int32_t*)dp; p_out = (int32_t*)o; p_vis = visible ? (const uint8_t*)vs : nullptr;
        }
    }
    int rc = B3D_OK;
    if (e == cudaSuccess) {
        const size_t smem = es * 3 * ((size_t)V + NN_THREADS * NN_LOAD) + 4 * (size_t)(2 * nparts + 3 + NN_THREADS * NN_LOAD);
        // one CTA per body when the batch fills the machine (the cloud is staged once), more chunks in flight when it does not
        const int64_t chunks = (maxnd + NN_THREADS * NN_LOAD - 1) / (NN_THREADS * NN_LOAD);
        const dim3 grid((unsigned)std::min<int64_t>(chunks, std::max(1, (2 * 148 + B - 1) / B)), B);
        if (smem > 232448) rc = set_error(B3D_EINVAL, "b3d_icp_nearest: model too large for the shared-memory cloud (%zu B)", smem);
        else if (dtype == B3D_F32) {
            e = cudaFuncSetAttribute(icp_nearest_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess)
                icp_nearest_kernel<float><<<grid, NN_THREADS, smem>>>(V, nparts, d_meta, d_meta + nparts + 1, (const float*)p_cloud, p_vis, d_dptr,
                                                                      (const float*)p_data, p_dpart, p_out);
        } else {
            e = cudaFuncSetAttribute(icp_nearest_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess)
                icp_nearest_kernel<double><<<grid, NN_THREADS, smem>>>(V, nparts, d_meta, d_meta + nparts + 1, (const double*)p_cloud, p_vis, d_dptr,
                                                                       (const double*)p_data, p_dpart, p_out);
        }
        if (!rc && e == cudaSuccess) { count_launch(); e = cudaGetLastError(); }
        if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(nullptr);
        if (!rc && e == cudaSuccess && !dev) e = cudaMemcpy(nn_out, p_out, 4 * (size_t)ndt, cudaMemcpyDeviceToHost);
    }
    cudaFree(d_meta); cudaFree(d_dptr); cudaFree(d_in);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_icp_nearest: %s", cudaGetErrorString(e));
    return rc;
}
