// Overlay rasteriser: smpl::render::renderSMPLMesh (reference include/RenderSMPLMesh.h:16-110) for B bodies on the device.
//
// Reference semantics kept: per-vertex projection (float u, v; vertices with Z <= 1e-6 invalid), per-face backface cull
// (n.z >= 0), flat shade round(220 * clamp(n^ . v^, 0, 1)), painter's order by mean Z (far first), integer pixel vertices
// (round half away from zero), one convex-polygon fill per face with OpenCV's coverage rule for LINE_8 (the three Bresenham
// edges of cv::LineIterator(leftToRight) after cv::clipLine, then the 16.16 fixed-point scanline fill of FillConvexPoly).
// What the painter's loop leaves in a pixel is the colour of the LAST face drawn over it = the covering face with the smallest
// mean depth (ties: highest face index, what a stable far-to-near sort draws last).  That makes the image order-free:
//   pass A  every face rasterises with atomicMin(depth bits) into a per-CTA depth plane,
//   pass B  faces whose depth equals the plane's value claim the pixel with atomicMax((face + 1) << 8 | colour),
//   pass C  the touched bounding box is resolved into the B x H x W x 3 image and both planes are reset.
// One CTA per body at a time (persistent over bodies); projected vertices, per-face depth and colour live in shared memory.
// Arithmetic that decides an integer (projection, shade, depth, clip) is done with explicitly rounded fp64 intrinsics so that no
// FMA contraction can move a rounding: the image is bit-identical to the CPU restatement (oracle/render.cpp), which is itself
// pinned against cv2.fillConvexPoly.  Not reproduced: cv::LINE_AA edge blending (order-dependent read-modify-write).
#include <climits>

#include "internal.hpp"

namespace b3d {
namespace {

constexpr int RT = 512;                    // threads per CTA
constexpr int COORD_LIMIT = 1 << 20;       // faces with a pixel coordinate beyond this are dropped (see b3d.h)
constexpr int XY_SHIFT = 16;
constexpr long long XY_ONE = 1ll << XY_SHIFT;

struct P2 { long long x, y; };

__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double ddiv(double a, double b) { return __ddiv_rn(a, b); }

// cv::clipLine(Size2l, Point2l&, Point2l&)
__device__ bool clip_line(long long width, long long height, P2& p1, P2& p2) {
    const long long right = width - 1, bottom = height - 1;
    long long &x1 = p1.x, &y1 = p1.y, &x2 = p2.x, &y2 = p2.y;
    int c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8;
    int c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8;
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
        long long a;
        if (c1 & 12) {
            a = c1 < 8 ? 0 : bottom;
            x1 += (long long)ddiv(dmul((double)(a - y1), (double)(x2 - x1)), (double)(y2 - y1));
            y1 = a;
            c1 = (x1 < 0) + (x1 > right) * 2;
        }
        if (c2 & 12) {
            a = c2 < 8 ? 0 : bottom;
            x2 += (long long)ddiv(dmul((double)(a - y2), (double)(x2 - x1)), (double)(y2 - y1));
            y2 = a;
            c2 = (x2 < 0) + (x2 > right) * 2;
        }
        if ((c1 & c2) == 0 && (c1 | c2) != 0) {
            if (c1) {
                a = c1 == 1 ? 0 : right;
                y1 += (long long)ddiv(dmul((double)(a - x1), (double)(y2 - y1)), (double)(x2 - x1));
                x1 = a;
                c1 = 0;
            }
            if (c2) {
                a = c2 == 1 ? 0 : right;
                y2 += (long long)ddiv(dmul((double)(a - x2), (double)(y2 - y1)), (double)(x2 - x1));
                x2 = a;
                c2 = 0;
            }
        }
    }
    return (c1 | c2) == 0;
}

// cv::Line -> LineIterator(img, pt1, pt2, 8, leftToRight = true)
template <class Put>
__device__ void line8(int W, int H, P2 p1, P2 p2, Put& put) {
    if ((unsigned long long)p1.x >= (unsigned long long)W || (unsigned long long)p2.x >= (unsigned long long)W ||
        (unsigned long long)p1.y >= (unsigned long long)H || (unsigned long long)p2.y >= (unsigned long long)H)
        if (!clip_line(W, H, p1, p2)) return;
    int dx = (int)(p2.x - p1.x), dy = (int)(p2.y - p1.y);
    int sy = 1;
    if (dx < 0) { dx = -dx; dy = -dy; p1 = p2; }
    if (dy < 0) { dy = -dy; sy = -1; }
    const bool vert = dy > dx;
    if (vert) { const int t = dx; dx = dy; dy = t; }
    int err = dx - (dy + dy);
    const int plus = dx + dx, minus = -(dy + dy);
    int x = (int)p1.x, y = (int)p1.y;
    for (int i = 0; i <= dx; ++i) {
        put(x, y);
        const bool m = err < 0;
        err += minus + (m ? plus : 0);
        if (vert) { y += sy; if (m) x += 1; }
        else { x += 1; if (m) y += sy; }
    }
}

// cv::FillConvexPoly(img, v, 3, colour, LINE_8, shift 0).  Rows above the image produce nothing and only advance the two
// edge accumulators linearly, so they are skipped in one step (OpenCV walks them one by one).
template <class Put>
__device__ void fill_tri(int W, int H, const P2 (&v)[3], Put& put) {
    int imin = 0, edges = 3;
    long long xmin = v[0].x, xmax = v[0].x, ymin = v[0].y, ymax = v[0].y;
    P2 p0 = v[2];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const P2 p = v[i];
        if (p.y < ymin) { ymin = p.y; imin = i; }
        ymax = max(ymax, p.y);
        xmax = max(xmax, p.x);
        xmin = min(xmin, p.x);
        line8(W, H, p0, p, put);
        p0 = p;
    }
    if (xmax < 0 || ymax < 0 || xmin >= W || ymin >= H) return;
    ymax = min(ymax, (long long)(H - 1));
    int e_idx[2] = {imin, imin}, e_ye[2];
    const int e_di[2] = {1, 2};
    long long e_x[2] = {-XY_ONE, -XY_ONE}, e_dx[2] = {0, 0};
    int y = (int)ymin;
    e_ye[0] = e_ye[1] = y;
    do {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            if (y >= e_ye[i]) {
                int idx0 = e_idx[i];
                int idx = idx0 + e_di[i];
                if (idx >= 3) idx -= 3;
                for (; edges-- > 0;) {
                    const int ty = (int)(idx == 0 ? v[0].y : idx == 1 ? v[1].y : v[2].y);
                    if (ty > y) {
                        const long long xs = (idx0 == 0 ? v[0].x : idx0 == 1 ? v[1].x : v[2].x) << XY_SHIFT;
                        const long long xe = (idx == 0 ? v[0].x : idx == 1 ? v[1].x : v[2].x) << XY_SHIFT;
                        e_ye[i] = ty;
                        e_dx[i] = ((xe - xs) * 2 + (ty - y)) / (2 * (ty - y));
                        e_x[i] = xs;
                        e_idx[i] = idx;
                        break;
                    }
                    idx0 = idx;
                    idx += e_di[i];
                    if (idx >= 3) idx -= 3;
                }
            }
        }
        if (edges < 0) break;
        int k = 1;
        if (y >= 0) {
            const bool sw = e_x[0] > e_x[1];
            int xx1 = (int)(((sw ? e_x[1] : e_x[0]) + (XY_ONE >> 1)) >> XY_SHIFT);
            int xx2 = (int)(((sw ? e_x[0] : e_x[1]) + (XY_ONE >> 1)) >> XY_SHIFT);
            if (xx2 >= 0 && xx1 < W) {
                if (xx1 < 0) xx1 = 0;
                if (xx2 >= W) xx2 = W - 1;
                for (int x = xx1; x <= xx2; ++x) put(x, y);
            }
        } else {
            const int stop = min(min(e_ye[0], e_ye[1]), 0);
            k = max(1, stop - y);
        }
        e_x[0] += e_dx[0] * k;
        e_x[1] += e_dx[1] * k;
        y += k;
    } while (y <= (int)ymax);
}


// shared memory: int2 pts[V] | u64 fdepth[F] | u8 fcolor[F] (colour 255 = not drawn)
template <typename T>
__global__ void __launch_bounds__(RT, 1)
render_overlay_kernel(int B, int V, int F, const T* __restrict__ cloud, const int32_t* __restrict__ faces, double fx, double fy, double cx,
                      double cy, int W, int H, int cull, unsigned long long* __restrict__ zplane, unsigned int* __restrict__ wplane,
                      uint8_t* __restrict__ images) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int2* pts = reinterpret_cast<int2*>(smem_raw);
    unsigned long long* fdepth = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)((V + 1) & ~1) * sizeof(int2));
    uint8_t* fcolor = reinterpret_cast<uint8_t*>(fdepth + F);
    __shared__ int bbox[4];
    const int tid = threadIdx.x;
    const size_t npix = (size_t)W * H;
    unsigned long long* Z = zplane + blockIdx.x * npix;
    unsigned int* Wn = wplane + blockIdx.x * npix;

    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const T* P = cloud + (size_t)b * V * 3;
        uint8_t* I = images + (size_t)b * npix * 3;
        if (tid == 0) { bbox[0] = W; bbox[1] = H; bbox[2] = -1; bbox[3] = -1; }
        // ---- vertices: projection (RenderSMPLMesh.h:40-47) and integer pixel (:79-84)
        for (int i = tid; i < V; i += RT) {
            const double X = (double)P[3 * i], Y = (double)P[3 * i + 1], Zc = (double)P[3 * i + 2];
            int2 q = make_int2(INT_MIN, INT_MIN);
            if (Zc > 1e-6) {
                const float u = __double2float_rn(dadd(ddiv(dmul(fx, X), Zc), cx));
                const float v = __double2float_rn(dadd(ddiv(dmul(fy, Y), Zc), cy));
                const float ru = roundf(u), rv = roundf(v);
                if (fabsf(ru) <= (float)COORD_LIMIT && fabsf(rv) <= (float)COORD_LIMIT) q = make_int2((int)ru, (int)rv);
                else q = make_int2(INT_MIN + 1, INT_MIN + 1);      // valid vertex, unusable pixel: its faces are dropped
            }
            pts[i] = q;
        }
        __syncthreads();
        // ---- pass A: face attributes (:50-77) + depth plane
        int bx0 = W, by0 = H, bx1 = -1, by1 = -1;
        for (int f = tid; f < F; f += RT) {
            const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
            const int2 q0 = pts[i0], q1 = pts[i1], q2 = pts[i2];
            uint8_t col = 255;
            unsigned long long dbits = ~0ull;
            if (q0.x > INT_MIN + 1 && q1.x > INT_MIN + 1 && q2.x > INT_MIN + 1) {
                const double a0 = (double)P[3 * i0], a1 = (double)P[3 * i0 + 1], a2 = (double)P[3 * i0 + 2];
                const double b0 = (double)P[3 * i1], b1 = (double)P[3 * i1 + 1], b2 = (double)P[3 * i1 + 2];
                const double c0 = (double)P[3 * i2], c1 = (double)P[3 * i2 + 1], c2 = (double)P[3 * i2 + 2];
                const double e10 = dsub(b0, a0), e11 = dsub(b1, a1), e12 = dsub(b2, a2);
                const double e20 = dsub(c0, a0), e21 = dsub(c1, a1), e22 = dsub(c2, a2);
                const double n0 = dsub(dmul(e11, e22), dmul(e12, e21));
                const double n1 = dsub(dmul(e12, e20), dmul(e10, e22));
                const double n2 = dsub(dmul(e10, e21), dmul(e11, e20));
                if (!(cull && n2 >= 0.0)) {
                    const double g0 = ddiv(dadd(dadd(a0, b0), c0), 3.0), g1 = ddiv(dadd(dadd(a1, b1), c1), 3.0);
                    const double g2 = ddiv(dadd(dadd(a2, b2), c2), 3.0);
                    const double cn2 = dadd(dadd(dmul(g0, g0), dmul(g1, g1)), dmul(g2, g2));
                    const double nn2 = dadd(dadd(dmul(n0, n0), dmul(n1, n1)), dmul(n2, n2));
                    const double cs = cn2 > 0.0 ? __dsqrt_rn(cn2) : 1.0, ns = nn2 > 0.0 ? __dsqrt_rn(nn2) : 1.0;
                    const double v0 = ddiv(-g0, cs), v1 = ddiv(-g1, cs), v2 = ddiv(-g2, cs);
                    const double h0 = ddiv(n0, ns), h1 = ddiv(n1, ns), h2 = ddiv(n2, ns);
                    double shade = dadd(dadd(dmul(h0, v0), dmul(h1, v1)), dmul(h2, v2));
                    shade = shade < 0.0 ? 0.0 : (shade > 1.0 ? 1.0 : shade);
                    col = (uint8_t)(int)round(dmul(220.0, shade));
                    dbits = (unsigned long long)__double_as_longlong(g2);       // mean Z > 0: bit pattern is order preserving
                    const P2 tri[3] = {{q0.x, q0.y}, {q1.x, q1.y}, {q2.x, q2.y}};
                    auto put = [&](int x, int y) {
                        atomicMin(&Z[(size_t)y * W + x], dbits);
                        bx0 = min(bx0, x); bx1 = max(bx1, x); by0 = min(by0, y); by1 = max(by1, y);
                    };
                    fill_tri(W, H, tri, put);
                }
            }
            fdepth[f] = dbits;
            fcolor[f] = col;
        }
        if (bx1 >= 0) { atomicMin(&bbox[0], bx0); atomicMin(&bbox[1], by0); atomicMax(&bbox[2], bx1); atomicMax(&bbox[3], by1); }
        __threadfence_block();
        __syncthreads();
        // ---- pass B: the nearest face of a pixel claims it (ties: highest face index = drawn last by a stable sort)
        for (int f = tid; f < F; f += RT) {
            const unsigned int col = fcolor[f];
            if (col == 255) continue;
            const unsigned long long dbits = fdepth[f];
            const int2 q0 = pts[faces[3 * f]], q1 = pts[faces[3 * f + 1]], q2 = pts[faces[3 * f + 2]];
            const P2 tri[3] = {{q0.x, q0.y}, {q1.x, q1.y}, {q2.x, q2.y}};
            const unsigned int tag = ((unsigned int)(f + 1) << 8) | col;
            auto put = [&](int x, int y) {
                const size_t p = (size_t)y * W + x;
                if (__ldcg(&Z[p]) == dbits) atomicMax(&Wn[p], tag);
            };
            fill_tri(W, H, tri, put);
        }
        __threadfence_block();
        __syncthreads();
        // ---- pass C: resolve the touched box into the image, reset the planes
        const int x0 = bbox[0], y0 = bbox[1], x1 = bbox[2], y1 = bbox[3];
        if (x1 >= x0) {
            const int bw = x1 - x0 + 1, n = bw * (y1 - y0 + 1);
            for (int k = tid; k < n; k += RT) {
                const int y = y0 + k / bw, x = x0 + k % bw;
                const size_t p = (size_t)y * W + x;
                const unsigned int w = __ldcg(&Wn[p]);
                if (w) {
                    const uint8_t c = (uint8_t)(w & 255u);
                    uint8_t* o = I + p * 3;
                    o[0] = c; o[1] = c; o[2] = c;
                    Wn[p] = 0u;
                    Z[p] = ~0ull;
                }
            }
        }
        __syncthreads();
    }
}

static size_t render_smem_bytes(int V, int F) { return (size_t)((V + 1) & ~1) * sizeof(int2) + (size_t)F * 9 + 16; }

}  // namespace
}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_render_overlay_dev(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, double fx, double fy, double cx, double cy,
                           int32_t W, int32_t H, uint32_t flags, uint8_t* images, void* stream) {
    if (!m || !cloud || !images) return set_error(B3D_EINVAL, "b3d_render_overlay: null argument");
    if (B < 0 || W <= 0 || H <= 0 || (dtype != B3D_F32 && dtype != B3D_F64)) return set_error(B3D_EINVAL, "b3d_render_overlay: bad size / dtype");
    if (m->F <= 0 || !m->d_faces) return set_error(B3D_EINVAL, "b3d_render_overlay: the model was created without faces");
    if (B == 0) return B3D_OK;
    if (int rc = require_device()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = render_smem_bytes(m->V, m->F);
    if (smem > 227 * 1024) return set_error(B3D_EINVAL, "b3d_render_overlay: mesh too large for the shared-memory face tables");
    int dev = 0, sms = 0;
    B3D_CUDA(cudaGetDevice(&dev));
    B3D_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = B < sms ? B : sms;
    const size_t npix = (size_t)W * H;
    unsigned long long* z = nullptr;
    B3D_CUDA(cudaMallocAsync((void**)&z, (size_t)grid * npix * 12, st));
    unsigned int* wn = reinterpret_cast<unsigned int*>(z + (size_t)grid * npix);
    cudaError_t e = cudaMemsetAsync(z, 0xff, (size_t)grid * npix * 8, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(wn, 0, (size_t)grid * npix * 4, st);
    if (e == cudaSuccess) {
        const int cull = (flags & B3D_RENDER_NO_CULL) ? 0 : 1;
        if (dtype == B3D_F32) {
            e = cudaFuncSetAttribute(render_overlay_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess)
                render_overlay_kernel<float><<<grid, RT, smem, st>>>(B, m->V, m->F, (const float*)cloud, m->d_faces, fx, fy, cx, cy, W, H, cull, z, wn, images);
        } else {
            e = cudaFuncSetAttribute(render_overlay_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess)
                render_overlay_kernel<double><<<grid, RT, smem, st>>>(B, m->V, m->F, (const double*)cloud, m->d_faces, fx, fy, cx, cy, W, H, cull, z, wn, images);
        }
        if (e == cudaSuccess) { count_launch(); e = cudaGetLastError(); }
    }
    cudaFreeAsync(z, st);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_render_overlay: %s", cudaGetErrorString(e));
    return B3D_OK;
}

int b3d_render_overlay(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, double fx, double fy, double cx, double cy,
                       int32_t W, int32_t H, uint32_t flags, uint8_t* images) {
    if (!m || !cloud || !images) return set_error(B3D_EINVAL, "b3d_render_overlay: null argument");
    if (B < 0 || W <= 0 || H <= 0 || (dtype != B3D_F32 && dtype != B3D_F64)) return set_error(B3D_EINVAL, "b3d_render_overlay: bad size / dtype");
    if (B == 0) return B3D_OK;
    if (int rc = require_device()) return rc;
    const size_t cb = (size_t)B * m->V * 3 * (dtype == B3D_F32 ? 4 : 8), ib = (size_t)B * W * H * 3;
    char* dc = nullptr;
    B3D_CUDA(cudaMalloc((void**)&dc, ((cb + 255) & ~(size_t)255) + ib));
    uint8_t* di = (uint8_t*)(dc + ((cb + 255) & ~(size_t)255));
    cudaError_t e = cudaMemcpy(dc, cloud, cb, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(di, images, ib, cudaMemcpyHostToDevice);
    int rc = B3D_OK;
    if (e == cudaSuccess) rc = b3d_render_overlay_dev(m, B, dtype, dc, fx, fy, cx, cy, W, H, flags, di, nullptr);
    if (e == cudaSuccess && rc == B3D_OK) e = cudaMemcpy(images, di, ib, cudaMemcpyDeviceToHost);
    cudaFree(dc);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_render_overlay: %s", cudaGetErrorString(e));
    return rc;
}

}  // extern "C"
