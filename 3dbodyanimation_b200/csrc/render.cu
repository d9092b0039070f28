// Overlay rasteriser: smpl::render::renderSMPLMesh (reference include/RenderSMPLMesh.h:16-110) for B bodies on the device.
//
// Reference semantics kept: per-vertex projection (float u, v; vertices with Z <= 1e-6 invalid), per-face backface cull
// (n.z >= 0), flat shade round(220 * clamp(n^ . v^, 0, 1)), painter's order by mean Z (far first), integer pixel vertices
// (round half away from zero), one convex-polygon fill per face with OpenCV's coverage rule for LINE_8 (the three Bresenham
// edges of cv::LineIterator(leftToRight) after cv::clipLine, then the 16.16 fixed-point scanline fill of FillConvexPoly).
// What the painter's loop leaves in a pixel is the colour of the LAST face drawn over it = the covering face with the smallest
// mean depth (ties: highest face index, what a stable far-to-near sort draws last).  That makes the image order-free:
//   pass A  every drawn face (one per lane, compacted after the cull) walks its pixels with atomicMin(depth bits) into a per-CTA depth plane,
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

// One triangle's pixel walk, flattened so that the 32 lanes of a warp (one face each) stay in ONE loop: a lane alternates
// between "set up the next run" and "emit the next pixel of the run".  Runs, in OpenCV's order:
//   three edge lines  cv::Line -> LineIterator(img, pt1, pt2, 8, leftToRight = true) after cv::clipLine,
//   then the rows of  cv::FillConvexPoly(img, v, 3, colour, LINE_8, shift 0): two edge accumulators in 16.16 fixed point.
// Rows above the image produce nothing and only advance the accumulators linearly, so they are skipped in one step
// (OpenCV walks them one by one).  The arithmetic is OpenCV's, statement for statement; only the loop nest is folded.
struct TriWalk {
    P2 v[3];
    int seg;                 // next run to set up: 0..2 edge lines (v2->v0, v0->v1, v1->v2), 3 = fill rows, 4 = finished
    int rem, x, y;           // pixels left in the current run, next pixel
    int err, plus, minus, sy;
    bool vert;
    // fill state
    int fy, ymax, edges;
    int e_idx[2], e_ye[2];
    long long e_x[2], e_dx[2];
};

__device__ __forceinline__ long long tw_vx(const TriWalk& t, int i) { return i == 0 ? t.v[0].x : i == 1 ? t.v[1].x : t.v[2].x; }
__device__ __forceinline__ long long tw_vy(const TriWalk& t, int i) { return i == 0 ? t.v[0].y : i == 1 ? t.v[1].y : t.v[2].y; }

__device__ void tw_setup_line(TriWalk& t, int W, int H, P2 p1, P2 p2) {
    t.rem = 0;
    if ((unsigned long long)p1.x >= (unsigned long long)W || (unsigned long long)p2.x >= (unsigned long long)W ||
        (unsigned long long)p1.y >= (unsigned long long)H || (unsigned long long)p2.y >= (unsigned long long)H)
        if (!clip_line(W, H, p1, p2)) return;
    int dx = (int)(p2.x - p1.x), dy = (int)(p2.y - p1.y);
    t.sy = 1;
    if (dx < 0) { dx = -dx; dy = -dy; p1 = p2; }
    if (dy < 0) { dy = -dy; t.sy = -1; }
    t.vert = dy > dx;
    if (t.vert) { const int q = dx; dx = dy; dy = q; }
    t.err = dx - (dy + dy);
    t.plus = dx + dx;
    t.minus = -(dy + dy);
    t.x = (int)p1.x;
    t.y = (int)p1.y;
    t.rem = dx + 1;
}

__device__ void tw_fill_init(TriWalk& t, int W, int H) {
    int imin = 0;
    long long xmin = t.v[0].x, xmax = t.v[0].x, ymin = t.v[0].y, ymax = t.v[0].y;
#pragma unroll
    for (int i = 1; i < 3; ++i) {
        if (t.v[i].y < ymin) { ymin = t.v[i].y; imin = i; }
        ymax = max(ymax, t.v[i].y);
        xmax = max(xmax, t.v[i].x);
        xmin = min(xmin, t.v[i].x);
    }
    if (xmax < 0 || ymax < 0 || xmin >= W || ymin >= H) { t.seg = 4; return; }
    t.ymax = (int)min(ymax, (long long)(H - 1));
    t.edges = 3;
    t.e_idx[0] = t.e_idx[1] = imin;
    t.fy = (int)ymin;
    t.e_ye[0] = t.e_ye[1] = t.fy;
    t.e_x[0] = t.e_x[1] = -XY_ONE;
    t.e_dx[0] = t.e_dx[1] = 0;
}

// one iteration of FillConvexPoly's do { } while (++y <= ymax): leaves the row's span as the current run (possibly empty)
__device__ void tw_fill_row(TriWalk& t, int W) {
    const int y = t.fy;
    t.rem = 0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        if (y >= t.e_ye[i]) {
            const int di = i == 0 ? 1 : 2;
            int idx0 = t.e_idx[i];
            int idx = idx0 + di;
            if (idx >= 3) idx -= 3;
            for (; t.edges-- > 0;) {
                const int ty = (int)tw_vy(t, idx);
                if (ty > y) {
                    const long long xs = tw_vx(t, idx0) << XY_SHIFT, xe = tw_vx(t, idx) << XY_SHIFT;
                    t.e_ye[i] = ty;
                    t.e_dx[i] = ((xe - xs) * 2 + (ty - y)) / (2 * (ty - y));
                    t.e_x[i] = xs;
                    t.e_idx[i] = idx;
                    break;
                }
                idx0 = idx;
                idx += di;
                if (idx >= 3) idx -= 3;
            }
        }
    }
    if (t.edges < 0) { t.seg = 4; return; }
    int k = 1;
    if (y >= 0) {
        const bool sw = t.e_x[0] > t.e_x[1];
        int xx1 = (int)(((sw ? t.e_x[1] : t.e_x[0]) + (XY_ONE >> 1)) >> XY_SHIFT);
        int xx2 = (int)(((sw ? t.e_x[0] : t.e_x[1]) + (XY_ONE >> 1)) >> XY_SHIFT);
        if (xx2 >= 0 && xx1 < W) {
            if (xx1 < 0) xx1 = 0;
            if (xx2 >= W) xx2 = W - 1;
            if (xx2 >= xx1) {       // a horizontal run: never steps in y
                t.x = xx1; t.y = y; t.rem = xx2 - xx1 + 1;
                t.vert = false; t.err = 0x3fffffff; t.plus = 0; t.minus = 0;
            }
        }
    } else {
        k = max(1, min(min(t.e_ye[0], t.e_ye[1]), 0) - y);
    }
    t.e_x[0] += t.e_dx[0] * k;
    t.e_x[1] += t.e_dx[1] * k;
    t.fy = y + k;
    if (t.fy > t.ymax) t.seg = 4;
}

// Warp-synchronous: every lane of the warp must call this (alive = false for lanes without a face).
template <class Put>
__device__ __forceinline__ void walk_triangle(int W, int H, bool alive, const P2 (&tri)[3], Put& put) {
    TriWalk t;
    t.v[0] = tri[0]; t.v[1] = tri[1]; t.v[2] = tri[2];
    t.seg = alive ? 0 : 4;
    t.rem = 0;
    while (__any_sync(0xffffffffu, t.seg < 4 || t.rem > 0)) {
        if (t.rem == 0 && t.seg < 4) {
            if (t.seg < 3) {
                const int s = t.seg++;
                tw_setup_line(t, W, H, s == 0 ? t.v[2] : s == 1 ? t.v[0] : t.v[1], s == 0 ? t.v[0] : s == 1 ? t.v[1] : t.v[2]);
                if (t.seg == 3) tw_fill_init(t, W, H);
            } else {
                tw_fill_row(t, W);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (t.rem > 0) {
                put(t.x, t.y);
                const bool m = t.err < 0;
                t.err += t.minus + (m ? t.plus : 0);
                if (t.vert) { t.y += t.sy; if (m) t.x += 1; }
                else { t.x += 1; if (m) t.y += t.sy; }
                --t.rem;
            }
        }
    }
}

// shared memory: int2 pts[V] | u64 cdepth[F] | u32 ctag[F]   (compact list of the faces that are drawn: depth bits, face << 8 | colour)
template <typename T>
__global__ void __launch_bounds__(RT, 1)
render_overlay_kernel(int B, int V, int F, const T* __restrict__ cloud, const int32_t* __restrict__ faces, double fx, double fy, double cx,
                      double cy, int W, int H, int cull, unsigned long long* __restrict__ zplane, unsigned int* __restrict__ wplane,
                      uint8_t* __restrict__ images) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int2* pts = reinterpret_cast<int2*>(smem_raw);
    unsigned long long* cdepth = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)((V + 1) & ~1) * sizeof(int2));
    unsigned int* ctag = reinterpret_cast<unsigned int*>(cdepth + F);
    __shared__ int bbox[4];
    __shared__ int ndrawn;
    const int tid = threadIdx.x, lane = tid & 31, wbase = tid & ~31;
    const size_t npix = (size_t)W * H;
    unsigned long long* Z = zplane + blockIdx.x * npix;
    unsigned int* Wn = wplane + blockIdx.x * npix;

    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const T* P = cloud + (size_t)b * V * 3;
        uint8_t* I = images + (size_t)b * npix * 3;
        if (tid == 0) { bbox[0] = W; bbox[1] = H; bbox[2] = -1; bbox[3] = -1; ndrawn = 0; }
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
        // ---- faces: cull, shade, depth (:50-77); the drawn ones are appended to the compact list
        for (int f0 = wbase; f0 < F; f0 += RT) {
            const int f = f0 + lane;
            bool draw = false;
            unsigned int tag = 0;
            unsigned long long dbits = 0;
            if (f < F) {
                const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
                if (pts[i0].x > INT_MIN + 1 && pts[i1].x > INT_MIN + 1 && pts[i2].x > INT_MIN + 1) {
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
                        tag = ((unsigned int)f << 8) | (unsigned int)(int)round(dmul(220.0, shade));
                        dbits = (unsigned long long)__double_as_longlong(g2);   // mean Z > 0: the bit pattern is order preserving
                        draw = true;
                    }
                }
            }
            const unsigned int vote = __ballot_sync(0xffffffffu, draw);
            int base = 0;
            if (lane == 0 && vote) base = atomicAdd(&ndrawn, __popc(vote));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (draw) {
                const int k = base + __popc(vote & ((1u << lane) - 1u));
                cdepth[k] = dbits;
                ctag[k] = tag;
            }
        }
        __syncthreads();
        const int nd = ndrawn;
        // ---- pass A: depth plane
        int bx0 = W, by0 = H, bx1 = -1, by1 = -1;
        for (int k0 = wbase; k0 < nd; k0 += RT) {
            const int k = k0 + lane;
            const bool alive = k < nd;
            const unsigned int tag = alive ? ctag[k] : 0u;
            const unsigned long long dbits = alive ? cdepth[k] : 0ull;
            const int f = (int)(tag >> 8);
            const int2 q0 = pts[faces[3 * f]], q1 = pts[faces[3 * f + 1]], q2 = pts[faces[3 * f + 2]];
            const P2 tri[3] = {{q0.x, q0.y}, {q1.x, q1.y}, {q2.x, q2.y}};
            auto put = [&](int x, int y) {
                atomicMin(&Z[(size_t)y * W + x], dbits);
                bx0 = min(bx0, x); bx1 = max(bx1, x); by0 = min(by0, y); by1 = max(by1, y);
            };
            walk_triangle(W, H, alive, tri, put);
        }
        if (bx1 >= 0) { atomicMin(&bbox[0], bx0); atomicMin(&bbox[1], by0); atomicMax(&bbox[2], bx1); atomicMax(&bbox[3], by1); }
        __threadfence_block();
        __syncthreads();
        // ---- pass B: the nearest face of a pixel claims it (ties: highest face index = drawn last by a stable sort)
        for (int k0 = wbase; k0 < nd; k0 += RT) {
            const int k = k0 + lane;
            const bool alive = k < nd;
            const unsigned int tag = alive ? ctag[k] : 0u;
            const unsigned long long dbits = alive ? cdepth[k] : 0ull;
            const int f = (int)(tag >> 8);
            const int2 q0 = pts[faces[3 * f]], q1 = pts[faces[3 * f + 1]], q2 = pts[faces[3 * f + 2]];
            const P2 tri[3] = {{q0.x, q0.y}, {q1.x, q1.y}, {q2.x, q2.y}};
            const unsigned int claim = tag + 256u;          // (face + 1) << 8 | colour: 0 stays "no face"
            auto put = [&](int x, int y) {
                const size_t p = (size_t)y * W + x;
                if (__ldcg(&Z[p]) == dbits) atomicMax(&Wn[p], claim);
            };
            walk_triangle(W, H, alive, tri, put);
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

static size_t render_smem_bytes(int V, int F) { return (size_t)((V + 1) & ~1) * sizeof(int2) + (size_t)F * 12 + 16; }

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
    static std::atomic<int> pool_kept{-1};      // keep the default pool's memory across calls (the planes are ~1.5 MB per SM)
    if (pool_kept.exchange(dev) != dev) {
        cudaMemPool_t pool;
        unsigned long long keep = ~0ull;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
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
