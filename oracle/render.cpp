// CPU ORACLE — TEST INFRASTRUCTURE ONLY.  Overlay renderer.
//
// Restates smpl::render::renderSMPLMesh (include/RenderSMPLMesh.h:16-110 of the reference): project, backface cull, flat shade
// 220 * clamp(n^ . v^), painter's order far to near, one cv::fillConvexPoly per face.  OpenCV is a third-party dependency of the
// reference that is not in its tree (CMake find_package(OpenCV), no pinned version); the polygon fill below restates the
// published algorithm of OpenCV's modules/imgproc/src/drawing.cpp (clipLine, LineIterator with leftToRight, Line,
// FillConvexPoly) for line_type = LINE_8, shift = 0, and is PINNED against the real library: tests/test_render_cpu.py compares it
// with cv2.fillConvexPoly (OpenCV 4.13 in this image) on random triangles in and across the image borders and on whole painter
// loops, and tests/golden/render_*.npz holds images produced by cv2 with the generating script (tools/make_render_golden.py).
//
// Deviation from the reference, stated once: the reference passes cv::LINE_AA.  Anti-aliased edges are read-modify-write blends
// whose result depends on everything drawn before them, face after face; the restatement (and the CUDA rasteriser it checks)
// uses LINE_8 coverage, i.e. the same triangles, order, colours and fill rule without the edge blend.  Ties in the painter's
// sort (std::sort is unstable, the reference leaves them unspecified) are broken by face index: the higher index is drawn later.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

struct Pt { int64_t x, y; };

// drawing.cpp clipLine(Size2l, Point2l&, Point2l&)
bool clip_line(int64_t width, int64_t height, Pt& p1, Pt& p2) {
    if (width <= 0 || height <= 0) return false;
    const int64_t right = width - 1, bottom = height - 1;
    int64_t &x1 = p1.x, &y1 = p1.y, &x2 = p2.x, &y2 = p2.y;
    int c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8;
    int c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8;
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
        int64_t a;
        if (c1 & 12) {
            a = c1 < 8 ? 0 : bottom;
            x1 += (int64_t)((double)(a - y1) * (x2 - x1) / (y2 - y1));
            y1 = a;
            c1 = (x1 < 0) + (x1 > right) * 2;
        }
        if (c2 & 12) {
            a = c2 < 8 ? 0 : bottom;
            x2 += (int64_t)((double)(a - y2) * (x2 - x1) / (y2 - y1));
            y2 = a;
            c2 = (x2 < 0) + (x2 > right) * 2;
        }
        if ((c1 & c2) == 0 && (c1 | c2) != 0) {
            if (c1) {
                a = c1 == 1 ? 0 : right;
                y1 += (int64_t)((double)(a - x1) * (y2 - y1) / (x2 - x1));
                x1 = a;
                c1 = 0;
            }
            if (c2) {
                a = c2 == 1 ? 0 : right;
                y2 += (int64_t)((double)(a - x2) * (y2 - y1) / (x2 - x1));
                x2 = a;
                c2 = 0;
            }
        }
    }
    return (c1 | c2) == 0;
}

template <class Put>
void line8(int W, int H, Pt p1, Pt p2, Put put) {   // Line() -> LineIterator(img, pt1, pt2, 8, leftToRight = true)
    if ((uint64_t)p1.x >= (uint64_t)W || (uint64_t)p2.x >= (uint64_t)W || (uint64_t)p1.y >= (uint64_t)H || (uint64_t)p2.y >= (uint64_t)H)
        if (!clip_line(W, H, p1, p2)) return;
    int dx = (int)(p2.x - p1.x), dy = (int)(p2.y - p1.y);
    int sx = 1, sy = 1;
    if (dx < 0) { dx = -dx; dy = -dy; p1 = p2; }   // leftToRight
    if (dy < 0) { dy = -dy; sy = -1; }
    const bool vert = dy > dx;
    if (vert) std::swap(dx, dy);
    int err = dx - (dy + dy);
    const int plus = dx + dx, minus = -(dy + dy);
    int x = (int)p1.x, y = (int)p1.y;
    for (int i = 0; i <= dx; ++i) {
        put(x, y);
        const bool m = err < 0;
        err += minus + (m ? plus : 0);
        if (vert) { y += sy; if (m) x += sx; }
        else { x += sx; if (m) y += sy; }
    }
}

// FillConvexPoly(img, v, 3, color, LINE_8, shift = 0)
template <class Put>
void fill_tri(int W, int H, const Pt v[3], Put put) {
    constexpr int XY_SHIFT = 16;
    constexpr int64_t XY_ONE = 1 << XY_SHIFT;
    const int npts = 3;
    struct { int idx, di; int64_t x, dx; int ye; } edge[2];
    int imin = 0, edges = npts;
    int64_t xmin = v[0].x, xmax = v[0].x, ymin = v[0].y, ymax = v[0].y;
    Pt p0 = v[npts - 1];
    for (int i = 0; i < npts; ++i) {
        const Pt p = v[i];
        if (p.y < ymin) { ymin = p.y; imin = i; }
        ymax = std::max(ymax, p.y);
        xmax = std::max(xmax, p.x);
        xmin = std::min(xmin, p.x);
        line8(W, H, p0, p, put);
        p0 = p;
    }
    if ((int)xmax < 0 || (int)ymax < 0 || (int)xmin >= W || (int)ymin >= H) return;
    ymax = std::min<int64_t>(ymax, H - 1);
    edge[0].idx = edge[1].idx = imin;
    int y = (int)ymin;
    edge[0].ye = edge[1].ye = y;
    edge[0].di = 1;
    edge[1].di = npts - 1;
    edge[0].x = edge[1].x = -XY_ONE;
    edge[0].dx = edge[1].dx = 0;
    do {
        for (int i = 0; i < 2; ++i) {
            if (y >= edge[i].ye) {
                int idx0 = edge[i].idx;
                const int di = edge[i].di;
                int idx = idx0 + di;
                if (idx >= npts) idx -= npts;
                for (; edges-- > 0;) {
                    const int ty = (int)v[idx].y;
                    if (ty > y) {
                        const int64_t xs = v[idx0].x << XY_SHIFT, xe = v[idx].x << XY_SHIFT;
                        edge[i].ye = ty;
                        edge[i].dx = ((xe - xs) * 2 + (ty - y)) / (2 * (ty - y));
                        edge[i].x = xs;
                        edge[i].idx = idx;
                        break;
                    }
                    idx0 = idx;
                    idx += di;
                    if (idx >= npts) idx -= npts;
                }
            }
        }
        if (edges < 0) break;
        if (y >= 0) {
            int left = 0, right = 1;
            if (edge[0].x > edge[1].x) { left = 1; right = 0; }
            int xx1 = (int)((edge[left].x + (XY_ONE >> 1)) >> XY_SHIFT);
            int xx2 = (int)((edge[right].x + (XY_ONE >> 1)) >> XY_SHIFT);
            if (xx2 >= 0 && xx1 < W) {
                if (xx1 < 0) xx1 = 0;
                if (xx2 >= W) xx2 = W - 1;
                for (int x = xx1; x <= xx2; ++x) put(x, y);
            }
        }
        edge[0].x += edge[0].dx;
        edge[1].x += edge[1].dx;
    } while (++y <= (int)ymax);
}

struct FaceItem { int face; double depth; Pt pts[3]; int color; };

}  // namespace

extern "C" {

// One triangle with integer vertices into a W x H x ch uint8 image (cv2.fillConvexPoly(img, pts, (c, c, c), LINE_8)).
void orc_fill_triangle(uint8_t* img, int W, int H, int ch, const int32_t* xy, int color) {
    const Pt v[3] = {{xy[0], xy[1]}, {xy[2], xy[3]}, {xy[4], xy[5]}};
    fill_tri(W, H, v, [&](int x, int y) { for (int c = 0; c < ch; ++c) img[((size_t)y * W + x) * ch + c] = (uint8_t)color; });
}

// renderSMPLMesh(cloud, faces, img, fx, fy, cx, cy, fill = true, backface_cull, wireframe = false) on B bodies.
//   cloud  B x V x 3 doubles (camera coordinates), faces F x 3, img B x H x W x 3 uint8 (in/out).
//   drawn (optional, B): faces that reached the draw list.  order (optional, B x F): the draw list (face ids, far to near), -1 padded.
void orc_render_overlay(int B, int V, int F, const double* cloud, const int32_t* faces, double fx, double fy, double cx, double cy,
                        int W, int H, int backface_cull, uint8_t* img, int32_t* drawn, int32_t* order) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        const double* P = cloud + (size_t)b * V * 3;
        uint8_t* I = img + (size_t)b * H * W * 3;
        std::vector<float> pu(V, -9999.f), pv(V, -9999.f);
        std::vector<char> valid(V, 0);
        for (int i = 0; i < V; ++i) {
            const double X = P[3 * i], Y = P[3 * i + 1], Z = P[3 * i + 2];
            if (Z <= 1e-6) continue;
            pu[i] = (float)(fx * X / Z + cx);
            pv[i] = (float)(fy * Y / Z + cy);
            valid[i] = 1;
        }
        std::vector<FaceItem> list;
        list.reserve(F);
        for (int f = 0; f < F; ++f) {
            const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
            if (!valid[i0] || !valid[i1] || !valid[i2]) continue;
            const double* a = P + 3 * i0; const double* bb = P + 3 * i1; const double* c = P + 3 * i2;
            const double e1[3] = {bb[0] - a[0], bb[1] - a[1], bb[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
            const double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            if (backface_cull && n[2] >= 0.0) continue;
            const double ce[3] = {(a[0] + bb[0] + c[0]) / 3.0, (a[1] + bb[1] + c[1]) / 3.0, (a[2] + bb[2] + c[2]) / 3.0};
            // Eigen normalized(): v / sqrt(squaredNorm) when the squared norm is > 0, else v itself
            const double cn2 = ce[0] * ce[0] + ce[1] * ce[1] + ce[2] * ce[2], nn2 = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];
            const double cs = cn2 > 0.0 ? std::sqrt(cn2) : 1.0, ns = nn2 > 0.0 ? std::sqrt(nn2) : 1.0;
            const double view[3] = {-ce[0] / cs, -ce[1] / cs, -ce[2] / cs}, nh[3] = {n[0] / ns, n[1] / ns, n[2] / ns};
            double shade = nh[0] * view[0] + nh[1] * view[1] + nh[2] * view[2];
            shade = shade < 0.0 ? 0.0 : (shade > 1.0 ? 1.0 : shade);
            FaceItem it;
            it.face = f;
            it.depth = (a[2] + bb[2] + c[2]) / 3.0;
            it.color = (int)std::round(220 * shade);
            const int idx[3] = {i0, i1, i2};
            bool sane = true;   // beyond +-2^20 px cv::Point overflows / FillConvexPoly walks millions of rows: such faces are dropped
            for (int k = 0; k < 3; ++k) {
                const float ru = std::round(pu[idx[k]]), rv = std::round(pv[idx[k]]);
                if (!(std::fabs(ru) <= 1048576.f && std::fabs(rv) <= 1048576.f)) { sane = false; break; }
                it.pts[k] = {(int64_t)(int)ru, (int64_t)(int)rv};
            }
            if (sane) list.push_back(it);
        }
        std::stable_sort(list.begin(), list.end(), [](const FaceItem& x, const FaceItem& y) { return x.depth > y.depth; });
        if (drawn) drawn[b] = (int)list.size();
        if (order) {
            for (int f = 0; f < F; ++f) order[(size_t)b * F + f] = f < (int)list.size() ? list[f].face : -1;
        }
        for (const FaceItem& it : list)
            fill_tri(W, H, it.pts, [&](int x, int y) { uint8_t* p = I + ((size_t)y * W + x) * 3; p[0] = p[1] = p[2] = (uint8_t)it.color; });
    }
}

}  // extern "C"
