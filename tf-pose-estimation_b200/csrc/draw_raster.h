// draw_raster.h -- integer rasterisation rules of the two primitives TfPoseEstimator.draw_humans uses
// (tf_pose/estimator.py:391-415): cv2.circle(center, 3, color, thickness=3, lineType=8) and
// cv2.line(p0, p1, color, 3).  The reference delegates both to OpenCV (third-party, not vendored; any 3.x/4.x:
// drawing.cpp has been stable there); this is a restatement of OpenCV's published algorithm for exactly these
// calls -- ThickLine = FillConvexPoly of the offset quad (scan conversion with 16-bit fixed-point edges, outline by
// the fixed-point line stepper) + filled end caps; a thick circle = ThickLine over the 12-gon ellipse2Poly yields
// for radius 3 -- so that the pixels match cv2's bit for bit (tests/test_draw.py compares with cv2 itself).
//
// Everything is a template over an Emit functor `void(int x, int y)` called once or more per covered pixel
// (inside the image only), so the same code runs in the CUDA kernel (csrc/k_draw.cu) and in the host-side
// checker build that the CPU tests compare with cv2.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TFPD_HD __host__ __device__ __forceinline__
#else
#define TFPD_HD inline
#endif

namespace tfpd_draw {

constexpr int XY_SHIFT = 16;
constexpr int64_t XY_ONE = 1 << XY_SHIFT;

struct P64 {
    int64_t x, y;
};

TFPD_HD int64_t round_half_even(double v) {
    // cvRound: nearest integer, ties to even (lrint under the default rounding mode)
    double f = v < 0 ? -v : v;
    int64_t i = (int64_t)f;
    double d = f - (double)i;
    if (d > 0.5 || (d == 0.5 && (i & 1))) ++i;
    return v < 0 ? -i : i;
}

TFPD_HD double sqrt_d(double v) {
#if defined(__CUDA_ARCH__)
    return sqrt(v);
#else
    return __builtin_sqrt(v);
#endif
}

// cv::clipLine(Size2l, Point2l&, Point2l&): Cohen-Sutherland with the intersection computed in double
TFPD_HD bool clip_line(int64_t width, int64_t height, P64 &pt1, P64 &pt2) {
    if (width <= 0 || height <= 0) return false;
    const int64_t right = width - 1, bottom = height - 1;
    int64_t &x1 = pt1.x, &y1 = pt1.y, &x2 = pt2.x, &y2 = pt2.y;
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

// Line2: 8-connected line between two fixed-point end points (the polygon outline FillConvexPoly draws)
template <class Emit>
TFPD_HD void line2(int W, int H, P64 pt1, P64 pt2, Emit &emit) {
    if (!clip_line((int64_t)W << XY_SHIFT, (int64_t)H << XY_SHIFT, pt1, pt2)) return;
    int64_t dx = pt2.x - pt1.x, dy = pt2.y - pt1.y;
    int64_t j = dx < 0 ? -1 : 0, ax = (dx ^ j) - j;
    int64_t i = dy < 0 ? -1 : 0, ay = (dy ^ i) - i;
    int64_t x_step, y_step;
    int ecount;
    if (ax > ay) {
        dy = (dy ^ j) - j;
        pt1.x ^= pt2.x & j;
        pt2.x ^= pt1.x & j;
        pt1.x ^= pt2.x & j;
        pt1.y ^= pt2.y & j;
        pt2.y ^= pt1.y & j;
        pt1.y ^= pt2.y & j;
        x_step = XY_ONE;
        y_step = (dy << XY_SHIFT) / (ax | 1);
        ecount = (int)((pt2.x - pt1.x) >> XY_SHIFT);
    } else {
        dx = (dx ^ i) - i;
        pt1.x ^= pt2.x & i;
        pt2.x ^= pt1.x & i;
        pt1.x ^= pt2.x & i;
        pt1.y ^= pt2.y & i;
        pt2.y ^= pt1.y & i;
        pt1.y ^= pt2.y & i;
        x_step = (dx << XY_SHIFT) / (ay | 1);
        y_step = XY_ONE;
        ecount = (int)((pt2.y - pt1.y) >> XY_SHIFT);
    }
    pt1.x += (XY_ONE >> 1);
    pt1.y += (XY_ONE >> 1);
    auto put = [&](int64_t x, int64_t y) {
        if (0 <= x && x < W && 0 <= y && y < H) emit((int)x, (int)y);
    };
    put((pt2.x + (XY_ONE >> 1)) >> XY_SHIFT, (pt2.y + (XY_ONE >> 1)) >> XY_SHIFT);
    while (ecount >= 0) {
        put(pt1.x >> XY_SHIFT, pt1.y >> XY_SHIFT);
        pt1.x += x_step;
        pt1.y += y_step;
        ecount--;
    }
}

// FillConvexPoly(v[npts], line_type = 8, shift = XY_SHIFT): outline with line2, then scan conversion with
// two active edges whose x advances by a rounded per-scanline increment
template <class Emit>
TFPD_HD void fill_convex_poly(int W, int H, const P64 *v, int npts, Emit &emit) {
    const int shift = XY_SHIFT;
    struct Edge {
        int idx, di;
        int64_t x, dx;
        int ye;
    } edge[2];
    const int64_t delta = (int64_t)1 << shift >> 1;
    const int64_t delta1 = XY_ONE >> 1, delta2 = XY_ONE >> 1;
    int i, y, imin = 0, edges = npts;
    int64_t xmin, xmax, ymin, ymax;
    P64 p0 = v[npts - 1];
    xmin = xmax = v[0].x;
    ymin = ymax = v[0].y;
    for (i = 0; i < npts; i++) {
        P64 p = v[i];
        if (p.y < ymin) {
            ymin = p.y;
            imin = i;
        }
        ymax = ymax > p.y ? ymax : p.y;
        xmax = xmax > p.x ? xmax : p.x;
        xmin = xmin < p.x ? xmin : p.x;
        line2(W, H, p0, p, emit);
        p0 = p;
    }
    xmin = (xmin + delta) >> shift;
    xmax = (xmax + delta) >> shift;
    ymin = (ymin + delta) >> shift;
    ymax = (ymax + delta) >> shift;
    if (npts < 3 || (int)xmax < 0 || (int)ymax < 0 || (int)xmin >= W || (int)ymin >= H) return;
    ymax = ymax < H - 1 ? ymax : H - 1;
    edge[0].idx = edge[1].idx = imin;
    edge[0].ye = edge[1].ye = y = (int)ymin;
    edge[0].di = 1;
    edge[1].di = npts - 1;
    edge[0].x = edge[1].x = -XY_ONE;
    edge[0].dx = edge[1].dx = 0;
    do {
        for (i = 0; i < 2; i++) {
            if (y >= edge[i].ye) {
                int idx0 = edge[i].idx, di = edge[i].di;
                int idx = idx0 + di;
                if (idx >= npts) idx -= npts;
                int ty = 0;
                for (; edges-- > 0;) {
                    ty = (int)((v[idx].y + delta) >> shift);
                    if (ty > y) {
                        const int64_t xs = v[idx0].x, xe = v[idx].x;
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
            if (edge[0].x > edge[1].x) {
                left = 1;
                right = 0;
            }
            int xx1 = (int)((edge[left].x + delta1) >> XY_SHIFT);
            int xx2 = (int)((edge[right].x + delta2) >> XY_SHIFT);
            if (xx2 >= 0 && xx1 < W) {
                if (xx1 < 0) xx1 = 0;
                if (xx2 >= W) xx2 = W - 1;
                for (int x = xx1; x <= xx2; ++x) emit(x, y);
            }
        }
        edge[0].x += edge[0].dx;
        edge[1].x += edge[1].dx;
    } while (++y <= (int)ymax);
}

// Circle(center, radius, fill = 1): midpoint circle, one or two spans per step, clipped to the image
template <class Emit>
TFPD_HD void fill_circle(int W, int H, int cx, int cy, int radius, Emit &emit) {
    int err = 0, dx = radius, dy = 0, plus = 1, minus = (radius << 1) - 1;
    auto hline = [&](int x1, int x2, int y) {
        if (y < 0 || y >= H) return;
        if (x1 < 0) x1 = 0;
        if (x2 >= W) x2 = W - 1;
        for (int x = x1; x <= x2; ++x) emit(x, y);
    };
    while (dx >= dy) {
        const int y11 = cy - dy, y12 = cy + dy, y21 = cy - dx, y22 = cy + dx;
        const int x11 = cx - dx, x12 = cx + dx, x21 = cx - dy, x22 = cx + dy;
        if (x11 < W && x12 >= 0 && y21 < H && y22 >= 0) {
            hline(x11, x12, y11);
            hline(x11, x12, y12);
            hline(x21, x22, y21);
            hline(x21, x22, y22);
        }
        dy++;
        err += plus;
        plus += 2;
        const int mask = (err <= 0) - 1;
        err -= minus & mask;
        dx += mask;
        minus -= mask & 2;
    }
}

// ThickLine(p0, p1 in XY_SHIFT fixed point, thickness, line_type = 8, flags): offset quad + end caps
template <class Emit>
TFPD_HD void thick_line(int W, int H, P64 p0, P64 p1, int thickness, int flags, Emit &emit) {
    const double INV_XY_ONE = 1.0 / (double)XY_ONE;
    P64 pt[4], dp = {0, 0};
    const double dx = (double)(p0.x - p1.x) * INV_XY_ONE, dy = (double)(p1.y - p0.y) * INV_XY_ONE;
    double r = dx * dx + dy * dy;
    const int odd = thickness & 1;
    const int64_t th = (int64_t)thickness << (XY_SHIFT - 1);
    if ((r < 0 ? -r : r) > 2.2204460492503131e-16) {
        r = ((double)th + odd * (double)XY_ONE * 0.5) / sqrt_d(r);
        dp.x = round_half_even(dy * r);
        dp.y = round_half_even(dx * r);
        pt[0].x = p0.x + dp.x;
        pt[0].y = p0.y + dp.y;
        pt[1].x = p0.x - dp.x;
        pt[1].y = p0.y - dp.y;
        pt[2].x = p1.x - dp.x;
        pt[2].y = p1.y - dp.y;
        pt[3].x = p1.x + dp.x;
        pt[3].y = p1.y + dp.y;
        fill_convex_poly(W, H, pt, 4, emit);
    }
    for (int i = 0; i < 2; i++) {
        if (flags & (i + 1)) {
            const int cx = (int)((p0.x + (XY_ONE >> 1)) >> XY_SHIFT), cy = (int)((p0.y + (XY_ONE >> 1)) >> XY_SHIFT);
            fill_circle(W, H, cx, cy, (int)((th + (XY_ONE >> 1)) >> XY_SHIFT), emit);
        }
        p0 = p1;
    }
}

// cv2.line(img, (x0,y0), (x1,y1), color, thickness >= 2): cv::line first clips the segment to the image
// rectangle grown by `thickness` on every side (integer clipLine), then ThickLine with both caps
template <class Emit>
TFPD_HD void cv_line(int W, int H, int x0, int y0, int x1, int y1, int thickness, Emit &emit) {
    P64 a = {(int64_t)x0 + thickness, (int64_t)y0 + thickness}, b = {(int64_t)x1 + thickness, (int64_t)y1 + thickness};
    if (!clip_line((int64_t)W + 2 * thickness, (int64_t)H + 2 * thickness, a, b)) return;
    a.x = (a.x - thickness) << XY_SHIFT;
    a.y = (a.y - thickness) << XY_SHIFT;
    b.x = (b.x - thickness) << XY_SHIFT;
    b.y = (b.y - thickness) << XY_SHIFT;
    thick_line(W, H, a, b, thickness, 3, emit);
}

// cv2.circle(img, (cx,cy), 3, color, thickness=3, lineType=8, shift=0): EllipseEx with delta = 30 degrees ->
// the open polyline over 13 vertices center + 3*65536*(cos, sin)(30k degrees), rounded; segment s joins
// vertex s and s+1 (caps at both ends of segment 0, at the far end of the others: PolyLine's flags)
TFPD_HD void circle3_vertex(int cx, int cy, int k, P64 &p) {
    // 196608 * cos / sin of multiples of 30 degrees, cvRound'ed (OpenCV's SinTable holds the float sines)
    const int64_t c[12] = {196608, 170268, 98304, 0, -98304, -170268, -196608, -170268, -98304, 0, 98304, 170268};
    const int kk = k % 12;
    p.x = ((int64_t)cx << XY_SHIFT) + c[kk];
    p.y = ((int64_t)cy << XY_SHIFT) + c[(kk + 9) % 12];        // sin(a) = cos(a - 90)
}
template <class Emit>
TFPD_HD void cv_circle3_segment(int W, int H, int cx, int cy, int seg, Emit &emit) {
    P64 a, b;
    circle3_vertex(cx, cy, seg, a);
    circle3_vertex(cx, cy, seg + 1, b);
    thick_line(W, H, a, b, 3, seg == 0 ? 3 : 2, emit);
}

}  // namespace tfpd_draw
