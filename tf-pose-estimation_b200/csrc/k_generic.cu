// k_generic.cu -- the general-size map path: any (h,w) -> (H,W), any odd smoother size.
// Used when the upsample is not the x4 replication handled by k1_fused.cu (e.g.
// resize_to_default=False, eval.py's ratio 8.0, ratio 1.0; tf_pose/estimator.py:522-525).
// Plain, coalesced, one-output-element-per-thread kernels; correctness first, the x4 fast
// path is the one the roofline is reported on.
//
//   k_resize_area  tf.image.resize_area(align_corners=False)   estimator.py:326-329, SURVEY.md A.1
//   k_blur_h/_v    Smoother depthwise conv SAME, separable      smoother.py:62-95
//   k_nms          max-pool 3x3 SAME + equal + where            estimator.py:335-337
//                  + the peak scan of pafprocess.cpp:24-37
#include <cstdlib>

#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

__device__ __forceinline__ float area_wt(int i, float a, float b, float s) {
    if ((float)i < a) return ((float)(i + 1) > b) ? s : __fsub_rn((float)(i + 1), a);
    return ((float)(i + 1) > b) ? __fsub_rn(b, (float)i) : 1.0f;
}

// out[n,Y,X,c] for channels [c_off, c_off+C) of the 57-channel input
__global__ void k_resize_area(const float *__restrict__ X, int N, int ih, int iw, int c_off, int C,
                              float *__restrict__ out, int oh, int ow) {
    const float hs = __fdiv_rn((float)ih, (float)oh), ws = __fdiv_rn((float)iw, (float)ow);
    const float scale = __fdiv_rn(1.0f, __fmul_rn(hs, ws));
    const size_t total = (size_t)N * oh * ow * C;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(idx % C);
        size_t pix = idx / C;
        const int x = (int)(pix % ow);
        pix /= ow;
        const int y = (int)(pix % oh);
        const int n = (int)(pix / oh);
        const float ya = __fmul_rn((float)y, hs), yb = __fmul_rn((float)(y + 1), hs);
        const float xa = __fmul_rn((float)x, ws), xb = __fmul_rn((float)(x + 1), ws);
        const int y0 = (int)floorf(ya), y1 = (int)ceilf(yb);
        const int x0 = (int)floorf(xa), x1 = (int)ceilf(xb);
        const float *img = X + (size_t)n * ih * iw * kNetC + c_off + c;
        float sum = 0.0f;
        for (int iy = y0; iy < y1; ++iy) {
            const float wy = area_wt(iy, ya, yb, hs);
            const float *row = img + (size_t)min(max(iy, 0), ih - 1) * iw * kNetC;
            float sx = 0.0f;
            for (int ix = x0; ix < x1; ++ix) {
                const float wx = area_wt(ix, xa, xb, ws);
                sx = __fadd_rn(sx, __fmul_rn(__ldg(row + (size_t)min(max(ix, 0), iw - 1) * kNetC), wx));
            }
            sum = __fadd_rn(sum, __fmul_rn(sx, wy));
        }
        out[idx] = __fmul_rn(sum, scale);
    }
}

// 1-D pass of the separable smoother over axis `stride_px` (1 = along X, W = along Y)
__global__ void k_blur_1d(const float *__restrict__ in, float *__restrict__ out, int N, int H, int W, int C,
                          const float *__restrict__ k1d, int ks, int along_y) {
    const int pad = ks / 2;
    const size_t total = (size_t)N * H * W * C;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        size_t pix = idx / C;
        const int x = (int)(pix % W);
        const int y = (int)((pix / W) % H);
        float acc = 0.0f;
        if (along_y) {
            for (int r = 0; r < ks; ++r) {
                int yy = y + r - pad;
                if (yy >= 0 && yy < H) acc = fmaf(__ldg(k1d + r), __ldg(in + idx + (ptrdiff_t)(r - pad) * W * C), acc);
            }
        } else {
            for (int q = 0; q < ks; ++q) {
                int xx = x + q - pad;
                if (xx >= 0 && xx < W) acc = fmaf(__ldg(k1d + q), __ldg(in + idx + (ptrdiff_t)(q - pad) * C), acc);
            }
        }
        out[idx] = acc;
    }
}

// heat score of a peak (pafprocess.cpp:32, HEAT(y,x,part)): read from `heat` = the dense heat_up tensor
// (hshift 0, hw = W, hc = 19) or, when the upsample is a power-of-two replication, from the low-res network
// output (hshift = log2 factor, hw = w, hc = 57; heat_up[y,x] == X[y>>hshift, x>>hshift] bit-exactly)
__global__ void k_nms(const float *__restrict__ g, const float *__restrict__ heat, int hshift, int hw, int hc,
                      size_t heat_frame_stride, float *__restrict__ peaks, int N, int H, int W, float thr, Workspace ws) {
    const int C = kHeatC;
    const size_t total = (size_t)N * H * W * C;
    const float NEG = -__int_as_float(0x7f800000);
    const float thr_lo = thr * (1.0f - 4e-5f);
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(idx % C);
        size_t pix = idx / C;
        const int x = (int)(pix % W);
        const int y = (int)((pix / W) % H);
        const int n = (int)(pix / ((size_t)W * H));
        const float v = g[idx];
        float m = NEG;
        for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
                int yy = y + dy, xx = x + dx;
                if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
                m = fmaxf(m, __ldg(g + idx + ((ptrdiff_t)dy * W + dx) * C));
            }
        const bool is_peak = (v == m);
        if (peaks) peaks[idx] = is_peak ? v : 0.0f;
        if (c < kParts) {
            if (is_peak) {
                if (v > thr) {
                    int k = atomicAdd(&ws.counters[n].n_peak_rec, 1);
                    if (k < ws.max_peaks) {
                        ws.rec_key[(size_t)n * ws.max_peaks + k] = peak_key(c, y, x);
                        ws.rec_val[(size_t)n * ws.max_peaks + k] =
                            heat[(size_t)n * heat_frame_stride + ((size_t)(y >> hshift) * hw + (x >> hshift)) * hc + c];
                    }
                }
                if (fabsf(v - thr) <= 2e-5f * thr) atomicAdd(&ws.counters[n].near_ties, 1);
            } else if (v > thr_lo && (m - v) <= 2e-5f * fabsf(m)) {
                atomicAdd(&ws.counters[n].near_ties, 1);
            }
        }
    }
}

// hb[n, iy, X, c] = sum_t wp[X % F][t] * heat[n, iy, (X >> s) + t - R, c]   (cells outside the row are zero:
// SAME padding of the upsampled row, whose edges coincide with cell edges)
__global__ void k_hblur_lowres(const float *__restrict__ X, int N, int h, int w, int s, const float *__restrict__ wp,
                               int T, float *__restrict__ hb) {
    const int W = w << s, F = 1 << s, R = T >> 1;
    const size_t total = (size_t)N * h * W * kHeatC;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(idx % kHeatC);
        size_t pix = idx / kHeatC;
        const int x = (int)(pix % W);
        const size_t row = pix / W;                       // n * h + iy
        const int cx = x >> s;
        const float *wrow = wp + (x & (F - 1)) * T;
        const float *src = X + row * w * kNetC + c;
        float acc = 0.0f;
        for (int t = 0; t < T; ++t) {
            const int cc = cx + t - R;
            if (cc >= 0 && cc < w) acc = fmaf(__ldg(wrow + t), __ldg(src + (size_t)cc * kNetC), acc);
        }
        hb[idx] = acc;
    }
}

// g[n, Y, X, c] = sum_t wp[Y % F][t] * hb[n, (Y >> s) + t - R, X, c]
__global__ void k_vblur_poly(const float *__restrict__ hb, int N, int h, int W, int s, const float *__restrict__ wp,
                             int T, float *__restrict__ g) {
    const int H = h << s, F = 1 << s, R = T >> 1;
    const size_t rowlen = (size_t)W * kHeatC;
    const size_t total = (size_t)N * H * rowlen;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t col = idx % rowlen;
        const size_t r = idx / rowlen;
        const int y = (int)(r % H);
        const int n = (int)(r / H);
        const int iy = y >> s;
        const float *wrow = wp + (y & (F - 1)) * T;
        const float *src = hb + (size_t)n * h * rowlen + col;
        float acc = 0.0f;
        for (int t = 0; t < T; ++t) {
            const int rr = iy + t - R;
            if (rr >= 0 && rr < h) acc = fmaf(__ldg(wrow + t), __ldg(src + (size_t)rr * rowlen), acc);
        }
        g[idx] = acc;
    }
}

// Dense upsample of a power-of-two replication: a CTA stages (a quarter of) one low-res row of the network output
// in shared memory and writes the F output rows it replicates to, 16 bytes per store; each float4 is composed once
// and stored F times.  Requires (W*19) % 4 == 0 so that every float4 stays inside one output row.
constexpr int kRepSplit = 4;
template <int C, int C_OFF>
__device__ __forceinline__ void replicate_rows(const float *s_row, float *out_row0, int row4, int F, int s,
                                               size_t rowlen) {
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < row4; q += kRepSplit * blockDim.x) {
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int e = 4 * q + k, x = e / C, c = e - x * C;
            v[k] = s_row[(x >> s) * kNetC + C_OFF + c];
        }
        const float4 o = make_float4(v[0], v[1], v[2], v[3]);
        float4 *dst = reinterpret_cast<float4 *>(out_row0) + q;
        for (int r = 0; r < F; ++r) __stcs(dst + (size_t)r * (rowlen >> 2), o);
    }
}

__global__ void __launch_bounds__(256) k_replicate_pow2(const float *__restrict__ X, int h, int w, int s,
                                                         float *__restrict__ heat_up, float *__restrict__ paf_up) {
    extern __shared__ float s_row[];                       // [w * 57]
    const int n = blockIdx.z, iy = blockIdx.y, F = 1 << s, W = w << s;
    const float *src = X + ((size_t)n * h + iy) * w * kNetC;
    for (int i = threadIdx.x; i < w * kNetC; i += blockDim.x) s_row[i] = __ldg(src + i);
    __syncthreads();
    const size_t row0 = ((size_t)n * h + iy) * F;          // first output row of this cell row
    const size_t hl = (size_t)W * kHeatC, pl = (size_t)W * kPafC;
    replicate_rows<kHeatC, 0>(s_row, heat_up + row0 * hl, (int)(hl >> 2), F, s, hl);
    replicate_rows<kPafC, kHeatC>(s_row, paf_up + row0 * pl, (int)(pl >> 2), F, s, pl);
}

// same replication one element per thread, for row lengths that are not a multiple of four floats
__global__ void k_replicate_scalar(const float *__restrict__ X, int N, int h, int w, int s, int c_off, int C,
                                   float *__restrict__ out) {
    const int H = h << s, W = w << s;
    const size_t total = (size_t)N * H * W * C;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(idx % C);
        size_t pix = idx / C;
        const int x = (int)(pix % W);
        pix /= W;
        const int y = (int)(pix % H);
        const size_t n = pix / H;
        out[idx] = __ldg(X + ((n * h + (y >> s)) * w + (x >> s)) * kNetC + c_off + c);
    }
}

// Smoother (both polyphase passes) + 3x3 NMS + peak records in one kernel: a CTA owns kTileX upsampled columns and a
// strip of low-res rows of one frame and walks down the strip.  One thread per (column, channel):
//   * per low-res row it forms ONE horizontally smoothed value from T cells of the network output (its x-phase
//     weights sit in registers) and keeps the last T+2 of them in registers (sliding window);
//   * the F + 2 smoothed output rows around the current cell row come from that window with the y-phase weights,
//     which arrive as kernel parameters (constant-bank operands of the FFMAs);
//   * the vertical 3-max is taken in registers, only it goes through shared memory for the horizontal 3-max.
// The smoothed map never goes to HBM.  S = log2 F, T = cell taps.  Same fmaf order as k_hblur_lowres/k_vblur_poly.
constexpr int kTileX = 32;
constexpr int kFusedCols = (kTileX + 2) * kHeatC;          // 646, halo column each side
constexpr int kFusedThreads = 672;
struct PolyWeights {
    float w[48];                                           // [F][T]
};
// HB: the horizontally smoothed low-res rows were formed by k_hblur_lowres (one load per row instead of T; pays
// when F is small, i.e. when the horizontal pass is a large share of the work).
template <int S, int T, bool HB>
__global__ void __launch_bounds__(kFusedThreads, (HB && S <= 3) ? 2 : 1) k_pow2_fused(const float *__restrict__ X, const float *__restrict__ hb,
                                                                const PolyWeights wp, const float *__restrict__ wp_g,
                                                                float *__restrict__ peaks, int h, int w, int strip,
                                                                float thr, Workspace ws) {
    constexpr int F = 1 << S, R = T / 2, NR = F + 2;
    static_assert(F * T <= 48, "weights must fit PolyWeights");
    __shared__ float s_vm[F][kFusedCols];
    const int H = h << S, W = w << S;
    const int n = blockIdx.z, iy0 = blockIdx.y * strip, iy1 = min(h, iy0 + strip), X0 = blockIdx.x * kTileX;
    const float NEG = -__int_as_float(0x7f800000);
    const int rowlen = W * kHeatC;
    const int j = threadIdx.x;
    const int e = (X0 - 1) * kHeatC + j;                   // element index inside the upsampled row
    const bool live = j < kFusedCols && e >= 0 && e < rowlen;
    const bool inner = live && j >= kHeatC && j < kFusedCols - kHeatC;
    const int xl = j / kHeatC, c = j - xl * kHeatC;        // xl counts from the halo column
    const int x = X0 - 1 + xl, cx = x >> S;
    const float *Xn = X + (size_t)n * h * w * kNetC + c;

    constexpr int NRAW = HB ? 1 : T;
    float wx[NRAW];                                        // x-phase weights of this column
    int off[NRAW];                                         // cell offsets inside a low-res row, -1 = outside (zero)
    if (!HB) {
#pragma unroll
        for (int t = 0; t < NRAW; ++t) {
            wx[t] = live ? __ldg(wp_g + (x & (F - 1)) * T + t) : 0.0f;
            const int cc = cx + t - R;
            off[t] = (live && cc >= 0 && cc < w) ? cc * kNetC : -1;
        }
    }
    const float *hb_n = HB ? hb + (size_t)n * h * rowlen + e : nullptr;
    // a low-res row is fetched one iteration before it is reduced into the window, so its latency hides behind
    // the previous row's arithmetic and barriers
    float raw[NRAW];
    auto fetch = [&](int rr) {
        const bool ok = live && rr >= 0 && rr < h;
        if (HB) {
            raw[0] = ok ? __ldg(hb_n + (size_t)rr * rowlen) : 0.0f;
        } else {
            const float *src = Xn + (size_t)rr * w * kNetC;
#pragma unroll
            for (int t = 0; t < NRAW; ++t) raw[t] = (ok && off[t] >= 0) ? __ldg(src + off[t]) : 0.0f;
        }
    };
    auto reduce = [&]() -> float {
        if (HB) return raw[0];
        float acc = 0.0f;
#pragma unroll
        for (int t = 0; t < NRAW; ++t) acc = fmaf(wx[t], raw[t], acc);
        return acc;
    };

    float col[T + 2];                                      // at the top of iteration iy: rows iy-R-1 .. iy+R+1
    col[0] = 0.0f;
#pragma unroll
    for (int r = 1; r < T + 2; ++r) {
        fetch(iy0 - R - 2 + r);
        col[r] = reduce();
    }
    fetch(iy0 + R + 1);
    float g[NR];
    const float thr_lo = thr * (1.0f - 4e-5f);

    for (int iy = iy0; iy < iy1; ++iy) {
#pragma unroll
        for (int r = 0; r < T + 1; ++r) col[r] = col[r + 1];
        col[T + 1] = reduce();
        fetch(iy + R + 2);
        if (live) {
            if (iy == iy0) {
#pragma unroll
                for (int yl = 0; yl < 2; ++yl) {           // last row of the cell above, first row of this cell
                    float acc = 0.0f;
#pragma unroll
                    for (int t = 0; t < T; ++t) acc = fmaf(wp.w[(yl == 0 ? F - 1 : 0) * T + t], col[yl + t], acc);
                    g[yl] = acc;
                }
                if (iy == 0) g[0] = NEG;
            } else {
                g[0] = g[F];
                g[1] = g[F + 1];
            }
#pragma unroll
            for (int yl = 2; yl < NR; ++yl) {
                const int cell = (yl == NR - 1) ? 2 : 1;   // col[] offset of the cell's first tap
                const int phase = (yl == NR - 1) ? 0 : yl - 1;
                float acc = 0.0f;
#pragma unroll
                for (int t = 0; t < T; ++t) acc = fmaf(wp.w[phase * T + t], col[cell + t], acc);
                g[yl] = acc;
            }
            if (iy == h - 1) g[NR - 1] = NEG;
        } else {
#pragma unroll
            for (int yl = 0; yl < NR; ++yl) g[yl] = NEG;
        }
        if (j < kFusedCols) {
#pragma unroll
            for (int yl = 0; yl < F; ++yl) s_vm[yl][j] = fmaxf(fmaxf(g[yl], g[yl + 1]), g[yl + 2]);
        }
        __syncthreads();
        if (inner) {
            // common case kept short: 3-max, equality, (dense) one store, one threshold test folded into a bit mask;
            // everything that happens only next to a blob (records, near-tie counts) runs after the loop
            unsigned hot = 0, pk = 0;
#pragma unroll
            for (int yl = 0; yl < F; ++yl) {
                const float v = g[yl + 1];
                const float own = fmaxf(fmaxf(g[yl], v), g[yl + 2]);
                const float m = fmaxf(fmaxf(s_vm[yl][j - kHeatC], s_vm[yl][j + kHeatC]), own);
                hot |= (v > thr_lo ? 1u : 0u) << yl;
                pk |= (v == m ? 1u : 0u) << yl;
            }
            if (peaks != nullptr) {
                float *prow = peaks + ((size_t)n * H + (size_t)iy * F) * rowlen + e;
#pragma unroll
                for (int yl = 0; yl < F; ++yl) {
                    __stcs(prow, ((pk >> yl) & 1) ? g[yl + 1] : 0.0f);
                    prow += rowlen;
                }
            }
            if (hot != 0 && c < kParts) {
#pragma unroll
                for (int yl = 0; yl < F; ++yl) {
                    if (!((hot >> yl) & 1)) continue;
                    const float v = g[yl + 1];
                    const int y = iy * F + yl;
                    if ((pk >> yl) & 1) {
                        if (v > thr) {
                            int k = atomicAdd(&ws.counters[n].n_peak_rec, 1);
                            if (k < ws.max_peaks) {
                                ws.rec_key[(size_t)n * ws.max_peaks + k] = peak_key(c, y, x);
                                ws.rec_val[(size_t)n * ws.max_peaks + k] =
                                    __ldg(Xn + ((size_t)iy * w + cx) * kNetC);
                            }
                        }
                        if (fabsf(v - thr) <= 2e-5f * thr) atomicAdd(&ws.counters[n].near_ties, 1);
                    } else {
                        const float own = fmaxf(fmaxf(g[yl], v), g[yl + 2]);
                        const float m = fmaxf(fmaxf(s_vm[yl][j - kHeatC], s_vm[yl][j + kHeatC]), own);
                        if ((m - v) <= 2e-5f * fabsf(m)) atomicAdd(&ws.counters[n].near_ties, 1);
                    }
                }
            }
        }
        __syncthreads();
    }
}

template <int S, int T>
void launch_pow2_fused(const float *X, float *hb, const float *wp_host, const float *wp_dev, float *peaks, int N, int h,
                       int w, float thr, const Workspace &ws, cudaStream_t s) {
    PolyWeights pw;
    for (int i = 0; i < (1 << S) * T; ++i) pw.w[i] = wp_host[i];
    const int tiles = ((w << S) + kTileX - 1) / kTileX;
    // strips as long as possible (each one re-forms T+1 window rows) while the grid still fills ~2 waves of 3 CTAs/SM
    int strips = (2 * 3 * 148 + tiles * N - 1) / (tiles * N);
    strips = strips < 1 ? 1 : (strips > (h + 3) / 4 ? (h + 3) / 4 : strips);
    const int strip = (h + strips - 1) / strips;
    dim3 grid(tiles, (h + strip - 1) / strip, N);
    if (hb)
        k_pow2_fused<S, T, true><<<grid, kFusedThreads, 0, s>>>(X, hb, pw, wp_dev, peaks, h, w, strip, thr, ws);
    else
        k_pow2_fused<S, T, false><<<grid, kFusedThreads, 0, s>>>(X, nullptr, pw, wp_dev, peaks, h, w, strip, thr, ws);
}

inline int grid_for(size_t total) {
    size_t want = (total + 255) / 256, cap = (size_t)148 * 16;
    return (int)(want < cap ? want : cap);
}

}  // namespace

// scratch: 2 * N*H*W*19 floats (horizontal pass, blurred map)
int launch_generic_maps(const float *X, int N, int h, int w, int H, int W, const float *d_k1d, int ksize,
                        float *heat_up, float *paf_up, float *peaks, float *scratch, const Workspace &ws, float thr,
                        cudaStream_t s) {
    const size_t nh = (size_t)N * H * W * kHeatC, np = (size_t)N * H * W * kPafC;
    float *tmp = scratch, *g = scratch + nh;
    k_resize_area<<<grid_for(nh), 256, 0, s>>>(X, N, h, w, 0, kHeatC, heat_up, H, W);
    k_resize_area<<<grid_for(np), 256, 0, s>>>(X, N, h, w, kHeatC, kPafC, paf_up, H, W);
    k_blur_1d<<<grid_for(nh), 256, 0, s>>>(heat_up, tmp, N, H, W, kHeatC, d_k1d, ksize, 0);
    k_blur_1d<<<grid_for(nh), 256, 0, s>>>(tmp, g, N, H, W, kHeatC, d_k1d, ksize, 1);
    k_nms<<<grid_for(nh), 256, 0, s>>>(g, heat_up, 0, W, kHeatC, (size_t)H * W * kHeatC, peaks, N, H, W, thr, ws);
    return cudaGetLastError() == cudaSuccess ? 5 : -1;
}

// Power-of-two factors other than the fused x4 kernel's (eval.py's default ratio 8.0, ratio 2.0, ...): the area
// resize is an exact replication, so the smoother collapses the same way as in k1_fused.cu -- a horizontal pass
// over the LOW-RES rows (one per h rows, not per H) with per-phase cell weights, then a vertical pass with per-phase
// row weights (T = 2*ceil(pad/F)+1 taps instead of ksize).  heat_up / paf_up are only materialised in dense mode.
int launch_pow2_maps(const float *X, int N, int h, int w, int log2f, const float *d_wp, const float *h_wp, int T,
                     int path, float *heat_up, float *paf_up, float *peaks, float *scratch, const Workspace &ws, float thr,
                     cudaStream_t s) {
    const int H = h << log2f, W = w << log2f;
    const size_t nh = (size_t)N * H * W * kHeatC, np = (size_t)N * H * W * kPafC;
    const size_t nhb = (size_t)N * h * W * kHeatC;
    float *hb = scratch, *g = scratch + nhb;
    int launches = 0;
    if (heat_up) {
        if ((W * kHeatC) % 4 == 0 && w * kNetC * 4 <= 48 * 1024 && N <= 65535 && h <= 65535) {
            k_replicate_pow2<<<dim3(kRepSplit, h, N), 256, w * kNetC * 4, s>>>(X, h, w, log2f, heat_up, paf_up);
            launches -= 1;
        } else {
            k_replicate_scalar<<<grid_for(nh), 256, 0, s>>>(X, N, h, w, log2f, 0, kHeatC, heat_up);
            k_replicate_scalar<<<grid_for(np), 256, 0, s>>>(X, N, h, w, log2f, kHeatC, kPafC, paf_up);
        }
        launches += 2;
    }
    // the usual smoother sizes have a fully fused kernel; anything else takes three passes through HBM
    const bool fused = path != 4 && N <= 65535 && h <= 65535 &&
                       ((log2f == 3 && T == 5) || (log2f == 1 && T == 13) || (log2f == 4 && T == 3) || (log2f == 2 && T == 7) ||
                        (log2f == 0 && T == 25));
    if (fused) {
        // measured (tools/generic_ratio_times.py): the separate horizontal pass pays for F = 2 and 16, not for F = 8
        const bool use_hb = path == 2 ? false : (path == 3 ? true : log2f != 3);
        if (use_hb) {
            k_hblur_lowres<<<grid_for(nhb), 256, 0, s>>>(X, N, h, w, log2f, d_wp, T, hb);
            ++launches;
        }
        float *hbp = use_hb ? hb : nullptr;
        if (log2f == 3) launch_pow2_fused<3, 5>(X, hbp, h_wp, d_wp, peaks, N, h, w, thr, ws, s);
        else if (log2f == 1) launch_pow2_fused<1, 13>(X, hbp, h_wp, d_wp, peaks, N, h, w, thr, ws, s);
        else if (log2f == 4) launch_pow2_fused<4, 3>(X, hbp, h_wp, d_wp, peaks, N, h, w, thr, ws, s);
        else if (log2f == 0) launch_pow2_fused<0, 25>(X, hbp, h_wp, d_wp, peaks, N, h, w, thr, ws, s);
        else launch_pow2_fused<2, 7>(X, hbp, h_wp, d_wp, peaks, N, h, w, thr, ws, s);
        return cudaGetLastError() == cudaSuccess ? launches + 1 : -1;
    }
    k_hblur_lowres<<<grid_for(nhb), 256, 0, s>>>(X, N, h, w, log2f, d_wp, T, hb);
    k_vblur_poly<<<grid_for(nh), 256, 0, s>>>(hb, N, h, W, log2f, d_wp, T, g);
    k_nms<<<grid_for(nh), 256, 0, s>>>(g, X, log2f, w, kNetC, (size_t)h * w * kNetC, peaks, N, H, W, thr, ws);
    return cudaGetLastError() == cudaSuccess ? launches + 3 : -1;
}

}  // namespace tfpd
