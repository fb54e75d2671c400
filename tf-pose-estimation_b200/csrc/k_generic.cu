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

__global__ void k_nms(const float *__restrict__ g, const float *__restrict__ heat_up, float *__restrict__ peaks,
                      int N, int H, int W, float thr, Workspace ws) {
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
                        ws.rec_val[(size_t)n * ws.max_peaks + k] = heat_up[idx];
                    }
                }
                if (fabsf(v - thr) <= 2e-5f * thr) atomicAdd(&ws.counters[n].near_ties, 1);
            } else if (v > thr_lo && (m - v) <= 2e-5f * fabsf(m)) {
                atomicAdd(&ws.counters[n].near_ties, 1);
            }
        }
    }
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
    k_nms<<<grid_for(nh), 256, 0, s>>>(g, heat_up, peaks, N, H, W, thr, ws);
    return cudaGetLastError() == cudaSuccess ? 5 : -1;
}

}  // namespace tfpd
