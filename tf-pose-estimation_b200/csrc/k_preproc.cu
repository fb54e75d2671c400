// Image pre-processing in front of the network:
//   * k_resize_cubic_u8: the cv2.resize(npimg, target_size, interpolation=cv2.INTER_CUBIC) of
//     TfPoseEstimator._get_scaled_img(npimg, None) (tf_pose/estimator.py:417-424, called by inference at :534-535);
//   * k_quantize_img:    TfPoseEstimator._quantize_img (estimator.py:383-388).
// The resize follows OpenCV's own 8-bit bicubic arithmetic statement by statement so that the bytes are the ones
// cv2 produces (IPP off): float32 interpolateCubic (A = -0.75) without FMA contraction, coefficients rounded to
// 11-bit fixed point, exact integer horizontal pass, vertical pass in float32 (separately rounded multiply and add,
// round-half-even) except for the last (dst_w*cn) % 8 elements of a row, which take the fixed-point cast of the
// scalar tail of OpenCV's 128-bit SIMD loop.
//
// Three forms with identical arithmetic (option "resize_path"): one thread per output pixel (the plain statement),
// the strip form (default) and a shared-memory tiled form that measured slower; the tests force each.
#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

struct CubicTap {
    int ofs;      // source index of tap 1 (tap k reads clamp(ofs - 1 + k))
    int c[4];     // 11-bit fixed-point coefficients
};

__device__ __forceinline__ CubicTap cubic_tap(int d, double scale) {
    // fx = (float)((d + 0.5) * scale - 0.5); s = floor(fx); fx -= s
    const double p = __dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
    float fx = __double2float_rn(p);
    const float fl = floorf(fx);
    fx = __fsub_rn(fx, fl);
    const float A = -0.75f;
    const float x1 = __fadd_rn(fx, 1.0f);
    float c0 = __fsub_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fsub_rn(__fmul_rn(A, x1), -3.75f), x1), -6.0f), x1), -3.0f);
    float c1 = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(1.25f, fx), 2.25f), fx), fx), 1.0f);
    const float xm = __fsub_rn(1.0f, fx);
    float c2 = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(1.25f, xm), 2.25f), xm), xm), 1.0f);
    float c3 = __fsub_rn(__fsub_rn(__fsub_rn(1.0f, c0), c1), c2);
    CubicTap t;
    t.ofs = (int)fl;
    const float cf[4] = {c0, c1, c2, c3};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int v = __float2int_rn(__fmul_rn(cf[k], 2048.0f));    // saturate_cast<short>(cvRound(c * 2^11))
        t.c[k] = max(-32768, min(32767, v));
    }
    return t;
}

template <int CN>
__global__ void __launch_bounds__(128) k_resize_cubic_u8(const unsigned char *__restrict__ src,
                                                          unsigned char *__restrict__ dst, int sh, int sw, int dh,
                                                          int dw, double scale_x, double scale_y) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, n = blockIdx.z;
    if (x >= dw) return;
    const CubicTap tx = cubic_tap(x, scale_x), ty = cubic_tap(y, scale_y);
    int col[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) col[k] = max(0, min(sw - 1, tx.ofs - 1 + k)) * CN;
    int hs[4][CN];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int row = max(0, min(sh - 1, ty.ofs - 1 + r));
        const unsigned char *p = src + ((size_t)n * sh + row) * (size_t)sw * CN;
#pragma unroll
        for (int c = 0; c < CN; ++c) {
            int v = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) v += (int)__ldg(p + col[k] + c) * tx.c[k];
            hs[r][c] = v;
        }
    }
    const float sc = 1.0f / (2048.0f * 2048.0f);
    const float b0 = (float)ty.c[0] * sc, b1 = (float)ty.c[1] * sc, b2 = (float)ty.c[2] * sc, b3 = (float)ty.c[3] * sc;
    const int simd_end = (dw * CN) & ~7;
    unsigned char *q = dst + (((size_t)n * dh + y) * dw + x) * CN;
#pragma unroll
    for (int c = 0; c < CN; ++c) {
        int v;
        if (x * CN + c < simd_end) {
            float acc = __fmul_rn((float)hs[3][c], b3);
            acc = __fadd_rn(__fmul_rn((float)hs[2][c], b2), acc);
            acc = __fadd_rn(__fmul_rn((float)hs[1][c], b1), acc);
            acc = __fadd_rn(__fmul_rn((float)hs[0][c], b0), acc);
            v = __float2int_rn(acc);
        } else {
            const int s = hs[0][c] * ty.c[0] + hs[1][c] * ty.c[1] + hs[2][c] * ty.c[2] + hs[3][c] * ty.c[3];
            v = (s + (1 << 21)) >> 22;
        }
        q[c] = (unsigned char)max(0, min(255, v));
    }
}

// Strip form (the default): a thread owns one output column of a strip of kStrip output rows and marches down it
// with a register window of the horizontal sums of four consecutive source rows.  Consecutive output rows share most
// of their source rows, so a source row's horizontal pass runs once per strip instead of once per output row that
// reads it, and the column taps are computed once per strip; the strip's row taps sit in shared memory.  The kernel
// is issue-bound (ALU, not L1 wavefronts: fetching the taps as realigned 32-bit words measured slower), so the
// instruction count per output pixel is what counts.  Measured per 64 frames 640x480x3 -> 432x368x3: per-pixel form
// 0.152 ms (0.170 with word loads), tiled 0.173, this one 0.084 (0.093 with word loads; CTA widths that leave fewer
// idle lanes across the 432 columns change nothing).
constexpr int kStrip = 16;

// int <-> float through the 1.5*2^23 constant: exact for |v| < 2^22 (horizontal sums stay below 2^20; the vertical
// sum is a pixel value), round-half-even like cvtps2dq / __float2int_rn, and full-rate FADD/IADD instead of the
// quarter-rate conversion pipe (ncu r01j: the conversions were a third of the strip kernel's issue slots).
__device__ __forceinline__ float exact_i2f(int v) { return __fadd_rn(__int_as_float(v + 0x4B400000), -12582912.0f); }
__device__ __forceinline__ int round_f2i(float v) { return __float_as_int(__fadd_rn(v, 12582912.0f)) - 0x4B400000; }

// TAIL: the row length dw*CN is not a multiple of 8, so the last elements take the fixed-point cast and the integer
// sums have to be kept beside their float copies.
template <int CN, bool TAIL>
__global__ void __launch_bounds__(128) k_resize_cubic_u8_strip(const unsigned char *__restrict__ src,
                                                                unsigned char *__restrict__ dst, int sh, int sw, int dh,
                                                                int dw, double scale_x, double scale_y) {
    __shared__ int yofs[kStrip];
    __shared__ int ycoef[kStrip][4];
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y0 = blockIdx.y * kStrip, n = blockIdx.z;
    const int th = min(kStrip, dh - y0);
    if (threadIdx.x < th) {
        const CubicTap t = cubic_tap(y0 + threadIdx.x, scale_y);
        yofs[threadIdx.x] = t.ofs;
#pragma unroll
        for (int k = 0; k < 4; ++k) ycoef[threadIdx.x][k] = t.c[k];
    }
    __syncthreads();
    if (x >= dw) return;
    const CubicTap tx = cubic_tap(x, scale_x);
    // the four tap columns as pointers into row 0 of this image; a source row adds one 32-bit offset
    const unsigned char *img = src + (size_t)n * sh * (size_t)sw * CN;
    const unsigned char *pc0 = img + max(0, min(sw - 1, tx.ofs - 1)) * CN, *pc1 = img + max(0, min(sw - 1, tx.ofs)) * CN,
                        *pc2 = img + max(0, min(sw - 1, tx.ofs + 1)) * CN, *pc3 = img + max(0, min(sw - 1, tx.ofs + 2)) * CN;
    const unsigned pitch = (unsigned)sw * CN;
    auto hrow = [&](int u, int (&hi)[CN], float (&hf)[CN]) {      // horizontal sums of source row clamp(u)
        const unsigned off = (unsigned)max(0, min(sh - 1, u)) * pitch;     // < 2^32: sh, sw <= 32767, CN <= 4 (checked by the API)
#pragma unroll
        for (int c = 0; c < CN; ++c) {
            const int v = (int)__ldg(pc0 + off + c) * tx.c[0] + (int)__ldg(pc1 + off + c) * tx.c[1] +
                          (int)__ldg(pc2 + off + c) * tx.c[2] + (int)__ldg(pc3 + off + c) * tx.c[3];
            hi[c] = v;
            hf[c] = exact_i2f(v);
        }
    };
    int i0[CN], i1[CN], i2[CN], i3[CN];
    float h0[CN], h1[CN], h2[CN], h3[CN];
    int base = yofs[0] - 1;                       // window = unclamped source rows base .. base+3
    hrow(base, i0, h0);
    hrow(base + 1, i1, h1);
    hrow(base + 2, i2, h2);
    hrow(base + 3, i3, h3);
    const float sc = 1.0f / (2048.0f * 2048.0f);
    const int simd_end = (dw * CN) & ~7;
    unsigned char *q = dst + (((size_t)n * dh + y0) * dw + x) * CN;
    for (int yl = 0; yl < th; ++yl, q += (size_t)dw * CN) {
        const int nb = yofs[yl] - 1;              // uniform over the CTA: no divergence in the window update
        if (nb - base >= 4) {
            base = nb;
            hrow(base, i0, h0);
            hrow(base + 1, i1, h1);
            hrow(base + 2, i2, h2);
            hrow(base + 3, i3, h3);
        } else {
            for (; base < nb; ++base) {
#pragma unroll
                for (int c = 0; c < CN; ++c) {
                    h0[c] = h1[c];
                    h1[c] = h2[c];
                    h2[c] = h3[c];
                    if (TAIL) {
                        i0[c] = i1[c];
                        i1[c] = i2[c];
                        i2[c] = i3[c];
                    }
                }
                hrow(base + 4, i3, h3);
            }
        }
        const int c0 = ycoef[yl][0], c1 = ycoef[yl][1], c2 = ycoef[yl][2], c3 = ycoef[yl][3];
        const float b0 = (float)c0 * sc, b1 = (float)c1 * sc, b2 = (float)c2 * sc, b3 = (float)c3 * sc;
#pragma unroll
        for (int c = 0; c < CN; ++c) {
            int v;
            if (!TAIL || x * CN + c < simd_end) {
                float acc = __fmul_rn(h3[c], b3);
                acc = __fadd_rn(__fmul_rn(h2[c], b2), acc);
                acc = __fadd_rn(__fmul_rn(h1[c], b1), acc);
                acc = __fadd_rn(__fmul_rn(h0[c], b0), acc);
                v = round_f2i(fminf(fmaxf(acc, -1024.0f), 1024.0f));      // clamp first: the rounding trick needs |acc| < 2^22
            } else {
                v = (i0[c] * c0 + i1[c] * c1 + i2[c] * c2 + i3[c] * c3 + (1 << 21)) >> 22;
            }
            q[c] = (unsigned char)max(0, min(255, v));
        }
    }
}

// Tiled form: one CTA = TY output rows x tile_w output columns.  The coefficient tables of the tile are computed once
// (not per pixel), the horizontal pass runs once per SOURCE row the tile touches (TY*scale + 3 rows instead of 4 per
// output row) into shared memory as exact int32, the vertical pass reads it back.  Same arithmetic as above.
constexpr int kTY = 16, kPreThreads = 256;

template <int CN>
__global__ void __launch_bounds__(kPreThreads) k_resize_cubic_u8_tiled(const unsigned char *__restrict__ src,
                                                                        unsigned char *__restrict__ dst, int sh, int sw,
                                                                        int dh, int dw, double scale_x, double scale_y,
                                                                        int tile_w, int max_rows) {
    extern __shared__ int smem_i[];
    const int x0 = blockIdx.x * tile_w, y0 = blockIdx.y * kTY, n = blockIdx.z;
    const int tw = min(tile_w, dw - x0), th = min(kTY, dh - y0), twe = tw * CN;
    int *xcol = smem_i;                       // [tile_w][4] clamped source columns * CN
    int *xc = xcol + tile_w * 4;              // [tile_w][4]
    int *yrow = xc + tile_w * 4;              // [kTY][4] clamped source rows
    int *yc = yrow + kTY * 4;                 // [kTY][4]
    int *hbuf = yc + kTY * 4;                 // [max_rows][tile_w*CN]
    const int tid = threadIdx.x;
    for (int i = tid; i < tw; i += kPreThreads) {
        const CubicTap t = cubic_tap(x0 + i, scale_x);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            xcol[i * 4 + k] = max(0, min(sw - 1, t.ofs - 1 + k)) * CN;
            xc[i * 4 + k] = t.c[k];
        }
    }
    if (tid < th) {
        const CubicTap t = cubic_tap(y0 + tid, scale_y);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            yrow[tid * 4 + k] = max(0, min(sh - 1, t.ofs - 1 + k));
            yc[tid * 4 + k] = t.c[k];
        }
    }
    __syncthreads();
    const int r_lo = yrow[0], nrows = yrow[(th - 1) * 4 + 3] - r_lo + 1;
    if (nrows > max_rows) __trap();           // the host sized the buffer from the same scale: cannot happen
    const unsigned char *base = src + ((size_t)n * sh + r_lo) * (size_t)sw * CN;
    for (int idx = tid; idx < nrows * twe; idx += kPreThreads) {
        const int r = idx / twe, e = idx - r * twe;
        const int i = e / CN, c = e - i * CN;
        const int4 col = *reinterpret_cast<const int4 *>(xcol + i * 4);
        const int4 w = *reinterpret_cast<const int4 *>(xc + i * 4);
        const unsigned char *p = base + (size_t)r * sw * CN + c;
        hbuf[idx] = (int)__ldg(p + col.x) * w.x + (int)__ldg(p + col.y) * w.y + (int)__ldg(p + col.z) * w.z +
                    (int)__ldg(p + col.w) * w.w;
    }
    __syncthreads();
    const float sc = 1.0f / (2048.0f * 2048.0f);
    const int simd_end = (dw * CN) & ~7;
    const int quads = (twe + 3) >> 2;
    for (int idx = tid; idx < th * quads; idx += kPreThreads) {
        const int yl = idx / quads, e0 = (idx - yl * quads) * 4;
        const int4 rr = *reinterpret_cast<const int4 *>(yrow + yl * 4);
        const int4 cc = *reinterpret_cast<const int4 *>(yc + yl * 4);
        const int *h0 = hbuf + (rr.x - r_lo) * twe, *h1 = hbuf + (rr.y - r_lo) * twe, *h2 = hbuf + (rr.z - r_lo) * twe,
                  *h3 = hbuf + (rr.w - r_lo) * twe;
        const float b0 = (float)cc.x * sc, b1 = (float)cc.y * sc, b2 = (float)cc.z * sc, b3 = (float)cc.w * sc;
        unsigned packed = 0;
        const int cnt = min(4, twe - e0);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int e = min(e0 + j, twe - 1);
            const int s0 = h0[e], s1 = h1[e], s2 = h2[e], s3 = h3[e];
            int v;
            if (x0 * CN + e < simd_end) {
                float acc = __fmul_rn((float)s3, b3);
                acc = __fadd_rn(__fmul_rn((float)s2, b2), acc);
                acc = __fadd_rn(__fmul_rn((float)s1, b1), acc);
                acc = __fadd_rn(__fmul_rn((float)s0, b0), acc);
                v = __float2int_rn(acc);
            } else {
                v = (s0 * cc.x + s1 * cc.y + s2 * cc.z + s3 * cc.w + (1 << 21)) >> 22;
            }
            packed |= (unsigned)max(0, min(255, v)) << (8 * j);
        }
        unsigned char *q = dst + (((size_t)n * dh + (y0 + yl)) * dw + x0) * CN + e0;
        if (cnt == 4 && (reinterpret_cast<uintptr_t>(q) & 3) == 0) {
            *reinterpret_cast<unsigned *>(q) = packed;
        } else {
            for (int j = 0; j < cnt; ++j) q[j] = (unsigned char)(packed >> (8 * j));
        }
    }
}

// (x + 1.0) / (2.0 / 2**8) in float32, truncated; values outside 0..255 keep the low byte of the int32 truncation
// (what numpy's astype(uint8) yields on x86-64: an input of exactly 1.0 gives 0)
__device__ __forceinline__ unsigned char quantize1(float x) {
    return (unsigned char)(__float2int_rz(__fdiv_rn(__fadd_rn(x, 1.0f), 0.0078125f)) & 0xff);
}

// four elements per thread (128-bit load, 32-bit store) when both pointers allow it; n4 = number of such groups
__global__ void k_quantize_img(const float *__restrict__ src, unsigned char *__restrict__ dst, size_t n, size_t n4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n4; i += stride) {
        const float4 v = __ldcs(reinterpret_cast<const float4 *>(src) + i);
        __stcs(reinterpret_cast<uchar4 *>(dst) + i, make_uchar4(quantize1(v.x), quantize1(v.y), quantize1(v.z), quantize1(v.w)));
    }
    for (size_t i = n4 * 4 + t0; i < n; i += stride) dst[i] = quantize1(src[i]);
}

}  // namespace

int launch_resize_cubic_u8(const unsigned char *src, unsigned char *dst, int N, int sh, int sw, int cn, int dh, int dw,
                           double fx, double fy, int path, cudaStream_t s) {
    // cv::resize: inv_scale = (double)dsize/ssize when dsize is given, fx / fy themselves when it is derived from them;
    // hal::resize: scale = 1./inv_scale
    const double scale_x = 1.0 / (fx > 0 ? fx : (double)dw / (double)sw), scale_y = 1.0 / (fy > 0 ? fy : (double)dh / (double)sh);
    // tiled form: tile width = dw cut into equal parts of <= 128 columns; source rows of a tile bounded from scale_y
    const int parts = (dw + 127) / 128, tile_w = (dw + parts - 1) / parts;
    const int max_rows = (int)ceil((kTY - 1) * scale_y) + 5;
    const size_t smem = ((size_t)tile_w * 8 + kTY * 8 + (size_t)max_rows * tile_w * cn) * sizeof(int);
    const bool tiled = path == 1;      // measured slower than the per-pixel kernel (0.17-0.28 ms against 0.15 per 64 frames)
    const bool strip = path == 0 || path == 3;
    if (tiled && smem > 200 * 1024) return -1;
#define TFPD_RESIZE_CASE(CN)                                                                                           \
    case CN:                                                                                                           \
        if (tiled) {                                                                                                   \
            if (smem > 48 * 1024 &&                                                                                    \
                cudaFuncSetAttribute(k_resize_cubic_u8_tiled<CN>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                                     (int)smem) != cudaSuccess)                                                        \
                return -1;                                                                                             \
            k_resize_cubic_u8_tiled<CN><<<dim3(parts, (dh + kTY - 1) / kTY, N), kPreThreads, smem, s>>>(                \
                src, dst, sh, sw, dh, dw, scale_x, scale_y, tile_w, max_rows);                                         \
        } else if (strip) {                                                                                            \
            if ((dw * CN) & 7)                                                                                         \
                k_resize_cubic_u8_strip<CN, true><<<dim3((dw + 127) / 128, (dh + kStrip - 1) / kStrip, N), 128, 0, s>>>( \
                    src, dst, sh, sw, dh, dw, scale_x, scale_y);                                                       \
            else                                                                                                       \
                k_resize_cubic_u8_strip<CN, false><<<dim3((dw + 127) / 128, (dh + kStrip - 1) / kStrip, N), 128, 0, s>>>( \
                    src, dst, sh, sw, dh, dw, scale_x, scale_y);                                                       \
        } else {                                                                                                       \
            k_resize_cubic_u8<CN><<<dim3((dw + 127) / 128, dh, N), 128, 0, s>>>(src, dst, sh, sw, dh, dw, scale_x,      \
                                                                               scale_y);                               \
        }                                                                                                              \
        break;
    switch (cn) {
        TFPD_RESIZE_CASE(1)
        TFPD_RESIZE_CASE(2)
        TFPD_RESIZE_CASE(3)
        TFPD_RESIZE_CASE(4)
        default: return -1;
    }
#undef TFPD_RESIZE_CASE
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

int launch_quantize_img(const float *src, unsigned char *dst, size_t n, int sm_count, cudaStream_t s) {
    if (n == 0) return 0;
    const bool vec = (reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0;
    const size_t n4 = vec ? n / 4 : 0, items = vec ? n4 + 3 : n;
    size_t blocks = (items + 255) / 256;
    const size_t cap = (size_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    k_quantize_img<<<(unsigned)blocks, 256, 0, s>>>(src, dst, n, n4);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
