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
// Byte work, HBM/L2-bound: one thread per output pixel (all channels), neighbouring threads read neighbouring
// source bytes; a frame (≈1 MB in, 0.5 MB out) stays L2-resident between the four rows an output row reads.
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

// (x + 1.0) / (2.0 / 2**8) in float32, truncated; values outside 0..255 keep the low byte of the int32 truncation
// (what numpy's astype(uint8) yields on x86-64: an input of exactly 1.0 gives 0)
__global__ void k_quantize_img(const float *__restrict__ src, unsigned char *__restrict__ dst, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float v = __fdiv_rn(__fadd_rn(src[i], 1.0f), 0.0078125f);
        dst[i] = (unsigned char)(__float2int_rz(v) & 0xff);
    }
}

}  // namespace

int launch_resize_cubic_u8(const unsigned char *src, unsigned char *dst, int N, int sh, int sw, int cn, int dh, int dw,
                           cudaStream_t s) {
    // cv::resize: inv_scale = (double)dsize/ssize; hal::resize: scale = 1./inv_scale
    const double scale_x = 1.0 / ((double)dw / (double)sw), scale_y = 1.0 / ((double)dh / (double)sh);
    const dim3 block(128), grid((dw + 127) / 128, dh, N);
    switch (cn) {
        case 1: k_resize_cubic_u8<1><<<grid, block, 0, s>>>(src, dst, sh, sw, dh, dw, scale_x, scale_y); break;
        case 2: k_resize_cubic_u8<2><<<grid, block, 0, s>>>(src, dst, sh, sw, dh, dw, scale_x, scale_y); break;
        case 3: k_resize_cubic_u8<3><<<grid, block, 0, s>>>(src, dst, sh, sw, dh, dw, scale_x, scale_y); break;
        case 4: k_resize_cubic_u8<4><<<grid, block, 0, s>>>(src, dst, sh, sw, dh, dw, scale_x, scale_y); break;
        default: return -1;
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

int launch_quantize_img(const float *src, unsigned char *dst, size_t n, int sm_count, cudaStream_t s) {
    if (n == 0) return 0;
    size_t blocks = (n + 255) / 256;
    const size_t cap = (size_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    k_quantize_img<<<(unsigned)blocks, 256, 0, s>>>(src, dst, n);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
