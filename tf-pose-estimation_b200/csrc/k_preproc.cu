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
// Four forms with identical arithmetic (option "resize_path"): one thread per output pixel (the plain statement),
// the strip form, its shared-memory staged version (default) and a tiled form that measured slower; the tests force each.
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

// Staged strip form (the default where the source rows are 4-byte aligned): the strip form's arithmetic in about half
// its instructions (64 frames 640x480x3 -> 432x368x3: 53.8 M -> 28 M warp instructions, 76 -> 39 us; ncu r02y / r03h).
//  * The source window of a CTA (the rows its strip reads x the columns its output columns read; out-of-image columns
//    and rows replicated) is staged in shared memory: one TMA bulk copy per row for the whole 16-byte units, issued by
//    warp 0 before anything else, word loads for the few units at the row ends, byte stores for the border columns.
//    A thread's 4 tap columns x CN channels are then 4*CN CONTIGUOUS bytes at a fixed offset of every staged row,
//    fetched as CN+1 aligned words (the strip form issues 4*CN byte loads with 64-bit address arithmetic each) and
//    realigned with funnel shifts.
//  * Byte permutes gather a channel's four taps into one register and two dp2a (s16 coefficient pairs x u8 taps)
//    make its exact integer sum; the accumulator starts at the bit pattern of 1.5*2^23, so the sum arrives as the
//    float's bits and one FADD finishes the exact int->float conversion.
//  * A thread computes TWO output columns (x and x + blockDim.x): the per-row bookkeeping is shared and the 2*CN
//    channel values always pair up for the packed vertical pass.
//  * The window of four source rows is a register ring rotated by unrolling (no moves): the outer loop walks SOURCE
//    rows and emits every output row whose fourth tap row was just computed.
//  * The vertical pass runs on packed FP32 (FMUL2 + an FFMA2 by 1.0: the same separately rounded multiplies and
//    adds, see pk_add); saturate_cast<uchar>(cvRound()) is one F2IP.U8.F32.
//  * Strip height and CTA width are chosen per launch (launch_resize_cubic_u8): whole waves for small batches, as
//    little total work as fits four CTAs per SM otherwise.
// Same bytes as the other forms (tests force each).  What is left (ncu r03h): 74 % of the issue slots busy on ~28 M
// instructions that are the formulation's own (per two columns and source row: 8 word loads, 6 funnel shifts,
// 12 permutes, 12 dp2a, 6 FADD; per output row 21 packed FP32, 6 F2IP, 6 byte stores).  TRIED, NOT KEPT (profiles/
// r03i_*): a producer / consumer ring version (one extra warp staging 8-row chunks through full / empty mbarriers,
// consumers walking 41-row segments so that a source row's sums are computed once: 1.05 passes over the source
// instead of 1.49) removed the staging waits and measured the same 39 us (its per-row chunk bookkeeping ate the
// gain); 720p sources gained 7 %.
constexpr int kStagedPx = 2;             // output columns per thread: x and x + blockDim.x
template <int CN>
struct HRow {                            // horizontal sums of one source row: CN channels of both columns, in pairs
    float2 p[CN * kStagedPx / 2];
};

__device__ __forceinline__ float2 pk_mul(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b), rd;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2 *>(&rd);
}
// a + b as fma(a, one, b) with one == 1.0f passed in at RUN time: ptxas contracts mul.rn.f32x2 feeding add.rn.f32x2
// into one FFMA2 (observed with CUDA 12.9 even under --fmad=false, .rn notwithstanding), which would round the product
// and the sum once instead of twice.  A product feeding the MULTIPLICAND of an fma with an opaque multiplier cannot be
// contracted, and a * 1.0 + b rounds exactly like a + b.
__device__ __forceinline__ float2 pk_add(float2 a, float2 one, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), ro = *reinterpret_cast<unsigned long long *>(&one),
                       rb = *reinterpret_cast<unsigned long long *>(&b), rd;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(ro), "l"(rb));
    return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ int dp2a_lo_su(unsigned a, unsigned b, int c) {     // c + s16(a.lo)*u8(b.0) + s16(a.hi)*u8(b.1)
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(unsigned a, unsigned b, int c) {     // c + s16(a.lo)*u8(b.2) + s16(a.hi)*u8(b.3)
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// saturate_cast<uchar>(cvRound(v)) in one instruction (F2IP.U8.F32): round half even, clamped to 0..255
__device__ __forceinline__ unsigned f2u8_rn_sat(float v) {
    unsigned r;
    asm("cvt.rni.u8.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}

// the four taps of channel C (bytes C, CN+C, 2CN+C, 3CN+C of the realigned tap bytes A[]) in one register
template <int CN, int C>
__device__ __forceinline__ unsigned gather_taps(const unsigned (&A)[CN]) {
    constexpr int i0 = C, i1 = CN + C, i2 = 2 * CN + C, i3 = 3 * CN + C;
    if constexpr (CN == 1) {
        return A[0];
    } else if constexpr (CN == 2) {
        return __byte_perm(A[0], A[1], i0 | (i1 << 4) | (i2 << 8) | (i3 << 12));
    } else if constexpr (CN == 3) {
        // taps below byte 8 come from (A0, A1); the rest from A2
        constexpr unsigned s1 = (unsigned)(i0 | (i1 << 4) | ((i2 < 8 ? i2 : 0) << 8));
        const unsigned P = __byte_perm(A[0], A[1], s1);
        constexpr unsigned s2 = 0u | (1u << 4) | ((i2 < 8 ? 2u : (unsigned)(4 + i2 - 8)) << 8) | ((unsigned)(4 + i3 - 8) << 12);
        return __byte_perm(P, A[2], s2);
    } else {
        const unsigned lo = __byte_perm(A[0], A[1], (i0 & 3) | ((4 + (i1 & 3)) << 4));
        const unsigned hi = __byte_perm(A[2], A[3], (i2 & 3) | ((4 + (i3 & 3)) << 4));
        return __byte_perm(lo, hi, 0x5410);
    }
}

// shared-memory loads through explicit 32-bit shared addresses: with generic pointers the compiler rebuilt the shared
// window base (S2UR SR_CgaCtaId, ULEA ...) in front of every group of accesses of the hot loop (ncu r02z: a quarter of
// its instructions were uniform-datapath bookkeeping).  "memory" keeps them behind the barrier that ends the staging.
template <int OFF>
__device__ __forceinline__ unsigned lds_u32(unsigned addr) {
    unsigned v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF) : "memory");
    return v;
}
__device__ __forceinline__ float4 lds_f4(unsigned addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ int4 lds_i4(unsigned addr) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// TMA bulk load global -> shared with an mbarrier counting the bytes (16-byte aligned source, destination and size)
__device__ __forceinline__ void st_mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void st_mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void st_mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n"
        " bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void st_bulk_load(unsigned sdst, const void *gsrc, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sdst),
                 "l"(gsrc), "r"(bytes), "r"(bar)
                 : "memory");
}

constexpr int kStagePad = 16;            // bytes in front of a staged row: room for replicated left border columns
constexpr int kStagedMaxTy = 32;         // most output rows per CTA (the strip height is a launch parameter)
constexpr int kStagedHeader = 36 * kStagedMaxTy;   // bytes in front of the staged rows

// TAIL: dw*CN is not a multiple of 8, the last elements of a row take the fixed-point cast (kept out of the common
// variant altogether: the compiler hoists the tail's float->int conversions out of their branch)
template <int CN, bool TAIL>
__global__ void __launch_bounds__(256) k_resize_cubic_u8_staged(const unsigned char *__restrict__ src,
                                                                 unsigned char *__restrict__ dst, int sh, int sw, int dh,
                                                                 int dw, double scale_x, double scale_y, int stride,
                                                                 int max_rows, float one, int TY) {
    constexpr int PX = kStagedPx, VC = CN * PX, NP = VC / 2;
    // dynamic shared memory: ylast[32] | ybf[32] | ybi[32] | the staged rows [max_rows][stride]   (TY <= 32 output rows)
    extern __shared__ __align__(16) unsigned char st_dyn[];
    int *ylast = reinterpret_cast<int *>(st_dyn);                    // source row (unclamped) of an output row's fourth tap
    float4 *ybf = reinterpret_cast<float4 *>(st_dyn + 4 * kStagedMaxTy);       // its coefficients / 2^22 as floats (vertical pass)
    int4 *ybi = reinterpret_cast<int4 *>(st_dyn + 20 * kStagedMaxTy);          // and as integers (fixed-point tail of a row)
    unsigned char *st_rows = st_dyn + kStagedHeader;
    __shared__ __align__(8) unsigned long long s_bar;      // mbarrier of the staging copies
    const int tid = threadIdx.x, T = blockDim.x;
    const unsigned bar32 = (unsigned)__cvta_generic_to_shared(&s_bar);
    const int x0 = blockIdx.x * (PX * T), tw = min(PX * T, dw - x0);
    const int y0 = blockIdx.y * TY, n = blockIdx.z;
    const int th = min(TY, dh - y0);
    const int pitch = sw * CN;
    const unsigned char *img = src + (size_t)n * sh * (size_t)pitch;
    const bool al16 = ((reinterpret_cast<uintptr_t>(img) | (unsigned)pitch) & 15) == 0;
    const unsigned st32 = (unsigned)__cvta_generic_to_shared(st_rows);
    // What the CTA stages, from its first / last source column (wl, wr) and row (u_first, u_last), all unclamped:
    // staged row r <-> source row clamp(u_first + r); staged byte s of a row <-> source byte s - kStagePad + g0 of that
    // row (g0: 16-byte aligned), columns clamped; units [un_a, un_b) of 16 bytes lie wholly inside the source row.
    struct Ext {
        int wl, wr, u_first, u_last, nrows, g0, units, un_a, un_b;
    };
    auto extents = [&](int wl, int wr, int u_first, int u_last) {
        Ext e;
        e.wl = wl, e.wr = wr, e.u_first = u_first, e.u_last = u_last;
        e.nrows = u_last - u_first + 1;
        e.g0 = (max(wl, 0) * CN) & ~15;
        const int s_end = (wr + 1) * CN - e.g0 + kStagePad;     // one past the last byte any thread reads
        e.units = (s_end + 4 + 15) >> 4;                         // + the extra word of the realigning loads
        e.un_a = e.g0 >= kStagePad ? 0 : (kStagePad - e.g0) >> 4;
        e.un_b = al16 ? max(e.un_a, min(e.units, ((pitch - e.g0 + kStagePad) >> 4))) : e.un_a;
        return e;
    };
    // Staging.  Every warp finds the extents for itself (four taps on four lanes, shuffled: no barrier in front of the
    // staging); warp 0 then issues the whole units of every row as ONE bulk copy (TMA) per row, lane = row (strided),
    // and everything else -- the edge units, the header, every thread's own column taps (float64 arithmetic) -- is
    // computed while the copies are in flight.  Out-of-image rows are copies of the first / last row (the vertical
    // taps clamp).  One barrier at the end.
    Ext ex;
    {
        const int l = tid & 31;
        const int d = l == 0 ? x0 : l == 1 ? x0 + tw - 1 : l == 2 ? y0 : y0 + th - 1;
        const CubicTap t = cubic_tap(d, l < 2 ? scale_x : scale_y);
        const int v = t.ofs + ((l & 1) ? 2 : -1);
        ex = extents(__shfl_sync(0xffffffffu, v, 0), __shfl_sync(0xffffffffu, v, 1), __shfl_sync(0xffffffffu, v, 2),
                     __shfl_sync(0xffffffffu, v, 3));
    }
    const int wl = ex.wl, wr = ex.wr, u_first = ex.u_first, u_last = ex.u_last, nrows = ex.nrows, g0 = ex.g0, units = ex.units,
              un_a = ex.un_a, un_b = ex.un_b;
    if (nrows > max_rows || units * 16 > stride) __trap();      // the host sized the buffer from the same scales
    const unsigned row_bytes = (unsigned)(un_b - un_a) * 16u;
    if (tid < 32) {
        if (tid == 0) {
            st_mbar_init(bar32, 1);
            if (row_bytes != 0u) st_mbar_expect_tx(bar32, row_bytes * (unsigned)nrows);
        }
        __syncwarp();
        if (row_bytes != 0u)
            for (int r = tid; r < nrows; r += 32)
                st_bulk_load(st32 + (unsigned)r * (unsigned)stride + 16u * (unsigned)un_a,
                             img + (size_t)max(0, min(sh - 1, u_first + r)) * pitch + (g0 - kStagePad) + 16 * un_a, row_bytes, bar32);
    }
    // The units that straddle or lie outside the source row (all of them when the rows are only 4-byte aligned), flat
    // over (row, unit): words inside the row are loaded (rows are 4-byte aligned: host check), the rest is zero; the
    // replicated border columns are patched in below, after a barrier.  (Filling them here, byte by byte, in the same
    // pass measured slower: 39 -> 46 us, the few threads with such units keep the whole CTA at the barrier.)
    const int n_edge = units - (un_b - un_a);
    for (int idx = tid; idx < nrows * n_edge; idx += T) {
        const int r = idx / n_edge, e = idx - r * n_edge;
        const int un = e < un_a ? e : e + (un_b - un_a);
        const unsigned char *row = img + (size_t)max(0, min(sh - 1, u_first + r)) * pitch;
        const int g = un * 16 - kStagePad + g0;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (g >= 0 && g + 4 <= pitch) v.x = __ldg(reinterpret_cast<const unsigned *>(row + g));
        if (g + 4 >= 0 && g + 8 <= pitch) v.y = __ldg(reinterpret_cast<const unsigned *>(row + g + 4));
        if (g + 8 >= 0 && g + 12 <= pitch) v.z = __ldg(reinterpret_cast<const unsigned *>(row + g + 8));
        if (g + 12 >= 0 && g + 16 <= pitch) v.w = __ldg(reinterpret_cast<const unsigned *>(row + g + 12));
        *reinterpret_cast<uint4 *>(st_rows + (size_t)r * stride + un * 16) = v;
    }
    // header: the row taps of the strip, by the last threads of the CTA (warp 0 has the copies to issue)
    const float sc = 1.0f / (2048.0f * 2048.0f);
    if (T - 1 - tid < th) {
        const int i = T - 1 - tid;
        const CubicTap t = cubic_tap(y0 + i, scale_y);
        ylast[i] = t.ofs + 2;
        ybf[i] = make_float4((float)t.c[0] * sc, (float)t.c[1] * sc, (float)t.c[2] * sc, (float)t.c[3] * sc);
        if (TAIL) ybi[i] = make_int4(t.c[0], t.c[1], t.c[2], t.c[3]);
    }
    // this thread's columns: x0 + tid and x0 + tid + T (taps of an out-of-range column are those of the last column)
    int xs[PX], ta[PX];                   // column, source byte offset of its first tap byte
    unsigned c01[PX], c23[PX];
#pragma unroll
    for (int j = 0; j < PX; ++j) {
        xs[j] = x0 + tid + j * T;
        const CubicTap tx = cubic_tap(min(xs[j], dw - 1), scale_x);
        ta[j] = (tx.ofs - 1) * CN;
        c01[j] = ((unsigned)tx.c[0] & 0xffffu) | ((unsigned)tx.c[1] << 16);
        c23[j] = ((unsigned)tx.c[2] & 0xffffu) | ((unsigned)tx.c[3] << 16);
    }
    // the replicated border columns a window can reach (at most two on either side), after the units they lie in
    const int nb_l = max(0, -wl) * CN, nb_r = max(0, wr - (sw - 1)) * CN;
    if (nb_l + nb_r > 0) {
        __syncthreads();
        const int nb = nb_l + nb_r;
        for (int idx = tid; idx < nrows * nb; idx += T) {
            const int r = idx / nb, i = idx - r * nb;
            const unsigned char *row = img + (size_t)max(0, min(sh - 1, u_first + r)) * pitch;
            unsigned char *srow = st_rows + (size_t)r * stride;
            if (i < nb_l) {           // byte i of the columns wl .. -1: channel i % CN of column 0
                srow[wl * CN + i - g0 + kStagePad] = __ldg(row + i % CN);
            } else {                  // byte j of the columns sw .. wr: channel j % CN of column sw-1
                const int j = i - nb_l;
                srow[sw * CN + j - g0 + kStagePad] = __ldg(row + (sw - 1) * CN + j % CN);
            }
        }
    }
    __syncthreads();              // staged bytes and header written; the mbarrier (initialised by thread 0) may be waited on
    if (row_bytes != 0u) st_mbar_wait(bar32, 0u);
    if (xs[0] >= dw) return;
    const bool valid1 = xs[PX - 1] < dw;

    // one 32-bit shared base, made opaque so that it stays in a register (the compiler otherwise rebuilds it from
    // SR_CgaCtaId at every use)
    unsigned sbase = (unsigned)__cvta_generic_to_shared(st_dyn);
    asm volatile("" : "+r"(sbase));
    const unsigned a_ylast = sbase, a_ybf = sbase + 4u * kStagedMaxTy, a_ybi = sbase + 20u * kStagedMaxTy;
    // a column's tap bytes: 4*CN contiguous bytes from staged byte a of every row, fetched as CN+1 aligned words
    unsigned tp32[PX], shift[PX];
#pragma unroll
    for (int j = 0; j < PX; ++j) {
        const int a = ta[j] - g0 + kStagePad;
        tp32[j] = sbase + (unsigned)kStagedHeader + (unsigned)(a & ~3);
        shift[j] = (unsigned)(a & 3) * 8u;
    }
    const int kMagicI = 0x4B400000;
    const float kMagicF = 12582912.0f;

    auto fetch = [&](unsigned roff, unsigned (&wv)[PX][CN + 1]) {
#pragma unroll
        for (int j = 0; j < PX; ++j) {
            const unsigned q = tp32[j] + roff;
            wv[j][0] = lds_u32<0>(q);
            wv[j][1] = lds_u32<4>(q);
            if (CN >= 2) wv[j][CN >= 2 ? 2 : 0] = lds_u32<8>(q);
            if (CN >= 3) wv[j][CN >= 3 ? 3 : 0] = lds_u32<12>(q);
            if (CN >= 4) wv[j][CN >= 4 ? 4 : 0] = lds_u32<16>(q);
        }
    };
    // exact integer sums of the fetched row: byte permutes gather a channel's taps, two dp2a sum them on top of the
    // bits of 1.5*2^23, one FADD finishes the int->float conversion
    auto reduce = [&](const unsigned (&wv)[PX][CN + 1], HRow<CN> &R) {
        float hf[VC];
#pragma unroll
        for (int j = 0; j < PX; ++j) {
            unsigned A[CN];
#pragma unroll
            for (int k = 0; k < CN; ++k) A[k] = __funnelshift_r(wv[j][k], wv[j][k + 1], shift[j]);
#pragma unroll
            for (int c = 0; c < CN; ++c) {
                unsigned Tp;
                if (c == 0) Tp = gather_taps<CN, 0>(A);
                else if (c == 1) Tp = gather_taps<CN, (CN > 1 ? 1 : 0)>(A);
                else if (c == 2) Tp = gather_taps<CN, (CN > 2 ? 2 : 0)>(A);
                else Tp = gather_taps<CN, (CN > 3 ? 3 : 0)>(A);
                hf[j * CN + c] = __fadd_rn(__int_as_float(dp2a_hi_su(c23[j], Tp, dp2a_lo_su(c01[j], Tp, kMagicI))), -kMagicF);
            }
        }
#pragma unroll
        for (int k = 0; k < NP; ++k) R.p[k] = make_float2(hf[2 * k], hf[2 * k + 1]);
    };

    const int simd_end = (dw * CN) & ~7;
    unsigned char *q0 = dst + (((size_t)n * dh + y0) * dw + xs[0]) * CN;       // column 0's pixel of the next output row
    const size_t opitch = (size_t)dw * CN;
    const int qstep = T * CN;                                                   // to column 1's pixel
    const float2 one2 = make_float2(one, one);
    auto emit = [&](int yl, const HRow<CN> &R0, const HRow<CN> &R1, const HRow<CN> &R2, const HRow<CN> &R3) {
        const float4 b = lds_f4(a_ybf + 16u * (unsigned)yl);
        float res[VC];
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            float2 acc = pk_mul(R3.p[k], make_float2(b.w, b.w));
            acc = pk_add(pk_mul(R2.p[k], make_float2(b.z, b.z)), one2, acc);
            acc = pk_add(pk_mul(R1.p[k], make_float2(b.y, b.y)), one2, acc);
            acc = pk_add(pk_mul(R0.p[k], make_float2(b.x, b.x)), one2, acc);
            res[2 * k] = acc.x;
            res[2 * k + 1] = acc.y;
        }
        unsigned out[VC];
#pragma unroll
        for (int v = 0; v < VC; ++v) out[v] = f2u8_rn_sat(res[v]);
        if (TAIL) {
            const int4 bi = lds_i4(a_ybi + 16u * (unsigned)yl);
#pragma unroll
            for (int v = 0; v < VC; ++v) {
                if (xs[v / CN] * CN + v % CN >= simd_end) {     // the horizontal sums are exact integers below 2^22
                    const float h0 = (v & 1) ? R0.p[v / 2].y : R0.p[v / 2].x, h1 = (v & 1) ? R1.p[v / 2].y : R1.p[v / 2].x,
                                h2 = (v & 1) ? R2.p[v / 2].y : R2.p[v / 2].x, h3 = (v & 1) ? R3.p[v / 2].y : R3.p[v / 2].x;
                    const int t = ((int)h0 * bi.x + (int)h1 * bi.y + (int)h2 * bi.z + (int)h3 * bi.w + (1 << 21)) >> 22;
                    out[v] = (unsigned)max(0, min(255, t));
                }
            }
        }
#pragma unroll
        for (int c = 0; c < CN; ++c) q0[c] = (unsigned char)out[c];
        if (valid1) {
#pragma unroll
            for (int c = 0; c < CN; ++c) q0[qstep + c] = (unsigned char)out[CN + c];
        }
        q0 += opitch;
    };

    HRow<CN> Ra, Rb, Rc, Rd;
    int yl = 0, u = u_first;
    int next_last = ylast[0];                   // source row that completes the window of output row yl
    unsigned roff = 0;                          // staged row of u
    unsigned wv[PX][CN + 1];
    fetch(roff, wv);
    // one source row per step: its sums go to S3 (the words of the NEXT row are requested before anything else uses
    // them), then every output row whose fourth tap row this is gets emitted
#define TFPD_STAGED_STEP(S0, S1, S2, S3)                                              \
    {                                                                                  \
        reduce(wv, S3);                                                                \
        if (u < u_last) roff += (unsigned)stride;                                      \
        fetch(roff, wv);                                                               \
        while (next_last == u) {                                                       \
            emit(yl, S0, S1, S2, S3);                                                  \
            ++yl;                                                                      \
            next_last = yl < th ? (int)lds_u32<0>(a_ylast + 4u * (unsigned)yl) : 0x7fffffff; \
        }                                                                              \
        if (u == u_last) break;                                                        \
        ++u;                                                                           \
    }
    for (;;) {
        TFPD_STAGED_STEP(Ra, Rb, Rc, Rd)
        TFPD_STAGED_STEP(Rb, Rc, Rd, Ra)
        TFPD_STAGED_STEP(Rc, Rd, Ra, Rb)
        TFPD_STAGED_STEP(Rd, Ra, Rb, Rc)
    }
#undef TFPD_STAGED_STEP
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
    // staged strip form: needs 4-byte aligned source rows (its edge loads are words) and a window that fits in
    // shared memory; otherwise the plain strip form
    if (path == 0 && ((reinterpret_cast<uintptr_t>(src) | (size_t)sw * cn) & 3) == 0 && scale_x < 64.0 && scale_y < 64.0) {
        static const int ty_env = getenv("TFPD_RESIZE_TY") ? atoi(getenv("TFPD_RESIZE_TY")) : 0;
        static const int th_env = getenv("TFPD_RESIZE_THREADS") ? atoi(getenv("TFPD_RESIZE_THREADS")) : 0;
        static int sm_of_dev[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        int &sms = sm_of_dev[dev & 63];
        if (sms == 0 && cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
        // CTA width t (a thread computes columns x and x + t) and strip height ty, by a model of the kernel's time:
        // waves of CTAs x the work of one CTA x the CTAs that share an SM.  Per thread a source row costs ~56
        // instructions, an output row ~41 (ncu r03c); taller strips re-do fewer source rows, shorter ones leave a
        // fuller last wave when there are only one or two.
        int threads = 0, ty = 0, stride = 0, rows = 0;
        double best = 0.0;
        for (int t = 64; t <= 256; t += 32) {
            if (th_env >= 32 && th_env <= 256 && th_env % 32 == 0 && t != th_env) continue;
            const int xt = (dw + kStagedPx * t - 1) / (kStagedPx * t);
            const int wcols = (int)ceil((kStagedPx * t - 1) * scale_x) + 6;
            const int st = (wcols * cn + 15 + kStagePad + 4 + 15) / 16 * 16;
            for (int y = 4; y <= kStagedMaxTy; ++y) {
                if (ty_env >= 4 && ty_env <= kStagedMaxTy && y != ty_env) continue;
                const int rw = (int)ceil((y - 1) * scale_y) + 6;
                const size_t sm_bytes = (size_t)st * rw + kStagedHeader;
                if (sm_bytes > 96 * 1024) continue;
                int per_sm = (int)((227 * 1024) / (sm_bytes + 1024));
                per_sm = per_sm < 2048 / t ? per_sm : 2048 / t;          // resident threads
                per_sm = per_sm < 65536 / (t * 64) ? per_sm : 65536 / (t * 64);   // 64 registers per thread
                if (per_sm < 1) continue;
                const double ctas = (double)xt * ((dh + y - 1) / y) * N;
                // whole waves only where there are few of them (small batches: 16 frames ran 25 % faster on strips
                // that fill exactly one wave); with several waves the measured time follows the total work
                // (64 frames: 16-row strips 39 us, 14-row strips at 2.92 waves 41 us)
                const double wv_exact = ctas / ((double)sms * per_sm), waves = wv_exact < 2.0 ? ceil(wv_exact) : wv_exact;
                const double work = ((y - 1) * scale_y + 4.0) * 56.0 + y * 41.0 + 250.0;
                // idle lanes of the last column tile and few resident warps (latency no longer hidden) cost time too
                const double lanes = (double)xt * kStagedPx * t / dw;
                const double warps = per_sm * (t / 32.0);
                const double cost = waves * work * per_sm * t * lanes * (warps < 24.0 ? 24.0 / warps : 1.0);
                if (threads == 0 || cost < best) best = cost, threads = t, ty = y, stride = st, rows = rw;
            }
        }
        if (threads != 0) {
            const size_t st_smem = (size_t)stride * rows + kStagedHeader;
            const dim3 grid((dw + kStagedPx * threads - 1) / (kStagedPx * threads), (dh + ty - 1) / ty, N);
#define TFPD_STAGED_CASE_T(CN, TAIL)                                                                                   \
    {                                                                                                                  \
        static size_t attr_set[64] = {}; /* largest size allowed so far, per device (function attributes are) */     \
        if (st_smem > 48 * 1024 && st_smem > attr_set[dev & 63]) {                                                     \
            if (cudaFuncSetAttribute(k_resize_cubic_u8_staged<CN, TAIL>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                     (int)st_smem) != cudaSuccess)                                                    \
                return -1;                                                                                             \
            attr_set[dev & 63] = st_smem;                                                                              \
        }                                                                                                              \
        k_resize_cubic_u8_staged<CN, TAIL><<<grid, threads, st_smem, s>>>(src, dst, sh, sw, dh, dw, scale_x, scale_y,  \
                                                                          stride, rows, 1.0f, ty);                     \
    }
#define TFPD_STAGED_CASE(CN)                                                                                           \
    if ((dw * CN) & 7) TFPD_STAGED_CASE_T(CN, true) else TFPD_STAGED_CASE_T(CN, false)
            switch (cn) {
                case 1: TFPD_STAGED_CASE(1) break;
                case 2: TFPD_STAGED_CASE(2) break;
                case 3: TFPD_STAGED_CASE(3) break;
                case 4: TFPD_STAGED_CASE(4) break;
                default: return -1;
            }
#undef TFPD_STAGED_CASE
#undef TFPD_STAGED_CASE_T
            return cudaGetLastError() == cudaSuccess ? 1 : -1;
        }
    }
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
