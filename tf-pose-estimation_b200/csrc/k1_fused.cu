// k1_fused.cu -- K1: fused channel split + x4 area upsample + 25x25 smoother + 3x3 NMS +
// peak records, the reference's TF graph ops at tf_pose/estimator.py:321-337 and
// tf_pose/tensblur/smoother.py:62-95, for the power-of-two factor every reference caller
// uses (factor 4: run.py:30, estimator.py:523).
//
// What the reference computes                       how this kernel gets it
//   heat_up/paf_up = resize_area(x4)                 pixel replication (bit-exact for f=4,
//     (estimator.py:326-329)                          SURVEY.md A.1): one low-res row chunk is
//                                                    expanded ONCE in shared memory and written
//                                                    to its 4 output rows by TMA bulk stores
//   g = depthwise 25x25 conv SAME (smoother.py:91)   separable (K = s s^T) and polyphase: a
//                                                    25-tap window over a x4-replicated signal
//                                                    is 7 low-res taps with phase weights, so
//                                                    g costs 7 + 7/4 FMA per pixel, not 625
//   peaks = where(g == maxpool3x3(g), g, 0)          registers + warp shuffles; the dense tensor
//     (estimator.py:335-337)                          row is a zeroed smem row with the sparse
//                                                    maxima poked in, then a TMA bulk store
//   peak listing (pafprocess.cpp:24-37)              (part,y,x) keyed records appended per frame,
//                                                    ordered afterwards by k_sort_peaks
//
// CTA = 19 "blur" warps (one heat channel each, lanes = low-res cells, marching down the
// output rows with a 7-row register window) + 5 "copy" warps (replication + staging).  The
// two roles never synchronise with each other; each has its own smem ring and named barrier.
// Work item = (frame, band of low-res rows, chunk of <=56 cells); grid = SM count, static
// stride over items.  Nothing here is a dense contraction: no tensor cores (north_star).
#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

constexpr int CPL = 2;                      // low-res cells per lane
constexpr int kBlurWarps = kHeatC;          // 19
constexpr int kCopyWarps = 5;
constexpr int kBlurThreads = kBlurWarps * 32;
constexpr int kCopyThreads = kCopyWarps * 32;
constexpr int kThreads = kBlurThreads + kCopyThreads;   // 768
// a warp pass owns at most 32*CPL-8 = 56 cells (see launch_k1_fused_f4: chunk_cells <= 56)
constexpr int NPX = 4 * CPL;                // output pixels per lane per row

__constant__ float c_wp[4][8];              // polyphase weights: phase p, low-res tap t (cell offset t-3)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void named_bar(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__host__ __device__ constexpr int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

struct Item {
    int n, iy0, nrows, c0, tc;
};
__device__ __forceinline__ Item decode_item(const K1Params &p, int item) {
    Item it;
    int per_frame = p.nbands * p.nchunks;
    it.n = item / per_frame;
    int rem = item - it.n * per_frame;
    int band = rem / p.nchunks;
    int chunk = rem - band * p.nchunks;
    it.iy0 = band * p.band_rows;
    it.nrows = min(p.band_rows, p.h - it.iy0);
    it.c0 = chunk * p.chunk_cells;
    it.tc = min(p.chunk_cells, p.w - it.c0);
    return it;
}

// ------------------------------------------------------------------------------------------
// copy role: heat_up / paf_up rows of one item.  One low-res row chunk -> smem (already in the
// NHWC order of an upsampled row) -> 4 + 4 bulk stores.  Two staging slots: while slot A drains
// through the TMA engine, slot B is being filled.
// ------------------------------------------------------------------------------------------
__device__ void copy_role(const K1Params &p, const Item &it, float *cslots, int slot_floats, int ctid,
                          unsigned &phase) {
    const int H = 4 * p.h, W = 4 * p.w;
    const int tc = it.tc;
    for (int r = 0; r < it.nrows; ++r) {
        const int iy = it.iy0 + r;
        float *sh = cslots + (size_t)(phase & 1u) * slot_floats;   // heat part: tc*76 floats
        float *sp = sh + tc * 4 * kHeatC;                          // paf part:  tc*152 floats
        ++phase;
        const float *src = p.X + ((size_t)(it.n * p.h + iy) * p.w + it.c0) * kNetC;
        // a cell of the upsampled row is its 19 (38) channel values repeated 4 times
        for (int q = ctid; q < tc * kHeatC; q += kCopyThreads) {
            int cell = q / kHeatC, m = q - cell * kHeatC;
            const float *b = src + cell * kNetC;
            int ch = 4 * m;
            float4 v;
            v.x = __ldg(b + (ch % kHeatC));
            v.y = __ldg(b + ((ch + 1) % kHeatC));
            v.z = __ldg(b + ((ch + 2) % kHeatC));
            v.w = __ldg(b + ((ch + 3) % kHeatC));
            reinterpret_cast<float4 *>(sh)[q] = v;
        }
        for (int q = ctid; q < tc * kPafC; q += kCopyThreads) {
            int cell = q / kPafC, m = q - cell * kPafC;
            const float *b = src + cell * kNetC + kHeatC;
            int ch = 4 * m;
            float4 v;
            v.x = __ldg(b + (ch % kPafC));
            v.y = __ldg(b + ((ch + 1) % kPafC));
            v.z = __ldg(b + ((ch + 2) % kPafC));
            v.w = __ldg(b + ((ch + 3) % kPafC));
            reinterpret_cast<float4 *>(sp)[q] = v;
        }
        fence_proxy_async();
        if (ctid == 0) bulk_wait_read0();    // the OTHER slot has drained -> it may be refilled after the barrier
        named_bar(2, kCopyThreads);
        if (ctid == 0) {
            const uint32_t hb = (uint32_t)tc * 4 * kHeatC * 4, pb = (uint32_t)tc * 4 * kPafC * 4;
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                size_t pix = ((size_t)it.n * H + (4 * iy + rr)) * W + 4 * it.c0;
                bulk_store(p.heat_up + pix * kHeatC, sh, hb);
                bulk_store(p.paf_up + pix * kPafC, sp, pb);
            }
            bulk_commit();
        }
    }
}

// ------------------------------------------------------------------------------------------
// blur role: warp = heat channel c, lane l owns low-res cells cb + 2l, cb + 2l + 1 (cb = c0-4).
// ------------------------------------------------------------------------------------------
template <int PR, bool DENSE>
__device__ void blur_role(const K1Params &p, const Item &it, float *pslots, int pslot_floats, int c, int lane,
                          unsigned &pgroup, uint32_t (&wmask)[2]) {
    const int h = p.h, w = p.w, H = 4 * h, W = 4 * w;
    const int tc = it.tc, c0 = it.c0, n = it.n;
    const int pitch_px = 4 * p.chunk_cells;     // slot row pitch: independent of the item's width
    const float NEG = -__int_as_float(0x7f800000);
    const unsigned FULL = 0xffffffffu;

    int cell[CPL];
    bool cin[CPL], own[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        cell[k] = c0 - 4 + lane * CPL + k;
        cin[k] = (cell[k] >= 0) && (cell[k] < w);
        own[k] = (cell[k] >= c0) && (cell[k] < c0 + tc);
    }
    const float *Xn = p.X + (size_t)n * h * w * kNetC + c;
    auto hv = [&](int r, int k) -> float {
        return (cin[k] && r >= 0 && r < h) ? __ldg(Xn + ((size_t)r * w + cell[k]) * kNetC) : 0.0f;
    };

    // 7-row register window of low-res heat values around the current low-res row iy
    float win[CPL][7], nxt[CPL];
    int iy = it.iy0 - 1;
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
#pragma unroll
        for (int t = 0; t < 7; ++t) win[k][t] = hv(iy - 3 + t, k);
        nxt[k] = hv(iy + 4, k);
    }
    float gp2[NPX], gp1[NPX], hc1[CPL];
#pragma unroll
    for (int i = 0; i < NPX; ++i) gp2[i] = gp1[i] = NEG;
#pragma unroll
    for (int k = 0; k < CPL; ++k) hc1[k] = 0.0f;

    const int Ys = 4 * it.iy0 - 1, Ye = 4 * (it.iy0 + it.nrows);   // rows whose g is needed
    const float thr = p.thresh;
    const float thr_lo = thr * (1.0f - 4e-5f);

    for (int grp = 0; grp < it.nrows + 2; ++grp, ++iy) {
#pragma unroll
        for (int py = 0; py < 4; ++py) {
            const int Y = 4 * iy + py;
            if (Y < Ys || Y > Ye) continue;     // warp-uniform: only the first/last group are partial
            float g[NPX];
            if (Y >= 0 && Y < H) {
                // vertical polyphase: gv[cell] = sum_t wp[py][t] * X[iy-3+t, cell, c]
                float gv[CPL];
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    float a = c_wp[py][0] * win[k][0];
#pragma unroll
                    for (int t = 1; t < 7; ++t) a = fmaf(c_wp[py][t], win[k][t], a);
                    gv[k] = a;
                }
                // window of gv over cells (own first cell - 3 .. own last cell + 3) via shuffles
                float Wn[CPL + 6];
#pragma unroll
                for (int j = 0; j < CPL + 6; ++j) {
                    const int d = j - 3;
                    const int dl = floordiv(d, CPL);
                    const int kk = d - dl * CPL;
                    if (dl == 0) Wn[j] = gv[kk];
                    else if (dl < 0) Wn[j] = __shfl_up_sync(FULL, gv[kk], -dl);
                    else Wn[j] = __shfl_down_sync(FULL, gv[kk], dl);
                }
                // horizontal polyphase: g[cell,px] = sum_t wp[px][t] * gv[cell-3+t]
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
#pragma unroll
                    for (int px = 0; px < 4; ++px) {
                        float a = c_wp[px][0] * Wn[k];
#pragma unroll
                        for (int t = 1; t < 7; ++t) a = fmaf(c_wp[px][t], Wn[k + t], a);
                        g[k * 4 + px] = cin[k] ? a : NEG;
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < NPX; ++i) g[i] = NEG;
            }

            // 3x3 NMS of row Yc = Y-1: vertical max in registers, horizontal neighbours by shuffle
            const int Yc = Y - 1;
            const int j = Yc - 4 * it.iy0;     // candidate row inside the band
            if (j >= 0 && j < 4 * it.nrows) {  // warp-uniform
                float v[NPX + 2];
#pragma unroll
                for (int i = 0; i < NPX; ++i) v[i + 1] = fmaxf(fmaxf(gp2[i], gp1[i]), g[i]);
                v[0] = __shfl_up_sync(FULL, v[NPX], 1);
                v[NPX + 1] = __shfl_down_sync(FULL, v[1], 1);

                float *slot = nullptr;
                int jr = 0;
                if (DENSE) {
                    jr = j % PR;
                    const int sidx = (int)((pgroup + (unsigned)(j / PR)) & 1u);
                    slot = pslots + (size_t)sidx * pslot_floats;
                    if (jr == 0 && wmask[sidx]) {
                        // lazily clear what this lane poked into the slot two groups ago
                        uint32_t mk = wmask[sidx];
                        while (mk) {
                            int b = __ffs(mk) - 1;
                            mk &= mk - 1;
                            int rr = b / NPX, i = b - rr * NPX;
                            int X = 4 * cell[i >> 2] + (i & 3);
                            slot[((size_t)rr * pitch_px + (X - 4 * c0)) * kHeatC + c] = 0.0f;
                        }
                        wmask[sidx] = 0;
                    }
                }
#pragma unroll
                for (int i = 0; i < NPX; ++i) {
                    const int k = i >> 2;
                    if (!own[k]) continue;
                    const float m = fmaxf(fmaxf(v[i], v[i + 1]), v[i + 2]);
                    const float gc = gp1[i];
                    const int X = 4 * cell[k] + (i & 3);
                    if (gc == m) {
                        if (DENSE) {
                            slot[((size_t)jr * pitch_px + (X - 4 * c0)) * kHeatC + c] = gc;
                            wmask[(pgroup + (unsigned)(j / PR)) & 1u] |= 1u << (jr * NPX + i);
                        }
                        if (c < kParts && gc > thr) {
                            int idx = atomicAdd(&p.counters[n].n_peak_rec, 1);
                            if (idx < p.max_peaks) {
                                p.rec_key[(size_t)n * p.max_peaks + idx] = peak_key(c, Yc, X);
                                p.rec_val[(size_t)n * p.max_peaks + idx] = hc1[k];
                            }
                        }
                        if (c < kParts && fabsf(gc - thr) <= 2e-5f * thr) atomicAdd(&p.counters[n].near_ties, 1);
                    } else if (c < kParts && gc > thr_lo && (m - gc) <= 2e-5f * fabsf(m)) {
                        atomicAdd(&p.counters[n].near_ties, 1);     // lost the 3x3 max by a hair
                    }
                }
                if (DENSE && (jr == PR - 1 || j == 4 * it.nrows - 1)) {
                    // the slot's rows are complete: hand them to the TMA engine
                    fence_proxy_async();
                    if (c == 0 && lane == 0) bulk_wait_read0();     // previous slot drained
                    named_bar(1, kBlurThreads);
                    if (c == 0 && lane == 0) {
                        const int jfirst = j - jr;
                        const uint32_t bytes = (uint32_t)tc * 4 * kHeatC * 4;
                        for (int rr = 0; rr <= jr; ++rr) {
                            size_t pix = ((size_t)n * H + (4 * it.iy0 + jfirst + rr)) * W + 4 * c0;
                            bulk_store(p.peaks + pix * kHeatC, slot + (size_t)rr * pitch_px * kHeatC, bytes);
                        }
                        bulk_commit();
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < NPX; ++i) {
                gp2[i] = gp1[i];
                gp1[i] = g[i];
            }
#pragma unroll
            for (int k = 0; k < CPL; ++k) hc1[k] = win[k][3];
        }
        // slide the low-res window down one row; prefetch the row after next
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
#pragma unroll
            for (int t = 0; t < 6; ++t) win[k][t] = win[k][t + 1];
            win[k][6] = nxt[k];
            nxt[k] = hv(iy + 5, k);
        }
    }
    if (DENSE) pgroup += (unsigned)((4 * it.nrows + PR - 1) / PR);
}

template <int PR, bool DENSE>
__global__ void __launch_bounds__(kThreads, 1) k1_fused_f4(const K1Params p) {
    extern __shared__ __align__(128) float smem[];
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int pslot_floats = PR * p.chunk_cells * 4 * kHeatC;
    const int cslot_floats = p.chunk_cells * 4 * kNetC;
    float *pslots = smem;                                 // 2 peak-row slots (blur role)
    float *cslots = smem + 2 * (size_t)pslot_floats;      // 2 replication slots (copy role)

    if (warp < kBlurWarps) {
        if (DENSE) {
            float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int q = tid; q < 2 * pslot_floats / 4; q += kBlurThreads) reinterpret_cast<float4 *>(pslots)[q] = z;
            named_bar(1, kBlurThreads);
        }
        unsigned pgroup = 0;
        uint32_t wmask[2] = {0u, 0u};
        for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
            Item it = decode_item(p, item);
            blur_role<PR, DENSE>(p, it, pslots, pslot_floats, warp, lane, pgroup, wmask);
        }
        if (DENSE && tid == 0) bulk_wait_all();
    } else {
        if (DENSE) {
            const int ctid = tid - kBlurThreads;
            unsigned phase = 0;
            for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                Item it = decode_item(p, item);
                copy_role(p, it, cslots, cslot_floats, ctid, phase);
            }
            if (ctid == 0) bulk_wait_all();
        }
    }
}

}  // namespace

int upload_polyphase_weights(const double *s1d, int ksize) {
    if (ksize != 25) return -1;
    float wp[4][8];
    for (int p = 0; p < 4; ++p) {
        double acc[7] = {0, 0, 0, 0, 0, 0, 0};
        for (int q = 0; q < 25; ++q) {
            int d = p + q - 12;                     // upsampled column offset relative to the cell start
            int cellofs = (d >= 0) ? d / 4 : -((-d + 3) / 4);
            acc[cellofs + 3] += s1d[q];
        }
        for (int t = 0; t < 7; ++t) wp[p][t] = (float)acc[t];
        wp[p][7] = 0.0f;
    }
    return cudaMemcpyToSymbol(c_wp, wp, sizeof(wp)) == cudaSuccess ? 0 : -2;
}

int launch_k1_fused_f4(const K1Params &p, bool dense, int sm_count, cudaStream_t s) {
    const int grid = min(p.nitems, sm_count);
    if (grid <= 0) return 0;
    if (!dense) {
        k1_fused_f4<4, false><<<grid, kBlurThreads, 0, s>>>(p);
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    const size_t cslot = (size_t)p.chunk_cells * 4 * kNetC * 4;
    const size_t pslot4 = (size_t)4 * p.chunk_cells * 4 * kHeatC * 4;
    const size_t limit = 227 * 1024 - 1024;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k1_fused_f4<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        cudaFuncSetAttribute(k1_fused_f4<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        attr_set = true;
    }
    if (2 * pslot4 + 2 * cslot <= limit) {
        k1_fused_f4<4, true><<<grid, kThreads, 2 * pslot4 + 2 * cslot, s>>>(p);
    } else {
        k1_fused_f4<2, true><<<grid, kThreads, pslot4 + 2 * cslot, s>>>(p);
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
