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
// output rows with a 7-row register window) + 1 "copy" warp (replication + staging).  The
// two roles never synchronise with each other; each has its own smem buffers.
// Scheduling: the (frame, chunk of <=56 cells, low-res row) space is flattened and cut into
// gridDim.x equal contiguous spans, one per CTA (grid = SM count): every SM gets the same
// number of rows (+-1) and pays the 8-row window prologue once per span / frame boundary.
// Nothing here is a dense contraction: no tensor cores (north_star).
#include <type_traits>

#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

constexpr int CPL = 2;                      // low-res cells per lane
constexpr int kBlurWarps = kHeatC;          // 19
constexpr int kCopyWarps = 1;                // dedicated copy warps (20 warps = 5 per SM sub-partition -> 96 registers)
constexpr int kHelperWarps = 2;              // blur warps that turn into copy warps once the blur role is done
constexpr int kBlurThreads = kBlurWarps * 32;
constexpr int kCopyThreads = kCopyWarps * 32;
constexpr int kThreads = kBlurThreads + kCopyThreads;   // 640 -> 96 registers per thread
// minimal mode (blur role alone): small CTAs, several per SM -- 6 warps x 4 CTAs = 24 warps per SM at <= 85 registers
#ifndef TFPD_MIN_WARPS
#define TFPD_MIN_WARPS 6                    // (a build with 3 warps x 8 CTAs per SM was measured against this one: profiles/README.md)
#endif
constexpr int kMinWarps = TFPD_MIN_WARPS;
constexpr int kMinThreads = kMinWarps * 32;
#ifndef TFPD_MIN_SM_WARPS
#define TFPD_MIN_SM_WARPS 24                // minimal-mode warps per SM the kernel is compiled for (24 -> 74 registers)
#endif
constexpr int kMinCtasPerSm = TFPD_MIN_SM_WARPS / kMinWarps;
// a warp pass owns at most 32*CPL-8 = 56 cells (see launch_k1_fused_f4: chunk_cells <= 56)
constexpr int NPX = 4 * CPL;                // output pixels per lane per row


__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
}
// same with an L2 eviction-priority hint (the dense outputs are written once and never re-read here)
__device__ __forceinline__ void bulk_store_hint(void *gdst, const void *ssrc, uint32_t bytes, uint64_t policy) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst),
                 "r"(smem_u32(ssrc)), "r"(bytes), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n"
        " bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// TMA bulk load global -> shared, completion counted in bytes on an mbarrier (16-byte aligned src/dst/size)
__device__ __forceinline__ void bulk_load(void *sdst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void named_bar(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// bytes of one raw input chunk buffer: the 16-byte aligned superset of copy_chunk cells, 128-byte padded
__host__ __device__ constexpr int k1_raw_bytes(int copy_chunk) { return ((copy_chunk * kNetC * 4 + 32) + 127) / 128 * 128; }

struct Item {
    int n, iy0, nrows, c0, tc;
};
// contiguous share [lo, hi) of `total` units for CTA b of g
__device__ __forceinline__ void span_of(long long total, int b, int g, long long &lo, long long &hi) {
    lo = total * b / g;
    hi = total * (b + 1) / g;
}

// ------------------------------------------------------------------------------------------
// copy role (one warp): heat_up / paf_up.  Item = one low-res row chunk (frame, iy, chunk of
// <= 32 cells).  Its tc*57 floats come in by ONE TMA bulk load (the 16-byte aligned superset of
// the chunk; the next item's load is in flight while this one is expanded).  Lane l owns cell l:
// it reads its 57 values (stride 57 words: conflict-free) and writes the cell's x4 replicated
// blocks -- 19 float4 of heat, 38 float4 of PAF, all register indices static -- into a staging
// slot that already has the NHWC layout of an upsampled row; the slot then goes out as 4 + 4 TMA
// bulk stores (x4 along Y).  Two raw buffers, two slots.
// ------------------------------------------------------------------------------------------
struct CopyItem {
    int n, iy, c0, tc;
    size_t src_off;     // byte offset of the chunk inside X
};
__device__ __forceinline__ CopyItem copy_item(const K1Params &p, long long R) {
    CopyItem ci;
    const long long grow = R / p.copy_nchunks;                       // grow = n*h + iy
    const int chunk = (int)(R - grow * p.copy_nchunks);
    ci.c0 = chunk * p.copy_chunk;
    ci.tc = min(p.copy_chunk, p.w - ci.c0);
    ci.n = (int)(grow / p.h);
    ci.iy = (int)(grow - (long long)ci.n * p.h);
    ci.src_off = ((size_t)grow * p.w + ci.c0) * (kNetC * 4);
    return ci;
}
// lane 0: start the bulk load of item R into raw (bytes are re-derived by the waiter)
__device__ __forceinline__ void copy_prefetch(const K1Params &p, long long R, unsigned char *raw, uint64_t *bar) {
    const CopyItem ci = copy_item(p, R);
    const size_t total = (size_t)p.N * p.h * p.w * (kNetC * 4);
    const size_t a0 = ci.src_off & ~(size_t)15;
    size_t a1 = (ci.src_off + (size_t)ci.tc * (kNetC * 4) + 15) & ~(size_t)15;
    if (a1 > total) a1 = total & ~(size_t)15;      // never read past the tensor; the tail comes by LDG
    const uint32_t bytes = (uint32_t)(a1 - a0);
    fence_proxy_async();                           // earlier generic reads of `raw` precede the async write
    mbar_expect_tx(bar, bytes);
    bulk_load(raw, reinterpret_cast<const unsigned char *>(p.X) + a0, bytes, bar);
}

// One copy warp.  Items come from a grid-wide queue (atomic counter in the launch's first FrameCounters):
// per-CTA finish times of a static split spread by +-20 % (measured, tools/k1_cta_times.py) because the
// warp's pace is set by how fast ITS stores drain, so the split has to follow the pace.  The warp always
// holds two grabbed items: the one it expands and the one whose bulk load is in flight.
__device__ void copy_role(const K1Params &p, float *cslots, int slot_floats, unsigned char *raws, int raw_bytes,
                          uint64_t *bars, int lane) {
    const int H = 4 * p.h, W = 4 * p.w;
    const size_t total_bytes = (size_t)p.N * p.h * p.w * (kNetC * 4);
    const long long Rend = (long long)p.N * p.h * p.copy_nchunks;
    int *queue = &p.counters[0].k1_queue;
    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const uint64_t pol_first = policy_evict_first();
    long long R = 0, Rn = 0;
    if (lane == 0) {
        R = atomicAdd(queue, 1);
        Rn = atomicAdd(queue, 1);
        if (R < Rend) copy_prefetch(p, R, raws, &bars[0]);
    }
    R = __shfl_sync(0xffffffffu, R, 0);
    Rn = __shfl_sync(0xffffffffu, Rn, 0);
    unsigned k = 0;
    for (; R < Rend; ++k) {
        const unsigned b = k & 1u;
        long long Rnn = 0;
        if (lane == 0) {
            if (Rn < Rend) copy_prefetch(p, Rn, raws + (size_t)(b ^ 1u) * raw_bytes, &bars[b ^ 1u]);
            Rnn = atomicAdd(queue, 1);                     // consumed at the end of this iteration
        }
        const CopyItem ci = copy_item(p, R);
        float *sh = cslots + (size_t)b * slot_floats;      // heat part: tc*76 floats
        float *sp = sh + ci.tc * 4 * kHeatC;               // paf part:  tc*152 floats
        if (lane == 0) bulk_wait_read1();                  // the stores that last used THIS slot have drained
        mbar_wait(&bars[b], (k >> 1) & 1u);                // the chunk has landed in raw[b]
        __syncwarp();
        const unsigned char *rawb = raws + (size_t)b * raw_bytes;
        const float *src = reinterpret_cast<const float *>(rawb + (ci.src_off & 15));
        const int ne = ci.tc * kNetC;
        // floats a clamped load left out (only at the very end of a tensor whose size is not a multiple of 16 B)
        const size_t a0 = ci.src_off & ~(size_t)15;
        const size_t got = ((total_bytes & ~(size_t)15) - a0);
        const int ne_loaded = (int)min((size_t)ne, (got - (ci.src_off & 15)) / 4);
        if (lane < ci.tc) {
            const float *sc = src + lane * kNetC;
            const bool whole = (lane + 1) * kNetC <= ne_loaded;
            const float *gc = p.X + ci.src_off / 4 + lane * kNetC;
            {
                float v[kHeatC];
#pragma unroll
                for (int c = 0; c < kHeatC; ++c) v[c] = sc[c];
                if (!whole) {      // rare: the clamped load left part of this cell out
#pragma unroll
                    for (int c = 0; c < kHeatC; ++c) v[c] = __ldg(gc + c);
                }
                float4 *d = reinterpret_cast<float4 *>(sh + lane * (4 * kHeatC));
#pragma unroll
                for (int j = 0; j < kHeatC; ++j)
                    d[j] = make_float4(v[(4 * j) % kHeatC], v[(4 * j + 1) % kHeatC], v[(4 * j + 2) % kHeatC],
                                       v[(4 * j + 3) % kHeatC]);
            }
            {
                float u[kPafC];
#pragma unroll
                for (int c = 0; c < kPafC; ++c) u[c] = sc[kHeatC + c];
                if (!whole) {
#pragma unroll
                    for (int c = 0; c < kPafC; ++c) u[c] = __ldg(gc + kHeatC + c);
                }
                float4 *d = reinterpret_cast<float4 *>(sp + lane * (4 * kPafC));
#pragma unroll
                for (int j = 0; j < kPafC; ++j)
                    d[j] = make_float4(u[(4 * j) % kPafC], u[(4 * j + 1) % kPafC], u[(4 * j + 2) % kPafC],
                                       u[(4 * j + 3) % kPafC]);
            }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
            const uint32_t hb = (uint32_t)ci.tc * 4 * kHeatC * 4, pb = (uint32_t)ci.tc * 4 * kPafC * 4;
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                size_t pix = ((size_t)ci.n * H + (4 * ci.iy + rr)) * W + 4 * ci.c0;
                if (p.l2_knobs & 2) {
                    bulk_store_hint(p.heat_up + pix * kHeatC, sh, hb, pol_first);
                    bulk_store_hint(p.paf_up + pix * kPafC, sp, pb, pol_first);
                } else {
                    bulk_store(p.heat_up + pix * kHeatC, sh, hb);
                    bulk_store(p.paf_up + pix * kPafC, sp, pb);
                }
            }
            bulk_commit();
        }
        R = Rn;
        Rn = __shfl_sync(0xffffffffu, Rnn, 0);
    }
}

// ------------------------------------------------------------------------------------------
// blur role: warp = heat channel c, lane l owns low-res cells cb + 2l, cb + 2l + 1 (cb = c0-4).
// ------------------------------------------------------------------------------------------
// A local maximum was found (rare: ~0.5 % of the pixels): dense poke + peak record + flags.
// nb = largest of the 8 neighbours (near-tie test from the winner's side).
__device__ __forceinline__ void emit_peak(const K1Params &p, int n, int c, int Yc, int X, float gc, float hc, float nb,
                                       float *slot_addr) {
    if (slot_addr) *slot_addr = gc;
    if (c < kParts) {
        const float thr = p.thresh;
        if (gc > thr) {                                   // PEAKS(y,x,part) > THRESH_HEAT, pafprocess.cpp:27
            int idx = atomicAdd(&p.counters[n].n_peak_rec, 1);
            if (idx < p.max_peaks) {
                p.rec_key[(size_t)n * p.max_peaks + idx] = peak_key(c, Yc, X);
                p.rec_val[(size_t)n * p.max_peaks + idx] = hc;     // score = un-blurred HEAT(y,x,part), :32
            }
        }
        const bool near_thr = fabsf(gc - thr) <= 2e-5f * thr;
        const bool near_nb = gc > thr * (1.0f - 4e-5f) && (gc - nb) <= 2e-5f * fabsf(gc);
        if (near_thr || near_nb) atomicAdd(&p.counters[n].near_ties, 1);
    }
}

// Packed FP32 (Blackwell): one FFMA2 / FMUL2 issue does two IEEE round-to-nearest FMAs.  ptxas broadcasts a scalar
// register or constant-bank operand to both halves by itself when both components of a float2 are the same value
// (SASS `R8.F32` / `UR9.F32` next to `.F32x2.HI_LO`), so no operand has to be duplicated in registers.  Every chain
// below keeps the t = 0..6 order of the scalar code it replaces: the blurred values are bit-identical to it.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b),
                       rc = *reinterpret_cast<unsigned long long *>(&c), rd;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b), rd;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 bcast2(float v) { return make_float2(v, v); }
// acc |= bit iff a == b, as a compare and a PREDICATED or (2 instructions per bit).  Written `acc |= c ? bit : 0`
// the compiler materialises the flag with a SEL first (3 per bit); sm_100a has no FSET with an integer result
// (set.eq.u32.f32 also becomes FSETP + SEL).
__device__ __forceinline__ uint32_t or_bit_if_eq(uint32_t acc, float a, float b, uint32_t bit) {
    asm("{\n .reg .pred q;\n setp.eq.f32 q, %1, %2;\n @q or.b32 %0, %0, %3;\n}" : "+r"(acc) : "f"(a), "f"(b), "r"(bit));
    return acc;
}

template <int PR, bool DENSE>
__device__ void blur_role(const K1Params &p, const Item &it, float *pslots, int pslot_floats, int c, int lane,
                          unsigned &pgroup, uint32_t &wmask0, uint32_t &wmask1) {
    const int h = p.h, w = p.w, H = 4 * h, W = 4 * w;
    const int tc = it.tc, c0 = it.c0, n = it.n;
    const int pitch_px = 4 * p.chunk_cells;     // slot row pitch: independent of the item's width
    const float NEG = -__int_as_float(0x7f800000);
    const unsigned FULL = 0xffffffffu;

    // lane -> its two low-res cells.  Per-lane border state is kept as VALUES, not predicates: nf is the start value
    // of a cell's horizontal accumulators (+0 inside the image, -inf outside: the 3x3 max then sees clipped border
    // windows, estimator.py:335), ownmask the pixels this lane decides.
    const int cell0 = c0 - 4 + lane * CPL;
    const bool cin0 = (cell0 >= 0) && (cell0 < w), cin1 = (cell0 + 1 >= 0) && (cell0 + 1 < w);
    const float2 nf0 = bcast2(cin0 ? 0.0f : NEG), nf1 = bcast2(cin1 ? 0.0f : NEG);
    uint32_t ownmask = 0;
    if ((cell0 >= c0) && (cell0 < c0 + tc)) ownmask |= 0x0fu;
    if ((cell0 + 1 >= c0) && (cell0 + 1 < c0 + tc)) ownmask |= 0xf0u;
    const int lx0 = 4 * (lane * CPL - 4);        // slot column of this lane's first pixel
    // low-res heat values of both cells in row r (zero outside the image: SAME zero padding, smoother.py:91)
    const float *Xc = p.X + (size_t)n * h * w * kNetC + c + (ptrdiff_t)cell0 * kNetC;
    // r * w * 57 fits 32 bits: h <= 2047 and w <= 4095 (the upsampled map is capped at 8191 x 16383, tfpd_create)
    const int rowstride = w * kNetC;
    auto hv2 = [&](int r) -> float2 {
        float2 v = make_float2(0.0f, 0.0f);
        if (r >= 0 && r < h) {
            const float *row = Xc + r * rowstride;
            if (cin0) v.x = __ldg(row);
            if (cin1) v.y = __ldg(row + kNetC);
        }
        return v;
    };

    // 7-row register window of low-res heat values around the current low-res row iy, both cells packed
    // (nxt / nxt2 = the next two rows, already in flight: global-load latency under a saturated write stream is
    // microseconds, one row of lead was measured not to cover it)
    float2 win[7], nxt, nxt2;
    int iy = it.iy0 - 1;
#pragma unroll
    for (int t = 0; t < 7; ++t) win[t] = hv2(iy - 3 + t);
    nxt = hv2(iy + 4);
    nxt2 = hv2(iy + 5);
    // NMS state: ga/gb ping-pong hold g of the previous / current row, vm = max(g[Y-2], g[Y-1])
    float ga[NPX], gb[NPX], vm[NPX];
#pragma unroll
    for (int i = 0; i < NPX; ++i) ga[i] = gb[i] = vm[i] = NEG;

    const int Ys = 4 * it.iy0 - 1;                       // first row whose g is needed; last = 4*(iy0+nrows)
    const float thr_lo = p.thresh * (1.0f - 4e-5f);
    const bool is_part = c < kParts;

    auto slide = [&]() {
        // slide the low-res window down one row; prefetch two rows ahead
        ++iy;
#pragma unroll
        for (int t = 0; t < 6; ++t) win[t] = win[t + 1];
        win[6] = nxt;
        nxt = nxt2;
        nxt2 = hv2(iy + 5);
    };

    // one output row: g(Y) -> cur ; NMS decision for row Y-1 whose g is in prev.  The phase
    // py = Y & 3 is a compile-time constant (rows are unrolled by 4), so every smoother weight is
    // a constant-bank operand.
    auto do_row = [&](auto py_tag, const int Y, float (&cur)[NPX], float (&prev)[NPX]) {
        constexpr int py = decltype(py_tag)::value;
        if (Y >= 0 && Y < H) {
            // vertical polyphase, both cells at once: gv = sum_t wp[py][t] * X[iy-3+t, cell, c]
            float2 gv = fmul2(win[0], bcast2(p.wp[py][0]));
#pragma unroll
            for (int t = 1; t < 7; ++t) gv = ffma2(win[t], bcast2(p.wp[py][t]), gv);
            // window of gv over cells (own first cell - 3 .. own last cell + 3) via shuffles
            float Wn[CPL + 6];
            Wn[3] = gv.x;
            Wn[4] = gv.y;
            Wn[0] = __shfl_up_sync(FULL, gv.y, 2);
            Wn[1] = __shfl_up_sync(FULL, gv.x, 1);
            Wn[2] = __shfl_up_sync(FULL, gv.y, 1);
            Wn[5] = __shfl_down_sync(FULL, gv.x, 1);
            Wn[6] = __shfl_down_sync(FULL, gv.y, 1);
            Wn[7] = __shfl_down_sync(FULL, gv.x, 2);
            // horizontal polyphase, two x phases at once: g[cell, 2q..2q+1] = sum_t wh[q][t] * gv[cell-3+t]; the
            // accumulator starts from nf (+0 / -inf), which is where a cell outside the image is excluded
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    float2 a = ffma2(bcast2(Wn[k]), p.wh[q][0], k == 0 ? nf0 : nf1);
#pragma unroll
                    for (int t = 1; t < 7; ++t) a = ffma2(bcast2(Wn[k + t]), p.wh[q][t], a);
                    cur[k * 4 + 2 * q] = a.x;
                    cur[k * 4 + 2 * q + 1] = a.y;
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < NPX; ++i) cur[i] = NEG;
        }

        // 3x3 NMS of row Yc = Y-1: vertical max in registers, horizontal neighbours by shuffle
        const int Yc = Y - 1;
        const int j = Yc - 4 * it.iy0;     // candidate row inside the band
        // Minimal mode writes no peaks tensor: a row in which no pixel of this warp's channel comes near the heat
        // threshold can neither yield a peak record (needs g > thresh) nor a near-tie count (both sides need
        // g > thr_lo), so its whole NMS step is skipped -- exact, and most rows of most channels are like that.
        bool row_live = true;
        if (!DENSE) {
            float hp = fmaxf(fmaxf(prev[0], prev[1]), prev[2]);
            hp = fmaxf(fmaxf(hp, prev[3]), prev[4]);
            hp = fmaxf(fmaxf(hp, prev[5]), fmaxf(prev[6], prev[7]));
            row_live = __any_sync(FULL, hp > thr_lo);
        }
        if (j >= 0 && j < 4 * it.nrows && row_live) {  // warp-uniform
            float v[NPX + 2];
#pragma unroll
            for (int i = 0; i < NPX; ++i) v[i + 1] = fmaxf(vm[i], cur[i]);
            v[0] = __shfl_up_sync(FULL, v[NPX], 1);
            v[NPX + 1] = __shfl_down_sync(FULL, v[1], 1);
            uint32_t pm = 0;
            float hot = NEG;
#pragma unroll
            for (int i = 0; i < NPX; ++i) {
                const float m = fmaxf(fmaxf(v[i], v[i + 1]), v[i + 2]);
                pm = or_bit_if_eq(pm, prev[i], m, 1u << i);
                hot = fmaxf(hot, prev[i]);
            }
            pm &= ownmask;
            if (is_part && hot > thr_lo) {
                // a pixel above the heat threshold that lost its 3x3 max by a hair: near-tie (loser side).  In minimal
                // mode every row that reaches this point has hot > thr_lo somewhere in the warp, so the exact test below
                // ran on every live row (15 % of the kernel's instructions, ncu r02i) for an event that is almost never
                // there.  Screen first, one FMA and one compare per pixel: the exact condition (m - prev) <= 2e-5*|m|
                // implies prev >= m - 3e-5*|m| (the 1e-5*|m| of slack is two orders of magnitude above the roundings of
                // either side), so a lane whose screen stays clear cannot have a near-tie; the exact test decides.
                bool cand = false;
#pragma unroll
                for (int i = 0; i < NPX; ++i) {
                    const float m = fmaxf(fmaxf(v[i], v[i + 1]), v[i + 2]);
                    cand |= (prev[i] != m) & (prev[i] >= fmaf(-3e-5f, fabsf(m), m));
                }
                if (cand) {
                    uint32_t nm = 0;
#pragma unroll
                    for (int i = 0; i < NPX; ++i) {
                        const float m = fmaxf(fmaxf(v[i], v[i + 1]), v[i + 2]);
                        const bool nr = (prev[i] > thr_lo) & (prev[i] != m) & ((m - prev[i]) <= 2e-5f * fabsf(m));
                        nm |= nr ? (1u << i) : 0u;
                    }
                    nm &= ownmask;
                    if (nm) atomicAdd(&p.counters[n].near_ties, __popc(nm));
                }
            }

            float *slot = nullptr;
            int jr = 0;
            if (DENSE) {
                jr = j % PR;
                const unsigned sidx = (pgroup + (unsigned)(j / PR)) & 1u;
                slot = pslots + (size_t)sidx * pslot_floats;
                uint32_t &wm = sidx ? wmask1 : wmask0;
                if (jr == 0 && wm) {
                    // lazily clear what this lane poked into the slot two groups ago
                    uint32_t mk = wm;
                    while (mk) {
                        const int b = __ffs(mk) - 1;
                        mk &= mk - 1;
                        const int rr = b / NPX, i = b - rr * NPX;
                        TFPD_CHK(lx0 + i >= 0 && ((size_t)rr * pitch_px + lx0 + i) * kHeatC + c < (size_t)pslot_floats);
                        slot[((size_t)rr * pitch_px + lx0 + i) * kHeatC + c] = 0.0f;
                    }
                    wm = 0;
                }
                wm |= pm << (jr * NPX);
            }
            // rare path (~0.5 % of the pixels are local maxima): one compact loop over the set bits,
            // operands picked out of the register arrays by predicated moves (no local memory, no call)
            for (uint32_t mk = pm; mk;) {
                const int i = __ffs(mk) - 1;
                mk &= mk - 1;
                float gc = 0.f, nb = 0.f, hc = 0.f;
#pragma unroll
                for (int q = 0; q < NPX; ++q)
                    if (i == q) {
                        gc = prev[q];
                        // largest neighbour except the one above (that pixel tests this one as its "below")
                        nb = fmaxf(fmaxf(v[q], v[q + 2]), cur[q]);
                        // heat score of row Yc: its low-res row is the window centre, or the one above it
                        // when Y just crossed into the next low-res row
                        const float2 hw = (py == 0) ? win[2] : win[3];
                        hc = (q >> 2) ? hw.y : hw.x;
                    }
                if (DENSE) TFPD_CHK(lx0 + i >= 0 && ((size_t)jr * pitch_px + lx0 + i) * kHeatC + c < (size_t)pslot_floats);
                TFPD_CHK(Yc >= 0 && Yc < H && 4 * (cell0 + (i >> 2)) + (i & 3) >= 0 && 4 * (cell0 + (i >> 2)) + (i & 3) < W);
                emit_peak(p, n, c, Yc, 4 * (cell0 + (i >> 2)) + (i & 3), gc, hc, nb,
                          DENSE ? slot + ((size_t)jr * pitch_px + lx0 + i) * kHeatC + c : nullptr);
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
                        if (p.l2_knobs & 2)
                            bulk_store_hint(p.peaks + pix * kHeatC, slot + (size_t)rr * pitch_px * kHeatC, bytes,
                                            policy_evict_first());
                        else
                            bulk_store(p.peaks + pix * kHeatC, slot + (size_t)rr * pitch_px * kHeatC, bytes);
                    }
                    bulk_commit();
                }
            }
        }
        // vm for the next row = max(g[Y-1], g[Y])
#pragma unroll
        for (int i = 0; i < NPX; ++i) vm[i] = fmaxf(prev[i], cur[i]);
        if (py == 3) slide();
    };

    // rows Ys .. Ys + 4*nrows + 1; Ys = 4*iy0 - 1 has phase 3, so a group of four rows has the
    // phases 3,0,1,2 and ga/gb swap roles without moves; the last two rows (phases 3, 0) only feed
    // the NMS of the band's last output rows.
    using P0 = std::integral_constant<int, 0>;
    using P1 = std::integral_constant<int, 1>;
    using P2 = std::integral_constant<int, 2>;
    using P3 = std::integral_constant<int, 3>;
    int Yr = Ys;
    bool last_dead = false;
#pragma unroll 1
    for (int g4 = 0; g4 < it.nrows; ++g4, Yr += 4) {
        if (!DENSE) {
            // Early-out (minimal mode only; dense mode must write every row of the peaks tensor).  This group
            // computes g of rows Yr .. Yr+3 and decides rows Yr-1 .. Yr+2; all of them read low-res rows
            // iy-3 .. iy+4 = win[0..6] and nxt, cells own-3 .. own+3 -- a subset of the cells this warp holds.
            // If no lane holds a heat value above dead_bound there, no g of those rows can come within reach of
            // the heat threshold (bound derived in api.cu): no peak record and no near-tie count can arise, and
            // for the rows decided later these rows are only ever the losing neighbour.  The group is skipped with
            // its g taken as -inf; exact.  (Ties with the group before: its last rows were decided against -inf
            // instead of a value below the bound, which cannot change a decision about a pixel above thr_lo.)
            float wm_ = fmaxf(fmaxf(win[0].x, win[0].y), fmaxf(nxt.x, nxt.y));
#pragma unroll
            for (int t = 1; t < 7; ++t) wm_ = fmaxf(wm_, fmaxf(win[t].x, win[t].y));
            if (!__any_sync(FULL, wm_ > p.dead_bound)) {
#pragma unroll
                for (int i = 0; i < NPX; ++i) ga[i] = gb[i] = vm[i] = NEG;
                slide();
                last_dead = true;
                continue;
            }
            last_dead = false;
        }
        do_row(P3(), Yr, ga, gb);
        do_row(P0(), Yr + 1, gb, ga);
        do_row(P1(), Yr + 2, ga, gb);
        do_row(P2(), Yr + 3, gb, ga);
    }
    // The two tail rows compute g of rows Yr, Yr+1 (low-res rows iy-3 .. iy+4 again) and decide rows Yr-1, Yr.  When the
    // band's last group was skipped (g of Yr-2, Yr-1 out of reach) and this window is dead as well, there is nothing to
    // decide: same argument as above.
    bool skip_tail = false;
    if (!DENSE && last_dead) {
        float wm_ = fmaxf(fmaxf(win[0].x, win[0].y), fmaxf(nxt.x, nxt.y));
#pragma unroll
        for (int t = 1; t < 7; ++t) wm_ = fmaxf(wm_, fmaxf(win[t].x, win[t].y));
        skip_tail = !__any_sync(FULL, wm_ > p.dead_bound);
    }
    if (!skip_tail) {
        do_row(P3(), Yr, ga, gb);
        do_row(P0(), Yr + 1, gb, ga);
    }
    if (DENSE) pgroup += (unsigned)((4 * it.nrows + PR - 1) / PR);
}

template <int PR, bool DENSE>
__global__ void __launch_bounds__(DENSE ? kThreads : kMinThreads, DENSE ? 1 : kMinCtasPerSm) k1_fused_f4(const K1Params p) {
    extern __shared__ __align__(128) float smem[];
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int pslot_floats = PR * p.chunk_cells * 4 * kHeatC;
    const int cslot_floats = p.copy_chunk * 4 * kNetC;
    const int raw_bytes = k1_raw_bytes(p.copy_chunk);
    float *pslots = smem;                                 // 2 peak-row slots (blur role)
    // per copy warp: 2 replication slots + 2 raw chunks; then 2 mbarriers per copy warp
    const size_t cw_bytes = 2 * (size_t)cslot_floats * 4 + 2 * (size_t)raw_bytes;
    // (TFPD_GUARD_BYTES is 0 in the product build; the checked build guards the two boundaries between the roles' areas)
    unsigned char *g_slots = reinterpret_cast<unsigned char *>(smem + 2 * (size_t)pslot_floats);
    unsigned char *cbase = g_slots + TFPD_GUARD_BYTES;
    unsigned char *g_copy = cbase + kCopyWarps * cw_bytes;
    uint64_t *bars_all = reinterpret_cast<uint64_t *>(g_copy + TFPD_GUARD_BYTES);
#ifdef TFPD_CHECKED
    if (DENSE) {
        guard_fill(g_slots, tid, (int)blockDim.x);
        guard_fill(g_copy, tid, (int)blockDim.x);
        __syncthreads();
    }
#endif

    if (p.dbg_times && tid == 0 && blockIdx.x < 1024) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.dbg_times[blockIdx.x * 4 + 0] = global_ns();
        p.dbg_times[blockIdx.x * 4 + 3] = smid;
    }
    if (!DENSE) {
        // ---- minimal mode: nothing dense is written, the kernel is the blur role alone ----
        // Only peak records and near-tie counts leave the kernel, so
        //  * channel 18 (the background heat map) is not computed: it yields neither (pafprocess.cpp:24 scans parts
        //    0..17), and its heat is ~1 everywhere, so it could never take the early-out;
        //  * warps need nothing from each other (no slots, no barriers).
        // Grid = SMALL, non-persistent CTAs: blockIdx.x = item (a band of `min_band` low-res rows of one frame and
        // column chunk), blockIdx.y = one of three groups of six part channels, warp = channel.  6 warps per CTA, four
        // CTAs per SM (24 warps per SM; the 19-warp persistent CTA of the dense kernel ran issue-bound at 65 % of the
        // issue slots, ncu r02e).  The six channels of a CTA share the 32-byte sectors of a cell (L1 locality), the
        // other two groups of the same band are the neighbouring CTAs.  Which rows and channels are skipped differs
        // from CTA to CTA; the hardware's CTA scheduler does the balancing, and with thousands of CTAs of a few
        // microseconds each the last wave is short (with one CTA per band: 768 CTAs on 592 slots, the SMs were
        // active 66 % of the kernel, ncu r02f).  Unlike persistent CTAs pulling from a software queue (tried: r02d),
        // CTAs that retire every few microseconds also let the scoring / assembly kernels of the batch before take SM
        // slots in between when several contexts are in flight.
        if (!(p.debug_roles & 1)) return;
        static_assert(kParts % kMinWarps == 0, "channel groups");
        const int band = p.min_band;
        const int nbands = (p.h + band - 1) / band;
        const long long g = blockIdx.x;
        const long long q = g / nbands;         // q = n * nchunks + chunk
        const int b = (int)(g - q * nbands);
        Item it;
        it.iy0 = b * band;
        it.nrows = min(band, p.h - it.iy0);
        it.n = (int)(q / p.nchunks);
        const int chunk = (int)(q - (long long)it.n * p.nchunks);
        it.c0 = chunk * p.chunk_cells;
        it.tc = min(p.chunk_cells, p.w - it.c0);
        unsigned pgroup = 0;
        uint32_t wmask0 = 0u, wmask1 = 0u;
        blur_role<PR, false>(p, it, nullptr, 0, (int)blockIdx.y * kMinWarps + warp, lane, pgroup, wmask0, wmask1);
        if (p.dbg_times && lane == 0 && warp == 0 && blockIdx.y == 0 && blockIdx.x < 1024)
            p.dbg_times[blockIdx.x * 4 + 1] = global_ns();
        return;
    }
    if (warp < kBlurWarps) {
        if (!(p.debug_roles & 1)) return;
        if (DENSE) {
            float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int q = tid; q < 2 * pslot_floats / 4; q += kBlurThreads) reinterpret_cast<float4 *>(pslots)[q] = z;
            named_bar(1, kBlurThreads);
        }
        unsigned pgroup = 0;
        uint32_t wmask0 = 0u, wmask1 = 0u;
        // this CTA's contiguous share of the flattened (frame, chunk, low-res row) space, walked as
        // maximal runs inside one (frame, chunk)
        long long u, uend;
        span_of((long long)p.N * p.nchunks * p.h, blockIdx.x, gridDim.x, u, uend);
        while (u < uend) {
            const long long q = u / p.h;                    // q = n * nchunks + chunk
            Item it;
            it.iy0 = (int)(u - q * p.h);
            it.nrows = (int)min((long long)(p.h - it.iy0), uend - u);
            it.n = (int)(q / p.nchunks);
            const int chunk = (int)(q - (long long)it.n * p.nchunks);
            it.c0 = chunk * p.chunk_cells;
            it.tc = min(p.chunk_cells, p.w - it.c0);
            blur_role<PR, DENSE>(p, it, pslots, pslot_floats, warp, lane, pgroup, wmask0, wmask1);
            u += it.nrows;
        }
        if (DENSE && tid == 0) bulk_wait_all();
#ifdef TFPD_CHECKED
        if (DENSE && lane == 0) TFPD_CHK(guard_ok(g_slots) && guard_ok(g_copy));
#endif
        if (p.dbg_times && tid == 0) p.dbg_times[blockIdx.x * 4 + 1] = global_ns();
        if (DENSE && (p.debug_roles & 2) && kHelperWarps > 0) {
            // The blur role of this CTA is done (measured: at ~65 % of the kernel) and its peak-row slots are
            // idle: the first kHelperWarps blur warps carve them into staging buffers and pull copy items from
            // the same queue as the copy warp until it is empty.
            named_bar(1, kBlurThreads);                   // every blur warp is past its last slot access; stores drained
            const size_t cw_bytes_h = 2 * (size_t)cslot_floats * 4 + 2 * (size_t)raw_bytes;
            const int fit = (int)((2 * (size_t)pslot_floats * 4) / cw_bytes_h);
            if (warp < kHelperWarps && warp < fit) {
                float *hslots = reinterpret_cast<float *>(reinterpret_cast<unsigned char *>(pslots) + warp * cw_bytes_h);
                unsigned char *hraws = reinterpret_cast<unsigned char *>(hslots + 2 * (size_t)cslot_floats);
                copy_role(p, hslots, cslot_floats, hraws, raw_bytes, bars_all + 2 * (kCopyWarps + warp), lane);
                if (lane == 0) bulk_wait_all();
            }
        }
    } else {
        const int cw = warp - kBlurWarps;
        if (DENSE && (p.debug_roles & 2)) {
            float *cslots = reinterpret_cast<float *>(cbase + cw * cw_bytes);
            unsigned char *raws = reinterpret_cast<unsigned char *>(cslots + 2 * (size_t)cslot_floats);
            copy_role(p, cslots, cslot_floats, raws, raw_bytes, bars_all + 2 * cw, lane);
            if (lane == 0) bulk_wait_all();
#ifdef TFPD_CHECKED
            if (lane == 0) TFPD_CHK(guard_ok(g_slots) && guard_ok(g_copy));
#endif
        }
        if (p.dbg_times && lane == 0 && cw == 0) p.dbg_times[blockIdx.x * 4 + 2] = global_ns();
    }
}

}  // namespace

// Polyphase weights of the 25-tap separable factor over a x4-replicated signal: wp[p][t] = sum of the taps that
// fall on low-res cell offset t-3 for output phase p, summed in float64.  They travel in the kernel parameters
// (constant bank, and per context: two contexts may have different smoother parameters).
int polyphase_weights(const double *s1d, int ksize, float wp[4][8]) {
    if (ksize != 25) return -1;
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
    return 0;
}

int launch_k1_fused_f4(const K1Params &p, bool dense, int sm_count, cudaStream_t s) {
    if (!dense) {
        // minimal mode: one small CTA per (frame, column chunk, band of rows); the grid size does not depend on sm_count
        const long long items = (long long)p.N * p.nchunks * ((p.h + p.min_band - 1) / p.min_band);
        if (items <= 0) return 0;
        if (items > 0x7fffffffLL) return -1;
        k1_fused_f4<4, false><<<dim3((unsigned)items, kParts / kMinWarps), kMinThreads, 0, s>>>(p);
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    // one CTA per SM; fewer only when there are fewer low-res rows than SMs
    const long long units = (long long)p.N * p.nchunks * p.h;
    const int grid = (int)min(units, (long long)sm_count);
    if (grid <= 0) return 0;
    const size_t cslot = (size_t)p.copy_chunk * 4 * kNetC * 4;
    const size_t raw = (size_t)k1_raw_bytes(p.copy_chunk);
    const size_t prow = (size_t)p.chunk_cells * 4 * kHeatC * 4;     // one peaks row of a blur chunk
    const size_t fixed = kCopyWarps * (2 * cslot + 2 * raw) + 16 * (kCopyWarps + kHelperWarps) + 2 * TFPD_GUARD_BYTES;
    const size_t limit = 227 * 1024 - 1024;
    // function attributes are per device: one flag per device a ctx of this process may live on
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63]) {
        cudaFuncSetAttribute(k1_fused_f4<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        cudaFuncSetAttribute(k1_fused_f4<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        cudaFuncSetAttribute(k1_fused_f4<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        attr_set[dev & 63] = true;
    }
    // peak rows per TMA store group: as many as fit next to the copy warps' staging
    int pr = 4;
    if (p.force_pr >= 2 && p.force_pr <= 4) pr = p.force_pr;
    while (pr > 2 && 2 * pr * prow + fixed > limit) --pr;
    const size_t smem = 2 * pr * prow + fixed;
    if (smem > limit) return -1;
    if (pr == 4) k1_fused_f4<4, true><<<grid, kThreads, smem, s>>>(p);
    else if (pr == 3) k1_fused_f4<3, true><<<grid, kThreads, smem, s>>>(p);
    else k1_fused_f4<2, true><<<grid, kThreads, smem, s>>>(p);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
