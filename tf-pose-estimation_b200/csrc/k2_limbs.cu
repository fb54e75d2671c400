// k2_limbs.cu -- K2: PAF line-integral scoring of every (peak a, peak b) pair of every limb,
// pafprocess.cpp:48-94 with get_paf_vectors (:216-233) and roundpaf (:235-237).
//
// A warp scores three candidate pairs at a time: 10 lanes per pair, one lane per PAF sample
// (STEP_PAF = 10, pafprocess.h:13).  The ten sample scores of a pair are then summed in the
// reference's i = 0..9 order by every lane of the group (shuffle broadcast), so the float32
// sum is bit-identical to the sequential C++ loop.  Arithmetic follows the C++ promotion
// rules statement by statement; products and sums use the _rn intrinsics so
// that nvcc cannot contract them into FMAs (the reference is built for baseline x86-64, which
// has no FMA).
//
// PAF vectors are gathered from the low-res network output X when the upsample is the x4
// replication (paf_up[y,x] == X[y>>2, x>>2] bit-exactly), else from the dense paf_up tensor.
#include "tfpd_internal.cuh"

namespace tfpd {

namespace {

constexpr int kK2Threads = 256;
constexpr int kPairsPerWarp = 3;
constexpr int kUnroll = 2;                 // candidate triples in flight per warp (gather latency overlap)

struct PairSetup {
    int limb, a, b;
    float vx, vy, norm;
    bool valid, skip;
    const float *pp;        // address of the PAF pixel this lane samples
    int ch0, ch1;
};

__global__ void __launch_bounds__(kK2Threads) k_limb_score(Workspace ws, PafSource src, size_t frame_stride,
                                                          Thresholds th) {
    extern __shared__ __align__(16) unsigned char k2_smem[];
    int16_t *s_px = reinterpret_cast<int16_t *>(k2_smem);          // [max_peaks] the frame's peak coordinates
    int16_t *s_py = s_px + ws.max_peaks;
    __shared__ int s_first[kHeatC + 1];
    __shared__ int s_tstart[kLimbs + 1];
    const int n = blockIdx.y;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid <= kHeatC) s_first[tid] = ws.part_first[n * 20 + tid];
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int l = 0; l < kLimbs; ++l) {
            s_tstart[l] = acc;
            int pa = c_limb_part[l][0], pb = c_limb_part[l][1];
            acc += (s_first[pa + 1] - s_first[pa]) * (s_first[pb + 1] - s_first[pb]);
        }
        s_tstart[kLimbs] = acc;
    }
    {
        const int npk = min(s_first[kParts], ws.max_peaks);
        const int16_t *gx = ws.pk_x + (size_t)n * ws.max_peaks, *gy = ws.pk_y + (size_t)n * ws.max_peaks;
        for (int i = tid; i < npk; i += kK2Threads) {
            s_px[i] = gx[i];
            s_py[i] = gy[i];
        }
    }
    __syncthreads();
    const int total = s_tstart[kLimbs];
    const float *paf = src.base + (size_t)n * frame_stride;

    const int slot = lane / kStepPaf;           // 0..2 ; lanes 30,31 -> 3 (idle)
    const int i = lane - slot * kStepPaf;       // sample index 0..9
    const int group_base = slot * kStepPaf;
    const int warps_total = gridDim.x * (kK2Threads / 32);
    const int stride = warps_total * kPairsPerWarp;

    auto setup = [&](int q0) -> PairSetup {
        PairSetup ps;
        const int q = q0 + slot;
        ps.valid = (slot < kPairsPerWarp) && (q < total);
        ps.limb = 0;
        ps.a = ps.b = 0;
        ps.vx = ps.vy = 0.f;
        ps.norm = 1.f;
        ps.skip = true;
        ps.pp = paf;
        ps.ch0 = ps.ch1 = 0;
        if (ps.valid) {
            int limb = 0;
            while (q >= s_tstart[limb + 1]) ++limb;
            const int pa = c_limb_part[limb][0], pb = c_limb_part[limb][1];
            const int nb = s_first[pb + 1] - s_first[pb];
            const int r = q - s_tstart[limb];
            const int a = r / nb;
            const int b = r - a * nb;
            const int ia = s_first[pa] + a, ib = s_first[pb] + b;
            const int x1 = s_px[ia], y1 = s_py[ia], x2 = s_px[ib], y2 = s_py[ib];
            float vx = (float)(x2 - x1);                                              // :63-64
            float vy = (float)(y2 - y1);
            const float norm = __fsqrt_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy))); // :65
            ps.limb = limb;
            ps.a = a;
            ps.b = b;
            ps.norm = norm;
            ps.skip = (norm < 1e-12f);                                             // :66 (norm is 0 or >= 1)
            if (!ps.skip) {
                ps.vx = __fdiv_rn(vx, norm);                                       // :67-68
                ps.vy = __fdiv_rn(vy, norm);
                const float stepx = __fdiv_rn((float)(x2 - x1), (float)kStepPaf);   // :219-220
                const float stepy = __fdiv_rn((float)(y2 - y1), (float)kStepPaf);
                // roundpaf(peak1.x + i*STEP_X): float product, float sum, then +0.5 in double (:223-224,:235-237)
                const float fx = __fadd_rn((float)x1, __fmul_rn((float)i, stepx));
                const float fy = __fadd_rn((float)y1, __fmul_rn((float)i, stepy));
                const int lx = (int)((double)fx + 0.5);
                const int ly = (int)((double)fy + 0.5);
                ps.pp = paf + ((size_t)(ly >> src.shift) * src.cols + (lx >> src.shift)) * src.chan + src.chan_off;
                ps.ch0 = c_limb_chan[limb][0];
                ps.ch1 = c_limb_chan[limb][1];
            }
        }
        return ps;
    };

    for (int q0 = (blockIdx.x * (kK2Threads / 32) + warp) * kPairsPerWarp; q0 < total; q0 += stride * kUnroll) {
        PairSetup ps[kUnroll];
        float pvx[kUnroll], pvy[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) ps[u] = setup(q0 + u * stride);       // q >= total -> invalid
        // all gathers of the kUnroll triples are issued before any is consumed
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            pvx[u] = pvy[u] = 0.f;
            if (ps[u].valid && !ps[u].skip) {
                pvx[u] = __ldg(ps[u].pp + ps[u].ch0);
                pvy[u] = __ldg(ps[u].pp + ps[u].ch1);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            if (q0 + u * stride >= total) break;                                   // warp-uniform
            float s = 0.f;
            if (ps[u].valid && !ps[u].skip)
                s = __fadd_rn(__fmul_rn(ps[u].vx, pvx[u]), __fmul_rn(ps[u].vy, pvy[u]));         // :77
            // scores += score, i = 0..9, in order (:78); criterion1 counts score > THRESH_VECTOR_SCORE (:80)
            float total_s = 0.0f;
            int hits = 0;
#pragma unroll
            for (int k = 0; k < kStepPaf; ++k) {
                const float sk = __shfl_sync(0xffffffffu, s, (group_base + k) & 31);
                total_s = __fadd_rn(total_s, sk);
                hits += (sk > th.vector_score) ? 1 : 0;
            }
            if (ps[u].valid && !ps[u].skip && i == 0) {
                // float criterion2 = scores / STEP_PAF + min(0.0, 0.5 * h1 / norm - 1.0);   (:83)
                const double pen = 0.5 * (double)src.H / (double)ps[u].norm - 1.0;
                const double d = (double)__fdiv_rn(total_s, (float)kStepPaf) + (pen < 0.0 ? pen : 0.0);
                const float crit2 = (float)d;
                if (hits > th.vector_cnt1 && crit2 > 0.0f) {                       // :85
                    int idx = atomicAdd(&ws.counters[n].n_cand, 1);
                    if (idx < ws.max_cands) {
                        // ascending key order == (limb asc, score desc, idx1 asc, idx2 asc)
                        unsigned long long key = ((unsigned long long)ps[u].limb << 56) |
                                                 ((unsigned long long)(0xffffffffu - __float_as_uint(crit2)) << 24) |
                                                 ((unsigned long long)ps[u].a << 12) | (unsigned long long)ps[u].b;
                        ws.cand_key[(size_t)n * ws.max_cands + idx] = key;
                    }
                }
            }
        }
    }
}

}  // namespace

int launch_limb_score(const Workspace &ws, int N, const PafSource &src, size_t frame_stride, const Thresholds &th,
                      int sm_count, cudaStream_t s) {
    // blocks per frame: ~6 CTAs (48 warps) per SM overall -- the kernel is gather-latency bound, occupancy is
    // what hides it; a CTA's fixed cost is the peak-coordinate staging
    int bpf = max(1, (6 * sm_count + N - 1) / N);
    if (bpf > 32) bpf = 32;
    dim3 grid(bpf, N);
    k_limb_score<<<grid, kK2Threads, (size_t)ws.max_peaks * 4, s>>>(ws, src, frame_stride, th);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
