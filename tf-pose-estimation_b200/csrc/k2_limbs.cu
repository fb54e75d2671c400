// k2_limbs.cu -- K2: PAF line-integral scoring of every (peak a, peak b) pair of every limb,
// pafprocess.cpp:48-94 with get_paf_vectors (:216-233) and roundpaf (:235-237).
//
// A warp takes 30 candidate pairs per pass: pair-level values are computed with one lane per pair, then
// three pairs at a time are scored with 10 lanes per pair, one lane per PAF sample
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
constexpr int kPairsPerWarp = 3;           // pairs scored at once: 10 lanes (PAF samples) each
constexpr int kPairsPerPass = 30;          // pairs set up at once: one lane each

__global__ void __launch_bounds__(kK2Threads) k_limb_score(Workspace ws, PafSource src, size_t frame_stride,
                                                          Thresholds th) {
    extern __shared__ __align__(16) unsigned char k2_smem[];
    int16_t *s_px = reinterpret_cast<int16_t *>(k2_smem);          // [max_peaks] the frame's peak coordinates
    int16_t *s_py = s_px + ws.max_peaks;
    __shared__ int s_first[kHeatC + 1];
    __shared__ int s_tstart[kLimbs + 1];
    __shared__ unsigned char s_owner[kK2Threads / 32][32];      // per warp: lane owning the k-th surviving pair of the pass
    const int n = blockIdx.y;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (ws.k2_fused_sort) {
        // The peak table in Peak.id order (pafprocess.cpp:24-37: part-major, row-major) straight from the unordered
        // records the map kernel left: keys (part, y, x) are unique, so a record's position is the number of
        // smaller keys.  Every CTA of the frame ranks for itself (a few hundred peaks: cheaper than a sort kernel
        // of its own in front of this one plus its launch gap); CTA 0 also writes the table for the assembly
        // kernel and tfpd_fetch_peaks.
        const size_t mp4 = ((size_t)ws.max_peaks * 4 + 15) & ~(size_t)15;
        uint32_t *s_key = reinterpret_cast<uint32_t *>(k2_smem + mp4);
        uint16_t *s_idx = reinterpret_cast<uint16_t *>(k2_smem + 2 * mp4);
        __shared__ int s_hist[kHeatC + 1];
        const int raw = ws.counters[n].n_peak_rec;
        const int cnt = min(raw, ws.max_peaks);
        const uint32_t *gk = ws.rec_key + (size_t)n * ws.max_peaks;
        const float *gv = ws.rec_val + (size_t)n * ws.max_peaks;
        int16_t *ox = ws.pk_x + (size_t)n * ws.max_peaks, *oy = ws.pk_y + (size_t)n * ws.max_peaks;
        float *os = ws.pk_score + (size_t)n * ws.max_peaks;
        if (tid <= kHeatC) s_hist[tid] = 0;
        __syncthreads();
        constexpr int kPerThread = 4;
        if (cnt <= kPerThread * kK2Threads) {
            // bucket by part first (18 shared-memory counters), then rank inside the bucket: n*n/18 compares
            uint32_t kk[kPerThread];
            int sl[kPerThread];
#pragma unroll
            for (int e = 0; e < kPerThread; ++e) {
                const int i = tid + e * kK2Threads;
                kk[e] = 0u;
                sl[e] = 0;
                if (i < cnt) {
                    kk[e] = gk[i];
                    sl[e] = atomicAdd(&s_hist[key_part(kk[e])], 1);
                }
            }
            __syncthreads();
            if (tid == 0) {                             // first[p] = index of the first peak with part >= p
                int acc = 0;
                for (int pp = 0; pp <= kHeatC; ++pp) {
                    s_first[pp] = acc;
                    acc += s_hist[pp];
                }
            }
            __syncthreads();
#pragma unroll
            for (int e = 0; e < kPerThread; ++e) {
                const int i = tid + e * kK2Threads;
                if (i < cnt) {
                    const int pos = s_first[key_part(kk[e])] + sl[e];
                    s_key[pos] = kk[e];
                    s_idx[pos] = (uint16_t)i;
                }
            }
            __syncthreads();
            for (int j = tid; j < cnt; j += kK2Threads) {
                const uint32_t k = s_key[j];
                const int lo = s_first[key_part(k)], hi = s_first[key_part(k) + 1];
                int r = lo;
                for (int t = lo; t < hi; ++t) r += (int)(s_key[t] < k);
                TFPD_CHK(r >= 0 && r < ws.max_peaks && s_idx[j] < cnt);
                s_px[r] = (int16_t)key_x(k);
                s_py[r] = (int16_t)key_y(k);
                if (blockIdx.x == 0) {
                    ox[r] = (int16_t)key_x(k);
                    oy[r] = (int16_t)key_y(k);
                    os[r] = gv[s_idx[j]];
                }
            }
        } else {
            // very many peaks: position = number of smaller keys among all of them
            const int cnt4 = (cnt + 3) & ~3;
            for (int i = tid; i < cnt4; i += kK2Threads) s_key[i] = (i < cnt) ? gk[i] : 0xffffffffu;
            __syncthreads();
            const uint4 *k4 = reinterpret_cast<const uint4 *>(s_key);
            for (int i = tid; i < cnt; i += kK2Threads) {
                const uint32_t k = s_key[i];
                int r = 0;
                for (int j = 0; j < (cnt4 >> 2); ++j) {
                    const uint4 v = k4[j];              // same address for the whole warp: one broadcast
                    r += (int)(v.x < k) + (int)(v.y < k) + (int)(v.z < k) + (int)(v.w < k);
                }
                s_px[r] = (int16_t)key_x(k);
                s_py[r] = (int16_t)key_y(k);
                atomicAdd(&s_hist[key_part(k)], 1);
                if (blockIdx.x == 0) {
                    ox[r] = (int16_t)key_x(k);
                    oy[r] = (int16_t)key_y(k);
                    os[r] = gv[i];
                }
            }
            __syncthreads();
            if (tid == 0) {
                int acc = 0;
                for (int pp = 0; pp <= kHeatC; ++pp) {
                    s_first[pp] = acc;
                    acc += s_hist[pp];
                }
            }
        }
        if (tid == 0 && blockIdx.x == 0) {
            ws.info[n].n_peaks = cnt;
            ws.info[n].near_ties = ws.counters[n].near_ties;
            if (raw > ws.max_peaks) atomicOr(&ws.counters[n].flags, TFPD_FLAG_PEAK_OVERFLOW);
        }
        __syncthreads();
        if (blockIdx.x == 0 && tid <= kHeatC) ws.part_first[n * 20 + tid] = s_first[tid];
    } else {
        if (tid <= kHeatC) s_first[tid] = ws.part_first[n * 20 + tid];
        const int16_t *gx = ws.pk_x + (size_t)n * ws.max_peaks, *gy = ws.pk_y + (size_t)n * ws.max_peaks;
        __syncthreads();
        const int npk = min(s_first[kParts], ws.max_peaks);
        for (int i = tid; i < npk; i += kK2Threads) {
            s_px[i] = gx[i];
            s_py[i] = gy[i];
        }
    }
    if (tid == 0) {
        int acc = 0;
        for (int l = 0; l < kLimbs; ++l) {
            s_tstart[l] = acc;
            int pa = c_limb_part[l][0], pb = c_limb_part[l][1];
            acc += (s_first[pa + 1] - s_first[pa]) * (s_first[pb + 1] - s_first[pb]);
        }
        s_tstart[kLimbs] = acc;
    }
    __syncthreads();
    const int total = s_tstart[kLimbs];
    const float *paf = src.base + (size_t)n * frame_stride;
    const unsigned FULL = 0xffffffffu;

    const int slot = lane / kStepPaf;           // 0..2 ; lanes 30,31 -> 3 (idle in the scoring phase)
    const int i = lane - slot * kStepPaf;       // sample index 0..9
    const int group_base = slot * kStepPaf;
    const unsigned group_mask = (slot < kPairsPerWarp) ? (0x3ffu << group_base) : 0u;
    const int warps_total = gridDim.x * (kK2Threads / 32);

    // Two phases per pass of 30 candidate pairs, because the per-PAIR arithmetic (limb lookup, index split,
    // sqrt, four IEEE divisions) was 3/4 of the instructions when every one of a pair's 10 sample lanes repeated
    // it (ncu: issue-bound, 427 warp instructions per pair triple):
    //   setup    lane p owns pair base+p and computes the pair-level values once;
    //   scoring  ten sub-steps, each takes three of the pairs, 10 lanes per pair (one per PAF sample); the
    //            pair-level values come by shuffle from the owning lane, the gathers of sub-step t+1 are issued
    //            before sub-step t is reduced.
    // pairs per pass: 30 when there is enough work, fewer (a multiple of 3) when the frame has so few pairs
    // that spreading them over all warps shortens the serial chain of sub-steps
    int per = (total + warps_total - 1) / warps_total;
    per = min(kPairsPerPass, max(kPairsPerWarp, (per + kPairsPerWarp - 1) / kPairsPerWarp * kPairsPerWarp));
    // leading samples the setup phase scores to reject a pair early: the fewest after which "no hit so far" proves that
    // the count cannot exceed vector_cnt1 (4 for the reference's cnt1 = 6); off for settings where that would be more
    // than 6 samples or none
    const int prefilter = (ws.k2_prefilter && th.vector_cnt1 >= 4 && th.vector_cnt1 < kStepPaf) ? kStepPaf - th.vector_cnt1 : 0;
    int limb_hint = 0;                          // pairs are visited in increasing order: resume the limb search
    for (int base = (blockIdx.x * (kK2Threads / 32) + warp) * per; base < total; base += warps_total * per) {
        // ---- setup (lane = pair) ----
        const int q = base + lane;
        const bool pvalid = (lane < per) && (q < total);
        int xy1 = 0, lab = 0, meta = 0;         // x1 | y1 << 16 ; limb << 24 | a << 12 | b ; chan0 | ok << 8
        float vx = 0.f, vy = 0.f, norm = 1.f, stepx = 0.f, stepy = 0.f;
        if (pvalid) {
            int limb = limb_hint;
            while (q >= s_tstart[limb + 1]) ++limb;
            limb_hint = limb;
            const int pa = c_limb_part[limb][0], pb = c_limb_part[limb][1];
            const int nb = s_first[pb + 1] - s_first[pb];
            const int r = q - s_tstart[limb];
            const int a = r / nb;
            const int b = r - a * nb;
            const int ia = s_first[pa] + a, ib = s_first[pb] + b;
            TFPD_CHK(ia >= 0 && ia < ws.max_peaks && ib >= 0 && ib < ws.max_peaks && limb < kLimbs);
            const int x1 = s_px[ia], y1 = s_py[ia], x2 = s_px[ib], y2 = s_py[ib];
            const float dx = (float)(x2 - x1), dy = (float)(y2 - y1);                     // :63-64
            norm = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));           // :65
            const bool ok = !(norm < 1e-12f);                                             // :66 (norm is 0 or >= 1)
            if (ok) {
                vx = __fdiv_rn(dx, norm);                                                 // :67-68
                vy = __fdiv_rn(dy, norm);
                stepx = __fdiv_rn(dx, (float)kStepPaf);                                   // :219-220
                stepy = __fdiv_rn(dy, (float)kStepPaf);
            }
            xy1 = (x1 & 0xffff) | (y1 << 16);
            lab = (limb << 24) | (a << 12) | b;
            bool keep = ok;
            if (ok && prefilter > 0) {
                // Early rejection, exact.  A pair becomes a candidate only if MORE than vector_cnt1 of its
                // STEP_PAF = 10 samples score above vector_score (pafprocess.cpp:80,85).  Samples 0 .. prefilter-1
                // are scored here by the pair's own lane with the statements of the scoring phase below (same
                // rounding, same loads, hence the same pass/fail per sample); when even a pass on every remaining
                // sample could not lift the count above vector_cnt1 the pair is dropped -- it could only have been
                // rejected later.  With the reference's constants (cnt1 = 6) that is: none of the first four samples
                // passes, which holds for the great majority of pairs that are not limbs.  Survivors go through the
                // unchanged 10-sample scoring, so accepted candidates and their scores are untouched.
                const int chan0 = c_limb_chan[limb][0];
                int hits0 = 0;
                for (int k = 0; k < prefilter; ++k) {
                    const float fx = __fadd_rn((float)x1, __fmul_rn((float)k, stepx));
                    const float fy = __fadd_rn((float)y1, __fmul_rn((float)k, stepy));
                    const int lx = (int)((double)fx + 0.5);
                    const int ly = (int)((double)fy + 0.5);
                    TFPD_CHK(lx >= 0 && ly >= 0 && (ly >> src.shift) < src.rows && (lx >> src.shift) < src.cols);
                    const float *pp = paf + ((size_t)(ly >> src.shift) * src.cols + (lx >> src.shift)) * src.chan +
                                      src.chan_off + chan0;
                    const float sk = __fadd_rn(__fmul_rn(vx, __ldg(pp)), __fmul_rn(vy, __ldg(pp + 1)));
                    hits0 += (sk > th.vector_score) ? 1 : 0;
                }
                keep = hits0 + (kStepPaf - prefilter) > th.vector_cnt1;
            }
            meta = c_limb_chan[limb][0] | (keep ? 0x100 : 0);
        }
        // the scoring phase walks the SURVIVORS only: sub-step t, slot s takes the (3t+s)-th lane whose pair is kept.
        // Every surviving lane publishes its lane id at its rank in a per-warp table (one store; __fns -- a software
        // loop -- looked the owner up twice per sub-step before and was 27 % of the kernel's instructions, ncu r02m).
        __syncwarp();                                                                     // last pass's readers are done
        const unsigned alive = __ballot_sync(FULL, pvalid && (meta & 0x100));
        if (pvalid && (meta & 0x100)) s_owner[warp][__popc(alive & ((1u << lane) - 1u))] = (unsigned char)lane;
        __syncwarp();
        const int npass = __popc(alive);                                                  // warp-uniform
        const int nsub = (npass + kPairsPerWarp - 1) / kPairsPerWarp;

        // ---- scoring (10 lanes per pair) ----
        auto fetch = [&](int t, float &o_vx, float &o_vy, float &o_pvx, float &o_pvy, bool &o_ok, int &o_owner) {
            const int rank = kPairsPerWarp * t + slot;
            const int owner = (slot < kPairsPerWarp && rank < npass) ? (int)s_owner[warp][rank] : 0;
            o_owner = owner;
            const int g_xy = __shfl_sync(FULL, xy1, owner);
            const int g_meta = __shfl_sync(FULL, meta, owner);
            const float g_sx = __shfl_sync(FULL, stepx, owner), g_sy = __shfl_sync(FULL, stepy, owner);
            o_vx = __shfl_sync(FULL, vx, owner);
            o_vy = __shfl_sync(FULL, vy, owner);
            o_ok = (slot < kPairsPerWarp) && (rank < npass) && (g_meta & 0x100);
            o_pvx = o_pvy = 0.f;
            if (o_ok) {
                const int x1 = (int)(int16_t)(g_xy & 0xffff), y1 = g_xy >> 16;
                // roundpaf(peak1.x + i*STEP_X): float product, float sum, then +0.5 in double (:223-224,:235-237)
                const float fx = __fadd_rn((float)x1, __fmul_rn((float)i, g_sx));
                const float fy = __fadd_rn((float)y1, __fmul_rn((float)i, g_sy));
                const int lx = (int)((double)fx + 0.5);
                const int ly = (int)((double)fy + 0.5);
                TFPD_CHK(lx >= 0 && ly >= 0 && (ly >> src.shift) < src.rows && (lx >> src.shift) < src.cols);
                const float *pp = paf + ((size_t)(ly >> src.shift) * src.cols + (lx >> src.shift)) * src.chan +
                                  src.chan_off + (g_meta & 0xff);
                o_pvx = __ldg(pp);              // the two PAF channels of a limb are adjacent (pafprocess.h:20-24)
                o_pvy = __ldg(pp + 1);
            }
        };
        float c_vx, c_vy, c_pvx, c_pvy;
        bool c_ok;
        int c_owner;
        fetch(0, c_vx, c_vy, c_pvx, c_pvy, c_ok, c_owner);
        for (int t = 0; t < nsub; ++t) {
            float n_vx = 0.f, n_vy = 0.f, n_pvx = 0.f, n_pvy = 0.f;
            bool n_ok = false;
            int n_owner = 0;
            if (t + 1 < nsub) fetch(t + 1, n_vx, n_vy, n_pvx, n_pvy, n_ok, n_owner);
            const float s = c_ok ? __fadd_rn(__fmul_rn(c_vx, c_pvx), __fmul_rn(c_vy, c_pvy)) : 0.f;   // :77
            // scores += score, i = 0..9, in order (:78); criterion1 counts score > THRESH_VECTOR_SCORE (:80)
            float total_s = 0.0f;
#pragma unroll
            for (int k = 0; k < kStepPaf; ++k) total_s = __fadd_rn(total_s, __shfl_sync(FULL, s, (group_base + k) & 31));
            const int hits = __popc(__ballot_sync(FULL, s > th.vector_score) & group_mask);
            const int owner = c_owner;
            const float g_norm = __shfl_sync(FULL, norm, owner);
            const int g_lab = __shfl_sync(FULL, lab, owner);
            bool accept = false;
            float crit2 = 0.0f;
            if (c_ok && i == 0 && hits > th.vector_cnt1) {                        // first half of :85
                // float criterion2 = scores / STEP_PAF + min(0.0, 0.5 * h1 / norm - 1.0);   (:83)
                const double pen = 0.5 * (double)src.H / (double)g_norm - 1.0;
                const double d = (double)__fdiv_rn(total_s, (float)kStepPaf) + (pen < 0.0 ? pen : 0.0);
                crit2 = (float)d;
                accept = crit2 > 0.0f;                                             // second half of :85
            }
            // one atomic per sub-step for all (up to three) accepted pairs of the warp: the list is unordered (the
            // assembly kernel sorts it by key), each accepted lane takes its offset by rank
            const unsigned am = __ballot_sync(FULL, accept);
            if (am) {
                const int leader = __ffs(am) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&ws.counters[n].n_cand, __popc(am));
                base = __shfl_sync(FULL, base, leader);
                if (accept) {
                    const int idx = base + __popc(am & ((1u << lane) - 1u));
                    if (idx < ws.max_cands) {
                        // ascending key order == (limb asc, score desc, idx1 asc, idx2 asc)
                        unsigned long long key = ((unsigned long long)(g_lab >> 24) << 56) |
                                                 ((unsigned long long)(0xffffffffu - __float_as_uint(crit2)) << 24) |
                                                 (unsigned long long)(g_lab & 0xffffff);
                        ws.cand_key[(size_t)n * ws.max_cands + idx] = key;
                    }
                }
            }
            c_vx = n_vx;
            c_vy = n_vy;
            c_pvx = n_pvx;
            c_pvy = n_pvy;
            c_ok = n_ok;
            c_owner = n_owner;
        }
    }
}

}  // namespace

int launch_limb_score(const Workspace &ws, int N, const PafSource &src, size_t frame_stride, const Thresholds &th,
                      int sm_count, cudaStream_t s) {
    // blocks per frame: ~6 CTAs (48 warps) per SM overall -- the kernel is gather-latency bound, occupancy is
    // what hides it; a CTA's fixed cost is the peak ranking / coordinate staging of its frame.  Alone, 8 CTAs per
    // frame at batch 64 are a little faster than these 14 (13.2 against 14.8 us), but with several batches in flight
    // (tail kernels squeezed onto the SMs the map kernel leaves free) the finer grid wins: 0.1433 against 0.1471 ms
    // per step.  "k2_blocks" overrides.
    int bpf = max(1, (6 * sm_count + N - 1) / N);
    if (bpf > 32) bpf = 32;
    if (ws.k2_blocks_per_frame > 0) bpf = ws.k2_blocks_per_frame;
    dim3 grid(bpf, N);
    const size_t smem = (size_t)ws.max_peaks * 10 + 64;     // sorted coordinates, keys, record indices
    static size_t attr_smem[64] = {};                      // per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > 48 * 1024 && smem > attr_smem[dev & 63]) {
        if (cudaFuncSetAttribute(k_limb_score, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return -1;
        attr_smem[dev & 63] = smem;
    }
    k_limb_score<<<grid, kK2Threads, smem, s>>>(ws, src, frame_stride, th);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
