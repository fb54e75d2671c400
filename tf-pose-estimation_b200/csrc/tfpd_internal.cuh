// tfpd_internal.cuh -- shared declarations of the decode kernels (not part of the C-ABI).
//
// Data layout in HBM (all per ctx, sized for max_batch frames):
//   net_out   X[N,h,w,57]  f32 NHWC   caller-owned   heat = ch 0..18, PAF = ch 19..56 (estimator.py:321-322)
//   dense     heat_up[N,H,W,19], paf_up[N,H,W,38], peaks[N,H,W,19]   caller-owned, optional
//   rec_key/rec_val [N,max_peaks]   unsorted peak records appended by K1 (key = part|y|x)
//   pk_x/pk_y/pk_score [N,max_peaks], part_first[N,20]   peak table, id = index (pafprocess.cpp:24-44)
//   cand_key [N,max_cands] u64      accepted limb candidates (limb|~score|a|b)
//   conn_* [N,max_cands]            accepted connections, limb-major (pafprocess.cpp:96-124)
//   humans [N,max_humans] tfpd_human, info[N] tfpd_frame_info
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "tfpose_decode.h"

// Checked build (make -C tf-pose-estimation_b200/csrc checked -> tools/ab/libtfpose_decode_checked.so, -DTFPD_CHECKED):
// TFPD_CHK(cond) traps with a message when cond is false; guard words sit between the shared-memory sub-arrays of the
// kernels and are verified when a role finishes.  Compiled out of the product build.  Stands in for compute-sanitizer,
// which this GPU pool does not allow (profiles/r02_sanitizers/README.md).
#ifdef TFPD_CHECKED
#include <cstdio>
#define TFPD_CHK(cond)                                                                                          \
    do {                                                                                                        \
        if (!(cond)) {                                                                                          \
            printf("TFPD_CHK failed: %s  at %s:%d  (block %d,%d thread %d)\n", #cond, __FILE__, __LINE__,          \
                   (int)blockIdx.x, (int)blockIdx.y, (int)threadIdx.x);                                         \
            __trap();                                                                                           \
        }                                                                                                       \
    } while (0)
#define TFPD_GUARD_BYTES 128
#else
#define TFPD_CHK(cond) ((void)0)
#define TFPD_GUARD_BYTES 0
#endif

namespace tfpd {

#ifdef TFPD_CHECKED
__device__ inline void guard_fill(void *g, int tid, int nthreads) {
    for (int i = tid; i < TFPD_GUARD_BYTES / 4; i += nthreads) reinterpret_cast<volatile uint32_t *>(g)[i] = 0xA5A5A5A5u;
}
__device__ inline bool guard_ok(const void *g) {
    bool ok = true;
    for (int i = 0; i < TFPD_GUARD_BYTES / 4; ++i) ok = ok && reinterpret_cast<const volatile uint32_t *>(g)[i] == 0xA5A5A5A5u;
    return ok;
}
#endif

constexpr int kParts = TFPD_NUM_PARTS;
constexpr int kLimbs = TFPD_NUM_LIMBS;
constexpr int kHeatC = TFPD_HEAT_CHANNELS;
constexpr int kPafC = TFPD_PAF_CHANNELS;
constexpr int kNetC = TFPD_NET_CHANNELS;
constexpr int kStepPaf = 10;               // pafprocess.h:13
constexpr int kMaxPeaksHard = 4096;        // 12-bit per-part indices inside the candidate key

// limb tables, pafprocess.h:16-24 (one copy per translation unit; no relocatable device code)
static __constant__ int c_limb_part[kLimbs][2] = {{1, 2},  {1, 5},   {2, 3},   {3, 4},  {5, 6},   {6, 7},  {1, 8},
                                                  {8, 9},  {9, 10},  {1, 11},  {11, 12}, {12, 13}, {1, 0},  {0, 14},
                                                  {14, 16}, {0, 15}, {15, 17}, {2, 16}, {5, 17}};
static __constant__ int c_limb_chan[kLimbs][2] = {{12, 13}, {20, 21}, {14, 15}, {16, 17}, {22, 23}, {24, 25}, {0, 1},
                                                  {2, 3},   {4, 5},   {6, 7},   {8, 9},   {10, 11}, {28, 29}, {30, 31},
                                                  {34, 35}, {32, 33}, {36, 37}, {18, 19}, {26, 27}};

// peak record key: part(5) | y(13) | x(14)  -> ascending key order == (part, y, x) order
__host__ __device__ inline uint32_t peak_key(int part, int y, int x) {
    return ((uint32_t)part << 27) | ((uint32_t)y << 14) | (uint32_t)x;
}
__host__ __device__ inline int key_part(uint32_t k) { return (int)(k >> 27); }
__host__ __device__ inline int key_y(uint32_t k) { return (int)((k >> 14) & 0x1FFF); }
__host__ __device__ inline int key_x(uint32_t k) { return (int)(k & 0x3FFF); }

struct Thresholds {
    float heat, vector_score, human_score;
    int vector_cnt1, part_cnt;
    int mode;
};

// per-frame counters zeroed before every call (one cudaMemsetAsync)
struct FrameCounters {
    int n_peak_rec;     // records appended by K1 / the dense scan (may exceed max_peaks)
    int near_ties;
    int n_cand;         // candidates appended by K2 (may exceed max_cands)
    int flags;
    int k1_queue;       // frame 0 of a launch only: head of K1's copy-item queue (dynamic distribution)
    int pad_[3];
};

struct Workspace {
    int max_batch, max_peaks, max_cands, max_rows, max_humans;
    int k3_warp_sort_max;       // longest limb segment sorted by one warp (1024; TFPD_K3_WARPSORT lowers it for tests)
    int k3_parallel_limbs;      // 1: whole-limb assembly steps where provably order-free (TFPD_K3_PARALLEL=0 turns them off)
    int k3_reg_rows;
    int k3_std_ties;            // 1: limb segments with bit-equal scores are ordered by a literal std::sort (the reference's order)
#ifdef TFPD_CHECKED
    int dbg_trip;            // checked build only, TFPD_CHECKED_TRIP=1: the assembly kernel writes one element past its key array (negative control)
#endif
    int k2_prefilter;        // 1: limb scoring rejects a pair after its first samples when it can no longer qualify (exact)
    int k2_blocks_per_frame; // > 0: CTAs of the scoring kernel per frame (default: ~6 CTAs per SM over the batch)
    int k2_fused_sort;      // the scoring kernel ranks the peak records itself (no k_sort_peaks launch)            // subset rows K3 keeps in registers before spilling to smem (128; TFPD_K3_REGROWS lowers it for tests)
    unsigned long long *dbg_times;  // null, or the ctx's timestamp buffer: K3 writes clock64 per phase at [2048 + frame*8 + phase], frame < 64
    FrameCounters *counters;    // [N]
    uint32_t *rec_key;          // [N,max_peaks]
    float *rec_val;             // [N,max_peaks]
    int16_t *pk_x, *pk_y;       // [N,max_peaks]
    float *pk_score;            // [N,max_peaks]
    int *part_first;            // [N,20]  first[part], first[18] = n_peaks (19 = pad)
    unsigned long long *cand_key;   // [N,max_cands]
    int16_t *conn_cid1, *conn_cid2; // [N,max_cands]
    float *conn_score;          // [N,max_cands]
    int *conn_start;            // [N,20]  segment start per limb inside conn_* ; [19] = unused
    int *conn_count;            // [N,20]
    tfpd_human *humans;         // [N,max_humans]
    tfpd_frame_info *info;      // [N]
};

// Where limb scoring reads PAF vectors from.
struct PafSource {
    const float *base;      // X (lowres, 57 ch) or paf_up (dense, 38 ch)
    int rows, cols;         // spatial dims of `base`
    int chan, chan_off;     // channels per pixel, offset of PAF channel 0
    int shift;              // lowres: log2(upsample factor) applied to sample coords; dense: 0
    int H;                  // upsampled height h1 used in the length penalty (pafprocess.cpp:83)
};

struct K1Params {
    const float *X;
    float *heat_up, *paf_up, *peaks;    // dense outputs or null
    uint32_t *rec_key;
    float *rec_val;
    FrameCounters *counters;
    int max_peaks;
    int N, h, w;                        // low-res dims; H = 4h, W = 4w
    int chunk_cells, nchunks;                              // blur role: <= 56-cell column chunks
    int copy_chunk, copy_nchunks;                          // copy role: <= 32-cell row chunks (one lane per cell)
    float thresh;
    float wp[4][8];                     // polyphase smoother weights: phase p, low-res tap t (cell offset t-3)
    float2 wh[2][8];                    // the same weights paired over two adjacent x phases: wh[q][t] = {wp[2q][t], wp[2q+1][t]}
                                        // (packed operand of the horizontal pass's FFMA2)
    int min_band;                       // minimal mode: low-res rows per work item of the dynamic queue
    float dead_bound;                   // minimal mode: a low-res window whose largest heat value is below this cannot
                                        // produce a blurred value near the heat threshold (k1_fused.cu, early-out)
    int l2_knobs;                       // bit0 up-front L2 prefetch of X, bit1 evict_first stores, bit2 evict_last prefetch
    int debug_roles;                    // bit0 blur role, bit1 copy role (3 = normal; tuning/profiling only)
    int force_pr;
    unsigned long long *dbg_times;      // null, or [grid][4]: CTA start / blur role end / copy role end (ns), SM id
};

// launchers (each returns the number of kernel launches it issued, <0 on launch error)
int launch_k1_fused_f4(const K1Params &p, bool dense, int sm_count, cudaStream_t s);
int polyphase_weights(const double *s1d, int ksize, float wp[4][8]);     // F=4, ksize=25 fast path
int launch_generic_maps(const float *X, int N, int h, int w, int H, int W, const float *d_k1d, int ksize,
                        float *heat_up, float *paf_up, float *peaks, float *scratch, const Workspace &ws, float thr,
                        cudaStream_t s);
int launch_pow2_maps(const float *X, int N, int h, int w, int log2f, const float *d_wp, const float *h_wp, int T,
                     int path, float *heat_up, float *paf_up, float *peaks, float *scratch, const Workspace &ws, float thr,
                     cudaStream_t s);
int launch_scan_dense_peaks(const float *peaks, const float *heat, int N, int H, int W, float thresh,
                            const Workspace &ws, cudaStream_t s);
int launch_sort_peaks(const Workspace &ws, int N, cudaStream_t s);
int launch_limb_score(const Workspace &ws, int N, const PafSource &src, size_t frame_stride,
                      const Thresholds &th, int sm_count, cudaStream_t s);
int launch_assemble(const Workspace &ws, int N, const Thresholds &th, cudaStream_t s);
size_t assemble_smem_bytes(const Workspace &ws);
int launch_draw(const int *d_centers, const int *d_counts, int max_h, unsigned char *d_img, int *d_ids, int N, int H,
                int W, cudaStream_t s);
int launch_centers_from_tables(const Workspace &ws, int N, int map_h, int map_w, int img_h, int img_w, int *d_centers,
                               int *d_counts, cudaStream_t s);
int launch_resize_cubic_u8(const unsigned char *src, unsigned char *dst, int N, int sh, int sw, int cn, int dh, int dw,
                           double fx, double fy, int path, cudaStream_t s);
int launch_quantize_img(const float *src, unsigned char *dst, size_t n, int sm_count, cudaStream_t s);

}  // namespace tfpd
