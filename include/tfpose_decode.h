/*
 * tfpose_decode.h -- C-ABI of libtfpose_decode.so: the B200 (sm_100a) heatmap/PAF decode
 * path of tf-pose-estimation.  Plain pointers and sizes only; no torch / CUDA types.
 *
 * What this boundary replaces in the reference (paths relative to the reference root):
 *
 *   seam 1  the post-processing half of TfPoseEstimator.inference
 *           (tf_pose/estimator.py:518-553): the TF1 graph ops built at
 *           tf_pose/estimator.py:321-337 (channel split, tf.image.resize_area x2,
 *           Smoother 25x25 depthwise conv -- tf_pose/tensblur/smoother.py:62-95 --,
 *           3x3 max-pool NMS) followed by PoseEstimator.estimate_paf
 *           (tf_pose/estimator.py:272-299).
 *           -> tfpd_decode_device / tfpd_decode_host + tfpd_fetch_humans.
 *
 *   seam 2  the swig-bound C++ module tf_pose/pafprocess (prototypes
 *           tf_pose/pafprocess/pafprocess.h:53-59, body pafprocess.cpp:19-214, binding
 *           tf_pose/pafprocess/pafprocess.i:1-15).
 *           -> tfpd_process_paf + tfpd_get_num_humans / tfpd_get_part_cid / tfpd_get_score /
 *              tfpd_get_part_x / tfpd_get_part_y / tfpd_get_part_score.
 *
 * Every function that can fail returns an int status (TFPD_OK == 0); tfpd_last_error()
 * gives the message.  A context (ctx) owns its device workspace and its results; results
 * stay valid until the next decode/process call on that ctx.  ONE call is in flight per ctx:
 * a decode/process call issued before the previous one was fetched is stream-ordered behind
 * it (whatever streams the two use) and replaces its results, as the next process_paf
 * replaces the reference's module globals (pafprocess.cpp:12-13,39,128); to overlap batches
 * use one ctx per batch in flight.  Calls on different contexts are independent (the
 * reference is not re-entrant).  There is no CPU fallback anywhere: without
 * a CUDA device tfpd_create fails with TFPD_ERR_CUDA.
 */
#ifndef TFPOSE_DECODE_H_
#define TFPOSE_DECODE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TFPD_API __attribute__((visibility("default")))
#else
#define TFPD_API
#endif

#define TFPD_NUM_PARTS 18          /* pafprocess.h:11 NUM_PART */
#define TFPD_NUM_LIMBS 19          /* pafprocess.h:15 COCOPAIRS_SIZE */
#define TFPD_HEAT_CHANNELS 19      /* estimator.py:321 */
#define TFPD_PAF_CHANNELS 38       /* estimator.py:322 */
#define TFPD_NET_CHANNELS 57

enum {
    TFPD_OK = 0,
    TFPD_ERR_INVALID = 1,    /* bad argument (shape, alignment, null pointer, batch > max_batch) */
    TFPD_ERR_CUDA = 2,       /* CUDA runtime error or no device */
    TFPD_ERR_CAPACITY = 3,   /* a per-frame table overflowed (see frame flags); nothing is silently truncated */
    TFPD_ERR_STATE = 4       /* fetch/getter without a preceding decode */
};

/* per-frame flag bits (tfpd_frame_info.flags) */
enum {
    TFPD_FLAG_SCORE_TIE = 1,       /* two candidates of one limb have bit-equal scores: the reference's order is then
                                      what libstdc++'s std::sort makes of the order the candidates were generated in
                                      (pafprocess.cpp:57-60,97).  Informational: the library reproduces that order
                                      (option "std_sort_ties", on by default); with the option off it orders
                                      (score desc, idx1, idx2) */
    TFPD_FLAG_FOUND_GT2 = 2,       /* a connection matched >2 subset rows and was dropped (pafprocess.cpp:136-141) */
    TFPD_FLAG_PEAK_OVERFLOW = 4,
    TFPD_FLAG_CAND_OVERFLOW = 8,
    TFPD_FLAG_ROW_OVERFLOW = 16,
    TFPD_FLAG_HUMAN_OVERFLOW = 32
};

enum { TFPD_MODE_ASBUILT = 0,     /* pafprocess.cpp:152 as compiled by GCC -O2: found==2 never merges */
       TFPD_MODE_UPSTREAM = 1 };  /* `int membership = 0;` (ildoonet upstream): merge disjoint rows */

typedef struct tfpd_ctx tfpd_ctx;

typedef struct tfpd_config {
    int device;               /* CUDA device ordinal */
    int max_batch;            /* frames per decode call */
    int max_h, max_w;         /* largest UPSAMPLED map (H, W) the ctx must handle */
    /* decode constants, pafprocess.h:6-13 */
    float thresh_heat;            /* 0.05 */
    float thresh_vector_score;    /* 0.05 */
    int thresh_vector_cnt1;       /* 6 */
    int thresh_part_cnt;          /* 4 */
    float thresh_human_score;     /* 0.3 */
    int mode;                     /* TFPD_MODE_* */
    /* smoother parameters, estimator.py:332 */
    int blur_ksize;               /* 25 */
    float blur_sigma;             /* 3.0 */
    /* per-frame table capacities (overflow is reported, never silent) */
    int max_peaks;                /* <= 4096 */
    int max_cands;                /* accepted limb candidates per frame */
    int max_rows;                 /* subset rows alive during assembly */
    int max_humans;               /* humans per frame after filtering */
} tfpd_config;

/* One human, as PoseEstimator.estimate_paf reads it through the getters
 * (estimator.py:277-297): cid < 0 means the part is absent. */
typedef struct tfpd_human {
    float score;                        /* get_score: subset[18]/subset[19] (pafprocess.cpp:202-204) */
    int16_t cid[TFPD_NUM_PARTS];        /* get_part_cid */
    int16_t x[TFPD_NUM_PARTS];          /* get_part_x(cid), upsampled pixels */
    int16_t y[TFPD_NUM_PARTS];          /* get_part_y(cid) */
    float part_score[TFPD_NUM_PARTS];   /* get_part_score(cid): un-blurred heat value */
} tfpd_human;

typedef struct tfpd_frame_info {
    int n_humans;
    int n_peaks;
    int n_cands;
    int flags;          /* TFPD_FLAG_* */
    int near_ties;      /* pixels whose peak / threshold decision is within 2e-5 relative of flipping */
    int n_rows;         /* subset rows before filtering (pafprocess.cpp:185) */
} tfpd_frame_info;

/* Optional dense outputs of seam 1, DEVICE pointers (16-byte aligned), NHWC float32,
 * the three tensors the reference's sess.run returns (estimator.py:536-539).
 * All three null = minimal mode (nothing dense is written). All three must be set or none. */
typedef struct tfpd_dense {
    float *heat_up;     /* [N,H,W,19] */
    float *paf_up;      /* [N,H,W,38] */
    float *peaks;       /* [N,H,W,19] */
} tfpd_dense;

TFPD_API void tfpd_default_config(tfpd_config *cfg);
TFPD_API int tfpd_create(const tfpd_config *cfg, tfpd_ctx **out);
TFPD_API void tfpd_destroy(tfpd_ctx *ctx);
/* Message of the last failing call on ctx (or of the last failed tfpd_create when ctx is null). */
TFPD_API const char *tfpd_last_error(const tfpd_ctx *ctx);

/* seam 1, device buffers: net_out is a DEVICE pointer to [N,h,w,57] float32 NHWC (the tensor
 * `Openpose/concat_stage7`, estimator.py:319); (H,W) is upsample_size (estimator.py:325,
 * 523-525).  Asynchronous on `stream` (a cudaStream_t passed as void*; null = default stream). */
TFPD_API int tfpd_decode_device(tfpd_ctx *ctx, const float *net_out, int N, int h, int w, int H, int W,
                       void *stream, const tfpd_dense *dense);

/* seam 1, host buffers: copies net_out host->device, decodes, and (when the pointers are
 * non-null) copies the dense tensors device->host -- what one persistent_sess.run +
 * estimate_paf does (estimator.py:536-551).  Synchronous. */
TFPD_API int tfpd_decode_host(tfpd_ctx *ctx, const float *net_out, int N, int h, int w, int H, int W,
                     float *heat_up, float *paf_up, float *peaks);

/* Same, asynchronous: enqueues the copies and kernels on the ctx's own streams and returns; net_out and
 * the dense output buffers must stay valid until tfpd_fetch_humans (which waits for everything) returns.
 * With page-locked host memory the input of the next batch (on a second ctx) crosses PCIe while this
 * batch is decoded. */
TFPD_API int tfpd_decode_host_async(tfpd_ctx *ctx, const float *net_out, int N, int h, int w, int H, int W,
                           float *heat_up, float *paf_up, float *peaks);

/* Waits for the decode, copies the per-frame tables to the host.
 * humans: [N * max_humans] (frame i at humans + i*max_humans; only the first info[i].n_humans records of a
 * frame are written, the rest of the caller's block is left as it was), info: [N].  Either may be null.
 * Returns TFPD_ERR_CAPACITY if any frame overflowed a table (tables are still filled). */
TFPD_API int tfpd_fetch_humans(tfpd_ctx *ctx, tfpd_human *humans, tfpd_frame_info *info);

/* Peak table of one frame (peak id == index == the reference's Peak.id, pafprocess.cpp:30):
 * arrays of length >= cap; *count receives the number of peaks.  Any array may be null. */
TFPD_API int tfpd_fetch_peaks(tfpd_ctx *ctx, int frame, int cap, int *part, int *x, int *y, float *score, int *count);

/* Accepted connections of one frame, limb-major in acceptance order (pafprocess.cpp:96-124):
 * arrays of length >= cap; *count = number written. */
TFPD_API int tfpd_fetch_connections(tfpd_ctx *ctx, int frame, int cap, int *limb, int *cid1, int *cid2,
                           float *score, int *count);

/* seam 2: same prototype as the reference's process_paf (pafprocess.h:53) plus ctx.
 * peaks [p1,p2,p3], heatmap [h1,h2,h3], pafmap [f1,f2,f3] are HOST float32 C-contiguous arrays
 * (what swig's IN_ARRAY3 typemap hands over, pafprocess_wrap.cpp:4930-4969); p3 = h3 = 19,
 * f3 = 38.  Returns 0 like the reference on success. */
TFPD_API int tfpd_process_paf(tfpd_ctx *ctx, int p1, int p2, int p3, const float *peaks, int h1, int h2, int h3,
                     const float *heatmap, int f1, int f2, int f3, const float *pafmap);
/* same, DEVICE pointers, asynchronous on stream */
TFPD_API int tfpd_process_paf_device(tfpd_ctx *ctx, int p1, int p2, int p3, const float *peaks, int h1, int h2,
                            int h3, const float *heatmap, int f1, int f2, int f3, const float *pafmap,
                            void *stream);
/* Peak table without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Needs no tfpd_ctx; calls no cudaMalloc / cudaFree / allocating library routine.
 *
 * Computes what tfpd_process_paf_device + tfpd_fetch_peaks compute (the reference's scan PEAKS(y,x,part) >
 * thresh_heat over part < 18 in (part, y, x) order, score = HEAT(y,x,part): pafprocess.cpp:22-44) on the CURRENT
 * device, asynchronously on `stream`:
 *   peaks, heatmap : DEVICE float32 [H,W,19]           H <= 8191, W <= 16383
 *   part, x, y     : DEVICE int32 [cap]                score : DEVICE float32 [cap]
 *   count          : DEVICE int32 [2] = { entries written (<= cap), peaks found }; found > cap means the table
 *                    overflowed and holds an arbitrary subset of `cap` peaks, in order
 *   err, err_len   : optional HOST buffer that receives the message of a non-zero return
 * Its scratch (append counter + unordered records) is a static __device__ array of TFPD_NOALLOC_SCRATCH_BYTES
 * that every GPU loading the library holds until the process exits; cap > TFPD_NOALLOC_MAX_PEAKS is rejected with
 * TFPD_ERR_CAPACITY.  All calls on a device SHARE that array: calls that overlap on the GPU (different streams or
 * threads) race -- keep them on one stream or order them with events. */
#define TFPD_NOALLOC_MAX_PEAKS 4096
#define TFPD_NOALLOC_SCRATCH_BYTES (256 + 8 * TFPD_NOALLOC_MAX_PEAKS) /* 33,024 bytes per GPU */
TFPD_API int tfpd_peak_table_noalloc(const float *peaks, const float *heatmap, int H, int W, float thresh_heat,
                            int cap, int32_t *part, int32_t *x, int32_t *y, float *score, int32_t *count,
                            void *stream, char *err, size_t err_len);

/* getters of pafprocess.h:54-59 (frame 0 of the last call).  Unlike the reference they are
 * bounds-checked: out-of-range arguments return -1 / 0.0f and set the error string. */
TFPD_API int tfpd_get_num_humans(tfpd_ctx *ctx);
TFPD_API int tfpd_get_part_cid(tfpd_ctx *ctx, int human_id, int part_id);
TFPD_API float tfpd_get_score(tfpd_ctx *ctx, int human_id);
TFPD_API int tfpd_get_part_x(tfpd_ctx *ctx, int cid);
TFPD_API int tfpd_get_part_y(tfpd_ctx *ctx, int cid);
TFPD_API float tfpd_get_part_score(tfpd_ctx *ctx, int cid);

/* draw_humans (tf_pose/estimator.py:391-415) on a headless buffer: for every human a radius-3, thickness-3
 * circle per present part and a thickness-3 line per CocoPairsRender pair, colours of common.py:108-110, later
 * primitives over earlier ones -- the pixels cv2.circle / cv2.line would set, bit for bit.  No window, no GL.
 *
 * tfpd_draw_humans: images [N,img_h,img_w,3] uint8 (HOST pointer when on_device == 0: copied in, drawn, copied
 *   back; DEVICE pointer otherwise, asynchronous on stream).  centers: HOST int32 [N][max_h][18][2] = (x, y) image
 *   pixels per part exactly as draw_humans computes them (int(part.x * w + 0.5)), x == INT32_MIN for an absent
 *   part; counts: HOST int32 [N] humans per image (<= max_h).
 * tfpd_draw_decoded: draws the humans of the LAST decode on ctx straight from its device tables into DEVICE
 *   images [N,img_h,img_w,3] (N = frames of that decode); nothing crosses PCIe. */
TFPD_API int tfpd_draw_humans(tfpd_ctx *ctx, uint8_t *images, int on_device, int N, int img_h, int img_w,
                     const int32_t *centers, const int32_t *counts, int max_h, void *stream);
TFPD_API int tfpd_draw_decoded(tfpd_ctx *ctx, uint8_t *images, int img_h, int img_w, void *stream);

/* Image pre-processing in front of the network (the step before the path).
 *
 * tfpd_resize_cubic_u8 replaces the cv2.resize(npimg, self.target_size, interpolation=cv2.INTER_CUBIC) of
 *   TfPoseEstimator._get_scaled_img(npimg, None) (tf_pose/estimator.py:417-424; inference calls it at :534-535):
 *   src [N,src_h,src_w,cn] uint8 -> dst [N,dst_h,dst_w,cn] uint8, cn = 1..4, the bytes OpenCV's own 8-bit bicubic
 *   code produces (cv2 with IPP disabled; the closed-source IPP variant differs by 1 LSB on a few percent of the
 *   pixels).  HOST pointers when on_device == 0 (copied in and out, synchronous); DEVICE pointers otherwise,
 *   asynchronous on stream.
 * tfpd_resize_cubic_u8_fxfy is the cv2.resize(npimg, dsize=None, fx=s, fy=s, interpolation=cv2.INTER_CUBIC) form of
 *   _get_scaled_img's scaled variants (estimator.py:428,452,482): the sampling step is 1/fx, 1/fy (not src/dst) and
 *   dst must have the size cv2 derives, dst_w = round_half_even(src_w*fx), dst_h likewise (else TFPD_ERR_INVALID).
 * tfpd_quantize_img replaces TfPoseEstimator._quantize_img (estimator.py:383-388) for float32 images:
 *   dst[i] = (uint8)((src[i] + 1.0f) / (2.0f / 256)), truncated; count elements. */
TFPD_API int tfpd_resize_cubic_u8(tfpd_ctx *ctx, const uint8_t *src, int on_device, int N, int src_h, int src_w, int cn,
                         uint8_t *dst, int dst_h, int dst_w, void *stream);
TFPD_API int tfpd_resize_cubic_u8_fxfy(tfpd_ctx *ctx, const uint8_t *src, int on_device, int N, int src_h, int src_w, int cn,
                              double fx, double fy, uint8_t *dst, int dst_h, int dst_w, void *stream);
TFPD_API int tfpd_quantize_img(tfpd_ctx *ctx, const float *src, int on_device, size_t count, uint8_t *dst, void *stream);

/* Tuning options, by name (unknown name or out-of-range value -> TFPD_ERR_INVALID):
 *   "k1_ctas"      CTAs of the fused upsample/smoother/NMS kernel; 0 = one per SM (default).  The kernel is
 *                  HBM-bound, so with several contexts in flight on different streams a smaller grid (128 of
 *                  148 SMs) leaves SMs on which the latency-bound scoring / assembly kernels of the previous
 *                  batch run concurrently.
 *   "host_chunks"  frame chunks tfpd_decode_host pipelines H2D / decode / D2H over; 0 = automatic.
 *   "k1_l2"        bit 1: evict-first stores (default 2); bit 0 is reserved (a bulk L2 prefetch of the input at
 *                  kernel start measured slower and was taken out again).
 *   "k2_blocks"    CTAs of the limb-scoring kernel per frame; 0 = automatic (about 6 CTAs per SM over the batch).
 *   "std_sort_ties" 1 (default): a limb whose candidates include bit-equal scores is ordered by a literal restatement of
 *                  libstdc++'s std::sort (introsort: median-of-3, unguarded partition, 16-element insertion-sort
 *                  threshold, heap-sort fallback) applied to the candidates in the reference's generation order -- the
 *                  order pafprocess.cpp:97 produces, so the greedy step accepts what the reference accepts.  0: the
 *                  total order (score desc, idx1 asc, idx2 asc).  Frames without such ties are unaffected.
 *   "k2_prefilter" 1 (default): the limb-scoring kernel scores the first 10 - thresh_vector_cnt1 samples of a candidate
 *                  pair (4 with the reference's constants) in its per-pair set-up and drops the pair when none of them
 *                  exceeds thresh_vector_score: it could no longer reach "more than thresh_vector_cnt1 of 10"
 *                  (pafprocess.cpp:80,85).  Exact: same candidates, same scores; 0 scores every pair in full.
 *   "fuse_sort"    1 (default): the limb-scoring kernel ranks the unordered peak records itself (every CTA of a frame
 *                  counts smaller keys in shared memory, CTA 0 writes the table); 0: a sort kernel of its own runs in
 *                  front of it.  Same tables.
 *   "graph"        1: tfpd_decode_device replays a CUDA graph of the whole decode (counter reset, the three kernels,
 *                  the table copy) captured once per distinct argument set (x4 path, profiling off, up to 16
 *                  cached): one launch instead of six stream operations.  Same results.
 *   "k1_turnstile" 1: the map kernels of all contexts on this device run one after the other (stream-ordered through
 *                  a per-device event), whatever their streams.  With several contexts in flight this keeps a
 *                  second HBM-bound map kernel from starting on the SMs "k1_ctas" leaves to the tail kernels.
 *   "k1_early_out" 1 (default): in minimal mode (no dense outputs) the fused map kernel skips a low-res row group of a
 *                  channel when the largest heat value in the 8 low-res rows that group reads is below
 *                  thresh_heat * (1 - 2e-4) / (sum of smoother weights)^2: every smoother weight is >= 0, so no
 *                  blurred value there can come within reach of the heat threshold or of a near-tie test.  Exact:
 *                  same tables, same near-tie counts; 0 computes every row.  Dense mode always computes every row.
 *   "k1_min_band"  minimal mode: low-res rows per CTA of the fused map kernel; 0 (default) = automatic, the largest of
 *                  16 / 8 / 4 that still gives about two waves of CTAs (what whole pipelined steps prefer).  In minimal mode the kernel
 *                  runs three small CTAs (6 warps = 6 part channels each, four CTAs per SM) per band of rows of a
 *                  frame.  Smaller bands balance better, larger ones re-read fewer window rows.  Same results.
 *                  ("k1_ctas" applies to dense mode only.)
 *   "tail_priority" 1: limb scoring and assembly run on a high-priority stream owned by the ctx, forked from and joined
 *                  back into the caller's stream by events (the caller's stream order is unchanged).  With several
 *                  contexts in flight their CTAs are then placed ahead of the pending CTAs of another context's map
 *                  kernel.  0 (default): everything on the caller's stream.  Same results.
 *   "early_fetch"  when the tables start their way to the host: 0 in tfpd_fetch_humans (default), 1 right behind
 *                  the last kernel on the same stream, 2 behind it on a side stream.
 *   "pow2_path"    kernels for power-of-two upsample factors other than the x4 fast path (2, 8, 16): 1 automatic
 *                  (default), 0 the general kernels, 2 fused, 3 fused with a separate horizontal pass, 4 unfused
 *                  polyphase passes.  All give the same maps up to the order of additions; for tests and tuning.
 *   "resize_path"  kernel of tfpd_resize_cubic_u8: 0 automatic (default: the staged strip form -- the source window of
 *                  a strip staged in shared memory by TMA, word loads + byte permutes + dp2a for the horizontal
 *                  sums, packed FP32 for the vertical pass, two columns per thread -- where the source rows are
 *                  4-byte aligned and the window fits, else the strip form), 1 tiled (horizontal pass once per
 *                  source row of a 16-row tile, in shared memory; measured slower), 2 one thread per output pixel,
 *                  3 strip (a thread marches down 16 output rows of a column with a register window of four
 *                  source rows' horizontal sums).  Same bytes every way.  */
TFPD_API int tfpd_set_option(tfpd_ctx *ctx, const char *name, int value);

/* Introspection used by the tests and the bench. */
TFPD_API int tfpd_version(void);
/* 25-tap separable factor and polyphase weights the kernels use (float64 -> float32):
 * k2d receives ksize*ksize float32 weights identical to Smoother.gauss_kernel (smoother.py:62-71). */
TFPD_API int tfpd_gauss_kernel(int ksize, float sigma, float *k2d);
/* Number of kernel launches issued by the last decode/process call, and the device time of
 * each stage of the last call in milliseconds (CUDA events on the caller's stream):
 * stage_ms[0]=K1 maps, [1]=peak sort, [2]=K2 limb scoring, [3]=K3 assembly.  Timing is only
 * recorded when tfpd_set_profiling(ctx, 1) was called. */
TFPD_API int tfpd_last_launch_count(const tfpd_ctx *ctx);
TFPD_API int tfpd_set_profiling(tfpd_ctx *ctx, int on);
TFPD_API int tfpd_last_stage_ms(tfpd_ctx *ctx, float stage_ms[4]);
/* Profiling aid (ctx created with the environment variable TFPD_K1_TIMES=1): per-CTA timestamps of the
 * last fused-map kernel launch, out[cta*4 + {0,1,2,3}] = CTA start / blur-role end / copy-role end
 * (globaltimer ns) / SM id, for the first max_ctas CTAs. */
TFPD_API int tfpd_debug_k1_times(tfpd_ctx *ctx, uint64_t *out, int max_ctas);

#ifdef __cplusplus
}
#endif
#endif /* TFPOSE_DECODE_H_ */
