// api.cu -- the C-ABI of include/tfpose_decode.h: context, workspace, stage sequencing.
// Host-side only; the kernels live in k1_fused.cu / k_generic.cu / k_peaks.cu / k2_limbs.cu /
// k3_assemble.cu.  There is deliberately no CPU implementation of any stage in this library.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <mutex>
#include <vector>

#include "tfpd_internal.cuh"

using namespace tfpd;

static_assert(sizeof(tfpd_human) == 184, "tfpd_human layout is part of the C-ABI (ctypes HumanRec, numpy HUMAN_DTYPE)");
static_assert(sizeof(tfpd_frame_info) == 24, "tfpd_frame_info layout is part of the C-ABI");

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    cudaError_t reserve(size_t need) {
        if (need <= bytes) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaMalloc(&p, need);
        if (e == cudaSuccess) bytes = need;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
};

}  // namespace

struct tfpd_ctx {
    tfpd_config cfg;
    int sm_count = 148;
    Workspace ws{};
    void *ws_block = nullptr;           // one allocation holding every workspace array
    tfpd_frame_info *d_info = nullptr;  // contiguous [info | humans] block (one D2H copy)
    size_t result_bytes = 0;
    unsigned char *h_result = nullptr;  // pinned mirror of the result block
    float *d_k1d = nullptr;             // float32 separable factor (generic path)
    std::vector<double> s1d;
    float wp[4][8] = {};                // polyphase weights of the x4 fast path
    std::vector<float> poly_host;       // host copy of the same (kernel-parameter weights of the fused pass)
    DevBuf poly_buf;                    // per-phase smoother weights of the power-of-two path, for factor poly_log2f
    int poly_log2f = -1, poly_T = 0;
    DevBuf draw_ids, draw_centers, draw_counts, draw_img;  // draw_humans scratch (lazily sized)
    int resize_path = 0;                // 0 auto (staged strip where it applies, else 3), 1 tiled, 2 one thread per output pixel, 3 strip
    DevBuf pre_src, pre_dst;            // host-pointer calls of the image pre-processing (lazily sized)
    DevBuf in_buf, dense_buf, gen_buf;  // lazily sized: host-call input, dense outputs, generic-path scratch
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    // host-buffer calls: H2D, kernels and D2H of successive frame chunks run on three streams
    cudaStream_t st_h2d = nullptr, st_work = nullptr, st_d2h = nullptr;
    std::vector<cudaEvent_t> ev_chunk;  // 2 per chunk: input landed / chunk decoded
    int pow2_path = 1;                  // 0 generic kernels, 1 automatic, 2 fused, 3 fused + separate h pass, 4 unfused
    int host_chunks = 0;                // 0 = automatic (TFPD_HOST_CHUNKS overrides: tuning knob)
    int k1_pr = 0;                      // TFPD_K1_PR
    int k1_grid = 0;                    // TFPD_K1_GRID: CTAs of the fused map kernel (0 = one per SM)
    int k1_l2_knobs = 2, k1_roles = 3;  // TFPD_K1_L2 / TFPD_K1_ROLES tuning and profiling knobs
    int k1_early_out = 1;               // minimal mode: skip low-res row groups that cannot reach the heat threshold
    int k1_min_band = 0;                // minimal mode: low-res rows per CTA of the map kernel (0 = automatic)
    unsigned long long *d_k1_times = nullptr;   // TFPD_K1_TIMES=1: per-CTA timestamps of the last K1 launch
    bool profiling = false, have_timing = false;
    cudaStream_t last_stream = nullptr;
    bool result_enqueued = false;       // the D2H of the tables is already enqueued (on last_stream or st_res)
    bool result_on_side = false;
    // CUDA graphs of whole decodes (x4 fast path): memset + 4 kernels + table D2H replayed with one launch
    struct GraphEntry {
        const void *net_out, *heat_up, *paf_up, *peaks;
        int N, h, w, H, W, k1_grid, k1_l2, launches;
        cudaGraphExec_t exec;
    };
    std::vector<GraphEntry> graphs;
    int use_graph = 0;
    int k1_turnstile = 0;               // serialise the map kernels of all contexts on this device
    int early_fetch = 0;                // 0 copy the tables at fetch time, 1 right behind the last kernel, 2 on st_res
    cudaStream_t st_res = nullptr;
    cudaEvent_t ev_res = nullptr;
    // "tail_priority": limb scoring and assembly run on a high-priority stream of the ctx, forked from and joined
    // back into the caller's stream by events, so that their CTAs are placed ahead of pending map-kernel CTAs of
    // other contexts
    // TFPD_CHECKED=1 in the environment at tfpd_create: a 256-byte guard band filled with 0xA5 follows every workspace
    // array; tfpd_fetch_humans verifies them (an out-of-bounds write of any kernel into the neighbouring array's
    // territory fails the call).  Stands in for compute-sanitizer memcheck, which this GPU pool does not allow.
    std::vector<size_t> guard_offsets;
    int tail_priority = 0;
    cudaStream_t st_tail = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_done = nullptr;      // behind the last operation of the latest decode/process call on last_stream
    int last_N = 0, last_H = 0, last_W = 0;
    int launches = 0;
    bool decoded = false, fetched = false, d2h_pending = false;
    std::string err;
    // host caches for the seam-2 getters
    std::vector<tfpd_human> g_humans;
    std::vector<int> g_px, g_py;
    std::vector<float> g_ps;
    bool getters_ready = false;
};

namespace {

std::mutex g_k1_mu;
cudaEvent_t g_k1_done[64] = {};        // per device: the last map kernel enqueued with k1_turnstile on

int fail(tfpd_ctx *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    else g_create_error = buf;
    return code;
}

#define CUDA_TRY(ctx, expr)                                                                        \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(ctx, TFPD_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__));      \
    } while (0)

// Makes the ctx's device current for the duration of an API call and puts the caller's device back afterwards.
struct DeviceGuard {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int dev) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
#define ON_CTX_DEVICE(ctx)                                                                          \
    DeviceGuard device_guard__((ctx)->cfg.device);                                                  \
    if (device_guard__.err != cudaSuccess)                                                          \
        return fail(ctx, TFPD_ERR_CUDA, "cudaSetDevice(%d): %s", (ctx)->cfg.device, cudaGetErrorString(device_guard__.err))

// Phi(x): scipy.stats.norm.cdf, used by Smoother.gauss_kernel (smoother.py:65)
double std_normal_cdf(double x) { return 0.5 * erfc(-x / sqrt(2.0)); }

// smoother.py:62-71: the 2-D float32 weights and the float64 separable factor s (K = s s^T)
void gauss_kernel_host(int ks, double nsig, std::vector<float> *k2d, std::vector<double> *s1d) {
    const double interval = (2.0 * nsig + 1.0) / ks;
    const double lo = -nsig - interval / 2.0, hi = nsig + interval / 2.0;
    const double step = (hi - lo) / ks;
    std::vector<double> cdf(ks + 1), k1(ks);
    for (int i = 0; i <= ks; ++i) cdf[i] = std_normal_cdf(i == ks ? hi : lo + i * step);
    for (int i = 0; i < ks; ++i) k1[i] = cdf[i + 1] - cdf[i];
    double total = 0.0, ssum = 0.0;
    for (int r = 0; r < ks; ++r)
        for (int q = 0; q < ks; ++q) total += sqrt(k1[r] * k1[q]);
    if (k2d) {
        k2d->resize((size_t)ks * ks);
        for (int r = 0; r < ks; ++r)
            for (int q = 0; q < ks; ++q) (*k2d)[(size_t)r * ks + q] = (float)(sqrt(k1[r] * k1[q]) / total);
    }
    if (s1d) {
        s1d->resize(ks);
        for (int i = 0; i < ks; ++i) ssum += sqrt(k1[i]);
        for (int i = 0; i < ks; ++i) (*s1d)[i] = sqrt(k1[i]) / ssum;
    }
}

int next_pow2(int v) {
    int p = 32;
    while (p < v) p <<= 1;
    return p;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// log2 of the upsample factor when (h,w)->(H,W) is a power-of-two replication the polyphase path handles
// (factor 1..16, at most 33 cell taps), else -1
int pow2_factor(int h, int w, int H, int W, int ksize) {
    for (int l = 0; l <= 4; ++l)
        if (H == (h << l) && W == (w << l)) {
            const int F = 1 << l, R = (ksize / 2 + F - 1) / F;
            return (2 * R + 1 <= 33) ? l : -1;
        }
    return -1;
}

Thresholds thresholds_of(const tfpd_config &c) {
    Thresholds t;
    t.heat = c.thresh_heat;
    t.vector_score = c.thresh_vector_score;
    t.human_score = c.thresh_human_score;
    t.vector_cnt1 = c.thresh_vector_cnt1;
    t.part_cnt = c.thresh_part_cnt;
    t.mode = c.mode;
    return t;
}

// zero the per-frame counters, then stages 2..4 shared by both seams
int run_tail_on(tfpd_ctx *ctx, const Workspace &ws, int N, const PafSource &src, size_t frame_stride, cudaStream_t s,
                bool timed);

int run_tail(tfpd_ctx *ctx, const Workspace &ws, int N, const PafSource &src, size_t frame_stride, cudaStream_t s,
             bool timed) {
    if (!ctx->tail_priority || !ctx->st_tail) return run_tail_on(ctx, ws, N, src, frame_stride, s, timed);
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_fork, s));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->st_tail, ctx->ev_fork, 0));
    const int rc = run_tail_on(ctx, ws, N, src, frame_stride, ctx->st_tail, timed);
    // join even after a launch error: the caller's stream must not be left forked (also ends a stream capture cleanly)
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_join, ctx->st_tail));
    CUDA_TRY(ctx, cudaStreamWaitEvent(s, ctx->ev_join, 0));
    return rc;
}

int run_tail_on(tfpd_ctx *ctx, const Workspace &ws, int N, const PafSource &src, size_t frame_stride, cudaStream_t s,
                bool timed) {
    int r;
    if (timed) cudaEventRecord(ctx->ev[1], s);
    if (!ws.k2_fused_sort) {
        if ((r = launch_sort_peaks(ws, N, s)) < 0) return fail(ctx, TFPD_ERR_CUDA, "k_sort_peaks launch failed");
        ctx->launches += r;
    }
    if (timed) cudaEventRecord(ctx->ev[2], s);
    if ((r = launch_limb_score(ws, N, src, frame_stride, thresholds_of(ctx->cfg), ctx->sm_count, s)) < 0)
        return fail(ctx, TFPD_ERR_CUDA, "k_limb_score launch failed");
    ctx->launches += r;
    if (timed) cudaEventRecord(ctx->ev[3], s);
    if ((r = launch_assemble(ws, N, thresholds_of(ctx->cfg), s)) < 0)
        return fail(ctx, TFPD_ERR_CUDA, "k_assemble launch failed: %s", cudaGetErrorString(cudaGetLastError()));
    ctx->launches += r;
    if (timed) {
        cudaEventRecord(ctx->ev[4], s);
        ctx->have_timing = true;
    }
    return TFPD_OK;
}

// the per-frame arrays of `w` starting at frame f0 (kernels index frames from 0)
Workspace ws_view(const Workspace &w, int f0) {
    Workspace v = w;
    const size_t f = (size_t)f0, MP = (size_t)w.max_peaks, MC = (size_t)w.max_cands;
    v.counters += f;
    v.rec_key += f * MP;
    v.rec_val += f * MP;
    v.pk_x += f * MP;
    v.pk_y += f * MP;
    v.pk_score += f * MP;
    v.part_first += f * 20;
    v.cand_key += f * MC;
    v.conn_cid1 += f * MC;
    v.conn_cid2 += f * MC;
    v.conn_score += f * MC;
    v.conn_start += f * 20;
    v.conn_count += f * 20;
    v.humans += f * (size_t)w.max_humans;
    v.info += f;
    return v;
}

// One call in flight per ctx.  A decode/process call that arrives before the previous one was fetched drops that
// call's results -- the reference overwrites its module globals the same way (pafprocess.cpp:39,128) -- but the
// previous call's kernels and copies still own the workspace until they finish: the new call's work on every stream
// it uses (`s`, and `s2` when non-null) is ordered behind them.
int order_after_previous(tfpd_ctx *ctx, cudaStream_t s, cudaStream_t s2 = nullptr) {
    if (!ctx->decoded || ctx->fetched) return TFPD_OK;
    CUDA_TRY(ctx, cudaStreamWaitEvent(s, ctx->ev_done, 0));
    if (s2) CUDA_TRY(ctx, cudaStreamWaitEvent(s2, ctx->ev_done, 0));
    if (ctx->result_on_side) CUDA_TRY(ctx, cudaStreamSynchronize(ctx->st_res));     // table copy into h_result
    if (ctx->d2h_pending) CUDA_TRY(ctx, cudaStreamSynchronize(ctx->st_d2h));        // dense maps leaving dense_buf
    return TFPD_OK;
}

void begin_call(tfpd_ctx *ctx, int N, int H, int W, cudaStream_t s) {
    ctx->launches = 0;
    ctx->last_stream = s;
    ctx->last_N = N;
    ctx->last_H = H;
    ctx->last_W = W;
    ctx->decoded = false;
    ctx->fetched = false;
    ctx->d2h_pending = false;
    ctx->result_enqueued = false;
    ctx->result_on_side = false;
    ctx->getters_ready = false;
    ctx->have_timing = false;
}

}  // namespace

extern "C" {

int tfpd_version(void) { return 100; }

void tfpd_default_config(tfpd_config *c) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->device = 0;
    c->max_batch = 64;
    c->max_h = 368;
    c->max_w = 656;
    c->thresh_heat = 0.05f;             // pafprocess.h:6-10
    c->thresh_vector_score = 0.05f;
    c->thresh_vector_cnt1 = 6;
    c->thresh_part_cnt = 4;
    c->thresh_human_score = 0.3f;
    c->mode = TFPD_MODE_ASBUILT;
    c->blur_ksize = 25;                 // estimator.py:332
    c->blur_sigma = 3.0f;
    c->max_peaks = 2048;
    c->max_cands = 4096;
    c->max_rows = 512;
    c->max_humans = 64;
}

const char *tfpd_last_error(const tfpd_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int tfpd_gauss_kernel(int ksize, float sigma, float *k2d) {
    if (ksize < 1 || (ksize & 1) == 0 || !k2d) return TFPD_ERR_INVALID;
    std::vector<float> k;
    gauss_kernel_host(ksize, (double)sigma, &k, nullptr);
    memcpy(k2d, k.data(), k.size() * sizeof(float));
    return TFPD_OK;
}

int tfpd_create(const tfpd_config *cfg, tfpd_ctx **out) {
    if (!cfg || !out) return fail(nullptr, TFPD_ERR_INVALID, "tfpd_create: null argument");
    *out = nullptr;
    // max_batch: frames index gridDim.y / gridDim.x of the per-frame kernels (<= 65535)
    if (cfg->max_batch < 1 || cfg->max_batch > 65535 || cfg->max_h < 1 || cfg->max_w < 1 || cfg->max_h > 8191 ||
        cfg->max_w > 16383)
        return fail(nullptr, TFPD_ERR_INVALID, "tfpd_create: bad max_batch (1..65535) / max_h (<=8191) / max_w (<=16383)");
    if (cfg->blur_ksize < 1 || (cfg->blur_ksize & 1) == 0 || cfg->blur_ksize > 127)
        return fail(nullptr, TFPD_ERR_INVALID, "tfpd_create: blur_ksize must be odd and <= 127");
    if (cfg->max_peaks < 1 || cfg->max_peaks > kMaxPeaksHard || cfg->max_cands < 1 || cfg->max_cands > 16384 ||
        cfg->max_rows < 1 || cfg->max_rows > 1024 || cfg->max_humans < 1 || cfg->max_humans > cfg->max_rows)
        return fail(nullptr, TFPD_ERR_INVALID,
                    "tfpd_create: capacities out of range (max_peaks<=4096, max_cands<=16384, "
                    "max_humans<=max_rows<=1024)");
    if (cfg->mode != TFPD_MODE_ASBUILT && cfg->mode != TFPD_MODE_UPSTREAM)
        return fail(nullptr, TFPD_ERR_INVALID, "tfpd_create: unknown mode %d", cfg->mode);
    {
        // K3 keeps a frame's candidate keys, subset rows and greedy bitmaps in shared memory
        Workspace probe{};
        probe.max_cands = next_pow2(cfg->max_cands);
        probe.max_rows = cfg->max_rows;
        probe.max_peaks = next_pow2(cfg->max_peaks);
        if (assemble_smem_bytes(probe) + 1024 > 227 * 1024)
            return fail(nullptr, TFPD_ERR_INVALID,
                        "tfpd_create: max_cands=%d with max_rows=%d (and the peak table) needs %zu bytes of shared memory in the assembly "
                        "kernel (limit 227 KB): lower one of them",
                        probe.max_cands, cfg->max_rows, assemble_smem_bytes(probe));
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(nullptr, TFPD_ERR_CUDA, "tfpd_create: no CUDA device (this library has no CPU path)");
    if (cfg->device < 0 || cfg->device >= ndev)
        return fail(nullptr, TFPD_ERR_INVALID, "tfpd_create: device %d of %d", cfg->device, ndev);
    DeviceGuard device_guard__(cfg->device);
    cudaError_t e = device_guard__.err;
    if (e != cudaSuccess) return fail(nullptr, TFPD_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));

    tfpd_ctx *ctx = new tfpd_ctx();
    ctx->cfg = *cfg;
    ctx->cfg.max_peaks = next_pow2(cfg->max_peaks);     // the in-smem sorts pad to a power of two
    ctx->cfg.max_cands = next_pow2(cfg->max_cands);
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, cfg->device);

    Workspace &w = ctx->ws;
    w.max_batch = cfg->max_batch;
    w.max_peaks = ctx->cfg.max_peaks;
    w.max_cands = ctx->cfg.max_cands;
    w.max_rows = cfg->max_rows;
    w.max_humans = cfg->max_humans;
    w.k3_reg_rows = 128;
    w.k2_fused_sort = 1;
    w.k2_prefilter = 1;
#ifdef TFPD_CHECKED
    w.dbg_trip = getenv("TFPD_CHECKED_TRIP") ? atoi(getenv("TFPD_CHECKED_TRIP")) : 0;
#endif
    w.k2_blocks_per_frame = 0;
    if (const char *e = getenv("TFPD_K2_BPF")) w.k2_blocks_per_frame = atoi(e);
    w.k3_warp_sort_max = 1024;
    w.k3_parallel_limbs = 1;
    w.k3_std_ties = 1;
    if (const char *e = getenv("TFPD_K3_PARALLEL")) w.k3_parallel_limbs = atoi(e) != 0;
    if (const char *e = getenv("TFPD_K3_WARPSORT")) w.k3_warp_sort_max = atoi(e);
    if (const char *e = getenv("TFPD_K3_REGROWS")) w.k3_reg_rows = atoi(e) < 0 ? 0 : atoi(e);
    const size_t N = (size_t)cfg->max_batch, MP = (size_t)w.max_peaks, MC = (size_t)w.max_cands;
    size_t off = 0;
    const bool checked = getenv("TFPD_CHECKED") && atoi(getenv("TFPD_CHECKED"));
    auto take = [&](size_t bytes) {
        size_t o = off;
        off = align_up(off + bytes, 256);
        if (checked && bytes) {
            ctx->guard_offsets.push_back(off);
            off += 256;
        }
        return o;
    };
    const size_t o_cnt = take(N * sizeof(FrameCounters));
    const size_t o_rk = take(N * MP * 4), o_rv = take(N * MP * 4);
    const size_t o_px = take(N * MP * 2), o_py = take(N * MP * 2), o_ps = take(N * MP * 4);
    const size_t o_pf = take(N * 20 * 4);
    const size_t o_ck = take(N * MC * 8);
    const size_t o_c1 = take(N * MC * 2), o_c2 = take(N * MC * 2), o_cs = take(N * MC * 4);
    const size_t o_st = take(N * 20 * 4), o_cc = take(N * 20 * 4);
    const size_t o_res = take(0);
    ctx->result_bytes = align_up(N * sizeof(tfpd_frame_info), 16) + N * w.max_humans * sizeof(tfpd_human);
    take(ctx->result_bytes);
    e = cudaMalloc(&ctx->ws_block, off);
    if (e != cudaSuccess) {
        delete ctx;
        return fail(nullptr, TFPD_ERR_CUDA, "cudaMalloc(%zu) workspace: %s", off, cudaGetErrorString(e));
    }
    cudaMemset(ctx->ws_block, 0, off);
    unsigned char *b = (unsigned char *)ctx->ws_block;
    for (size_t g : ctx->guard_offsets) cudaMemset(b + g, 0xA5, 256);
    w.counters = (FrameCounters *)(b + o_cnt);
    w.rec_key = (uint32_t *)(b + o_rk);
    w.rec_val = (float *)(b + o_rv);
    w.pk_x = (int16_t *)(b + o_px);
    w.pk_y = (int16_t *)(b + o_py);
    w.pk_score = (float *)(b + o_ps);
    w.part_first = (int *)(b + o_pf);
    w.cand_key = (unsigned long long *)(b + o_ck);
    w.conn_cid1 = (int16_t *)(b + o_c1);
    w.conn_cid2 = (int16_t *)(b + o_c2);
    w.conn_score = (float *)(b + o_cs);
    w.conn_start = (int *)(b + o_st);
    w.conn_count = (int *)(b + o_cc);
    w.info = (tfpd_frame_info *)(b + o_res);
    w.humans = (tfpd_human *)(b + o_res + align_up(N * sizeof(tfpd_frame_info), 16));
    ctx->d_info = w.info;

    e = cudaMallocHost((void **)&ctx->h_result, ctx->result_bytes);
    if (e != cudaSuccess) {
        cudaFree(ctx->ws_block);
        delete ctx;
        return fail(nullptr, TFPD_ERR_CUDA, "cudaMallocHost: %s", cudaGetErrorString(e));
    }
    gauss_kernel_host(cfg->blur_ksize, (double)cfg->blur_sigma, nullptr, &ctx->s1d);
    std::vector<float> s32(ctx->s1d.size());
    for (size_t i = 0; i < s32.size(); ++i) s32[i] = (float)ctx->s1d[i];
    e = cudaMalloc((void **)&ctx->d_k1d, s32.size() * 4);
    if (e == cudaSuccess) e = cudaMemcpy(ctx->d_k1d, s32.data(), s32.size() * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess || (cfg->blur_ksize == 25 && polyphase_weights(ctx->s1d.data(), 25, ctx->wp) != 0)) {
        tfpd_destroy(ctx);
        return fail(nullptr, TFPD_ERR_CUDA, "smoother weight upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    for (auto &ev : ctx->ev) cudaEventCreate(&ev);
    cudaStreamCreateWithFlags(&ctx->st_h2d, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&ctx->st_work, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&ctx->st_d2h, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&ctx->st_res, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&ctx->ev_res, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ctx->ev_done, cudaEventDisableTiming);
    {
        int lo = 0, hi = 0;                              // "greatest" priority is the numerically lowest value
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        cudaStreamCreateWithPriority(&ctx->st_tail, cudaStreamNonBlocking, hi);
        cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming);
    }
    if (const char *e = getenv("TFPD_HOST_CHUNKS")) ctx->host_chunks = atoi(e);
    if (const char *e = getenv("TFPD_K1_L2")) ctx->k1_l2_knobs = atoi(e);
    if (const char *e = getenv("TFPD_K1_ROLES")) ctx->k1_roles = atoi(e);
    if (const char *e = getenv("TFPD_K1_PR")) ctx->k1_pr = atoi(e);
    if (const char *e = getenv("TFPD_K1_GRID")) ctx->k1_grid = atoi(e);
    if (const char *e = getenv("TFPD_K1_TIMES"))
        if (atoi(e) && cudaMalloc((void **)&ctx->d_k1_times, 1024 * 4 * 8) == cudaSuccess) {
            cudaMemset(ctx->d_k1_times, 0, 1024 * 4 * 8);
            ctx->ws.dbg_times = ctx->d_k1_times;
        }
    // the memsets / uploads above ran on the legacy default stream, which the non-blocking streams of this ctx and
    // of the callers do not wait for: finish them before the first decode can be enqueued
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        tfpd_destroy(ctx);
        return fail(nullptr, TFPD_ERR_CUDA, "tfpd_create: %s", cudaGetErrorString(e));
    }
    *out = ctx;
    return TFPD_OK;
}

void tfpd_destroy(tfpd_ctx *ctx) {
    if (!ctx) return;
    DeviceGuard device_guard__(ctx->cfg.device);
    if (ctx->decoded) cudaStreamSynchronize(ctx->last_stream);
    for (auto &ev : ctx->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto &ev : ctx->ev_chunk) cudaEventDestroy(ev);
    for (auto &g : ctx->graphs) cudaGraphExecDestroy(g.exec);
    if (ctx->ev_res) cudaEventDestroy(ctx->ev_res);
    if (ctx->ev_done) cudaEventDestroy(ctx->ev_done);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->st_tail) cudaStreamDestroy(ctx->st_tail);
    for (cudaStream_t st : {ctx->st_h2d, ctx->st_work, ctx->st_d2h, ctx->st_res})
        if (st) cudaStreamDestroy(st);
    if (ctx->ws_block) cudaFree(ctx->ws_block);
    if (ctx->h_result) cudaFreeHost(ctx->h_result);
    if (ctx->d_k1d) cudaFree(ctx->d_k1d);
    if (ctx->d_k1_times) cudaFree(ctx->d_k1_times);
    ctx->in_buf.release();
    ctx->dense_buf.release();
    ctx->gen_buf.release();
    ctx->poly_buf.release();
    ctx->draw_ids.release();
    ctx->draw_centers.release();
    ctx->draw_counts.release();
    ctx->draw_img.release();
    ctx->pre_src.release();
    ctx->pre_dst.release();
    delete ctx;
}

int tfpd_set_profiling(tfpd_ctx *ctx, int on) {
    if (!ctx) return TFPD_ERR_INVALID;
    ctx->profiling = on != 0;
    return TFPD_OK;
}

static int draw_reserve_ids(tfpd_ctx *ctx, size_t npix) {
    const size_t had = ctx->draw_ids.bytes;
    CUDA_TRY(ctx, ctx->draw_ids.reserve(npix * 4));
    if (ctx->draw_ids.bytes != had) CUDA_TRY(ctx, cudaMemset(ctx->draw_ids.p, 0, ctx->draw_ids.bytes));   // the paint pass keeps it zero
    return TFPD_OK;
}

int tfpd_draw_humans(tfpd_ctx *ctx, uint8_t *images, int on_device, int N, int img_h, int img_w,
                     const int32_t *centers, const int32_t *counts, int max_h, void *stream) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!images || !centers || !counts) return fail(ctx, TFPD_ERR_INVALID, "draw_humans: null argument");
    if (N < 1 || N > 65535 || img_h < 1 || img_w < 1 || max_h < 1 || img_h > 32767 || img_w > 32767)     // N is gridDim.y of the draw kernels
        return fail(ctx, TFPD_ERR_INVALID, "draw_humans: bad dimensions (N 1..65535, image sides 1..32767)");
    for (int i = 0; i < N; ++i)
        if (counts[i] < 0 || counts[i] > max_h) return fail(ctx, TFPD_ERR_INVALID, "draw_humans: counts[%d]=%d outside 0..%d", i, counts[i], max_h);
    ON_CTX_DEVICE(ctx);
    cudaStream_t s = on_device ? (cudaStream_t)stream : ctx->st_work;
    const size_t npix = (size_t)N * img_h * img_w, cbytes = (size_t)N * max_h * 36 * 4;
    int rc = draw_reserve_ids(ctx, npix);
    if (rc != TFPD_OK) return rc;
    CUDA_TRY(ctx, ctx->draw_centers.reserve(cbytes));
    CUDA_TRY(ctx, ctx->draw_counts.reserve((size_t)N * 4));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->draw_centers.p, centers, cbytes, cudaMemcpyHostToDevice, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->draw_counts.p, counts, (size_t)N * 4, cudaMemcpyHostToDevice, s));
    unsigned char *d_img = images;
    if (!on_device) {
        CUDA_TRY(ctx, ctx->draw_img.reserve(npix * 3));
        d_img = (unsigned char *)ctx->draw_img.p;
        CUDA_TRY(ctx, cudaMemcpyAsync(d_img, images, npix * 3, cudaMemcpyHostToDevice, s));
    }
    if (launch_draw((const int *)ctx->draw_centers.p, (const int *)ctx->draw_counts.p, max_h, d_img, (int *)ctx->draw_ids.p,
                    N, img_h, img_w, s) < 0)
        return fail(ctx, TFPD_ERR_CUDA, "draw kernels failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (!on_device) CUDA_TRY(ctx, cudaMemcpyAsync(images, d_img, npix * 3, cudaMemcpyDeviceToHost, s));
    // the host tables were read by asynchronous copies: they may be reused once this returns
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    return TFPD_OK;
}

int tfpd_draw_decoded(tfpd_ctx *ctx, uint8_t *images, int img_h, int img_w, void *stream) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!ctx->decoded) return fail(ctx, TFPD_ERR_STATE, "draw_decoded: no decode on this ctx yet");
    if (!images || img_h < 1 || img_w < 1 || img_h > 32767 || img_w > 32767)
        return fail(ctx, TFPD_ERR_INVALID, "draw_decoded: bad argument");
    ON_CTX_DEVICE(ctx);
    cudaStream_t s = (cudaStream_t)stream;
    const int N = ctx->last_N, max_h = ctx->ws.max_humans;
    int rc = draw_reserve_ids(ctx, (size_t)N * img_h * img_w);
    if (rc != TFPD_OK) return rc;
    CUDA_TRY(ctx, ctx->draw_centers.reserve((size_t)N * max_h * 36 * 4));
    CUDA_TRY(ctx, ctx->draw_counts.reserve((size_t)N * 4));
    if (s != ctx->last_stream) CUDA_TRY(ctx, cudaStreamSynchronize(ctx->last_stream));     // the tables must exist
    if (launch_centers_from_tables(ctx->ws, N, ctx->last_H, ctx->last_W, img_h, img_w, (int *)ctx->draw_centers.p,
                                   (int *)ctx->draw_counts.p, s) < 0 ||
        launch_draw((const int *)ctx->draw_centers.p, (const int *)ctx->draw_counts.p, max_h, images,
                    (int *)ctx->draw_ids.p, N, img_h, img_w, s) < 0)
        return fail(ctx, TFPD_ERR_CUDA, "draw kernels failed: %s", cudaGetErrorString(cudaGetLastError()));
    return TFPD_OK;
}

static int resize_cubic_impl(tfpd_ctx *ctx, const uint8_t *src, int on_device, int N, int src_h, int src_w, int cn,
                             uint8_t *dst, int dst_h, int dst_w, double fx, double fy, void *stream);

int tfpd_resize_cubic_u8(tfpd_ctx *ctx, const uint8_t *src, int on_device, int N, int src_h, int src_w, int cn,
                         uint8_t *dst, int dst_h, int dst_w, void *stream) {
    return resize_cubic_impl(ctx, src, on_device, N, src_h, src_w, cn, dst, dst_h, dst_w, 0.0, 0.0, stream);
}

int tfpd_resize_cubic_u8_fxfy(tfpd_ctx *ctx, const uint8_t *src, int on_device, int N, int src_h, int src_w, int cn,
                              double fx, double fy, uint8_t *dst, int dst_h, int dst_w, void *stream) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!(fx > 0) || !(fy > 0)) return fail(ctx, TFPD_ERR_INVALID, "resize_cubic_u8_fxfy: fx, fy must be positive");
    // cv::resize with an empty dsize: dsize = saturate_cast<int>(src * f) = round half to even
    const double ew = nearbyint(src_w * fx), eh = nearbyint(src_h * fy);
    if ((double)dst_w != ew || (double)dst_h != eh)
        return fail(ctx, TFPD_ERR_INVALID, "resize_cubic_u8_fxfy: dst is %dx%d, cv2 makes %.0fx%.0f from fx=%g fy=%g", dst_h, dst_w, eh, ew, fx, fy);
    return resize_cubic_impl(ctx, src, on_device, N, src_h, src_w, cn, dst, dst_h, dst_w, fx, fy, stream);
}

static int resize_cubic_impl(tfpd_ctx *ctx, const uint8_t *src, int on_device, int N, int src_h, int src_w, int cn,
                             uint8_t *dst, int dst_h, int dst_w, double fx, double fy, void *stream) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!src || !dst) return fail(ctx, TFPD_ERR_INVALID, "resize_cubic_u8: null argument");
    if (N < 1 || N > 65535 || src_h < 1 || src_w < 1 || dst_h < 1 || dst_w < 1 || dst_h > 65535 || src_h > 32767 ||
        src_w > 32767 || dst_w > 32767)
        return fail(ctx, TFPD_ERR_INVALID, "resize_cubic_u8: bad dimensions %dx%d -> %dx%d, N=%d", src_h, src_w, dst_h, dst_w, N);
    if (cn < 1 || cn > 4) return fail(ctx, TFPD_ERR_INVALID, "resize_cubic_u8: %d channels (1..4 supported)", cn);
    ON_CTX_DEVICE(ctx);
    cudaStream_t s = on_device ? (cudaStream_t)stream : ctx->st_work;
    const size_t in_bytes = (size_t)N * src_h * src_w * cn, out_bytes = (size_t)N * dst_h * dst_w * cn;
    const unsigned char *d_src = src;
    unsigned char *d_dst = dst;
    if (!on_device) {
        CUDA_TRY(ctx, ctx->pre_src.reserve(in_bytes));
        CUDA_TRY(ctx, ctx->pre_dst.reserve(out_bytes));
        d_src = (const unsigned char *)ctx->pre_src.p;
        d_dst = (unsigned char *)ctx->pre_dst.p;
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pre_src.p, src, in_bytes, cudaMemcpyHostToDevice, s));
    }
    if (launch_resize_cubic_u8(d_src, d_dst, N, src_h, src_w, cn, dst_h, dst_w, fx, fy, ctx->resize_path, s) < 0)
        return fail(ctx, TFPD_ERR_CUDA, "resize kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (!on_device) {
        CUDA_TRY(ctx, cudaMemcpyAsync(dst, d_dst, out_bytes, cudaMemcpyDeviceToHost, s));
        CUDA_TRY(ctx, cudaStreamSynchronize(s));
    }
    return TFPD_OK;
}

int tfpd_quantize_img(tfpd_ctx *ctx, const float *src, int on_device, size_t count, uint8_t *dst, void *stream) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!src || !dst) return fail(ctx, TFPD_ERR_INVALID, "quantize_img: null argument");
    if (count == 0) return TFPD_OK;
    ON_CTX_DEVICE(ctx);
    cudaStream_t s = on_device ? (cudaStream_t)stream : ctx->st_work;
    const float *d_src = src;
    unsigned char *d_dst = dst;
    if (!on_device) {
        CUDA_TRY(ctx, ctx->pre_src.reserve(count * 4));
        CUDA_TRY(ctx, ctx->pre_dst.reserve(count));
        d_src = (const float *)ctx->pre_src.p;
        d_dst = (unsigned char *)ctx->pre_dst.p;
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pre_src.p, src, count * 4, cudaMemcpyHostToDevice, s));
    }
    if (launch_quantize_img(d_src, d_dst, count, ctx->sm_count, s) < 0)
        return fail(ctx, TFPD_ERR_CUDA, "quantize kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (!on_device) {
        CUDA_TRY(ctx, cudaMemcpyAsync(dst, d_dst, count, cudaMemcpyDeviceToHost, s));
        CUDA_TRY(ctx, cudaStreamSynchronize(s));
    }
    return TFPD_OK;
}

int tfpd_set_option(tfpd_ctx *ctx, const char *name, int value) {
    if (!ctx || !name) return TFPD_ERR_INVALID;
    if (!strcmp(name, "k1_ctas")) {
        if (value < 0 || value > ctx->sm_count) return fail(ctx, TFPD_ERR_INVALID, "k1_ctas=%d out of range", value);
        ctx->k1_grid = value;
    } else if (!strcmp(name, "host_chunks")) {
        if (value < 0 || value > 4096) return fail(ctx, TFPD_ERR_INVALID, "host_chunks=%d out of range", value);
        ctx->host_chunks = value;
    } else if (!strcmp(name, "k2_blocks")) {
        if (value < 0 || value > 1024) return fail(ctx, TFPD_ERR_INVALID, "k2_blocks=%d out of range", value);
        ctx->ws.k2_blocks_per_frame = value;
    } else if (!strcmp(name, "fuse_sort")) {
        ctx->ws.k2_fused_sort = value != 0;
    } else if (!strcmp(name, "std_sort_ties")) {
        ctx->ws.k3_std_ties = value != 0;
    } else if (!strcmp(name, "k2_prefilter")) {
        ctx->ws.k2_prefilter = value != 0;
    } else if (!strcmp(name, "graph")) {
        ctx->use_graph = value != 0;
    } else if (!strcmp(name, "k1_turnstile")) {
        ctx->k1_turnstile = value != 0;
    } else if (!strcmp(name, "early_fetch")) {
        if (value < 0 || value > 2) return fail(ctx, TFPD_ERR_INVALID, "early_fetch=%d out of range", value);
        ctx->early_fetch = value;
    } else if (!strcmp(name, "pow2_path")) {
        if (value < 0 || value > 4) return fail(ctx, TFPD_ERR_INVALID, "pow2_path=%d out of range", value);
        ctx->pow2_path = value;
    } else if (!strcmp(name, "resize_path")) {
        if (value < 0 || value > 3) return fail(ctx, TFPD_ERR_INVALID, "resize_path=%d out of range", value);
        ctx->resize_path = value;
    } else if (!strcmp(name, "tail_priority")) {
        ctx->tail_priority = value != 0;
    } else if (!strcmp(name, "k1_early_out")) {
        ctx->k1_early_out = value != 0;
    } else if (!strcmp(name, "k1_min_band")) {
        if (value < 0 || value > 8191) return fail(ctx, TFPD_ERR_INVALID, "k1_min_band=%d out of range", value);
        ctx->k1_min_band = value;
    } else if (!strcmp(name, "k1_l2")) {
        if (value < 0 || value > 7) return fail(ctx, TFPD_ERR_INVALID, "k1_l2=%d out of range", value);
        ctx->k1_l2_knobs = value;
    } else {
        return fail(ctx, TFPD_ERR_INVALID, "unknown option '%s'", name);
    }
    // captured graphs bake in grid sizes, kernel choices and the copy placement: any option change drops them
    if (!ctx->graphs.empty()) {
        DeviceGuard device_guard__(ctx->cfg.device);
        if (ctx->decoded) cudaStreamSynchronize(ctx->last_stream);
        for (auto &g : ctx->graphs) cudaGraphExecDestroy(g.exec);
        ctx->graphs.clear();
    }
    return TFPD_OK;
}

int tfpd_last_launch_count(const tfpd_ctx *ctx) { return ctx ? ctx->launches : 0; }

#ifdef TFPD_CHECKED
// checked build only (negative control of tools/checked_run.sh): overwrite one byte of workspace guard band `which`
extern "C" TFPD_API int tfpd_debug_poke_guard(tfpd_ctx *ctx, int which) {
    if (!ctx || which < 0 || (size_t)which >= ctx->guard_offsets.size()) return TFPD_ERR_INVALID;
    DeviceGuard device_guard__(ctx->cfg.device);
    return cudaMemset((unsigned char *)ctx->ws_block + ctx->guard_offsets[which] + 17, 0, 1) == cudaSuccess ? TFPD_OK : TFPD_ERR_CUDA;
}
#endif

int tfpd_debug_k1_times(tfpd_ctx *ctx, uint64_t *out, int max_ctas) {
    if (!ctx || !out || max_ctas < 1) return TFPD_ERR_INVALID;
    if (!ctx->d_k1_times) return fail(ctx, TFPD_ERR_STATE, "per-CTA timestamps are off (create the ctx with TFPD_K1_TIMES=1)");
    ON_CTX_DEVICE(ctx);
    CUDA_TRY(ctx, cudaDeviceSynchronize());
    const int n = max_ctas < 1024 ? max_ctas : 1024;
    CUDA_TRY(ctx, cudaMemcpy(out, ctx->d_k1_times, (size_t)n * 4 * 8, cudaMemcpyDeviceToHost));
    return TFPD_OK;
}

int tfpd_last_stage_ms(tfpd_ctx *ctx, float stage_ms[4]) {
    if (!ctx || !stage_ms) return TFPD_ERR_INVALID;
    if (!ctx->have_timing) return fail(ctx, TFPD_ERR_STATE, "no timing recorded (tfpd_set_profiling + decode first)");
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev[4]));
    for (int i = 0; i < 4; ++i) CUDA_TRY(ctx, cudaEventElapsedTime(&stage_ms[i], ctx->ev[i], ctx->ev[i + 1]));
    return TFPD_OK;
}

// The tables of a decode start their way back to the page-locked result block as soon as the last kernel is
// enqueued, so that tfpd_fetch_humans only has to wait for them (it used to start the copy itself).
static int enqueue_result_copy(tfpd_ctx *ctx, int N, cudaStream_t s, bool at_fetch = false) {
    if (!at_fetch && ctx->early_fetch == 0) return TFPD_OK;           // tfpd_fetch_humans starts the copy
    cudaStream_t cs = s;
    if (!at_fetch && ctx->early_fetch == 2) {                         // on a side stream, behind the last kernel
        CUDA_TRY(ctx, cudaEventRecord(ctx->ev_res, s));
        CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->st_res, ctx->ev_res, 0));
        cs = ctx->st_res;
        ctx->result_on_side = true;
    }
    const size_t info_bytes = align_up((size_t)ctx->cfg.max_batch * sizeof(tfpd_frame_info), 16);
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_result, ctx->d_info, (size_t)N * sizeof(tfpd_frame_info),
                                  cudaMemcpyDeviceToHost, cs));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_result + info_bytes, ctx->ws.humans,
                                  (size_t)N * ctx->ws.max_humans * sizeof(tfpd_human), cudaMemcpyDeviceToHost, cs));
    ctx->result_enqueued = true;
    return TFPD_OK;
}

// Enqueue the whole decode of N frames on stream s into the workspace view `ws` (arguments validated by the caller).
static int enqueue_decode(tfpd_ctx *ctx, const Workspace &ws, const float *net_out, int N, int h, int w, int H, int W,
                          cudaStream_t s, const tfpd_dense *dense, bool timed) {
    const bool want_dense = dense && dense->heat_up;
    CUDA_TRY(ctx, cudaMemsetAsync(ws.counters, 0, (size_t)N * sizeof(FrameCounters), s));
    if (timed) cudaEventRecord(ctx->ev[0], s);

    PafSource src;
    size_t frame_stride;
    const bool fast = (H == 4 * h && W == 4 * w && ctx->cfg.blur_ksize == 25);
    if (fast) {
        K1Params p;
        p.X = net_out;
        p.heat_up = want_dense ? dense->heat_up : nullptr;
        p.paf_up = want_dense ? dense->paf_up : nullptr;
        p.peaks = want_dense ? dense->peaks : nullptr;
        p.rec_key = ws.rec_key;
        p.rec_val = ws.rec_val;
        p.counters = ws.counters;
        p.max_peaks = ws.max_peaks;
        p.N = N;
        p.h = h;
        p.w = w;
        p.nchunks = (w + 55) / 56;
        p.chunk_cells = (w + p.nchunks - 1) / p.nchunks;
        static const int copy_cells = getenv("TFPD_K1_COPY_CELLS") ? atoi(getenv("TFPD_K1_COPY_CELLS")) : 32;
        p.copy_nchunks = (w + copy_cells - 1) / copy_cells;  // <= 32 cells (29 KB staged) per copy item
        p.copy_chunk = (w + p.copy_nchunks - 1) / p.copy_nchunks;
        p.thresh = ctx->cfg.thresh_heat;
        p.min_band = ctx->k1_min_band;
        if (p.min_band <= 0) {
            // automatic: the largest band of 16 / 8 / 4 low-res rows that still gives ~1.9 waves of CTAs (three CTAs per
            // band, four resident per SM).  Larger bands cost less SM time in total (fewer halo rows and window
            // prologues), smaller ones balance the last wave.  Measured on whole steps (profiles/r02s_band_sweep.log, 64
            // frames, band 4 | 8 | 16): 46x54 maps serial 118.0 | 119.3 | 124.9 us, four contexts in flight 60.9 | 57.1 |
            // 56.5; 46x82 serial 142.6 | 140.0 | 142.5, in flight 89.4 | 79.2 | 73.6; 92x164 serial 219 | 196 | 195, in
            // flight 159 | 132 | 126.  Round 2 first asked for ~6 waves (K1 timed alone: 50.0 | 51.5 | 54.9 us), which
            // left 5-18 % of the pipelined rate on the table for 1-3 % of the serial step.
            const long long enough = 19LL * 4 * ctx->sm_count / 10;
            p.min_band = 4;
            for (int cand = 16; cand >= 8; cand /= 2)
                if ((long long)N * p.nchunks * ((h + cand - 1) / cand) * (kParts / 6) >= enough) {
                    p.min_band = cand;
                    break;
                }
        }
        memcpy(p.wp, ctx->wp, sizeof(p.wp));
        for (int q = 0; q < 2; ++q)
            for (int t = 0; t < 8; ++t) p.wh[q][t] = make_float2(ctx->wp[2 * q][t], ctx->wp[2 * q + 1][t]);
        // Early-out bound of minimal mode.  Every polyphase weight is >= 0 and each phase's seven weights sum to
        // sum_w <= 1 + 1e-6, so a blurred value is at most max(0, largest heat value in its 7x7 low-res window) *
        // sum_w^2, up to the rounding of 14 FMAs (< 2e-6 relative).  Nothing downstream looks at a pixel unless its
        // blurred value exceeds thresh * (1 - 4e-5) (peak records and both near-tie tests), and the winner-side
        // near-tie test compares against neighbours within 2e-5 relative: a window below thresh * (1 - 2e-4) is
        // provably out of reach of all of them.  Disabled (bound = -inf: never dead) for a non-positive threshold.
        {
            double sw = 0.0;
            bool nonneg = true;                 // the bound needs non-negative weights (true for any Gaussian factor)
            for (int ph = 0; ph < 4; ++ph) {
                double a = 0.0;
                for (int t = 0; t < 7; ++t) {
                    a += (double)ctx->wp[ph][t];
                    nonneg = nonneg && ctx->wp[ph][t] >= 0.0f;
                }
                sw = a > sw ? a : sw;
            }
            const double bound = (double)ctx->cfg.thresh_heat * (1.0 - 2e-4) / (sw * sw);
            p.dead_bound = (ctx->cfg.thresh_heat > 0.0f && ctx->k1_early_out && nonneg) ? (float)bound : -INFINITY;
        }
        p.l2_knobs = ctx->k1_l2_knobs;
        p.debug_roles = ctx->k1_roles;
        p.dbg_times = ctx->d_k1_times;
        p.force_pr = ctx->k1_pr;
        int r;
        if (ctx->k1_turnstile) {
            // One map kernel at a time per device, whatever the ctx and stream: it is HBM-bound, so two of them only
            // dilute each other, and the early CTAs of the second one would take the SMs a reduced grid (k1_ctas)
            // leaves to the scoring / assembly kernels of the batches before it.
            std::lock_guard<std::mutex> lock(g_k1_mu);
            cudaEvent_t &done = g_k1_done[ctx->cfg.device & 63];
            if (done) CUDA_TRY(ctx, cudaStreamWaitEvent(s, done, 0));
            else CUDA_TRY(ctx, cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
            r = launch_k1_fused_f4(p, want_dense, ctx->k1_grid > 0 ? ctx->k1_grid : ctx->sm_count, s);
            if (r >= 0) CUDA_TRY(ctx, cudaEventRecord(done, s));
        } else {
            r = launch_k1_fused_f4(p, want_dense, ctx->k1_grid > 0 ? ctx->k1_grid : ctx->sm_count, s);
        }
        if (r < 0) return fail(ctx, TFPD_ERR_CUDA, "k1_fused_f4 launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        ctx->launches += r;
        src.base = net_out;
        src.rows = h;
        src.cols = w;
        src.chan = kNetC;
        src.chan_off = kHeatC;
        src.shift = 2;
        src.H = H;
        frame_stride = (size_t)h * w * kNetC;
    } else if (const int log2f = ctx->pow2_path ? pow2_factor(h, w, H, W, ctx->cfg.blur_ksize) : -1; log2f >= 0) {
        // power-of-two replication other than the fused kernel's case: polyphase passes on the low-res rows
        if (ctx->poly_log2f != log2f) {
            const int F = 1 << log2f, pad = ctx->cfg.blur_ksize / 2, R = (pad + F - 1) / F, T = 2 * R + 1;
            std::vector<float> wp((size_t)F * T);
            for (int p = 0; p < F; ++p) {
                std::vector<double> acc(T, 0.0);
                for (int q = 0; q < ctx->cfg.blur_ksize; ++q) {
                    const int d = p + q - pad;                       // upsampled offset relative to the cell start
                    const int cell = (d >= 0) ? d / F : -((-d + F - 1) / F);
                    acc[cell + R] += ctx->s1d[q];
                }
                for (int t = 0; t < T; ++t) wp[(size_t)p * T + t] = (float)acc[t];
            }
            CUDA_TRY(ctx, ctx->poly_buf.reserve(wp.size() * 4));
            CUDA_TRY(ctx, cudaMemcpyAsync(ctx->poly_buf.p, wp.data(), wp.size() * 4, cudaMemcpyHostToDevice, s));
            CUDA_TRY(ctx, cudaStreamSynchronize(s));                 // wp is a local; once per factor
            ctx->poly_host = wp;
            ctx->poly_log2f = log2f;
            ctx->poly_T = T;
        }
        const size_t px = (size_t)N * H * W;
        CUDA_TRY(ctx, ctx->gen_buf.reserve(((size_t)N * h * W + px) * kHeatC * 4));
        int r = launch_pow2_maps(net_out, N, h, w, log2f, (const float *)ctx->poly_buf.p, ctx->poly_host.data(), ctx->poly_T,
                                 ctx->pow2_path,
                                 want_dense ? dense->heat_up : nullptr, want_dense ? dense->paf_up : nullptr,
                                 want_dense ? dense->peaks : nullptr, (float *)ctx->gen_buf.p, ws, ctx->cfg.thresh_heat, s);
        if (r < 0) return fail(ctx, TFPD_ERR_CUDA, "power-of-two map kernels failed: %s", cudaGetErrorString(cudaGetLastError()));
        ctx->launches += r;
        src.base = net_out;
        src.rows = h;
        src.cols = w;
        src.chan = kNetC;
        src.chan_off = kHeatC;
        src.shift = log2f;
        src.H = H;
        frame_stride = (size_t)h * w * kNetC;
    } else {
        const size_t px = (size_t)N * H * W;
        float *heat_up, *paf_up, *peaks;
        size_t need = 2 * px * kHeatC * 4;
        if (!want_dense) need += px * (kHeatC + kPafC) * 4;
        CUDA_TRY(ctx, ctx->gen_buf.reserve(need));
        float *scratch = (float *)ctx->gen_buf.p;
        if (want_dense) {
            heat_up = dense->heat_up;
            paf_up = dense->paf_up;
            peaks = dense->peaks;
        } else {
            heat_up = scratch + 2 * px * kHeatC;
            paf_up = heat_up + px * kHeatC;
            peaks = nullptr;
        }
        int r = launch_generic_maps(net_out, N, h, w, H, W, ctx->d_k1d, ctx->cfg.blur_ksize, heat_up, paf_up, peaks,
                                    scratch, ws, ctx->cfg.thresh_heat, s);
        if (r < 0) return fail(ctx, TFPD_ERR_CUDA, "generic map kernels failed: %s", cudaGetErrorString(cudaGetLastError()));
        ctx->launches += r;
        src.base = paf_up;
        src.rows = H;
        src.cols = W;
        src.chan = kPafC;
        src.chan_off = 0;
        src.shift = 0;
        src.H = H;
        frame_stride = (size_t)H * W * kPafC;
    }
    return run_tail(ctx, ws, N, src, frame_stride, s, timed);
}

int tfpd_decode_device(tfpd_ctx *ctx, const float *net_out, int N, int h, int w, int H, int W, void *stream,
                       const tfpd_dense *dense) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!net_out) return fail(ctx, TFPD_ERR_INVALID, "decode: net_out is null");
    if (((uintptr_t)net_out) & 15) return fail(ctx, TFPD_ERR_INVALID, "decode: net_out must be 16-byte aligned");
    if (N < 1 || N > ctx->cfg.max_batch) return fail(ctx, TFPD_ERR_INVALID, "decode: batch %d outside 1..%d", N, ctx->cfg.max_batch);
    if (h < 1 || w < 1 || H < 1 || W < 1) return fail(ctx, TFPD_ERR_INVALID, "decode: non-positive dimension");
    if (H > ctx->cfg.max_h || W > ctx->cfg.max_w)
        return fail(ctx, TFPD_ERR_INVALID, "decode: upsample %dx%d exceeds ctx max %dx%d", H, W, ctx->cfg.max_h, ctx->cfg.max_w);
    const bool want_dense = dense && (dense->heat_up || dense->paf_up || dense->peaks);
    if (want_dense && !(dense->heat_up && dense->paf_up && dense->peaks))
        return fail(ctx, TFPD_ERR_INVALID, "decode: dense outputs must be all set or all null");
    if (want_dense && ((((uintptr_t)dense->heat_up) | ((uintptr_t)dense->paf_up) | ((uintptr_t)dense->peaks)) & 15))
        return fail(ctx, TFPD_ERR_INVALID, "decode: dense output pointers must be 16-byte aligned");
    ON_CTX_DEVICE(ctx);
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if ((rc = order_after_previous(ctx, s)) != TFPD_OK) return rc;
    begin_call(ctx, N, H, W, s);
    const bool fast = (H == 4 * h && W == 4 * w && ctx->cfg.blur_ksize == 25);
    if (ctx->use_graph && fast && !ctx->profiling && !ctx->k1_turnstile && !ctx->d_k1_times) {
        // the launch-bound part of a decode (seven stream operations) as one graph launch; captured once per
        // distinct argument set on an internal stream (never executed there), replayed on the caller's stream
        const void *hu = want_dense ? dense->heat_up : nullptr, *pu = want_dense ? dense->paf_up : nullptr;
        const void *pk = want_dense ? dense->peaks : nullptr;
        tfpd_ctx::GraphEntry *hit = nullptr;
        for (auto &g : ctx->graphs)
            if (g.net_out == net_out && g.heat_up == hu && g.paf_up == pu && g.peaks == pk && g.N == N && g.h == h &&
                g.w == w && g.H == H && g.W == W && g.k1_grid == ctx->k1_grid && g.k1_l2 == ctx->k1_l2_knobs) {
                hit = &g;
                break;
            }
        if (!hit) {
            CUDA_TRY(ctx, cudaStreamBeginCapture(ctx->st_work, cudaStreamCaptureModeRelaxed));
            rc = enqueue_decode(ctx, ctx->ws, net_out, N, h, w, H, W, ctx->st_work, want_dense ? dense : nullptr, false);
            if (rc == TFPD_OK) rc = enqueue_result_copy(ctx, N, ctx->st_work, true);
            cudaGraph_t graph = nullptr;
            cudaError_t ce = cudaStreamEndCapture(ctx->st_work, &graph);
            if (rc != TFPD_OK) {
                if (graph) cudaGraphDestroy(graph);
                return rc;
            }
            if (ce != cudaSuccess) return fail(ctx, TFPD_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(ce));
            cudaGraphExec_t exec = nullptr;
            ce = cudaGraphInstantiate(&exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ce != cudaSuccess) return fail(ctx, TFPD_ERR_CUDA, "graph instantiation failed: %s", cudaGetErrorString(ce));
            if (ctx->graphs.size() >= 16) {                  // oldest out
                cudaGraphExecDestroy(ctx->graphs.front().exec);
                ctx->graphs.erase(ctx->graphs.begin());
            }
            ctx->graphs.push_back({net_out, hu, pu, pk, N, h, w, H, W, ctx->k1_grid, ctx->k1_l2_knobs, ctx->launches, exec});
            hit = &ctx->graphs.back();
        }
        ctx->launches = hit->launches;
        CUDA_TRY(ctx, cudaGraphLaunch(hit->exec, s));
        CUDA_TRY(ctx, cudaEventRecord(ctx->ev_done, s));
        ctx->result_enqueued = true;
        ctx->decoded = true;
        return TFPD_OK;
    }
    rc = enqueue_decode(ctx, ctx->ws, net_out, N, h, w, H, W, s, want_dense ? dense : nullptr, ctx->profiling);
    if (rc != TFPD_OK) return rc;
    if ((rc = enqueue_result_copy(ctx, N, s)) != TFPD_OK) return rc;
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_done, s));
    ctx->decoded = true;
    return TFPD_OK;
}

int tfpd_fetch_humans(tfpd_ctx *ctx, tfpd_human *humans, tfpd_frame_info *info) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!ctx->decoded) return fail(ctx, TFPD_ERR_STATE, "fetch: no decode on this ctx yet");
    const int N = ctx->last_N;
    const size_t info_bytes = align_up((size_t)ctx->cfg.max_batch * sizeof(tfpd_frame_info), 16);
    if (!ctx->fetched) {
        ON_CTX_DEVICE(ctx);
        if (!ctx->result_enqueued) {
            int rc = enqueue_result_copy(ctx, N, ctx->last_stream, true);
            if (rc != TFPD_OK) return rc;
        }
        CUDA_TRY(ctx, cudaStreamSynchronize(ctx->last_stream));
        if (ctx->result_on_side) CUDA_TRY(ctx, cudaStreamSynchronize(ctx->st_res));
        if (ctx->d2h_pending) {         // dense maps of a host-buffer call still on their way back
            CUDA_TRY(ctx, cudaStreamSynchronize(ctx->st_d2h));
            ctx->d2h_pending = false;
        }
        ctx->fetched = true;
        if (!ctx->guard_offsets.empty()) {
            unsigned char probe[256];
            for (size_t gi = 0; gi < ctx->guard_offsets.size(); ++gi) {
                CUDA_TRY(ctx, cudaMemcpy(probe, (unsigned char *)ctx->ws_block + ctx->guard_offsets[gi], 256, cudaMemcpyDeviceToHost));
                for (int k = 0; k < 256; ++k)
                    if (probe[k] != 0xA5)
                        return fail(ctx, TFPD_ERR_CUDA, "TFPD_CHECKED: guard band %zu of the workspace was overwritten at byte %d "
                                                        "(an out-of-bounds write behind workspace array %zu)", gi, k, gi);
            }
        }
    }
    const tfpd_frame_info *hi = (const tfpd_frame_info *)ctx->h_result;
    if (info) memcpy(info, hi, (size_t)N * sizeof(tfpd_frame_info));
    if (humans) {                       // only the records that exist: the rest of the caller's block stays untouched
        const tfpd_human *src = (const tfpd_human *)(ctx->h_result + info_bytes);
        const size_t MH = (size_t)ctx->ws.max_humans;
        for (int i = 0; i < N; ++i) {
            const size_t n = (size_t)std::min(std::max(hi[i].n_humans, 0), (int)MH);
            if (n) memcpy(humans + i * MH, src + i * MH, n * sizeof(tfpd_human));
        }
    }
    const int ovf = TFPD_FLAG_PEAK_OVERFLOW | TFPD_FLAG_CAND_OVERFLOW | TFPD_FLAG_ROW_OVERFLOW | TFPD_FLAG_HUMAN_OVERFLOW;
    for (int i = 0; i < N; ++i)
        if (hi[i].flags & ovf)
            return fail(ctx, TFPD_ERR_CAPACITY,
                        "frame %d overflowed a table (flags=0x%x: 4=peaks>%d 8=cands>%d 16=rows>%d 32=humans>%d); "
                        "raise the capacity in tfpd_config",
                        i, hi[i].flags, ctx->ws.max_peaks, ctx->ws.max_cands, ctx->ws.max_rows, ctx->ws.max_humans);
    return TFPD_OK;
}

int tfpd_fetch_peaks(tfpd_ctx *ctx, int frame, int cap, int *part, int *x, int *y, float *score, int *count) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!ctx->decoded) return fail(ctx, TFPD_ERR_STATE, "fetch_peaks: no decode on this ctx yet");
    if (frame < 0 || frame >= ctx->last_N || cap < 0) return fail(ctx, TFPD_ERR_INVALID, "fetch_peaks: bad frame/cap");
    ON_CTX_DEVICE(ctx);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->last_stream));
    int first[20];
    CUDA_TRY(ctx, cudaMemcpy(first, ctx->ws.part_first + frame * 20, sizeof(first), cudaMemcpyDeviceToHost));
    const int n = first[kParts];
    if (count) *count = n;
    const int m = n < cap ? n : cap;
    if (m > 0) {
        std::vector<int16_t> hx(m), hy(m);
        const size_t o = (size_t)frame * ctx->ws.max_peaks;
        CUDA_TRY(ctx, cudaMemcpy(hx.data(), ctx->ws.pk_x + o, (size_t)m * 2, cudaMemcpyDeviceToHost));
        CUDA_TRY(ctx, cudaMemcpy(hy.data(), ctx->ws.pk_y + o, (size_t)m * 2, cudaMemcpyDeviceToHost));
        if (score) CUDA_TRY(ctx, cudaMemcpy(score, ctx->ws.pk_score + o, (size_t)m * 4, cudaMemcpyDeviceToHost));
        int p = 0;
        for (int i = 0; i < m; ++i) {
            while (p < kParts - 1 && i >= first[p + 1]) ++p;
            if (part) part[i] = p;
            if (x) x[i] = hx[i];
            if (y) y[i] = hy[i];
        }
    }
    return TFPD_OK;
}

int tfpd_fetch_connections(tfpd_ctx *ctx, int frame, int cap, int *limb, int *cid1, int *cid2, float *score,
                           int *count) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!ctx->decoded) return fail(ctx, TFPD_ERR_STATE, "fetch_connections: no decode on this ctx yet");
    if (frame < 0 || frame >= ctx->last_N || cap < 0) return fail(ctx, TFPD_ERR_INVALID, "fetch_connections: bad frame/cap");
    ON_CTX_DEVICE(ctx);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->last_stream));
    int st[20], cc[20];
    CUDA_TRY(ctx, cudaMemcpy(st, ctx->ws.conn_start + frame * 20, sizeof(st), cudaMemcpyDeviceToHost));
    CUDA_TRY(ctx, cudaMemcpy(cc, ctx->ws.conn_count + frame * 20, sizeof(cc), cudaMemcpyDeviceToHost));
    const size_t MC = (size_t)ctx->ws.max_cands;
    std::vector<int16_t> h1(MC), h2(MC);
    std::vector<float> hs(MC);
    CUDA_TRY(ctx, cudaMemcpy(h1.data(), ctx->ws.conn_cid1 + frame * MC, MC * 2, cudaMemcpyDeviceToHost));
    CUDA_TRY(ctx, cudaMemcpy(h2.data(), ctx->ws.conn_cid2 + frame * MC, MC * 2, cudaMemcpyDeviceToHost));
    CUDA_TRY(ctx, cudaMemcpy(hs.data(), ctx->ws.conn_score + frame * MC, MC * 4, cudaMemcpyDeviceToHost));
    int k = 0;
    for (int l = 0; l < kLimbs; ++l)
        for (int i = 0; i < cc[l]; ++i, ++k) {
            if (k >= cap) continue;
            if (limb) limb[k] = l;
            if (cid1) cid1[k] = h1[st[l] + i];
            if (cid2) cid2[k] = h2[st[l] + i];
            if (score) score[k] = hs[st[l] + i];
        }
    if (count) *count = k;
    return TFPD_OK;
}

int tfpd_decode_host_async(tfpd_ctx *ctx, const float *net_out, int N, int h, int w, int H, int W, float *heat_up,
                           float *paf_up, float *peaks) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!net_out) return fail(ctx, TFPD_ERR_INVALID, "decode_host: net_out is null");
    if (N < 1 || N > ctx->cfg.max_batch || h < 1 || w < 1 || H < 1 || W < 1)
        return fail(ctx, TFPD_ERR_INVALID, "decode_host: bad batch or dimension");
    if (H > ctx->cfg.max_h || W > ctx->cfg.max_w)
        return fail(ctx, TFPD_ERR_INVALID, "decode_host: upsample %dx%d exceeds ctx max %dx%d", H, W, ctx->cfg.max_h,
                    ctx->cfg.max_w);
    const bool want_dense = heat_up || paf_up || peaks;
    if (want_dense && !(heat_up && paf_up && peaks))
        return fail(ctx, TFPD_ERR_INVALID, "decode_host: dense outputs must be all set or all null");
    ON_CTX_DEVICE(ctx);
    if (int rc = order_after_previous(ctx, ctx->st_work, ctx->st_h2d)) return rc;
    const size_t frame_in = (size_t)h * w * kNetC * 4;      // multiple of 4; chunk offsets are kept 16-byte aligned below
    const size_t fpx = (size_t)H * W;
    CUDA_TRY(ctx, ctx->in_buf.reserve((size_t)N * frame_in));
    float *d_heat = nullptr, *d_paf = nullptr, *d_peaks = nullptr;
    if (want_dense) {
        const size_t px = (size_t)N * fpx;
        CUDA_TRY(ctx, ctx->dense_buf.reserve(px * (2 * kHeatC + kPafC) * 4));
        d_heat = (float *)ctx->dense_buf.p;
        d_paf = d_heat + px * kHeatC;
        d_peaks = d_paf + px * kPafC;
    }
    // Frame chunks: while chunk k is decoded, chunk k+1 crosses PCIe and (dense mode) the maps of chunk k-1 go
    // back on the other DMA direction.  The x4 fast path only (the generic path has whole-batch scratch), chunk
    // boundaries on 16-byte aligned input offsets, at least ~4 MB of input per chunk.
    const bool fast = (H == 4 * h && W == 4 * w && ctx->cfg.blur_ksize == 25);
    int nchunk = 1;
    if (fast && N > 1) {
        int want = ctx->host_chunks > 0 ? ctx->host_chunks : 4;
        const size_t min_bytes = ctx->host_chunks > 0 ? 1 : ((size_t)4 << 20);
        while (want > 1 && ((size_t)N * frame_in) / want < min_bytes) --want;
        nchunk = want < N ? want : N;
    }
    int align_frames = 1;
    while ((align_frames * frame_in) & 15) align_frames *= 2;          // <= 4
    int per = (N + nchunk - 1) / nchunk;
    per = (per + align_frames - 1) / align_frames * align_frames;
    nchunk = (N + per - 1) / per;
    while ((int)ctx->ev_chunk.size() < 2 * nchunk) {
        cudaEvent_t e;
        CUDA_TRY(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->ev_chunk.push_back(e);
    }
    begin_call(ctx, N, H, W, ctx->st_work);
    const bool timed = ctx->profiling && nchunk == 1;
    for (int k = 0; k < nchunk; ++k) {
        const int f0 = k * per, n = (f0 + per <= N) ? per : N - f0;
        const unsigned char *src = (const unsigned char *)net_out + (size_t)f0 * frame_in;
        unsigned char *dst = (unsigned char *)ctx->in_buf.p + (size_t)f0 * frame_in;
        CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, (size_t)n * frame_in, cudaMemcpyHostToDevice, ctx->st_h2d));
        CUDA_TRY(ctx, cudaEventRecord(ctx->ev_chunk[2 * k], ctx->st_h2d));
        CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->st_work, ctx->ev_chunk[2 * k], 0));
        tfpd_dense d = {nullptr, nullptr, nullptr};
        if (want_dense) {
            d.heat_up = d_heat + (size_t)f0 * fpx * kHeatC;
            d.paf_up = d_paf + (size_t)f0 * fpx * kPafC;
            d.peaks = d_peaks + (size_t)f0 * fpx * kHeatC;
        }
        int rc = enqueue_decode(ctx, ws_view(ctx->ws, f0), (const float *)dst, n, h, w, H, W, ctx->st_work,
                                want_dense ? &d : nullptr, timed);
        if (rc != TFPD_OK) return rc;
        if (want_dense) {
            CUDA_TRY(ctx, cudaEventRecord(ctx->ev_chunk[2 * k + 1], ctx->st_work));
            CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->st_d2h, ctx->ev_chunk[2 * k + 1], 0));
            const size_t o19 = (size_t)f0 * fpx * kHeatC, o38 = (size_t)f0 * fpx * kPafC;
            CUDA_TRY(ctx, cudaMemcpyAsync(heat_up + o19, d.heat_up, (size_t)n * fpx * kHeatC * 4, cudaMemcpyDeviceToHost, ctx->st_d2h));
            CUDA_TRY(ctx, cudaMemcpyAsync(paf_up + o38, d.paf_up, (size_t)n * fpx * kPafC * 4, cudaMemcpyDeviceToHost, ctx->st_d2h));
            CUDA_TRY(ctx, cudaMemcpyAsync(peaks + o19, d.peaks, (size_t)n * fpx * kHeatC * 4, cudaMemcpyDeviceToHost, ctx->st_d2h));
        }
    }
    if (int rc = enqueue_result_copy(ctx, N, ctx->st_work)) return rc;
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_done, ctx->st_work));
    ctx->decoded = true;
    ctx->d2h_pending = want_dense;
    return TFPD_OK;
}

int tfpd_decode_host(tfpd_ctx *ctx, const float *net_out, int N, int h, int w, int H, int W, float *heat_up,
                     float *paf_up, float *peaks) {
    int rc = tfpd_decode_host_async(ctx, net_out, N, h, w, H, W, heat_up, paf_up, peaks);
    if (rc != TFPD_OK) return rc;
    return tfpd_fetch_humans(ctx, nullptr, nullptr);    // synchronises; reports capacity overflow
}

int tfpd_process_paf_device(tfpd_ctx *ctx, int p1, int p2, int p3, const float *peaks, int h1, int h2, int h3,
                            const float *heatmap, int f1, int f2, int f3, const float *pafmap, void *stream) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!peaks || !heatmap || !pafmap) return fail(ctx, TFPD_ERR_INVALID, "process_paf: null array");
    if (p3 != kHeatC || h3 != kHeatC || f3 != kPafC)
        return fail(ctx, TFPD_ERR_INVALID, "process_paf: channel counts must be 19/19/38, got %d/%d/%d", p3, h3, f3);
    if (p1 < 1 || p2 < 1 || h1 != p1 || h2 != p2 || f1 != p1 || f2 != p2)
        return fail(ctx, TFPD_ERR_INVALID, "process_paf: the three arrays must share their spatial shape");
    if (p1 > 8191 || p2 > 16383) return fail(ctx, TFPD_ERR_INVALID, "process_paf: map larger than 8191x16383");
    ON_CTX_DEVICE(ctx);
    cudaStream_t s = (cudaStream_t)stream;
    if (int rc = order_after_previous(ctx, s)) return rc;
    begin_call(ctx, 1, p1, p2, s);
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->ws.counters, 0, sizeof(FrameCounters), s));
    if (ctx->profiling) cudaEventRecord(ctx->ev[0], s);
    int r = launch_scan_dense_peaks(peaks, heatmap, 1, p1, p2, ctx->cfg.thresh_heat, ctx->ws, s);
    if (r < 0) return fail(ctx, TFPD_ERR_CUDA, "k_scan_dense_peaks launch failed");
    ctx->launches += r;
    PafSource src;
    src.base = pafmap;
    src.rows = p1;
    src.cols = p2;
    src.chan = kPafC;
    src.chan_off = 0;
    src.shift = 0;
    src.H = h1;
    int rc = run_tail(ctx, ctx->ws, 1, src, (size_t)p1 * p2 * kPafC, s, ctx->profiling);
    if (rc != TFPD_OK) return rc;
    if ((rc = enqueue_result_copy(ctx, 1, s)) != TFPD_OK) return rc;
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_done, s));
    ctx->decoded = true;
    return 0;
}

int tfpd_process_paf(tfpd_ctx *ctx, int p1, int p2, int p3, const float *peaks, int h1, int h2, int h3,
                     const float *heatmap, int f1, int f2, int f3, const float *pafmap) {
    if (!ctx) return TFPD_ERR_INVALID;
    if (!peaks || !heatmap || !pafmap) return fail(ctx, TFPD_ERR_INVALID, "process_paf: null array");
    if (p1 < 1 || p2 < 1 || p3 < 1 || h1 < 1 || h2 < 1 || h3 < 1 || f1 < 1 || f2 < 1 || f3 < 1)
        return fail(ctx, TFPD_ERR_INVALID, "process_paf: empty array");
    ON_CTX_DEVICE(ctx);
    const size_t np = (size_t)p1 * p2 * p3, nh = (size_t)h1 * h2 * h3, nf = (size_t)f1 * f2 * f3;
    CUDA_TRY(ctx, ctx->dense_buf.reserve((np + nh + nf) * 4));
    float *dp = (float *)ctx->dense_buf.p, *dh = dp + np, *df = dh + nh;
    CUDA_TRY(ctx, cudaMemcpyAsync(dp, peaks, np * 4, cudaMemcpyHostToDevice, nullptr));
    CUDA_TRY(ctx, cudaMemcpyAsync(dh, heatmap, nh * 4, cudaMemcpyHostToDevice, nullptr));
    CUDA_TRY(ctx, cudaMemcpyAsync(df, pafmap, nf * 4, cudaMemcpyHostToDevice, nullptr));
    int rc = tfpd_process_paf_device(ctx, p1, p2, p3, dp, h1, h2, h3, dh, f1, f2, f3, df, nullptr);
    if (rc != TFPD_OK) return rc;
    rc = tfpd_fetch_humans(ctx, nullptr, nullptr);
    return rc == TFPD_OK ? 0 : rc;
}

static int ensure_getters(tfpd_ctx *ctx) {
    if (ctx->getters_ready) return TFPD_OK;
    if (!ctx->decoded) return fail(ctx, TFPD_ERR_STATE, "getter: no process_paf/decode on this ctx yet");
    int rc = tfpd_fetch_humans(ctx, nullptr, nullptr);
    if (rc != TFPD_OK && rc != TFPD_ERR_CAPACITY) return rc;
    const tfpd_frame_info *hi = (const tfpd_frame_info *)ctx->h_result;
    const size_t info_bytes = align_up((size_t)ctx->cfg.max_batch * sizeof(tfpd_frame_info), 16);
    const tfpd_human *hh = (const tfpd_human *)(ctx->h_result + info_bytes);
    ctx->g_humans.assign(hh, hh + hi[0].n_humans);
    const int np = hi[0].n_peaks;
    ctx->g_px.resize(np);
    ctx->g_py.resize(np);
    ctx->g_ps.resize(np);
    int cnt = 0;
    if (np > 0) {
        rc = tfpd_fetch_peaks(ctx, 0, np, nullptr, ctx->g_px.data(), ctx->g_py.data(), ctx->g_ps.data(), &cnt);
        if (rc != TFPD_OK) return rc;
    }
    ctx->getters_ready = true;
    return TFPD_OK;
}

int tfpd_get_num_humans(tfpd_ctx *ctx) {
    if (!ctx || ensure_getters(ctx) != TFPD_OK) return -1;
    return (int)ctx->g_humans.size();
}

int tfpd_get_part_cid(tfpd_ctx *ctx, int human_id, int part_id) {
    if (!ctx || ensure_getters(ctx) != TFPD_OK) return -1;
    if (human_id < 0 || human_id >= (int)ctx->g_humans.size() || part_id < 0 || part_id >= kParts) {
        fail(ctx, TFPD_ERR_INVALID, "get_part_cid(%d,%d) out of range", human_id, part_id);
        return -1;
    }
    return ctx->g_humans[human_id].cid[part_id];
}

float tfpd_get_score(tfpd_ctx *ctx, int human_id) {
    if (!ctx || ensure_getters(ctx) != TFPD_OK) return 0.0f;
    if (human_id < 0 || human_id >= (int)ctx->g_humans.size()) {
        fail(ctx, TFPD_ERR_INVALID, "get_score(%d) out of range", human_id);
        return 0.0f;
    }
    return ctx->g_humans[human_id].score;
}

int tfpd_get_part_x(tfpd_ctx *ctx, int cid) {
    if (!ctx || ensure_getters(ctx) != TFPD_OK) return -1;
    if (cid < 0 || cid >= (int)ctx->g_px.size()) {
        fail(ctx, TFPD_ERR_INVALID, "get_part_x(%d) out of range", cid);
        return -1;
    }
    return ctx->g_px[cid];
}

int tfpd_get_part_y(tfpd_ctx *ctx, int cid) {
    if (!ctx || ensure_getters(ctx) != TFPD_OK) return -1;
    if (cid < 0 || cid >= (int)ctx->g_py.size()) {
        fail(ctx, TFPD_ERR_INVALID, "get_part_y(%d) out of range", cid);
        return -1;
    }
    return ctx->g_py[cid];
}

float tfpd_get_part_score(tfpd_ctx *ctx, int cid) {
    if (!ctx || ensure_getters(ctx) != TFPD_OK) return 0.0f;
    if (cid < 0 || cid >= (int)ctx->g_ps.size()) {
        fail(ctx, TFPD_ERR_INVALID, "get_part_score(%d) out of range", cid);
        return 0.0f;
    }
    return ctx->g_ps[cid];
}

}  // extern "C"
