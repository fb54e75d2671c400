// k_noalloc.cu -- tfpd_peak_table_noalloc: the peak table of pafprocess.cpp:22-44 (what tfpd_process_paf_device
// builds first and tfpd_fetch_peaks returns) for callers that may not allocate: no context, no cudaMalloc / cudaFree,
// no allocating library call.  No counterpart in the reference.
//
// The scratch the two kernels need -- the append counter and the unordered (key, score) records -- lives in ONE
// file-scope __device__ array of TFPD_NOALLOC_SCRATCH_BYTES.  Every GPU that loads the library holds its own copy until
// the process exits; every call on a device shares it, and it keeps the last call's contents: the counter is cleared
// at the start of each call (stream-ordered), the records are only read up to the counter.  Calls that overlap on one
// GPU (two streams, two threads) race on it.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the address is resolved once per
// device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels receive.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

constexpr int kNaMaxPeaks = TFPD_NOALLOC_MAX_PEAKS;
constexpr int kNaHeader = 256;                           // counter block, keeps the record arrays 256-byte aligned
constexpr int kNaThreads = 1024;
constexpr int kNaMaxDevices = 64;
static_assert(kNaHeader + 8 * kNaMaxPeaks == TFPD_NOALLOC_SCRATCH_BYTES, "header constant out of step");
static_assert((kNaMaxPeaks & (kNaMaxPeaks - 1)) == 0 && kNaMaxPeaks * 8 <= 48 * 1024, "sort uses static shared memory");

static __device__ __align__(256) unsigned char g_na_scratch[TFPD_NOALLOC_SCRATCH_BYTES];

struct NaScratch {
    int *counter;        // [0] peaks found (may exceed cap)
    uint32_t *rec_key;   // [kNaMaxPeaks]
    float *rec_val;      // [kNaMaxPeaks]
};

__host__ __device__ inline NaScratch carve(unsigned char *base) {
    NaScratch s;
    s.counter = reinterpret_cast<int *>(base);
    s.rec_key = reinterpret_cast<uint32_t *>(base + kNaHeader);
    s.rec_val = reinterpret_cast<float *>(base + kNaHeader + 4 * kNaMaxPeaks);
    return s;
}

// the reference's scan `PEAKS(y,x,part) > THRESH_HEAT` over part < 18 (pafprocess.cpp:24-37), one coalesced pass;
// records are appended in arbitrary order and ordered by k_na_sort
__global__ void k_na_scan(const float *__restrict__ peaks, const float *__restrict__ heat, int H, int W, float thresh,
                          int cap, unsigned char *scratch) {
    const NaScratch s = carve(scratch);
    const size_t total = (size_t)H * W * kHeatC;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const float v = __ldg(peaks + i);
        if (v > thresh) {
            const int c = (int)(i % kHeatC);
            if (c < kParts) {
                const size_t pix = i / kHeatC;
                const int y = (int)(pix / W), x = (int)(pix - (size_t)y * W);
                const int idx = atomicAdd(s.counter, 1);
                if (idx < cap) {
                    s.rec_key[idx] = peak_key(c, y, x);
                    s.rec_val[idx] = __ldg(heat + i);      // score = HEAT(y,x,part), pafprocess.cpp:32
                }
            }
        }
    }
}

// one block: (part, y, x) order == the order in which the reference assigns Peak.id (pafprocess.cpp:30)
__global__ void __launch_bounds__(kNaThreads) k_na_sort(int cap, unsigned char *scratch, int32_t *o_part, int32_t *o_x,
                                                        int32_t *o_y, float *o_score, int32_t *o_count) {
    __shared__ uint32_t keys[kNaMaxPeaks];
    __shared__ float vals[kNaMaxPeaks];
    const NaScratch s = carve(scratch);
    const int tid = threadIdx.x;
    const int raw = *s.counter;
    const int cnt = min(raw, cap);
    int np2 = 32;
    while (np2 < cnt) np2 <<= 1;
    for (int i = tid; i < np2; i += kNaThreads) {
        keys[i] = i < cnt ? s.rec_key[i] : 0xffffffffu;
        vals[i] = i < cnt ? s.rec_val[i] : 0.0f;
    }
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += kNaThreads) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const bool up = ((i & k) == 0);
                    const uint32_t a = keys[i], b = keys[ixj];
                    if ((a > b) == up) {
                        keys[i] = b;
                        keys[ixj] = a;
                        const float t = vals[i];
                        vals[i] = vals[ixj];
                        vals[ixj] = t;
                    }
                }
            }
            __syncthreads();
        }
    }
    for (int i = tid; i < cnt; i += kNaThreads) {
        o_part[i] = key_part(keys[i]);
        o_x[i] = key_x(keys[i]);
        o_y[i] = key_y(keys[i]);
        o_score[i] = vals[i];
    }
    if (tid == 0) {
        o_count[0] = cnt;
        o_count[1] = raw;
    }
}

std::mutex g_na_mu;
unsigned char *g_na_ptr[kNaMaxDevices] = {};

int na_fail(char *err, size_t err_len, int code, const char *fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace
}  // namespace tfpd

using namespace tfpd;

extern "C" TFPD_API int tfpd_peak_table_noalloc(const float *peaks, const float *heatmap, int H, int W, float thresh_heat,
                                                int cap, int32_t *part, int32_t *x, int32_t *y, float *score,
                                                int32_t *count, void *stream, char *err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (!peaks || !heatmap || !part || !x || !y || !score || !count)
        return na_fail(err, err_len, TFPD_ERR_INVALID, "peak_table_noalloc: null array");
    if (H < 1 || W < 1 || H > 8191 || W > 16383)
        return na_fail(err, err_len, TFPD_ERR_INVALID, "peak_table_noalloc: map %dx%d outside 1x1 .. 8191x16383", H, W);
    if (cap < 1 || cap > kNaMaxPeaks)
        return na_fail(err, err_len, TFPD_ERR_CAPACITY,
                       "peak_table_noalloc: cap=%d outside 1 .. TFPD_NOALLOC_MAX_PEAKS (%d): the static scratch of %d "
                       "bytes holds no more; use tfpd_process_paf_device for larger tables",
                       cap, kNaMaxPeaks, TFPD_NOALLOC_SCRATCH_BYTES);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return na_fail(err, err_len, TFPD_ERR_CUDA, "peak_table_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kNaMaxDevices)
        return na_fail(err, err_len, TFPD_ERR_INVALID, "peak_table_noalloc: device %d outside 0 .. %d", dev, kNaMaxDevices - 1);
    unsigned char *scratch = nullptr;
    {
        // once per device, with that device current: the device address of this GPU's copy of the array
        std::lock_guard<std::mutex> lock(g_na_mu);
        if (!g_na_ptr[dev]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_na_scratch);
            if (e != cudaSuccess || !p)
                return na_fail(err, err_len, TFPD_ERR_CUDA, "peak_table_noalloc: cudaGetSymbolAddress on device %d: %s", dev,
                               cudaGetErrorString(e));
            g_na_ptr[dev] = static_cast<unsigned char *>(p);
        }
        scratch = g_na_ptr[dev];
    }
    cudaStream_t s = (cudaStream_t)stream;
    // the counter must start at zero; the records of an earlier call are never read beyond it
    e = cudaMemsetAsync(scratch, 0, kNaHeader, s);
    if (e != cudaSuccess) return na_fail(err, err_len, TFPD_ERR_CUDA, "peak_table_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
    const size_t total = (size_t)H * W * kHeatC;
    const size_t want = (total + 255) / 256;
    const int blocks = (int)(want < (size_t)(148 * 8) ? want : (size_t)(148 * 8));
    k_na_scan<<<blocks, 256, 0, s>>>(peaks, heatmap, H, W, thresh_heat, cap, scratch);
    k_na_sort<<<1, kNaThreads, 0, s>>>(cap, scratch, part, x, y, score, count);
    e = cudaGetLastError();
    if (e != cudaSuccess) return na_fail(err, err_len, TFPD_ERR_CUDA, "peak_table_noalloc: launch: %s", cudaGetErrorString(e));
    return TFPD_OK;
}
