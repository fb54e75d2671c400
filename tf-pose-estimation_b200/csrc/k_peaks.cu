// k_peaks.cu -- the peak table of pafprocess.cpp:22-44.
//
//  k_scan_dense_peaks : seam 2 (process_paf called with the dense `peaks` tensor): the
//                       reference's part/y/x scan `PEAKS(y,x,part) > THRESH_HEAT`
//                       (pafprocess.cpp:24-37) as one coalesced pass, appending (key, score)
//                       records in arbitrary order.
//  k_sort_peaks       : one block per frame orders the records by (part, y, x) -- the order in
//                       which the reference assigns Peak.id (pafprocess.cpp:30) -- with a
//                       bitonic sort in shared memory, and writes the peak table + part offsets.
#include "tfpd_internal.cuh"

namespace tfpd {
namespace {

__global__ void k_scan_dense_peaks(const float *__restrict__ peaks, const float *__restrict__ heat, int H, int W,
                                   float thresh, Workspace ws) {
    const int n = blockIdx.y;
    const size_t total = (size_t)H * W * kHeatC;
    const float *pk = peaks + (size_t)n * total;
    const float *ht = heat + (size_t)n * total;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        float v = __ldg(pk + i);
        if (v > thresh) {
            int c = (int)(i % kHeatC);
            if (c < kParts) {
                size_t pix = i / kHeatC;
                int y = (int)(pix / W), x = (int)(pix - (size_t)y * W);
                int idx = atomicAdd(&ws.counters[n].n_peak_rec, 1);
                if (idx < ws.max_peaks) {
                    ws.rec_key[(size_t)n * ws.max_peaks + idx] = peak_key(c, y, x);
                    ws.rec_val[(size_t)n * ws.max_peaks + idx] = __ldg(ht + i);   // HEAT(y,x,part), pafprocess.cpp:32
                }
            }
        }
    }
}

constexpr int kSortThreads = 1024;

__global__ void __launch_bounds__(kSortThreads) k_sort_peaks(Workspace ws) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    uint32_t *keys = reinterpret_cast<uint32_t *>(sm_raw);
    float *vals = reinterpret_cast<float *>(keys + ws.max_peaks);
    __shared__ int s_first[kHeatC + 1];

    const int n = blockIdx.x, tid = threadIdx.x;
    const int raw = ws.counters[n].n_peak_rec;
    const int cnt = min(raw, ws.max_peaks);
    int np2 = 32;
    while (np2 < cnt) np2 <<= 1;
    for (int i = tid; i < np2; i += kSortThreads) {
        if (i < cnt) {
            keys[i] = ws.rec_key[(size_t)n * ws.max_peaks + i];
            vals[i] = ws.rec_val[(size_t)n * ws.max_peaks + i];
        } else {
            keys[i] = 0xffffffffu;
            vals[i] = 0.0f;
        }
    }
    if (tid <= kHeatC) s_first[tid] = cnt;
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += kSortThreads) {
                int ixj = i ^ j;
                if (ixj > i) {
                    bool up = ((i & k) == 0);
                    uint32_t a = keys[i], b = keys[ixj];
                    if ((a > b) == up) {
                        keys[i] = b;
                        keys[ixj] = a;
                        float t = vals[i];
                        vals[i] = vals[ixj];
                        vals[ixj] = t;
                    }
                }
            }
            __syncthreads();
        }
    }
    // part offsets: first[p] = index of the first peak with part >= p
    for (int i = tid; i < cnt; i += kSortThreads) {
        int part = key_part(keys[i]);
        int prev = (i == 0) ? -1 : key_part(keys[i - 1]);
        for (int q = prev + 1; q <= part; ++q) s_first[q] = i;
        ws.pk_x[(size_t)n * ws.max_peaks + i] = (int16_t)key_x(keys[i]);
        ws.pk_y[(size_t)n * ws.max_peaks + i] = (int16_t)key_y(keys[i]);
        ws.pk_score[(size_t)n * ws.max_peaks + i] = vals[i];
    }
    __syncthreads();
    if (tid <= kHeatC) ws.part_first[n * 20 + tid] = s_first[tid];
    if (tid == 0) {
        ws.info[n].n_peaks = cnt;
        ws.info[n].near_ties = ws.counters[n].near_ties;
        if (raw > ws.max_peaks) atomicOr(&ws.counters[n].flags, TFPD_FLAG_PEAK_OVERFLOW);
    }
}

}  // namespace

int launch_scan_dense_peaks(const float *peaks, const float *heat, int N, int H, int W, float thresh,
                            const Workspace &ws, cudaStream_t s) {
    size_t total = (size_t)H * W * kHeatC;
    size_t want = (total + 255) / 256;
    int bx = (int)(want < (size_t)(148 * 8) ? want : (size_t)(148 * 8));
    dim3 grid(bx, N);
    k_scan_dense_peaks<<<grid, 256, 0, s>>>(peaks, heat, H, W, thresh, ws);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

int launch_sort_peaks(const Workspace &ws, int N, cudaStream_t s) {
    size_t smem = (size_t)ws.max_peaks * 8;
    k_sort_peaks<<<N, kSortThreads, smem, s>>>(ws);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
