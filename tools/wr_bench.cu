// wr_bench.cu -- developer microbenchmark (not product, not a test): what a WRITE-dominated stream
// can reach on this GPU, to calibrate K1's dense-mode roofline (K1 writes 773 MB and reads 36 MB per
// launch; MEASURED_PEAKS.json's hbm_gbs is a read+write copy).
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/wr_bench tools/wr_bench.cu
//   run  : tools/wr_bench [MB]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// A. plain 128-bit stores, grid-stride
__global__ void k_stg(float4 *dst, size_t n4) {
    float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) dst[i] = v;
}
// A2. each CTA owns a contiguous span
__global__ void k_stg_span(float4 *dst, size_t n4) {
    float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    size_t per = (n4 + gridDim.x - 1) / gridDim.x;
    size_t lo = per * blockIdx.x, hi = lo + per < n4 ? lo + per : n4;
    for (size_t i = lo + threadIdx.x; i < hi; i += blockDim.x) dst[i] = v;
}

// B. TMA bulk stores out of shared memory.  chunk bytes per store, `depth` stores in flight per CTA
// (one thread issues), interleaved (chunk c -> CTA c % grid) or span (CTA owns a contiguous range).
template <int SPAN>
__global__ void k_bulk(unsigned char *dst, size_t bytes, int chunk, int depth) {
    extern __shared__ __align__(128) unsigned char sm[];
    for (int i = threadIdx.x; i < chunk / 4; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x != 0) return;
    const size_t nchunks = bytes / chunk;
    size_t c, step, end;
    if (SPAN) {
        size_t per = (nchunks + gridDim.x - 1) / gridDim.x;
        c = per * blockIdx.x;
        end = c + per < nchunks ? c + per : nchunks;
        step = 1;
    } else {
        c = blockIdx.x;
        end = nchunks;
        step = gridDim.x;
    }
    int inflight = 0;
    for (; c < end; c += step) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * (size_t)chunk),
                     "r"(smem_u32(sm)), "r"(chunk)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (++inflight >= depth) {
            // allow depth-1 groups to stay pending
            switch (depth) {
                case 1: asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); break;
                case 2: asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); break;
                case 4: asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); break;
                default: asm volatile("cp.async.bulk.wait_group.read 7;" ::: "memory"); break;
            }
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// C. copy with a read fraction: reads `rd4` float4 per written float4 block of 16 (emulates 5 % reads)
__global__ void k_mix(float4 *dst, const float4 *src, size_t n4, int every) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
        if ((i / 32) % every == 0) v = __ldg(src + i / every);
        dst[i] = v;
    }
}

template <typename F>
float time_ms(F f, int reps = 8) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) f();
    CK(cudaDeviceSynchronize());
    float best = 1e9f;
    for (int i = 0; i < reps; ++i) {
        cudaEventRecord(a);
        f();
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main(int argc, char **argv) {
    size_t mb = argc > 1 ? atol(argv[1]) : 773;
    size_t bytes = mb * 1000 * 1000 / 65536 * 65536;
    unsigned char *d, *s;
    CK(cudaMalloc(&d, bytes));
    CK(cudaMalloc(&s, bytes));
    CK(cudaMemset(s, 1, bytes));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("buffer %.1f MB, %d SMs\n", bytes / 1e6, sms);
    float ms = time_ms([&] { cudaMemsetAsync(d, 0, bytes); });
    printf("cudaMemsetAsync            %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    for (int mult : {2, 4, 8, 16}) {
        ms = time_ms([&] { k_stg<<<sms * mult, 512>>>((float4 *)d, bytes / 16); });
        printf("STG.128 grid-stride x%-2d     %.3f ms  %.0f GB/s\n", mult, ms, bytes / ms / 1e6);
    }
    ms = time_ms([&] { k_stg_span<<<sms * 4, 512>>>((float4 *)d, bytes / 16); });
    printf("STG.128 span x4            %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    CK(cudaFuncSetAttribute(k_bulk<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CK(cudaFuncSetAttribute(k_bulk<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    for (int chunk : {4096, 8192, 16384, 32768, 65536})
        for (int depth : {2, 4, 8})
            for (int ctas : {1, 2}) {
                ms = time_ms([&] { k_bulk<0><<<sms * ctas, 64, chunk>>>(d, bytes, chunk, depth); });
                float ms2 = time_ms([&] { k_bulk<1><<<sms * ctas, 64, chunk>>>(d, bytes, chunk, depth); });
                printf("TMA bulk chunk %5d depth %d ctas/SM %d  interleaved %.3f ms %.0f GB/s | span %.3f ms %.0f GB/s\n",
                       chunk, depth, ctas, ms, bytes / ms / 1e6, ms2, bytes / ms2 / 1e6);
            }
    for (int every : {20, 4, 2}) {
        ms = time_ms([&] { k_mix<<<sms * 8, 512>>>((float4 *)d, (const float4 *)s, bytes / 16, every); });
        printf("STG + 1/%d reads            %.3f ms  %.0f GB/s (writes only: %.0f)\n", every, ms,
               (bytes + bytes / every) / ms / 1e6, bytes / ms / 1e6);
    }
    ms = time_ms([&] { cudaMemcpyAsync(d, s, bytes, cudaMemcpyDeviceToDevice); });
    printf("cudaMemcpy D2D             %.3f ms  %.0f GB/s (r+w)\n", ms, 2.0 * bytes / ms / 1e6);
    return 0;
}
