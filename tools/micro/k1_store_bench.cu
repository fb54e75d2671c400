// k1_store_bench.cu -- developer microbenchmark (not product, not a test): the STORE side of K1's copy role in
// isolation.  Work = K1's on the bench workload: 64 frames, 46 low-res rows, 2 chunks of 27 cells -> 5888 items from a
// grid-wide atomic queue; an item writes one staged upsampled row (heat 108 px x 76 B = 8208 B, paf 108 x 152 B =
// 16416 B) to 4 output rows of heat_up and 4 of paf_up (98.5 KB per item, 580 MB per launch).  The staged slot is
// constant: only the way the bytes leave the SM differs.
//   mode 0  one TMA bulk store per output row (what K1 does)
//   mode 1  TMA bulk stores of <= PIECE bytes
//   mode 2  the warp's own 128-bit streaming stores (LDS.128 from the slot, 4 x STG.128 .cs)
//   mode 3  same with default-policy stores
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/k1_store_bench tools/micro/k1_store_bench.cu
//   run  : tools/micro/k1_store_bench
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int N = 64, h = 46, w = 54, H = 4 * h, W = 4 * w, CHUNKS = 2, CELLS = 27;
constexpr int HEAT_B = CELLS * 4 * 19 * 4, PAF_B = CELLS * 4 * 38 * 4;      // 8208, 16416
constexpr int SLOT_B = HEAT_B + PAF_B;                                       // 24624
constexpr int ITEMS = N * h * CHUNKS;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_store(void *dst, const void *src, int bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}

template <int MODE>
__global__ void k_store(unsigned char *heat, unsigned char *paf, int *queue, int piece) {
    extern __shared__ __align__(128) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *slot = sm + (size_t)warp * SLOT_B;
    for (int i = lane; i < SLOT_B / 4; i += 32) reinterpret_cast<float *>(slot)[i] = (float)i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    for (;;) {
        int it = 0;
        if (lane == 0) it = atomicAdd(queue, 1);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= ITEMS) break;
        const int ch = it % CHUNKS, iy = (it / CHUNKS) % h, n = it / (CHUNKS * h);
        const size_t pix = ((size_t)n * H + 4 * iy) * W + 4 * CELLS * ch;
        unsigned char *dh = heat + pix * 76, *dp = paf + pix * 152;
        if (MODE <= 1) {
            if (lane == 0) {
                for (int r = 0; r < 4; ++r) {
                    unsigned char *a = dh + (size_t)r * W * 76, *b = dp + (size_t)r * W * 152;
                    if (MODE == 0) {
                        bulk_store(a, slot, HEAT_B);
                        bulk_store(b, slot + HEAT_B, PAF_B);
                    } else {
                        for (int o = 0; o < HEAT_B; o += piece) bulk_store(a + o, slot + o, min(piece, HEAT_B - o));
                        for (int o = 0; o < PAF_B; o += piece) bulk_store(b + o, slot + HEAT_B + o, min(piece, PAF_B - o));
                    }
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");     // two items in flight per warp
            }
            __syncwarp();
        } else {
            for (int u = lane; u < HEAT_B / 16; u += 32) {
                const float4 v = *reinterpret_cast<const float4 *>(slot + 16 * u);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    float4 *q = reinterpret_cast<float4 *>(dh + (size_t)r * W * 76) + u;
                    if (MODE == 2) __stcs(q, v); else *q = v;
                }
            }
            for (int u = lane; u < PAF_B / 16; u += 32) {
                const float4 v = *reinterpret_cast<const float4 *>(slot + HEAT_B + 16 * u);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    float4 *q = reinterpret_cast<float4 *>(dp + (size_t)r * W * 152) + u;
                    if (MODE == 2) __stcs(q, v); else *q = v;
                }
            }
        }
    }
    if (MODE <= 1 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int MODE>
float run(unsigned char *heat, unsigned char *paf, int *queue, int ctas, int warps, int piece) {
    const size_t smem = (size_t)warps * SLOT_B;
    CK(cudaFuncSetAttribute(k_store<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float sum = 0.f;
    const int reps = 10;
    for (int i = 0; i < reps + 3; ++i) {
        CK(cudaMemsetAsync(queue, 0, 4));
        cudaEventRecord(a);
        k_store<MODE><<<ctas, warps * 32, smem>>>(heat, paf, queue, piece);
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (i >= 3) sum += ms;
    }
    CK(cudaGetLastError());
    return sum / reps;
}

int main() {
    const size_t hb = (size_t)N * H * W * 76, pb = (size_t)N * H * W * 152;
    unsigned char *heat, *paf, *other;
    int *queue;
    CK(cudaMalloc(&heat, hb));
    CK(cudaMalloc(&paf, pb));
    CK(cudaMalloc(&other, 256u << 20));
    CK(cudaMalloc(&queue, 4));
    const double mb = (hb + pb) / 1e6;
    printf("K1 copy-role store pattern: %d items x %.1f KB = %.1f MB per launch (mean of 10 launches)\n", ITEMS,
           (4.0 * SLOT_B) / 1e3, mb);
    for (int ctas : {128, 148})
        for (int warps : {1, 2, 3, 4, 6, 8}) {
            if ((size_t)warps * SLOT_B > 220 * 1024) continue;
            float t0 = run<0>(heat, paf, queue, ctas, warps, 0);
            float t1a = run<1>(heat, paf, queue, ctas, warps, 4096);
            float t1b = run<1>(heat, paf, queue, ctas, warps, 2048);
            float t2 = run<2>(heat, paf, queue, ctas, warps, 0);
            float t3 = run<3>(heat, paf, queue, ctas, warps, 0);
            printf("%3d CTAs x %d copy warps: TMA per row %.1f us %.0f GB/s | TMA <=4 KB %.1f us %.0f | TMA <=2 KB %.1f us %.0f | "
                   "STG.128 .cs %.1f us %.0f | STG.128 %.1f us %.0f\n",
                   ctas, warps, t0 * 1e3, mb / t0, t1a * 1e3, mb / t1a, t1b * 1e3, mb / t1b, t2 * 1e3, mb / t2,
                   t3 * 1e3, mb / t3);
        }
    return 0;
}
