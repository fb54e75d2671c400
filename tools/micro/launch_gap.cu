// Launch-to-launch time of dependent kernels in one stream: small kernels, kernels with a large dynamic shared
// memory size (like k1_fused_f4), and the two alternating (does the shared-memory carve-out switch cost time?).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o launch_gap launch_gap.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_small(int *p) { if (threadIdx.x == 0 && blockIdx.x == 0 && p) p[0] += 1; }
__global__ void k_big(int *p) {
    extern __shared__ int sm[];
    sm[threadIdx.x] = threadIdx.x;
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0 && p) p[0] += sm[1];
}

int main() {
    int *d;
    cudaMalloc(&d, 4);
    cudaMemset(d, 0, 4);
    const int big = 200 * 1024;
    cudaFuncSetAttribute(k_big, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int R = 2000;
    for (int mode = 0; mode < 5; ++mode) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            for (int i = 0; i < R; ++i) {
                switch (mode) {
                case 0: k_small<<<64, 128>>>(d); break;
                case 1: k_big<<<128, 672, big>>>(d); break;
                case 2: if (i & 1) k_big<<<128, 672, big>>>(d); else k_small<<<64, 128>>>(d); break;
                case 3: if (i & 1) k_big<<<128, 672, 48 * 1024>>>(d); else k_small<<<64, 128>>>(d); break;
                case 4: if (i & 1) cudaMemsetAsync(d, 0, 4); else k_small<<<64, 128>>>(d); break;
                }
            }
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep) {
                const char *names[] = {"small only", "big-smem only", "small / big-smem (200 KB) alternating",
                                       "small / 48 KB-smem alternating", "small / memset alternating"};
                printf("%-40s %.2f us per launch\n", names[mode], ms * 1e3 / R);
            }
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
