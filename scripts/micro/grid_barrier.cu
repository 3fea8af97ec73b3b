// Micro-benchmark: latency of the hand-written grid barrier of kr_chain (one CTA of 512 threads per SM, cooperative
// launch), and of a dependent chain of L2 / DRAM loads between barriers -- the building blocks of a thin linkage round.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/micro/grid_barrier scripts/micro/grid_barrier.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
template <int MODE>
__device__ __forceinline__ void grid_barrier(unsigned int *bar, unsigned int &target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += gridDim.x;
        __threadfence();
        atomicAdd(bar, 1u);
        if (MODE == 0) while (*(volatile unsigned int *)bar < target) __nanosleep(64);
        if (MODE == 1) while (*(volatile unsigned int *)bar < target) {}
        if (MODE == 2) while (*(volatile unsigned int *)bar < target) __nanosleep(20);
        if (MODE == 3) { unsigned v; do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar)); } while (v < target); }
        __threadfence();
    }
    __syncthreads();
}
template <int MODE>
__global__ void __launch_bounds__(512, 1) k_bar(unsigned int *bar, int iters) {
    unsigned int target = *(volatile unsigned int *)bar;
    target -= target % gridDim.x;
    for (int i = 0; i < iters; ++i) grid_barrier<MODE>(bar, target);
}
// dependent pointer chase: every thread of the first `nthreads` global threads follows `depth` dependent loads through a
// random permutation table (table bytes decide L2 vs DRAM), then a barrier; repeated
__global__ void __launch_bounds__(512, 1) k_chase(unsigned int *bar, const unsigned int *tab, unsigned int mask, int depth, int nthreads,
                                                  int iters, unsigned int *sink) {
    unsigned int target = *(volatile unsigned int *)bar;
    target -= target % gridDim.x;
    const unsigned g = (threadIdx.x >> 5) * gridDim.x * 32 + blockIdx.x * 32 + (threadIdx.x & 31);
    unsigned int x = g * 2654435761u;
    for (int i = 0; i < iters; ++i) {
        if ((int)g < nthreads)
            for (int d = 0; d < depth; ++d) { unsigned int v; asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(tab + ((x + i * 40503u) & mask) * 8u)); x = v; }
        grid_barrier<0>(bar, target);
    }
    if (x == 0xdeadbeefu) sink[0] = x;
}
template <int MODE>
float run_bar(unsigned int *bar, int sms, int iters) {
    void *args[] = {&bar, &iters};
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int it1 = 10; void *a1[] = {&bar, &it1};
    cudaLaunchCooperativeKernel((const void *)k_bar<MODE>, dim3(sms), dim3(512), a1, 0, 0);
    cudaEventRecord(e0);
    cudaLaunchCooperativeKernel((const void *)k_bar<MODE>, dim3(sms), dim3(512), args, 0, 0);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms * 1e3f / iters;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    unsigned int *bar; cudaMalloc(&bar, 256); cudaMemset(bar, 0, 256);
    printf("grid barrier, %d CTAs x 512 threads: nanosleep(64) %.2f us, tight spin %.2f us, nanosleep(20) %.2f us, ld.acquire spin %.2f us\n", sms,
           run_bar<0>(bar, sms, 2000), run_bar<1>(bar, sms, 2000), run_bar<2>(bar, sms, 2000), run_bar<3>(bar, sms, 2000));
    // chase tables: 1 MB (L2 resident) and 2 GB (DRAM), entries 32 bytes apart
    for (int big = 0; big < 2; ++big) {
        const size_t entries = big ? (size_t)1 << 26 : (size_t)1 << 15;   // x 32 B = 2 GB / 1 MB
        unsigned int *tab, *sink; cudaMalloc(&tab, entries * 32); cudaMalloc(&sink, 4);
        unsigned int *h = (unsigned int *)malloc(entries * 32);
        unsigned int s = 12345;
        for (size_t k = 0; k < entries; ++k) { s = s * 1664525u + 1013904223u; h[k * 8] = s; }
        cudaMemcpy(tab, h, entries * 32, cudaMemcpyHostToDevice);
        for (int nthreads : {32, 14000, 75776})
            for (int depth : {1, 2, 4}) {
                int iters = 300; unsigned int mask = (unsigned int)(entries - 1);
                void *args[] = {&bar, &tab, &mask, &depth, &nthreads, &iters, &sink};
                cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
                cudaLaunchCooperativeKernel((const void *)k_chase, dim3(sms), dim3(512), args, 0, 0);
                cudaEventRecord(e0);
                cudaLaunchCooperativeKernel((const void *)k_chase, dim3(sms), dim3(512), args, 0, 0);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                printf("%s table, %6d threads, depth %d: %.2f us per (chase + barrier)\n", big ? "2 GB" : "1 MB", nthreads, depth, ms * 1e3f / iters);
            }
        cudaFree(tab); free(h);
    }
    return 0;
}
