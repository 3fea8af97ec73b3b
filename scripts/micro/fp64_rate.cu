// Micro-benchmark: FP64 instruction rates of one SM (DFMA, DMUL, DADD, __ddiv_rn, f32->f64 conversion, FP32 FMA for scale).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/micro/fp64_rate scripts/micro/fp64_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(double *out, float *outf, int iters, double seed) {
    double a[8];
    float f[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { a[u] = seed + threadIdx.x * 1e-9 + u; f[u] = (float)a[u]; }
    const double m = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (OP == 0) a[u] = __fma_rn(a[u], m, c);
            if (OP == 1) a[u] = __dmul_rn(a[u], m);
            if (OP == 2) a[u] = __dadd_rn(a[u], c);
            if (OP == 3) a[u] = __ddiv_rn(a[u], m);
            if (OP == 4) { f[u] = __fmaf_rn(f[u], 1.0000001f, 1e-9f); }
            if (OP == 5) { a[u] = __dadd_rn(a[u], (double)f[u]); f[u] = __fmaf_rn(f[u], 1.0000001f, 1e-9f); }  // F2F.F64.F32 + DADD + FFMA
            if (OP == 6) a[u] = __drcp_rn(a[u]) ;
        }
    }
    double s = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) s += a[u] + f[u];
    if (s == 12345.678) { out[0] = s; outf[0] = (float)s; }
}
template <int OP>
void run(const char *name, int sms, double clk_ghz) {
    double *o; float *of;
    cudaMalloc(&o, 8); cudaMalloc(&of, 4);
    const int iters = 2048, blocks = sms * 2, threads = 1024;
    k<OP><<<blocks, threads>>>(o, of, 16, 1.0);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(o, of, iters, 1.0);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)blocks * threads * iters * 8;
    printf("%-28s %8.3f ms  %8.2f Gop/s  %6.2f op/clk/SM (at %.3f GHz)\n", name, ms, ops / ms * 1e-6, ops / (ms * 1e-3) / sms / (clk_ghz * 1e9), clk_ghz);
}
int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double ghz = clk * 1e-6;
    printf("%s, %d SMs, %.3f GHz max\n", p.name, p.multiProcessorCount, ghz);
    run<0>("DFMA", p.multiProcessorCount, ghz);
    run<1>("DMUL", p.multiProcessorCount, ghz);
    run<2>("DADD", p.multiProcessorCount, ghz);
    run<3>("__ddiv_rn", p.multiProcessorCount, ghz);
    run<6>("__drcp_rn", p.multiProcessorCount, ghz);
    run<4>("FFMA", p.multiProcessorCount, ghz);
    run<5>("F2F.F64.F32+DADD+FFMA", p.multiProcessorCount, ghz);
    return 0;
}
