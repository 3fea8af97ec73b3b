// Shared helpers for the hicembler_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "hicembler_b200.h"

namespace hc {

void set_error(const char *fmt, ...);
int sm_count();
double wall_seconds();
// merge records -> scipy linkage array (dendrogram.cu)
void records_to_Z(int n, bool ward, const int *ra, const int *rb, double *rd, double *h_Z, double rel_tol,
                  long long *near_ties);

#define HC_CUDA_OK(expr)                                                                  \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            hc::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return HC_ECUDA;                                                              \
        }                                                                                 \
    } while (0)

#define HC_CHECK_LAUNCH() HC_CUDA_OK(cudaGetLastError())

#define HC_REQUIRE(cond, msg)                                      \
    do {                                                           \
        if (!(cond)) {                                             \
            hc::set_error("%s: %s", __func__, msg);                \
            return HC_EINVAL;                                      \
        }                                                          \
    } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// carve typed arrays out of a caller-provided workspace
struct Carver {
    char *base;
    size_t off, cap;
    Carver(void *p, size_t bytes) : base((char *)p), off(0), cap(bytes) {}
    template <typename T>
    T *take(size_t count) {
        off = align_up(off, 256);
        T *p = (T *)(base + off);
        off += count * sizeof(T);
        return p;
    }
    bool ok() const { return off <= cap; }
};

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// streaming 16-byte load that does not pollute L1
__device__ __forceinline__ uint4 ld_stream_u4(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

}  // namespace hc
