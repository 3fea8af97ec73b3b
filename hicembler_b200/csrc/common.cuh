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

// a / b for a loop-invariant divisor: the correctly rounded reciprocal is taken once, each element then
// costs FMUL + 2 FFMA (q = a*y; e = a - b*q exactly; q + e*y rounds to RN(a/b) when y = RN(1/b) and no
// intermediate leaves the normal range).  A group of operands is tested together: if any of them (or b) is
// outside [2^-60, 2^60) -- zero, negative, subnormal, inf/NaN included -- the whole group takes the plain
// IEEE division, so the result is a / b bit for bit in every case (checked against the IEEE division over
// whole binades by hc_selftest_scaled_division).
struct ScaledDivF {
    float b, y;
    bool ok;
    __device__ __forceinline__ explicit ScaledDivF(float b_) : b(b_), y(1.0f / b_) {
        ok = (__float_as_uint(b_) - 0x21800000u) < 0x3C000000u;
    }
    template <int N>
    __device__ __forceinline__ void apply(const float (&a)[N], float (&p)[N]) const {
        unsigned m = 0;
#pragma unroll
        for (int k = 0; k < N; ++k) m = max(m, __float_as_uint(a[k]) - 0x21800000u);
        if (ok && m < 0x3C000000u) {
#pragma unroll
            for (int k = 0; k < N; ++k) {
                const float q = __fmul_rn(a[k], y);
                const float e = __fmaf_rn(-b, q, a[k]);
                p[k] = __fmaf_rn(e, y, q);
            }
        } else {
#pragma unroll
            for (int k = 0; k < N; ++k) p[k] = a[k] / b;
        }
    }
};
struct ScaledDivD {     // float64 matrices are not on the hot path
    double b;
    __device__ __forceinline__ explicit ScaledDivD(double b_) : b(b_) {}
    template <int N>
    __device__ __forceinline__ void apply(const double (&a)[N], double (&p)[N]) const {
#pragma unroll
        for (int k = 0; k < N; ++k) p[k] = a[k] / b;
    }
};
template <typename T> struct ScaledDiv;
template <> struct ScaledDiv<float> { typedef ScaledDivF type; };
template <> struct ScaledDiv<double> { typedef ScaledDivD type; };

// streaming 16-byte load that does not pollute L1
__device__ __forceinline__ uint4 ld_stream_u4(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

}  // namespace hc
