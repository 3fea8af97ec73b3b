// K3-K5: Sinkhorn-Knopp balancing as HBM-bound matvec / rescale passes.
//
// Replaces SinkhornKnopp.fit (reference bin/sinkhorn_knopp/sinkhorn_knopp/sinkhorn_knopp.py:93-163)
// as called by normalize() (bam2clusters.py:36-49).  The reference spends O(N^3) per iteration in
// D1.dot(P).dot(D2) with dense diagonal matrices (:146-148); the arithmetic that matters is the two
// matvecs c = 1/(P^T r), r = 1/(P c) (:143-144).  Here each is one streaming read of the float32
// matrix (4 N^2 bytes) with float64 accumulation (numpy upcasts P to float64 inside dot), the
// division by vmax of normalize() (:46) is folded into the load, and the stopping test of
// :138-141 costs nothing: the column sums of D1.P.D2 are c * (P^T r), i.e. the next column pass.
//
// Row-sharded multi-GPU use: colpass works on a row block and produces partial column sums; the
// caller all-reduces y (N float64) over NCCL and continues locally.
#include "common.cuh"

namespace {

constexpr int RC = 256;  // rows per column-pass chunk
constexpr int RB = 8;    // rows per row-pass block
constexpr int ST = 256;  // tile edge of the symmetric pass

template <typename T>
struct Vec4;
template <>
struct Vec4<float> {
    typedef float4 type;
};
template <>
struct Vec4<double> {
    typedef double4 type;
};

template <typename T>
__device__ __forceinline__ void load4(const T *p, T v[4], bool vec, int valid) {
    if (vec) {
        if (sizeof(T) == 4) {
            float4 t = __ldcs((const float4 *)p);
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
            double2 a = __ldcs((const double2 *)p), b = __ldcs((const double2 *)p + 1);
            v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = k < valid ? __ldcs(p + k) : (T)0;
    }
}

// ypart[chunk][j] = sum_{i in chunk} (X[i][j]/vmax) * r[i]
template <typename T, bool CHECK>
__global__ void __launch_bounds__(256) k_colpass(const T *X, int nrows, int n, T vmax, const double *r,
                                                  double *ypart, int32_t *flags) {
    const int j0 = (blockIdx.x * 256 + threadIdx.x) * 4;
    const int i0 = blockIdx.y * RC, i1 = min(i0 + RC, nrows);
    if (j0 >= n) return;
    const bool vec = (n % 4 == 0) && (((uintptr_t)X & 15) == 0);
    const int valid = min(4, n - j0);
    double acc[4] = {0, 0, 0, 0};
    bool neg = false;
    const bool scale = vmax != (T)1;
    const typename hc::ScaledDiv<T>::type div(vmax);
#pragma unroll 4
    for (int i = i0; i < i1; ++i) {
        T v[4];
        load4(X + (size_t)i * n + j0, v, vec, valid);
        double ri = r ? __ldg(r + i) : 1.0;
        T p[4];
        if (scale) div.apply(v, p);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (CHECK) neg |= !(v[k] >= (T)0);
            acc[k] = fma((double)(scale ? p[k] : v[k]), ri, acc[k]);
        }
    }
    double *dst = ypart + (size_t)blockIdx.y * n + j0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (k < valid) dst[k] = acc[k];
    if (CHECK && neg) atomicOr(flags, 1);
}

__global__ void k_colfinish(const double *ypart, int n_chunks, int n, double *y) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double s = 0;
    for (int b = 0; b < n_chunks; ++b) s += ypart[(size_t)b * n + j];
    y[j] = s;
}

// c = 1/y; flags[2] |= any(y==0); flags[3] |= any(c_prev*y outside 1 +- eps)
__global__ void k_update_c(const double *y, const double *c_prev, int n, double eps, double *c, int32_t *flags) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double v = y[j];
    if (v == 0.0) atomicOr(flags + 2, 1);
    if (c_prev) {
        double s = c_prev[j] * v;
        if (s < 1.0 - eps || s > 1.0 + eps) atomicOr(flags + 3, 1);
    }
    c[j] = 1.0 / v;
}

__global__ void k_outside(const double *v, int n, double eps, int32_t *flag) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double s = v[j];
    if (s < 1.0 - eps || s > 1.0 + eps) atomicOr(flag, 1);
}

// x[i] = sum_j (X[i][j]/vmax) * c[j];  r[i] = 1/x[i]
template <typename T>
__global__ void __launch_bounds__(256) k_rowpass(const T *X, int nrows, int n, T vmax, const double *c,
                                                  double *x, double *r, int32_t *flags) {
    const int i0 = blockIdx.x * RB;
    const bool vec = (n % 4 == 0) && (((uintptr_t)X & 15) == 0);
    const bool scale = vmax != (T)1;
    const typename hc::ScaledDiv<T>::type div(vmax);
    double acc[RB];
#pragma unroll
    for (int k = 0; k < RB; ++k) acc[k] = 0;
    for (int j0 = threadIdx.x * 4; j0 < n; j0 += 1024) {
        const int valid = min(4, n - j0);
        double cj[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) cj[k] = c ? (k < valid ? __ldg(c + j0 + k) : 0.0) : 1.0;
        T v[RB][4];
#pragma unroll
        for (int rr = 0; rr < RB; ++rr) {
            if (i0 + rr < nrows) load4(X + (size_t)(i0 + rr) * n + j0, v[rr], vec, valid);
            else v[rr][0] = v[rr][1] = v[rr][2] = v[rr][3] = (T)0;
        }
#pragma unroll
        for (int rr = 0; rr < RB; ++rr) {
            T p[4];
            if (scale) div.apply(v[rr], p);
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[rr] = fma((double)(scale ? p[k] : v[rr][k]), cj[k], acc[rr]);
        }
    }
    __shared__ double sh[RB][8];
#pragma unroll
    for (int rr = 0; rr < RB; ++rr) {
        double s = hc::warp_sum_f64(acc[rr]);
        if ((threadIdx.x & 31) == 0) sh[rr][threadIdx.x >> 5] = s;
    }
    __syncthreads();
    if (threadIdx.x < RB && i0 + threadIdx.x < nrows) {
        double s = 0;
        for (int w = 0; w < 8; ++w) s += sh[threadIdx.x][w];
        x[i0 + threadIdx.x] = s;
        if (s == 0.0) atomicOr(flags + 4, 1);
        if (r) r[i0 + threadIdx.x] = 1.0 / s;
    }
}

// Symmetric matrix: y = P v from the upper triangle of 256 x 256 tiles only (2 N^2 bytes per pass instead
// of 4 N^2).  Tile (I, J), I < J, gives the rows of block I their partial over the columns of block J and,
// because P[j][i] == P[i][j], the rows of block J their partial over the columns of block I; a diagonal
// tile gives its rows their partial over its own columns.  part[b][k] (k in block K) holds the contribution
// of tile (min(K,b), max(K,b)): every slot is written exactly once, k_symfinish sums the nb slots of an
// element in a fixed order, so the result does not depend on scheduling.
template <typename T, bool CHECK>
__global__ void __launch_bounds__(256, 2) k_sympass(const T *X, int n, T vmax, const double *v, double *part,
                                                    int32_t *flags, int nb) {
    const long long t = blockIdx.x;
    const double e = 2.0 * nb + 1.0;
    int I = (int)((e - sqrt(e * e - 8.0 * (double)t)) * 0.5);
    auto first = [&](int i) { return (long long)i * nb - (long long)i * (i - 1) / 2; };   // linear index of tile (i, i)
    while (I > 0 && first(I) > t) --I;
    while (I + 1 < nb && first(I + 1) <= t) ++I;
    const int J = I + (int)(t - first(I));
    const bool diag = I == J;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int r0 = I * ST + w * 32, c0 = J * ST + lane * 4;
    const bool vec = (n % 4 == 0) && (((uintptr_t)X & 15) == 0);
    const bool scale = vmax != (T)1;
    const typename hc::ScaledDiv<T>::type div(vmax);
    double vc[2][4], cacc[2][4];
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int j = c0 + h * 128 + k;
            vc[h][k] = j < n ? (v ? __ldg(v + j) : 1.0) : 0.0;
            cacc[h][k] = 0;
        }
    bool neg = false;
    for (int b = 0; b < 4; ++b) {
        const int rb = r0 + b * 8;
        if (rb >= n) break;                      // warp-uniform
        double racc[8], vr[8];
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
            racc[rr] = 0;
            vr[rr] = rb + rr < n ? (v ? __ldg(v + rb + rr) : 1.0) : 0.0;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int j0 = c0 + h * 128, valid = min(4, n - j0);
            T x[8][4];
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
                if (rb + rr < n && valid > 0) load4(X + (size_t)(rb + rr) * n + j0, x[rr], vec, valid);
                else x[rr][0] = x[rr][1] = x[rr][2] = x[rr][3] = (T)0;
            }
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
                T p[4];
                if (scale) div.apply(x[rr], p);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (CHECK) neg |= !(x[rr][k] >= (T)0);
                    const double pd = (double)(scale ? p[k] : x[rr][k]);
                    racc[rr] = fma(pd, vc[h][k], racc[rr]);
                    if (!diag) cacc[h][k] = fma(pd, vr[rr], cacc[h][k]);
                }
            }
        }
        // 8 row sums over the 32 lanes: exchange 4, 2, 1 values, then two plain steps (9 shuffles, fixed order)
        double b4[4], b2[2], s;
        {
            const bool up = lane & 16;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double send = up ? racc[k] : racc[k + 4], keep = up ? racc[k + 4] : racc[k];
                b4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
        }
        {
            const bool up = lane & 8;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const double send = up ? b4[k] : b4[k + 2], keep = up ? b4[k + 2] : b4[k];
                b2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
        }
        {
            const bool up = lane & 4;
            const double send = up ? b2[0] : b2[1], keep = up ? b2[1] : b2[0];
            s = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        const int rr = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
        if ((lane & 3) == 0 && rb + rr < n) part[(size_t)J * n + rb + rr] = s;
    }
    if (!diag) {                                  // CTA-uniform
        __shared__ double sh[8][ST];
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int k = 0; k < 4; ++k) sh[w][h * 128 + lane * 4 + k] = cacc[h][k];
        __syncthreads();
        const int j = J * ST + threadIdx.x;
        if (j < n) {
            double s = 0;
#pragma unroll
            for (int ww = 0; ww < 8; ++ww) s += sh[ww][threadIdx.x];
            part[(size_t)I * n + j] = s;
        }
    }
    if (CHECK && neg) atomicOr(flags, 1);
}

// y[j] = sum_b part[b][j]; with r: flags[4] |= any(y == 0), r = 1/y
__global__ void k_symfinish(const double *part, int nb, int n, double *y, double *r, int32_t *flags) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double s = 0;
    for (int b = 0; b < nb; ++b) s += part[(size_t)b * n + j];
    y[j] = s;
    if (r) {
        if (s == 0.0) atomicOr(flags + 4, 1);
        r[j] = 1.0 / s;
    }
}

template <typename T, typename TO>
__global__ void __launch_bounds__(256) k_apply(const T *X, int nrows, int n, T vmax, const double *r,
                                                const double *c, TO *out) {
    const int i = blockIdx.y;
    const double ri = r[i];
    const bool scale = vmax != (T)1;
    const typename hc::ScaledDiv<T>::type div(vmax);
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        T v[1] = {X[(size_t)i * n + j]}, q[1];
        if (scale) div.apply(v, q);
        T p = scale ? q[0] : v[0];
        out[(size_t)i * n + j] = (TO)__dmul_rn(__dmul_rn(ri, (double)p), c[j]);
    }
}

// every float a of the binades [1,2) and [2,4) against each divisor b: ScaledDivF vs the IEEE division
__global__ void k_divcheck(const float *b, int nb, unsigned long long *mismatches) {
    const hc::ScaledDivF div(b[blockIdx.y]);
    const float bb = b[blockIdx.y];
    unsigned long long bad = 0;
    for (unsigned m = blockIdx.x * blockDim.x + threadIdx.x; m < (1u << 24); m += gridDim.x * blockDim.x) {
        const float a[1] = {__uint_as_float(0x3F800000u + m)};
        float q[1];
        div.apply(a, q);
        bad += __float_as_uint(q[0]) != __float_as_uint(a[0] / bb);
    }
    if (bad) atomicAdd(mismatches, bad);
}

}  // namespace

extern "C" int32_t hc_sk_num_row_chunks(int32_t nrows) { return (nrows + RC - 1) / RC; }

extern "C" int hc_sk_colpass(const void *X, int dtype, int32_t nrows, int32_t n, double vmax, const double *r,
                             double *ypart, int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && ypart && nrows > 0 && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    dim3 grid((n + 1023) / 1024, (nrows + RC - 1) / RC);
    if (dtype == 0) {
        if (d_flags) k_colpass<float, true><<<grid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, r, ypart, d_flags);
        else k_colpass<float, false><<<grid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, r, ypart, d_flags);
    } else {
        if (d_flags) k_colpass<double, true><<<grid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, r, ypart, d_flags);
        else k_colpass<double, false><<<grid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, r, ypart, d_flags);
    }
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_sk_colfinish(const double *ypart, int32_t n_chunks, int32_t n, double *y, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(ypart && y && n_chunks > 0 && n > 0, "bad arguments");
    k_colfinish<<<(n + 255) / 256, 256, 0, stream>>>(ypart, n_chunks, n, y);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int32_t hc_sk_sym_num_blocks(int32_t n) { return (n + ST - 1) / ST; }

extern "C" int hc_sk_sympass(const void *X, int dtype, int32_t n, double vmax, const double *v, double *part,
                             int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && part && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    const int nb = (n + ST - 1) / ST;
    const long long tiles = (long long)nb * (nb + 1) / 2;
    HC_REQUIRE(tiles < (1ll << 31), "matrix too large for the symmetric pass");
    if (dtype == 0) {
        if (d_flags) k_sympass<float, true><<<(unsigned)tiles, 256, 0, stream>>>((const float *)X, n, (float)vmax, v, part, d_flags, nb);
        else k_sympass<float, false><<<(unsigned)tiles, 256, 0, stream>>>((const float *)X, n, (float)vmax, v, part, d_flags, nb);
    } else {
        if (d_flags) k_sympass<double, true><<<(unsigned)tiles, 256, 0, stream>>>((const double *)X, n, vmax, v, part, d_flags, nb);
        else k_sympass<double, false><<<(unsigned)tiles, 256, 0, stream>>>((const double *)X, n, vmax, v, part, d_flags, nb);
    }
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_sk_symfinish(const double *part, int32_t n, double *y, double *r, int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(part && y && n > 0 && (!r || d_flags), "bad arguments");
    k_symfinish<<<(n + 255) / 256, 256, 0, stream>>>(part, (n + ST - 1) / ST, n, y, r, d_flags);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_selftest_scaled_division(const float *d_b, int32_t nb, unsigned long long *d_mismatches, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(d_b && d_mismatches && nb > 0 && nb < 65536, "bad arguments");
    HC_CUDA_OK(cudaMemsetAsync(d_mismatches, 0, 8, stream));
    k_divcheck<<<dim3(296, nb), 256, 0, stream>>>(d_b, nb, d_mismatches);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_sk_update_c(const double *y, const double *c_prev, int32_t n, double eps, double *c,
                              int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(y && c && d_flags && n > 0, "bad arguments");
    k_update_c<<<(n + 255) / 256, 256, 0, stream>>>(y, c_prev, n, eps, c, d_flags);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_sk_outside(const double *v, int32_t n, double eps, int32_t *d_flag, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(v && d_flag && n > 0, "bad arguments");
    k_outside<<<(n + 255) / 256, 256, 0, stream>>>(v, n, eps, d_flag);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_sk_rowpass(const void *X, int dtype, int32_t nrows, int32_t n, double vmax, const double *c,
                             double *x, double *r, int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && x && d_flags && nrows > 0 && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    int grid = (nrows + RB - 1) / RB;
    if (dtype == 0) k_rowpass<float><<<grid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, c, x, r, d_flags);
    else k_rowpass<double><<<grid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, c, x, r, d_flags);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_sk_apply(const void *X, int dtype, int32_t nrows, int32_t n, double vmax, const double *r,
                           const double *c, void *out, int out_dtype, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && r && c && out && nrows > 0 && n > 0, "bad arguments");
    dim3 grid(min((n + 255) / 256, 32), nrows);
    if (dtype == 0 && out_dtype == 0) k_apply<float, float><<<grid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, r, c, (float *)out);
    else if (dtype == 0 && out_dtype == 1) k_apply<float, double><<<grid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, r, c, (double *)out);
    else if (dtype == 1 && out_dtype == 1) k_apply<double, double><<<grid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, r, c, (double *)out);
    else if (dtype == 1 && out_dtype == 0) k_apply<double, float><<<grid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, r, c, (float *)out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}
