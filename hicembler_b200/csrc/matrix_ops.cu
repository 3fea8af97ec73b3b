// K2 (fold / symmetrise), K6 (contacts -> distances into the linkage working matrix),
// condensed <-> square conversions and K8 block reductions.
//
// All N x N passes here are tile-PAIR kernels: one CTA owns tiles (I,J) and (J,I), I <= J, stages
// both through shared memory and writes both.  That makes "d += d.T - diag(d)" (reference
// bam2clusters.py:42,511) a single read + single write of the matrix (8 N^2 bytes for float32 =
// the algorithmic minimum), lets it run in place, and lets the transform read only the upper
// triangle that d[np.triu_indices(n,1)] (bam2clusters.py:385) selects.
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace {

constexpr int TS = 32;  // tile side; blocks are (32, 8)

__device__ __forceinline__ void atomic_max_nonneg(float *addr, float v) {
    atomicMax((int *)addr, __float_as_int(v));
}
__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v) {
    atomicMax((long long *)addr, __double_as_longlong(v));
}

template <typename T>
__device__ __forceinline__ T block_max_2d(T v, T *sh) {
    for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (threadIdx.x == 0) sh[threadIdx.y] = v;
    __syncthreads();
    T r = sh[0];
    for (int k = 1; k < blockDim.y; ++k) r = fmax(r, sh[k]);
    return r;
}

// counts (u32, full orientation) -> float32 upper-triangular (mode 0) or symmetrised (mode 1)
__global__ void __launch_bounds__(256) k_fold_counts(const uint32_t *C, int n, int mode, float add,
                                                      float *out, float *vmax) {
    const int I = blockIdx.y, J = blockIdx.x;
    if (I > J) return;
    __shared__ uint32_t A[TS][TS + 1], B[TS][TS + 1];
    __shared__ float shm[8];
    const int tx = threadIdx.x;
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        A[r][tx] = (i < n && j < n) ? C[(size_t)i * n + j] : 0u;
        int i2 = J * TS + r, j2 = I * TS + tx;
        B[r][tx] = (i2 < n && j2 < n) ? C[(size_t)i2 * n + j2] : 0u;
    }
    __syncthreads();
    float m = 0.f;
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        if (i < n && j < n) {
            uint32_t s = (i == j) ? A[r][tx] : A[r][tx] + B[tx][r];
            float v;
            if (mode == 0) v = (i <= j) ? (float)s : 0.f;
            else v = (float)s + add;
            out[(size_t)i * n + j] = v;
            m = fmaxf(m, v);
        }
        if (I != J) {
            int i2 = J * TS + r, j2 = I * TS + tx;
            if (i2 < n && j2 < n) {
                float v = (mode == 0) ? 0.f : (float)(B[r][tx] + A[tx][r]) + add;
                out[(size_t)i2 * n + j2] = v;
                m = fmaxf(m, v);
            }
        }
    }
    if (vmax) {
        m = block_max_2d(m, shm);
        if (threadIdx.x == 0 && threadIdx.y == 0) atomic_max_nonneg(vmax, m);
    }
}

// in place  d += d.T - diag(d.diagonal())  (+ add)
template <typename T>
__global__ void __launch_bounds__(256) k_symmetrize(T *d, int n, T add, T *vmax) {
    const int I = blockIdx.y, J = blockIdx.x;
    if (I > J) return;
    __shared__ T A[TS][TS + 1], B[TS][TS + 1];
    __shared__ T shm[8];
    const int tx = threadIdx.x;
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        A[r][tx] = (i < n && j < n) ? d[(size_t)i * n + j] : (T)0;
        int i2 = J * TS + r, j2 = I * TS + tx;
        B[r][tx] = (i2 < n && j2 < n) ? d[(size_t)i2 * n + j2] : (T)0;
    }
    __syncthreads();
    T m = 0;
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        if (i < n && j < n) {
            T v = ((i == j) ? A[r][tx] : A[r][tx] + B[tx][r]) + add;
            d[(size_t)i * n + j] = v;
            m = fmax(m, v);
        }
        if (I != J) {
            int i2 = J * TS + r, j2 = I * TS + tx;
            if (i2 < n && j2 < n) {
                T v = (B[r][tx] + A[tx][r]) + add;
                d[(size_t)i2 * n + j2] = v;
                m = fmax(m, v);
            }
        }
    }
    if (vmax) {
        m = block_max_2d(m, shm);
        if (threadIdx.x == 0 && threadIdx.y == 0) atomic_max_nonneg(vmax, m);
    }
}

// ---- 64 x 64 tile-pair kernels for 4-byte element types (n % 4 == 0, 16-byte aligned base) ----
// One CTA per unordered tile pair {(I,J),(J,I)}, I <= J (linear grid over the upper triangle of the
// tile grid, no idle CTAs): 16-byte coalesced loads of both tiles into shared memory, one
// barrier, 16-byte coalesced stores of both results.  u32 counts -> f32 (fold) or f32 in place.
constexpr int T64 = 64;

__device__ __forceinline__ void tile_pair_from_linear(long long t, int T, int &I, int &J) {
    const double d = 2.0 * T + 1.0;
    long long i = (long long)((d - sqrt(d * d - 8.0 * (double)t)) * 0.5);
    if (i < 0) i = 0;
    if (i > T - 1) i = T - 1;
    // start(i) = i*T - i*(i-1)/2 = index of pair (i,i)
    while (i > 0 && i * T - i * (i - 1) / 2 > t) --i;
    while (i + 1 < T && (i + 1) * T - (i + 1) * i / 2 <= t) ++i;
    I = (int)i;
    J = (int)(i + (t - (i * T - i * (i - 1) / 2)));
}

template <typename TIn>
__device__ __forceinline__ float pair_value(TIn a, TIn b, bool diag, float add);
template <>
__device__ __forceinline__ float pair_value<uint32_t>(uint32_t a, uint32_t b, bool diag, float add) {
    return (float)(diag ? a : a + b) + add;
}
template <>
__device__ __forceinline__ float pair_value<float>(float a, float b, bool diag, float add) {
    return (diag ? a : a + b) + add;
}

// mode 0 (counts only): upper triangle incl. diagonal, zeros below; mode 1: symmetrised (+add)
template <typename TIn>
__global__ void __launch_bounds__(256) k_tilepair64(const TIn *C, int n, int T, int mode, float add, float *out,
                                                    float *vmax) {
    __shared__ TIn A[T64][T64 + 1], B[T64][T64 + 1];
    __shared__ float shm[8];
    int I, J;
    tile_pair_from_linear(blockIdx.x, T, I, J);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const uint4 zero4 = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;
        const int i = I * T64 + r, j = J * T64 + tx * 4;
        uint4 v = zero4, w = zero4;
        if (i < n && j < n) v = *(const uint4 *)(C + (size_t)i * n + j);  // n % 4 == 0: j + 3 < n
        const int i2 = J * T64 + r, j2 = I * T64 + tx * 4;
        if (i2 < n && j2 < n) w = *(const uint4 *)(C + (size_t)i2 * n + j2);
        const TIn *pv = (const TIn *)&v, *pw = (const TIn *)&w;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            A[r][tx * 4 + q] = pv[q];
            B[r][tx * 4 + q] = pw[q];
        }
    }
    __syncthreads();
    float m = 0.f;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;
        const int i = I * T64 + r, j = J * T64 + tx * 4;
        if (i < n && j < n) {
            float o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = tx * 4 + q;
                float v = pair_value<TIn>(A[r][c], B[c][r], i == j + q, add);
                if (mode == 0) v = (i <= j + q) ? v : 0.f;
                o[q] = v;
                m = fmaxf(m, v);
            }
            *(float4 *)(out + (size_t)i * n + j) = make_float4(o[0], o[1], o[2], o[3]);
        }
        if (I != J) {
            const int i2 = J * T64 + r, j2 = I * T64 + tx * 4;
            if (i2 < n && j2 < n) {
                float o[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = tx * 4 + q;
                    float v = (mode == 0) ? 0.f : pair_value<TIn>(B[r][c], A[c][r], false, add);
                    o[q] = v;
                    m = fmaxf(m, v);
                }
                *(float4 *)(out + (size_t)i2 * n + j2) = make_float4(o[0], o[1], o[2], o[3]);
            }
        }
    }
    if (vmax) {
        for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0) shm[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int k = 1; k < 8; ++k) m = fmaxf(m, shm[k]);
            atomic_max_nonneg(vmax, m);
        }
    }
}

// Row blocks of one matrix living in different buffers (the row-sharded build keeps block k in the HBM of rank k;
// the linkage rank reads them over NVLink): row i = base[i / per] + (i % per) * n.  One buffer: per >= n.
struct RowBlocks {
    const void *base[HC_MAX_RANKS];
    int per;
    template <typename T>
    __device__ __forceinline__ const T *row(int i, int n) const {
        const int b = i / per;
        return (const T *)base[b] + (size_t)(i - b * per) * n;
    }
};

struct Xform {
    int kind;
    double p0, p1;
    const double *r, *c;
    double vmax;
    int square;
    double inv_p1, K;  // 1/p1 and (1/p0)^(1/p1), set by make_xform
};

static Xform make_xform(int kind, double p0, double p1, const double *r, const double *c, double vmax, int square) {
    Xform t{kind, p0, p1, r, c, vmax, square, 0.0, 0.0};
    if (kind == 1 || kind == 3) {
        t.inv_p1 = 1.0 / p1;
        t.K = pow(1.0 / p0, t.inv_p1);
    }
    return t;
}

// value of the distance for upper-triangle entry (i,j), i<j, given raw X[i][j]
template <typename T>
__device__ __forceinline__ double xform(const Xform &t, T raw, int i, int j) {
    double v;
    if (t.r) {
        T p = raw / (T)t.vmax;                       // d/vmax in the matrix dtype (bam2clusters.py:46)
        v = __dmul_rn(__dmul_rn(t.r[i], (double)p), t.c[j]);  // (D1.P).D2  (sinkhorn_knopp.py:161)
        if (t.kind == 1) v = v + 1.0;
    } else if (t.kind == 1 || t.kind == 2) {
        v = (double)(T)(raw + (T)1);                 // x+1 in the matrix dtype
    } else {
        v = (double)raw;
    }
    double out;
    if (t.kind == 1 || t.kind == 3) {
        // (1/(a x))^(1/b) (bam2clusters.py:248-251) = K * exp(-log(x)/b), K = (1/a)^(1/b).  Nearly every entry of a
        // balanced (or empty) Hi-C matrix is x = 1 + y with y <= 2^-10 (y = v - 1 is exact): two short series replace
        // the division, log and exp (~120 FP64 instructions -> ~15), truncation error < 2^-60.  Elsewhere pow() is
        // ~3x the FP64 work of exp(e*log(u)); that form is used only while |e*log(u)| < 1, where the error of the
        // exponent is not amplified (<= 3 ulp in all).
        const double e = t.inv_p1, y = v - 1.0;
        if (sizeof(T) == 4 && !t.r) {
            // float32 matrix without balancing vectors (contigs2scaffold: distance_func on the float32 block means,
            // bam2scaffolds.py:265): numpy evaluates (1./(a*x))**(1./b) in float32 -- a and 1/b are scalars and do
            // not promote the array -- so a*x, the division and the power are float32 operations.
            const float u = __fdiv_rn(1.0f, __fmul_rn((float)t.p0, (float)v));
            out = (double)powf(u, (float)e);
        } else if (y >= 0.0 && y <= 0x1p-10 && fabs(e) <= 4.0) {   // |z| <= 2^-8: exp series truncation < 2^-57
            double L = fma(y, -1.0 / 6.0, 0.2);
            L = fma(y, L, -0.25);
            L = fma(y, L, 1.0 / 3.0);
            L = fma(y, L, -0.5);
            L = fma(y, L, 1.0);
            const double z = -__dmul_rn(e, __dmul_rn(y, L));  // -log1p(y)/b
            double E = fma(z, 1.0 / 120.0, 1.0 / 24.0);
            E = fma(z, E, 1.0 / 6.0);
            E = fma(z, E, 0.5);
            E = fma(z, E, 1.0);
            out = fma(t.K, __dmul_rn(z, E), t.K);  // K * (1 + expm1(z))
        } else {
            const double u = 1.0 / (t.p0 * v);
            const double lg = log(u), z = __dmul_rn(e, lg), zl = fma(e, lg, -z);  // z + zl = e*lg exactly
            if (fabs(z) < 1.0) {
                out = exp(z);
                out = fma(out, zl, out);
            } else {
                out = pow(u, e);
            }
        }
    }
    else if (t.kind == 2) {
        if (sizeof(T) == 4 && !t.r) out = (double)(logf((float)t.p0) - logf((float)v));
        else out = log(t.p0) - log(t.r ? v + 1.0 : v);
    } else out = v;
    return t.square ? __dmul_rn(out, out) : out;
}

template <typename T>
__global__ void __launch_bounds__(256) k_transform_working(RowBlocks X, int n, Xform t, double *W,
                                                            int32_t *flags) {
    const int I = blockIdx.y, J = blockIdx.x;
    if (I > J) return;
    __shared__ double A[TS][TS + 1];
    const int tx = threadIdx.x;
    bool bad = false;
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        double v = 0;
        if (i < n && j < n && i < j) {
            v = xform<T>(t, X.row<T>(i, n)[j], i, j);
            bad |= isnan(v);
        }
        A[r][tx] = v;
    }
    __syncthreads();
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        if (i < n && j < n) {
            double v = (i < j) ? A[r][tx] : (i == j ? inf : A[tx][r]);
            W[(size_t)i * n + j] = v;
        }
        if (I != J) {
            int i2 = J * TS + r, j2 = I * TS + tx;  // lower tile: W[i2][j2] = f(X[j2][i2])
            if (i2 < n && j2 < n) W[(size_t)i2 * n + j2] = A[tx][r];
        }
    }
    if (bad) atomicOr(flags, 1);
}

// 64 x 64 tile version for the float matrix (n % 4 == 0): one CTA per tile pair of the upper triangle, float4
// loads, the transformed tile staged once in shared memory as f64, both mirror tiles written with 16-byte stores.
__global__ void __launch_bounds__(256) k_transform_tile64(RowBlocks X, int n, int T, Xform t, double *W,
                                                           int32_t *flags) {
    __shared__ double A[T64][T64 + 1];
    int I, J;
    tile_pair_from_linear(blockIdx.x, T, I, J);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    bool bad = false;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;
        const int i = I * T64 + r, j = J * T64 + tx * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < n && j < n && i < j + 3) v = __ldcs((const float4 *)(X.row<float>(i, n) + j));
        const float *pv = (const float *)&v;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            double o = 0;
            if (i < n && j + q < n && i < j + q) {
                o = xform<float>(t, pv[q], i, j + q);
                bad |= isnan(o);
            }
            A[r][tx * 4 + q] = o;
        }
    }
    __syncthreads();
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;
        const int i = I * T64 + r, j = J * T64 + tx * 4;
        if (i < n && j < n) {
            double o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = tx * 4 + q;
                o[q] = (i < j + q) ? A[r][c] : (i == j + q ? inf : A[c][r]);  // (I == J only below the diagonal)
            }
            double2 *dst = (double2 *)(W + (size_t)i * n + j);
            dst[0] = make_double2(o[0], o[1]);
            dst[1] = make_double2(o[2], o[3]);
        }
        if (I != J) {
            const int i2 = J * T64 + r, j2 = I * T64 + tx * 4;
            if (i2 < n && j2 < n) {
                double2 *dst = (double2 *)(W + (size_t)i2 * n + j2);
                dst[0] = make_double2(A[tx * 4 + 0][r], A[tx * 4 + 1][r]);
                dst[1] = make_double2(A[tx * 4 + 2][r], A[tx * 4 + 3][r]);
            }
        }
    }
    if (bad) atomicOr(flags, 1);
}

// ---- the hot case of K6: float matrix, Sinkhorn vectors given, power law (bam2clusters.py:516 on the balanced matrix) ----
// k_transform_tile64 inlines the general xform() sixteen times (every kind, pow/log/exp fall-backs: 13 600 SASS lines,
// ~120 instructions per cell, 4-way shared-memory bank conflicts from its float4 thread layout) and ran at 42 % of the
// HBM peak, issue-bound.  Here: one lane per COLUMN (a warp reads / writes 128 / 256 contiguous bytes per instruction and
// walks shared memory conflict-free in both orientations, row stride 65 doubles), all 16 loads of a thread in flight
// before the first use, d/vmax by the exact reciprocal form (ScaledDivF), the two short series straight-line, and the
// rare cell outside the series range in ONE out-of-line call.  Same arithmetic as xform(): identical bits.
__device__ __noinline__ double power_outside_series(double v, double e, double p0) {
    const double u = 1.0 / (p0 * v);
    const double lg = log(u), z = __dmul_rn(e, lg), zl = fma(e, lg, -z);  // z + zl = e*lg exactly
    if (fabs(z) < 1.0) {
        const double out = exp(z);
        return fma(out, zl, out);
    }
    return pow(u, e);
}

// INTERIOR: a tile strictly above the diagonal with all 64 x 64 cells inside the matrix (all but ~2/T of the tiles):
// no guards, pointer increments only.
template <bool SQUARE, bool INTERIOR>
__device__ __forceinline__ void transform_bal64_tile(const RowBlocks &X, int n, int I, int J, const Xform &t,
                                                     double *__restrict__ W, int32_t *flags, double (*A)[T64 + 1]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const hc::ScaledDivF dv((float)t.vmax);
    const double e = t.inv_p1, K = t.K;
    const bool series_ok = fabs(e) <= 4.0;
    const int j0 = J * T64 + lane, j1 = j0 + 32;
    const double cj[2] = {INTERIOR || j0 < n ? __ldg(t.c + j0) : 0.0, INTERIOR || j1 < n ? __ldg(t.c + j1) : 0.0};
    float x[16], pq[16];
    double ri[8];
    // all rows of a tile live in the same row block (blocks are multiples of 64 rows, or one block holds everything)
    const int i_base = I * T64, blk = i_base / X.per;
    const float *Xb = (const float *)X.base[blk] + (size_t)(i_base - blk * X.per + warp) * n;
    unsigned valid = INTERIOR ? 0xffffu : 0u;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const int i = i_base + warp + 8 * p;
        const float *row = Xb + (size_t)(8 * p) * n;
        if (INTERIOR) {
            ri[p] = __ldg(t.r + i);
            x[2 * p] = __ldcs(row + j0);
            x[2 * p + 1] = __ldcs(row + j1);
        } else {
            x[2 * p] = x[2 * p + 1] = 1.0f;   // masked cells: a harmless operand for the grouped division
            ri[p] = 0.0;
            if (i < n) {
                ri[p] = __ldg(t.r + i);
                if (j0 < n && i < j0) { x[2 * p] = __ldcs(row + j0); valid |= 1u << (2 * p); }
                if (j1 < n && i < j1) { x[2 * p + 1] = __ldcs(row + j1); valid |= 2u << (2 * p); }
            }
        }
    }
    dv.apply(x, pq);                      // d/vmax in the matrix dtype (bam2clusters.py:46)
    // sixteen independent straight-line chains (the series of xform()); cells outside the series range are redone below
    double o[16];
    unsigned slow = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const double v = __dadd_rn(__dmul_rn(__dmul_rn(ri[k >> 1], (double)pq[k]), cj[k & 1]), 1.0);  // (D1.P).D2 + 1
        const double y = __dsub_rn(v, 1.0);
        if (!(y >= 0.0 && y <= 0x1p-10)) slow |= 1u << k;
        double L = fma(y, -1.0 / 6.0, 0.2);
        L = fma(y, L, -0.25);
        L = fma(y, L, 1.0 / 3.0);
        L = fma(y, L, -0.5);
        L = fma(y, L, 1.0);
        const double z = -__dmul_rn(e, __dmul_rn(y, L));
        double E = fma(z, 1.0 / 120.0, 1.0 / 24.0);
        E = fma(z, E, 1.0 / 6.0);
        E = fma(z, E, 0.5);
        E = fma(z, E, 1.0);
        o[k] = fma(K, __dmul_rn(z, E), K);
    }
    if (!series_ok) slow = 0xffffu;
    slow &= valid;
    if (slow) {
#pragma unroll
        for (int k = 0; k < 16; ++k)
            if (slow >> k & 1) {
                const double v = __dadd_rn(__dmul_rn(__dmul_rn(ri[k >> 1], (double)pq[k]), cj[k & 1]), 1.0);
                o[k] = power_outside_series(v, e, t.p0);
            }
    }
    bool bad = false;
    double *As = &A[warp][lane];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        double q = o[k];
        if (INTERIOR || (valid >> k & 1)) bad |= q != q;
        if (SQUARE) q = __dmul_rn(q, q);
        As[(k >> 1) * 8 * (T64 + 1) + (k & 1) * 32] = (INTERIOR || (valid >> k & 1)) ? q : 0.0;
    }
    __syncthreads();
    if (INTERIOR) {
        double *d1 = W + (size_t)(i_base + warp) * n + j0;            // W[i][j]
        double *d2 = W + (size_t)(J * T64 + warp) * n + i_base + lane;  // mirror tile: W[j][i] = the same value
        const double *a2 = &A[lane][warp];
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            const size_t g = (size_t)(8 * p) * n;
            d1[g] = As[8 * p * (T64 + 1)];
            d1[g + 32] = As[8 * p * (T64 + 1) + 32];
            d2[g] = a2[8 * p];
            d2[g + 32] = a2[32 * (T64 + 1) + 8 * p];
        }
    } else {
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            const int r = warp + 8 * p, i = I * T64 + r;
            if (i < n) {
                double *dst = W + (size_t)i * n;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c = lane + 32 * h, j = J * T64 + c;
                    if (j < n) dst[j] = (I != J || i < j) ? A[r][c] : (i == j ? inf : A[c][r]);
                }
            }
            if (I != J) {
                const int i2 = J * T64 + r;   // mirror tile: W[i2][j2] = f(X[j2][i2])
                if (i2 < n) {
                    double *dst = W + (size_t)i2 * n;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int c = lane + 32 * h, j2 = I * T64 + c;
                        if (j2 < n) dst[j2] = A[c][r];
                    }
                }
            }
        }
    }
    if (bad) atomicOr(flags, 1);
}

template <bool SQUARE>
__global__ void __launch_bounds__(256, 3) k_transform_bal64(RowBlocks X, int n, int T, Xform t, double *__restrict__ W,
                                                             int32_t *flags) {
    __shared__ double A[T64][T64 + 1];
    int I, J;
    tile_pair_from_linear(blockIdx.x, T, I, J);
    if (I != J && (J + 1) * T64 <= n) transform_bal64_tile<SQUARE, true>(X, n, I, J, t, W, flags, A);
    else transform_bal64_tile<SQUARE, false>(X, n, I, J, t, W, flags, A);
}

template <typename TI, typename TO>
__global__ void k_transform_flat(const TI *X, long long count, Xform t, TO *out) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < count;
         k += (long long)gridDim.x * blockDim.x)
        out[k] = (TO)xform<TI>(t, X[k], 0, 0);
}

__device__ __forceinline__ long long condensed_row_offset(long long i, long long n) {
    return i * (2 * n - i - 1) / 2;  // index of (i, i+1)
}

template <typename T>
__global__ void k_condense(const T *X, int n, double *y) {
    int i = blockIdx.x;
    long long off = condensed_row_offset(i, n);
    for (int j = i + 1 + threadIdx.x; j < n; j += blockDim.x) y[off + (j - i - 1)] = (double)X[(size_t)i * n + j];
}

__global__ void __launch_bounds__(256) k_expand_condensed(const double *y, int n, int square, double *W,
                                                           int32_t *flags) {
    const int I = blockIdx.y, J = blockIdx.x;
    if (I > J) return;
    __shared__ double A[TS][TS + 1];
    const int tx = threadIdx.x;
    bool bad = false;
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        double v = 0;
        if (i < n && j < n && i < j) {
            v = y[condensed_row_offset(i, n) + (j - i - 1)];
            bad |= isnan(v);
            if (square) v = __dmul_rn(v, v);
        }
        A[r][tx] = v;
    }
    __syncthreads();
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    for (int r = threadIdx.y; r < TS; r += 8) {
        int i = I * TS + r, j = J * TS + tx;
        if (i < n && j < n) W[(size_t)i * n + j] = (i < j) ? A[r][tx] : (i == j ? inf : A[tx][r]);
        if (I != J) {
            int i2 = J * TS + r, j2 = I * TS + tx;
            if (i2 < n && j2 < n) W[(size_t)i2 * n + j2] = A[tx][r];
        }
    }
    if (bad) atomicOr(flags, 1);
}

// ---- K8 ----------------------------------------------------------------------------------
// stage 1: tmp[k][j] = mean over rows of segment k (sequential float32 adds in row order, the
// order np.mean(A[indices], axis=0) uses for a C-contiguous selection), bam2scaffolds.py:63
__global__ void k_segmean_rows(const float *A, int n, const int32_t *seg_begin, const int32_t *seg_end,
                               int n_seg, float *tmp) {
    int k = blockIdx.y;
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int b = seg_begin[k], e = seg_end[k];
    float s = 0.f;
    for (int i = b; i < e; ++i) s = __fadd_rn(s, A[(size_t)i * n + j]);
    tmp[(size_t)k * n + j] = __fdiv_rn(s, (float)(e - b));
}
// stage 2: out[k1][k2] = mean over j in segment k2 of tmp[k1][j]
__global__ void k_segmean_cols(const float *tmp, int n, const int32_t *seg_begin, const int32_t *seg_end,
                               int n_seg, float *out) {
    int k2 = blockIdx.x * blockDim.x + threadIdx.x, k1 = blockIdx.y;
    if (k2 >= n_seg) return;
    int b = seg_begin[k2], e = seg_end[k2];
    float s = 0.f;
    for (int j = b; j < e; ++j) s = __fadd_rn(s, tmp[(size_t)k1 * n + j]);
    out[(size_t)k1 * n_seg + k2] = __fdiv_rn(s, (float)(e - b));
}

// normalize_rows (array2scaffolds.py:33-41): column sums, row sums and row maxima in one read
__global__ void k_rowcol_sums(const float *A, int n, double *colsum, double *rowsum, float *rowmax) {
    // block (256): rows [blockIdx.y*64, +64), columns blockIdx.x*256 + tx
    int j = blockIdx.x * 256 + threadIdx.x;
    int i0 = blockIdx.y * 64, i1 = min(i0 + 64, n);
    double cs = 0;
    for (int i = i0; i < i1; ++i) {
        float v = (j < n) ? A[(size_t)i * n + j] : 0.f;
        cs += v;
        float mx = v;
        double rs = hc::warp_sum_f64((double)v);
        for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(rowsum + i, rs);
            atomicMax((int *)(rowmax + i), __float_as_int(mx));
        }
    }
    if (j < n) atomicAdd(colsum + j, cs);
}
__global__ void k_colsum_max(const double *colsum, int n, float *maxv) {
    __shared__ float sh[8];
    float m = 0.f;
    for (int j = threadIdx.x; j < n; j += blockDim.x) m = fmaxf(m, (float)colsum[j]);
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) m = fmaxf(m, sh[k]);
        *maxv = fmaxf(m, sh[0]);
    }
}
__global__ void k_scale_rows(float *A, int n, const double *rowsum, const float *rowmax, const float *maxv) {
    int i = blockIdx.y;
    if (rowmax[i] == 0.f) return;  // "only if any signal"
    float f = __fdiv_rn(*maxv, (float)rowsum[i]);  // 1.*maxv/a[i].sum() in float32 (NEP 50)
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
        A[(size_t)i * n + j] = __fmul_rn(A[(size_t)i * n + j], f);
}

// bam2array's tail (bam2clusters.py:244): array = (array / sizes).T / sizes on the float32
// upper-triangular count matrix -> float64, transposed (lower triangular)
__global__ void k_contig_normalize(const uint32_t *C, int n, const int32_t *sizes, double *out) {
    int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= n) return;
    double v = 0.0;
    if (i >= j) {
        uint32_t u = (i == j) ? C[(size_t)i * n + i] : C[(size_t)j * n + i] + C[(size_t)i * n + j];
        v = __ddiv_rn(__ddiv_rn((double)(float)u, (double)sizes[i]), (double)sizes[j]);
    }
    out[(size_t)i * n + j] = v;
}

// ---- alternative normalisations ---------------------------------------------------
// d[indices, :][:, indices] (bam2clusters.py:55-57)
template <typename T>
__global__ void k_gather_square(const T *X, int n, const int32_t *keep, int m, T *out) {
    const int i = blockIdx.y;
    const T *src = X + (size_t)keep[i] * n;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < m; j += gridDim.x * blockDim.x)
        out[(size_t)i * m + j] = src[keep[j]];
}

// d *= (d*n2).T*n2 with n2 = mean(diag)/diag (bam2clusters.py:64-65), all in the matrix dtype like numpy:
// out[i][j] = d[i][j] * ((d[j][i] * n2[i]) * n2[j])
template <typename T>
__global__ void k_diag_normalize(const T *D, int n, T mean, T *out) {
    __shared__ T tile[TS][TS + 1];
    const int I = blockIdx.y, J = blockIdx.x, tx = threadIdx.x;
    for (int r = threadIdx.y; r < TS; r += 8) {  // tile (J, I) of D, read along rows
        const int i = J * TS + r, j = I * TS + tx;
        tile[r][tx] = (i < n && j < n) ? D[(size_t)i * n + j] : (T)0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < TS; r += 8) {
        const int i = I * TS + r, j = J * TS + tx;
        if (i < n && j < n) {
            const T n2i = mean / D[(size_t)i * n + i], n2j = mean / D[(size_t)j * n + j];
            const T f = (tile[tx][r] * n2i) * n2j;  // d[j][i] * n2[i] * n2[j]
            out[(size_t)i * n + j] = D[(size_t)i * n + j] * f;
        }
    }
}

// d *= 1. * windowSize**2 / (sizes * sizes.T) (bam2scaffolds.py:40): float64 factor, product cast back to T
template <typename T>
__global__ void k_window_size_normalize(T *D, int n, const long long *sizes, double ws2) {
    const int i = blockIdx.y;
    const long long si = sizes[i];
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const double f = ws2 / (double)(si * sizes[j]);
        D[(size_t)i * n + j] = (T)((double)D[(size_t)i * n + j] * f);
    }
}

}  // namespace

extern "C" int hc_fold_counts(const uint32_t *counts, int32_t n, int mode, float add, float *out,
                              float *d_vmax, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(counts && out && n > 0 && (mode == 0 || mode == 1), "bad arguments");
    if (d_vmax) HC_CUDA_OK(cudaMemsetAsync(d_vmax, 0, sizeof(float), stream));
    if ((n & 3) == 0 && (((uintptr_t)counts | (uintptr_t)out) & 15) == 0 && !getenv("HC_FOLD_TS32")) {
        const int T = (n + T64 - 1) / T64;
        const long long pairs = (long long)T * (T + 1) / 2;
        k_tilepair64<uint32_t><<<(unsigned)pairs, 256, 0, stream>>>(counts, n, T, mode, add, out, d_vmax);
    } else {
        int T = (n + TS - 1) / TS;
        k_fold_counts<<<dim3(T, T), dim3(TS, 8), 0, stream>>>(counts, n, mode, add, out, d_vmax);
    }
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_symmetrize(void *d, int32_t n, int dtype, double add, void *d_vmax, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(d && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    int T = (n + TS - 1) / TS;
    if (d_vmax) HC_CUDA_OK(cudaMemsetAsync(d_vmax, 0, dtype ? 8 : 4, stream));
    if (dtype == 0 && (n & 3) == 0 && ((uintptr_t)d & 15) == 0 && !getenv("HC_FOLD_TS32")) {
        const int T6 = (n + T64 - 1) / T64;
        const long long pairs = (long long)T6 * (T6 + 1) / 2;
        k_tilepair64<float><<<(unsigned)pairs, 256, 0, stream>>>((const float *)d, n, T6, 1, (float)add, (float *)d,
                                                                   (float *)d_vmax);
    } else if (dtype == 0)
        k_symmetrize<float><<<dim3(T, T), dim3(TS, 8), 0, stream>>>((float *)d, n, (float)add, (float *)d_vmax);
    else
        k_symmetrize<double><<<dim3(T, T), dim3(TS, 8), 0, stream>>>((double *)d, n, add, (double *)d_vmax);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

static int transform_blocks(const RowBlocks &rb, bool aligned16, int dtype, int32_t n, int kind, double p0, double p1,
                            const double *r, const double *c, double vmax, int square, double *W, int32_t *d_flags,
                            cudaStream_t stream) {
    Xform t = make_xform(kind, p0, p1, r, c, vmax, square);
    int T = (n + TS - 1) / TS;
    HC_CUDA_OK(cudaMemsetAsync(d_flags, 0, sizeof(int32_t), stream));
    if (dtype == 0 && kind == 1 && r && !getenv("HC_TRANSFORM_GENERIC")) {   // the balanced power-law case: any n, any alignment
        const int T6 = (n + T64 - 1) / T64;
        const long long pairs = (long long)T6 * (T6 + 1) / 2;
        if (square) k_transform_bal64<true><<<(unsigned)pairs, 256, 0, stream>>>(rb, n, T6, t, W, d_flags);
        else k_transform_bal64<false><<<(unsigned)pairs, 256, 0, stream>>>(rb, n, T6, t, W, d_flags);
    } else if (dtype == 0 && (n & 3) == 0 && aligned16 && (((uintptr_t)W) & 15) == 0 && !getenv("HC_FOLD_TS32")) {
        const int T6 = (n + T64 - 1) / T64;
        const long long pairs = (long long)T6 * (T6 + 1) / 2;
        k_transform_tile64<<<(unsigned)pairs, 256, 0, stream>>>(rb, n, T6, t, W, d_flags);
    } else if (dtype == 0)
        k_transform_working<float><<<dim3(T, T), dim3(TS, 8), 0, stream>>>(rb, n, t, W, d_flags);
    else
        k_transform_working<double><<<dim3(T, T), dim3(TS, 8), 0, stream>>>(rb, n, t, W, d_flags);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_transform_to_working(const void *X, int dtype, int32_t n, int kind, double p0, double p1,
                                       const double *r, const double *c, double vmax, int square,
                                       double *W, int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && W && d_flags && n > 0 && kind >= 0 && kind <= 3 && (dtype == 0 || dtype == 1), "bad arguments");
    HC_REQUIRE((r == nullptr) == (c == nullptr), "r and c must be given together");
    RowBlocks rb;
    memset(&rb, 0, sizeof rb);
    rb.base[0] = X;
    rb.per = n;
    return transform_blocks(rb, (((uintptr_t)X) & 15) == 0, dtype, n, kind, p0, p1, r, c, vmax, square, W, d_flags, stream);
}

extern "C" int hc_transform_gather(const void *const *h_blocks, int32_t nblocks, int32_t per, int dtype, int32_t n, int kind,
                                   double p0, double p1, const double *r, const double *c, double vmax, int square,
                                   double *W, int32_t *d_flags, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(h_blocks && W && d_flags && n > 0 && kind >= 0 && kind <= 3 && (dtype == 0 || dtype == 1), "bad arguments");
    HC_REQUIRE(nblocks >= 1 && nblocks <= HC_MAX_RANKS && per > 0 && per % T64 == 0 && (long long)per * nblocks >= n,
               "rows per block must be a multiple of 64 and cover the matrix");
    HC_REQUIRE((r == nullptr) == (c == nullptr), "r and c must be given together");
    RowBlocks rb;
    memset(&rb, 0, sizeof rb);
    bool aligned = true;
    for (int k = 0; k < nblocks; ++k) {
        HC_REQUIRE(h_blocks[k] || (long long)per * k >= n, "null block pointer");
        rb.base[k] = h_blocks[k];
        aligned = aligned && (((uintptr_t)h_blocks[k]) & 15) == 0;
    }
    rb.per = per;
    return transform_blocks(rb, aligned, dtype, n, kind, p0, p1, r, c, vmax, square, W, d_flags, stream);
}

// max over a flat array (np.max of array2scaffolds.py:31 / bam2clusters.py:45): block maxima, then one block
template <typename T>
__global__ void __launch_bounds__(256) k_max_partial(const T *X, long long count, T *part) {
    __shared__ T sh[8];
    T m = -INFINITY;
    bool nan = false;
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < count; k += (long long)gridDim.x * blockDim.x) {
        const T v = X[k];
        nan |= v != v;
        m = v > m ? v : m;
    }
    if (nan) m = (T)NAN;   // numpy's max propagates NaN
    for (int o = 16; o; o >>= 1) {
        const T w = __shfl_xor_sync(0xffffffffu, m, o);
        m = (m != m || w != w) ? (T)NAN : (w > m ? w : m);
    }
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) m = (m != m || sh[k] != sh[k]) ? (T)NAN : (sh[k] > m ? sh[k] : m);
        part[blockIdx.x] = m;
    }
}

extern "C" int hc_matrix_max(const void *X, int dtype, int64_t count, void *d_partial, int32_t n_partial, void *d_max,
                             void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && d_partial && d_max && count > 0 && n_partial >= 1 && (dtype == 0 || dtype == 1), "bad arguments");
    const int grid = (int)std::min<long long>(n_partial, (count + 255) / 256);
    if (dtype == 0) {
        k_max_partial<float><<<grid, 256, 0, stream>>>((const float *)X, count, (float *)d_partial);
        k_max_partial<float><<<1, 256, 0, stream>>>((const float *)d_partial, grid, (float *)d_max);
    } else {
        k_max_partial<double><<<grid, 256, 0, stream>>>((const double *)X, count, (double *)d_partial);
        k_max_partial<double><<<1, 256, 0, stream>>>((const double *)d_partial, grid, (double *)d_max);
    }
    HC_CHECK_LAUNCH();
    return HC_OK;
}

__global__ void k_gather_elements(const float *A, int n, const int32_t *rows, const int32_t *cols, long long m, float *out) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < m; k += (long long)gridDim.x * blockDim.x)
        out[k] = A[(size_t)rows[k] * n + cols[k]];
}

extern "C" int hc_gather_elements(const float *A, int32_t n, const int32_t *rows, const int32_t *cols, int64_t m,
                                  float *out, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(A && n > 0 && m >= 0 && (m == 0 || (rows && cols && out)), "bad arguments");
    if (m == 0) return HC_OK;
    k_gather_elements<<<(int)std::min<long long>((m + 255) / 256, 4096), 256, 0, stream>>>(A, n, rows, cols, m, out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_transform(const void *X, int dtype, int64_t count, int kind, double p0, double p1, void *out,
                            int out_dtype, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && out && count >= 0 && kind >= 0 && kind <= 3, "bad arguments");
    Xform t = make_xform(kind, p0, p1, nullptr, nullptr, 1.0, 0);
    if (count == 0) return HC_OK;
    int blocks = 8 * hc::sm_count();
    if (dtype == 0 && out_dtype == 0)
        k_transform_flat<float, float><<<blocks, 256, 0, stream>>>((const float *)X, count, t, (float *)out);
    else if (dtype == 0 && out_dtype == 1)
        k_transform_flat<float, double><<<blocks, 256, 0, stream>>>((const float *)X, count, t, (double *)out);
    else if (dtype == 1 && out_dtype == 1)
        k_transform_flat<double, double><<<blocks, 256, 0, stream>>>((const double *)X, count, t, (double *)out);
    else {
        hc::set_error("hc_transform: unsupported dtype combination");
        return HC_EINVAL;
    }
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_condense(const void *X, int dtype, int32_t n, double *y, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && y && n > 0, "bad arguments");
    if (n < 2) return HC_OK;
    if (dtype == 0) k_condense<float><<<n - 1, 256, 0, stream>>>((const float *)X, n, y);
    else k_condense<double><<<n - 1, 256, 0, stream>>>((const double *)X, n, y);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_expand_condensed(const double *y, int32_t n, int square, double *W, int32_t *d_flags,
                                   void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(y && W && d_flags && n > 0, "bad arguments");
    int T = (n + TS - 1) / TS;
    HC_CUDA_OK(cudaMemsetAsync(d_flags, 0, sizeof(int32_t), stream));
    k_expand_condensed<<<dim3(T, T), dim3(TS, 8), 0, stream>>>(y, n, square, W, d_flags);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_segment_mean_2d(const float *A, int32_t n, const int32_t *seg_begin, const int32_t *seg_end,
                                  int32_t n_seg, float *tmp, float *out, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(A && seg_begin && seg_end && tmp && out && n > 0 && n_seg > 0, "bad arguments");
    k_segmean_rows<<<dim3((n + 255) / 256, n_seg), 256, 0, stream>>>(A, n, seg_begin, seg_end, n_seg, tmp);
    k_segmean_cols<<<dim3((n_seg + 127) / 128, n_seg), 128, 0, stream>>>(tmp, n, seg_begin, seg_end, n_seg, out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

// hc_segment_mean_2d without an allocation anywhere (for callers that cannot cudaMalloc: graph capture, tight
// serving loops): the n_seg x n intermediate lives in a file-scope device array that every GPU loading the library
// holds until exit (HC_SEGMEAN_STATIC_MAX_ELEMS floats = 8 MiB).  The name below is only a host-side stand-in in
// host code; the device address is taken per device with cudaGetSymbolAddress, with that device current.
static __device__ __align__(256) float g_segmean_scratch[HC_SEGMEAN_STATIC_MAX_ELEMS];

static int segmean_scratch_ptr(float **out) {
    constexpr int MAX_DEV = 64;
    static float *ptr[MAX_DEV];       // written once per device; a race between two first calls stores the same value
    int dev = -1;
    HC_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) {
        hc::set_error("hc_segment_mean_2d_static: device ordinal %d not supported (max %d)", dev, MAX_DEV - 1);
        return HC_EINVAL;
    }
    float *p = __atomic_load_n(&ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        HC_CUDA_OK(cudaGetSymbolAddress((void **)&p, g_segmean_scratch));
        if (!p) {
            hc::set_error("hc_segment_mean_2d_static: cudaGetSymbolAddress returned a null pointer");
            return HC_ECUDA;
        }
        __atomic_store_n(&ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return HC_OK;
}

extern "C" int hc_segment_mean_2d_static(const float *A, int32_t n, const int32_t *seg_begin, const int32_t *seg_end,
                                         int32_t n_seg, float *out, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(A && seg_begin && seg_end && out && n > 0 && n_seg > 0, "bad arguments");
    if ((long long)n * n_seg > (long long)HC_SEGMEAN_STATIC_MAX_ELEMS) {
        hc::set_error("hc_segment_mean_2d_static: n * n_seg = %lld exceeds the static scratch of %lld floats "
                      "(HC_SEGMEAN_STATIC_MAX_ELEMS); use hc_segment_mean_2d with a caller-provided tmp",
                      (long long)n * n_seg, (long long)HC_SEGMEAN_STATIC_MAX_ELEMS);
        return HC_EWORKSPACE;
    }
    float *tmp = nullptr;
    int rc = segmean_scratch_ptr(&tmp);
    if (rc != HC_OK) return rc;
    // every element of tmp[0 .. n_seg*n) that stage 2 reads is written by stage 1 of this call: nothing has to
    // start at zero, whatever the previous call left behind
    k_segmean_rows<<<dim3((n + 255) / 256, n_seg), 256, 0, stream>>>(A, n, seg_begin, seg_end, n_seg, tmp);
    k_segmean_cols<<<dim3((n_seg + 127) / 128, n_seg), 128, 0, stream>>>(tmp, n, seg_begin, seg_end, n_seg, out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

// join_scaffolds (bam2scaffolds.py:142-174): the four corner-block sums that decide how two scaffolds are
// oriented.  One CTA per (job, corner); a job's windows are idx[off .. off+n1) (scaffold 1) followed by
// idx[off+n1 .. off+n1+n2) (scaffold 2).  Rows: first n1/2 (corners 0,1) or last n1/2 (2,3) windows of
// scaffold 1; columns: first n2/2 (0,2) or last n2/2 (1,3) windows of scaffold 2 (all of them when n2/2 == 0, as the
// reference's [-0:] slice does).  Sums are taken in double
// in a fixed order (exact for count matrices).
__global__ void __launch_bounds__(256) k_block_sums4(const float *A, int32_t n, const int32_t *idx,
                                                      const int64_t *job_off, const int32_t *job_n1,
                                                      const int32_t *job_n2, double *out) {
    __shared__ double sh[8];
    const int job = blockIdx.x >> 2, q = blockIdx.x & 3;
    const int32_t n1 = job_n1[job], n2 = job_n2[job];
    const int32_t *rows = idx + job_off[job], *cols = rows + n1;
    const int32_t h = n1 / 2;
    int32_t w = n2 / 2;
    if (q & 2) rows += n1 - h;
    if (q & 1) {
        if (w == 0) w = n2;  // Python's d[:, -0:] is the whole axis (n2 == 1, reachable with minWindows = 1)
        cols += n2 - w;
    }
    double acc = 0.0;
    const long long cells = (long long)h * w;
    for (long long i = threadIdx.x; i < cells; i += blockDim.x) {
        const int32_t r = (int32_t)(i / w), c = (int32_t)(i - (long long)r * w);
        acc += (double)A[(size_t)rows[r] * n + cols[c]];
    }
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += sh[k];
        out[blockIdx.x] = t;
    }
}

extern "C" int hc_block_sums4(const float *A, int32_t n, const int32_t *idx, const int64_t *job_off,
                              const int32_t *job_n1, const int32_t *job_n2, int32_t n_jobs, double *out,
                              void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(A && idx && job_off && job_n1 && job_n2 && out && n > 0 && n_jobs >= 0, "bad arguments");
    if (n_jobs == 0) return HC_OK;
    k_block_sums4<<<4 * n_jobs, 256, 0, stream>>>(A, n, idx, job_off, job_n1, job_n2, out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_normalize_rows(float *A, int32_t n, void *workspace, size_t workspace_bytes, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(A && workspace && n > 0, "bad arguments");
    size_t need = (size_t)n * 8 * 2 + (size_t)n * 4 + 1024;
    if (workspace_bytes < need) {
        hc::set_error("hc_normalize_rows: workspace too small");
        return HC_EWORKSPACE;
    }
    hc::Carver ws(workspace, workspace_bytes);
    double *colsum = ws.take<double>(n), *rowsum = ws.take<double>(n);
    float *rowmax = ws.take<float>(n), *maxv = ws.take<float>(1);
    HC_CUDA_OK(cudaMemsetAsync(workspace, 0, ws.off, stream));
    k_rowcol_sums<<<dim3((n + 255) / 256, (n + 63) / 64), 256, 0, stream>>>(A, n, colsum, rowsum, rowmax);
    k_colsum_max<<<1, 256, 0, stream>>>(colsum, n, maxv);
    k_scale_rows<<<dim3(min((n + 255) / 256, 64), n), 256, 0, stream>>>(A, n, rowsum, rowmax, maxv);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_contig_normalize(const uint32_t *counts, int32_t n, const int32_t *sizes, double *out,
                                   void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(counts && sizes && out && n > 0, "bad arguments");
    k_contig_normalize<<<dim3((n + 255) / 256, n), 256, 0, stream>>>(counts, n, sizes, out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_gather_square(const void *X, int dtype, int32_t n, const int32_t *keep, int32_t m, void *out,
                                void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && out && n > 0 && m >= 0 && (m == 0 || keep) && (dtype == 0 || dtype == 1), "bad arguments");
    if (m == 0) return HC_OK;
    const dim3 grid(min((m + 255) / 256, 64), m);
    if (dtype == 0) k_gather_square<float><<<grid, 256, 0, stream>>>((const float *)X, n, keep, m, (float *)out);
    else k_gather_square<double><<<grid, 256, 0, stream>>>((const double *)X, n, keep, m, (double *)out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_diag_normalize(const void *D, int dtype, int32_t n, double mean, void *out, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(D && out && D != out && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    const int T = (n + TS - 1) / TS;
    if (dtype == 0)
        k_diag_normalize<float><<<dim3(T, T), dim3(TS, 8), 0, stream>>>((const float *)D, n, (float)mean, (float *)out);
    else
        k_diag_normalize<double><<<dim3(T, T), dim3(TS, 8), 0, stream>>>((const double *)D, n, mean, (double *)out);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" int hc_window_size_normalize(void *D, int dtype, int32_t n, const int64_t *sizes, double ws2,
                                        void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(D && sizes && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    const dim3 grid(min((n + 255) / 256, 64), n);
    if (dtype == 0)
        k_window_size_normalize<float><<<grid, 256, 0, stream>>>((float *)D, n, (const long long *)sizes, ws2);
    else
        k_window_size_normalize<double><<<grid, 256, 0, stream>>>((double *)D, n, (const long long *)sizes, ws2);
    HC_CHECK_LAUNCH();
    return HC_OK;
}
