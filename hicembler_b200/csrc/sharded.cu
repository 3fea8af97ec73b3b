// Row-sharded build of ONE contact matrix over the GPUs of an NVLink / NVSwitch node (north star:
// "the matrix build and Sinkhorn iterations are sharded by row block").  The reference's analogue is the split of
// the contigs over Pool workers with a reduction in the parent (bam2clusters.py:224-230, bam2plot.py:138-161).
//
// One process per GPU.  Rank k owns the row block [k*per, (k+1)*per) of every N x N object.  The exchange steps are
// not separate collectives: the kernels read (and, for the 400 KB column-sum vector, signal) PEER MEMORY directly
// over NVLink, so the transfer overlaps the arithmetic tile by tile:
//   hc_fold_block   S_k[i][j] = C_k[i][j] + C_owner(j)[j][i]  -- the transposed tile comes straight out of the HBM
//                   of the rank that owns rows j (the symmetrisation d += d.T - diag(d), bam2clusters.py:42,511)
//   hc_sk_fit_block Sinkhorn-Knopp on the row block (sinkhorn_knopp.py:114-163): the partial column sums of every
//                   rank are summed by a one-shot all-reduce fused with the column-chunk reduction before it and the
//                   c = 1/y update and stopping test after it; the whole fit is queued without host round trips
//                   (the loop condition lives on the device, kernels of finished iterations return at once)
//   transform       hc_transform_gather (matrix_ops.cu) builds the linkage working matrix on one rank from the row
//                   blocks of all ranks.
// Peer memory = cudaMalloc + cudaIpc handles (hc_shared_*): the handles travel through torch.distributed, the
// pointers arrive here as plain tables.  On ONE device the same kernels run on several streams with local pointers
// in the tables -- that is how the parity tests emulate 2..8 row blocks.
#include <algorithm>

#include "common.cuh"

namespace {

constexpr int MAXR = HC_MAX_RANKS;
constexpr int T64 = 64;
constexpr int RC = 256;  // rows per column-pass chunk (as in sinkhorn.cu)

__device__ __forceinline__ void atomic_max_nonneg_f(float *addr, float v) {
    atomicMax((int *)addr, __float_as_int(v));  // non-negative floats order like their bit patterns
}

struct CountBlocks {
    const uint32_t *base[MAXR];  // counts block of every rank: (rows of that rank) x n, full orientation
    int per;                     // rows per rank (multiple of 64), the last block may be shorter
};

// out[i - row0][j] = C[i][j] + (i != j ? C[j][i] : 0) + add for the rows of this rank; C[j][i] is read from the
// block of the rank that owns row j.  One CTA per 64 x 64 tile: the mirrored tile is staged through shared memory.
__global__ void __launch_bounds__(256) k_fold_block(CountBlocks cb, int rank, int row0, int nrows, int n, float add,
                                                    float *out, float *vmax) {
    __shared__ uint32_t B[T64][T64 + 1];
    __shared__ float shm[8];
    // tile row inside the block (fastest), tile column of the matrix: consecutive CTAs read ADJACENT column ranges of
    // the same 64 rows of the peer's block, so the NVLink reads stream whole rows (with J fastest every CTA touched 64
    // other remote rows, 256 bytes each: 34 GB/s at two GPUs)
    const int I = blockIdx.x, J = blockIdx.y;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int gi0 = row0 + I * T64, gj0 = J * T64;
    // mirrored tile: rows gj0.. (owned by rank gj0 / per), columns gi0..
    const int owner = gj0 / cb.per;
    const uint32_t *peer = cb.base[owner] + (size_t)(gj0 - owner * cb.per) * n;
    const bool vec = (n & 3) == 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;              // row of the mirrored tile = column j of the output
        const int j = gj0 + r, i = gi0 + tx * 4;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (j < n && i < n) {
            const uint32_t *src = peer + (size_t)r * n + i;
            // peer memory: a plain streaming load (ld.global.nc on an NVLink mapping measured 34 GB/s at two GPUs)
            if (vec) v = __ldcs((const uint4 *)src);
            else {
                v.x = src[0];
                if (i + 1 < n) v.y = src[1];
                if (i + 2 < n) v.z = src[2];
                if (i + 3 < n) v.w = src[3];
            }
        }
        B[r][tx * 4 + 0] = v.x; B[r][tx * 4 + 1] = v.y; B[r][tx * 4 + 2] = v.z; B[r][tx * 4 + 3] = v.w;
    }
    __syncthreads();
    const uint32_t *mine = cb.base[rank];
    float m = 0.f;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;
        const int li = I * T64 + r, i = row0 + li, j = gj0 + tx * 4;
        if (li < nrows && j < n) {
            const uint32_t *src = mine + (size_t)li * n + j;
            uint32_t a[4] = {0, 0, 0, 0};
            if (vec) {
                const uint4 v = hc::ld_stream_u4(src);
                a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w;
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (j + q < n) a[q] = src[q];
            }
            float o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint32_t t = (i == j + q) ? a[q] : a[q] + B[tx * 4 + q][r];
                o[q] = (float)t + add;           // exact below 2^24 per cell (the reference's float32 += 1)
                if (j + q < n) m = fmaxf(m, o[q]);
            }
            float *dst = out + (size_t)li * n + j;
            if (vec) *(float4 *)dst = make_float4(o[0], o[1], o[2], o[3]);
            else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (j + q < n) dst[q] = o[q];
            }
        }
    }
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) shm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) m = fmaxf(m, shm[k]);
        atomic_max_nonneg_f(vmax, m);
    }
}

// ---- Sinkhorn-Knopp on a row block --------------------------------------------------------------------------------
// Per-rank exchange region (peer visible, zeroed once when it is allocated):
//   double y[2][ylen]            two slots (epoch parity) of this rank's partial column sums + 8 flag words
//   int    flag[2][nblk][MAXR]   flag[s][b][r] = last epoch for which rank r has published block b in slot s
struct SkPeers {
    double *y[MAXR];
    int *flag[MAXR];
    int nranks, rank;
};

struct SkState {
    int cond, next_cond, cur, iterations, stop, outside_c;
    int gflags[8];   // all-reduced: 0 negative / NaN entry, 1 zero column sum, 2 zero row sum, 3 raw row sum outside 1 +- eps
    int lflags[8];   // this rank's own, published with the next exchange
};

__device__ __forceinline__ void st_release_sys(int *p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int *p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double ld_volatile_f64(const double *p) {
    double v;
    asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

// ypart[chunk][j] = sum over the rows of the chunk of (X[i][j]/vmax) * r[i]  (as k_colpass of sinkhorn.cu, with the
// device-side loop condition and r == NULL meaning ones)
template <typename T, bool CHECK>
__global__ void __launch_bounds__(256) kb_colpass(const T *X, int nrows, int n, T vmax, const double *r, double *ypart,
                                                  SkState *st, int guarded) {
    if (guarded && !st->cond) return;
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
        const T *p = X + (size_t)i * n + j0;
        if (vec) {
            if (sizeof(T) == 4) {
                const float4 t = __ldcs((const float4 *)p);
                v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
            } else {
                const double2 a = __ldcs((const double2 *)p), b = __ldcs((const double2 *)p + 1);
                v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = k < valid ? __ldcs(p + k) : (T)0;
        }
        const double ri = r ? __ldg(r + i) : 1.0;
        T q[4];
        if (scale) div.apply(v, q);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (CHECK) neg |= !(v[k] >= (T)0);
            acc[k] = fma((double)(scale ? q[k] : v[k]), ri, acc[k]);
        }
    }
    double *dst = ypart + (size_t)blockIdx.y * n + j0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (k < valid) dst[k] = acc[k];
    if (CHECK && neg) atomicOr(&st->lflags[0], 1);
}

// x[i] = sum_j (X[i][j]/vmax) * c[j], r[i] = 1/x[i]; RAW also evaluates the row sums of P itself (c = ones) in the
// same read of the matrix: the first `while` test of sinkhorn_knopp.py:138-141 on P_eps = copy(P)
template <typename T, bool RAW>
__global__ void __launch_bounds__(256) kb_rowpass(const T *X, int nrows, int n, T vmax, const double *c0, const double *c1,
                                                  double *r, double eps, SkState *st, int guarded) {
    if (guarded && !st->cond) return;
    constexpr int RB = 8;
    const double *c = st->cur ? c1 : c0;
    const int i0 = blockIdx.x * RB;
    const bool vec = (n % 4 == 0) && (((uintptr_t)X & 15) == 0);
    const bool scale = vmax != (T)1;
    const typename hc::ScaledDiv<T>::type div(vmax);
    double acc[RB], raw[RB];
#pragma unroll
    for (int k = 0; k < RB; ++k) acc[k] = raw[k] = 0;
    for (int j0 = threadIdx.x * 4; j0 < n; j0 += 1024) {
        const int valid = min(4, n - j0);
        double cj[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) cj[k] = k < valid ? __ldg(c + j0 + k) : 0.0;
        T v[RB][4];
#pragma unroll
        for (int rr = 0; rr < RB; ++rr) {
            v[rr][0] = v[rr][1] = v[rr][2] = v[rr][3] = (T)0;
            if (i0 + rr < nrows) {
                const T *p = X + (size_t)(i0 + rr) * n + j0;
                if (vec) {
                    if (sizeof(T) == 4) {
                        const float4 t = __ldcs((const float4 *)p);
                        v[rr][0] = t.x; v[rr][1] = t.y; v[rr][2] = t.z; v[rr][3] = t.w;
                    } else {
                        const double2 a = __ldcs((const double2 *)p), b = __ldcs((const double2 *)p + 1);
                        v[rr][0] = a.x; v[rr][1] = a.y; v[rr][2] = b.x; v[rr][3] = b.y;
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (k < valid) v[rr][k] = __ldcs(p + k);
                }
            }
        }
#pragma unroll
        for (int rr = 0; rr < RB; ++rr) {
            T q[4];
            if (scale) div.apply(v[rr], q);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double pd = (double)(scale ? q[k] : v[rr][k]);
                acc[rr] = fma(pd, cj[k], acc[rr]);
                if (RAW) raw[rr] += (k < valid) ? pd : 0.0;
            }
        }
    }
    __shared__ double sh[RB][8], shr[RB][8];
#pragma unroll
    for (int rr = 0; rr < RB; ++rr) {
        const double s = hc::warp_sum_f64(acc[rr]);
        if ((threadIdx.x & 31) == 0) sh[rr][threadIdx.x >> 5] = s;
        if (RAW) {
            const double t = hc::warp_sum_f64(raw[rr]);
            if ((threadIdx.x & 31) == 0) shr[rr][threadIdx.x >> 5] = t;
        }
    }
    __syncthreads();
    if (threadIdx.x < RB && i0 + threadIdx.x < nrows) {
        double s = 0, t = 0;
        for (int w = 0; w < 8; ++w) { s += sh[threadIdx.x][w]; if (RAW) t += shr[threadIdx.x][w]; }
        if (s == 0.0) atomicOr(&st->lflags[2], 1);
        r[i0 + threadIdx.x] = 1.0 / s;
        if (RAW && (t < 1.0 - eps || t > 1.0 + eps)) atomicOr(&st->lflags[3], 1);
    }
}

// Fused: column-chunk reduction of this rank -> publish -> wait for the same block of every peer -> sum the partial
// vectors of all ranks in rank order (every rank gets the same bits) -> c_alt = 1/y and the stopping test.
// mode 0: first column pass (r = ones): c = 1/y into the CURRENT buffer, column sums of P tested against 1 +- eps
// mode 1: first pass of the loop: c_alt = 1/y, no test (the loop condition is the raw test, applied by k_advance)
// mode 2: later passes: test c_cur * y (= column sums of D1.P.D2, sinkhorn_knopp.py:138-141), c_alt = 1/y
__global__ void __launch_bounds__(256) kb_exchange(const double *ypart, int nchunks, int n, SkPeers pe, int epoch, int nblk,
                                                   double *y, double *c0, double *c1, double eps, SkState *st, int mode) {
    if (mode != 0 && !st->cond) return;
    const int j = blockIdx.x * 256 + threadIdx.x, ylen = n + 8;
    const int slot = epoch & 1, R = pe.nranks;
    double s = 0;
    if (j < n) {
        for (int b = 0; b < nchunks; ++b) s += ypart[(size_t)b * n + j];
    } else if (j < ylen) {
        s = (double)st->lflags[j - n];
    }
    if (R > 1) {
        if (j < ylen) pe.y[pe.rank][(size_t)slot * ylen + j] = s;
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x < R)
            st_release_sys(pe.flag[threadIdx.x] + ((size_t)slot * nblk + blockIdx.x) * MAXR + pe.rank, epoch);
        if (threadIdx.x < R) {
            const int *f = pe.flag[pe.rank] + ((size_t)slot * nblk + blockIdx.x) * MAXR + threadIdx.x;
            while (ld_acquire_sys(f) < epoch) { }
        }
        __syncthreads();
        if (j < ylen) {
            s = 0;
            for (int q = 0; q < R; ++q) s += ld_volatile_f64(pe.y[q] + (size_t)slot * ylen + j);
        }
    }
    if (j >= ylen) return;
    if (j >= n) {                       // all-reduced flags: every rank raises / warns alike
        if (s != 0.0) atomicOr(&st->gflags[j - n], 1);
        return;
    }
    y[j] = s;
    double *ccur = st->cur ? c1 : c0, *calt = st->cur ? c0 : c1;
    if (mode == 0) {
        if (s == 0.0) atomicOr(&st->gflags[1], 1);
        if (s < 1.0 - eps || s > 1.0 + eps) atomicOr(&st->outside_c, 1);
        ccur[j] = 1.0 / s;
    } else {
        if (mode == 2) {
            const double t = ccur[j] * s;
            if (t < 1.0 - eps || t > 1.0 + eps) atomicOr(&st->next_cond, 1);
        }
        calt[j] = 1.0 / s;
    }
}

// after an exchange of the loop: take the loop decision, swap in the new c only when the loop goes on
__global__ void kb_advance(SkState *st, int mode) {
    if (!st->cond && mode != 1) return;
    if (mode == 1) {
        if (!st->cond) return;           // (cond was preset to 1 for the first pass; cleared below when P needs no work)
        st->cond = (st->outside_c | st->gflags[3]) ? 1 : 0;
    } else {
        st->cond = st->next_cond;
    }
    st->next_cond = 0;
    if (st->cond) st->cur ^= 1;
}

// after a row pass of the loop: count the iteration, stop at max_iter (sinkhorn_knopp.py:150-154)
__global__ void kb_post_row(SkState *st, int max_iter) {
    if (!st->cond) return;
    st->iterations += 1;
    if (st->iterations >= max_iter) { st->stop = 1; st->cond = 0; }
}

__global__ void kb_init_state(SkState *st, int iterations0) {
    SkState s;
    memset(&s, 0, sizeof s);
    s.iterations = iterations0;
    s.cond = 1;
    *st = s;
}

struct FitLayout {
    SkState *st;
    double *ypart, *y, *c0, *c1;
};

bool carve_fit(void *ws, size_t bytes, int nrows, int n, FitLayout &L) {
    hc::Carver c(ws, bytes);
    L.st = c.take<SkState>(1);
    L.ypart = c.take<double>((size_t)((nrows + RC - 1) / RC) * n);
    L.y = c.take<double>(n + 8);
    L.c0 = c.take<double>(n);
    L.c1 = c.take<double>(n);
    return c.ok();
}

}  // namespace

// Loads every kernel of this file into the context now.  With CUDA's lazy module loading the FIRST launch of a kernel
// may wait for running kernels to finish; a rank that enters the exchange kernel spins until its peers have launched
// theirs, so with several (emulated) ranks on one device that first launch must not happen inside the fit.
extern "C" int hc_sharded_preload(void) {
    cudaFuncAttributes at;
    HC_CUDA_OK(cudaFuncGetAttributes(&at, k_fold_block));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_colpass<float, true>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_colpass<float, false>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_colpass<double, true>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_colpass<double, false>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_rowpass<float, true>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_rowpass<float, false>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_rowpass<double, true>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_rowpass<double, false>));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_exchange));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_advance));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_post_row));
    HC_CUDA_OK(cudaFuncGetAttributes(&at, kb_init_state));
    return HC_OK;
}

extern "C" int hc_shared_alloc(size_t bytes, void **h_dptr, void *h_handle64) {
    HC_REQUIRE(h_dptr && h_handle64 && bytes > 0, "bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    void *p = nullptr;
    HC_CUDA_OK(cudaMalloc(&p, bytes));
    HC_CUDA_OK(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {   // no IPC on this system: the buffer still works for single-process use
        cudaGetLastError();
        memset(&h, 0, sizeof h);
    }
    memcpy(h_handle64, &h, 64);
    *h_dptr = p;
    return HC_OK;
}

extern "C" int hc_shared_free(void *dptr) {
    if (dptr) HC_CUDA_OK(cudaFree(dptr));
    return HC_OK;
}

extern "C" int hc_shared_open(const void *h_handle64, void **h_dptr) {
    HC_REQUIRE(h_handle64 && h_dptr, "bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, h_handle64, 64);
    HC_CUDA_OK(cudaIpcOpenMemHandle(h_dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return HC_OK;
}

extern "C" int hc_shared_close(void *dptr) {
    if (dptr) HC_CUDA_OK(cudaIpcCloseMemHandle(dptr));
    return HC_OK;
}

extern "C" int hc_fold_block(const uint32_t *const *h_counts_blocks, int32_t nranks, int32_t rank, int32_t per, int32_t n,
                             float add, float *out, float *d_vmax, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(h_counts_blocks && out && d_vmax && nranks >= 1 && nranks <= MAXR && rank >= 0 && rank < nranks, "bad arguments");
    HC_REQUIRE(per > 0 && per % T64 == 0 && (long long)per * (nranks - 1) < n && (long long)per * nranks >= n,
               "rows per rank must be a multiple of 64 and cover the matrix");
    CountBlocks cb;
    memset(&cb, 0, sizeof cb);
    for (int k = 0; k < nranks; ++k) {
        HC_REQUIRE(h_counts_blocks[k], "null block pointer");
        cb.base[k] = h_counts_blocks[k];
    }
    cb.per = per;
    const int row0 = rank * per, nrows = std::min(per, n - row0);
    HC_REQUIRE(((uintptr_t)out & 15) == 0, "out must be 16-byte aligned");
    dim3 grid((nrows + T64 - 1) / T64, (n + T64 - 1) / T64);
    k_fold_block<<<grid, 256, 0, stream>>>(cb, rank, row0, nrows, n, add, out, d_vmax);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

extern "C" size_t hc_sk_exchange_bytes(int32_t n) {
    const size_t ylen = (size_t)n + 8, nblk = (ylen + 255) / 256;
    return hc::align_up(2 * ylen * sizeof(double), 256) + 2 * nblk * MAXR * sizeof(int) + 256;
}

extern "C" size_t hc_sk_fit_block_workspace_bytes(int32_t nrows, int32_t n) {
    return 4096 + (size_t)((nrows + RC - 1) / RC + 4) * ((size_t)n + 8) * sizeof(double) + 5 * 256;
}

extern "C" int hc_sk_fit_block(const void *X, int dtype, int32_t nrows, int32_t n, double vmax, double eps, int32_t max_iter,
                               int32_t iterations0, void *const *h_exchange, int32_t nranks, int32_t rank,
                               int32_t *h_epoch, void *workspace, size_t workspace_bytes, double *r, double *c,
                               int32_t *h_result, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(X && r && c && h_result && h_epoch && workspace && nrows > 0 && n > 0 && (dtype == 0 || dtype == 1), "bad arguments");
    HC_REQUIRE(nranks >= 1 && nranks <= MAXR && rank >= 0 && rank < nranks && (nranks == 1 || h_exchange), "bad rank table");
    FitLayout L;
    if (workspace_bytes < hc_sk_fit_block_workspace_bytes(nrows, n) || !carve_fit(workspace, workspace_bytes, nrows, n, L)) {
        hc::set_error("hc_sk_fit_block: workspace too small");
        return HC_EWORKSPACE;
    }
    const int ylen = n + 8, nblk = (ylen + 255) / 256, nchunks = (nrows + RC - 1) / RC;
    SkPeers pe;
    memset(&pe, 0, sizeof pe);
    pe.nranks = nranks;
    pe.rank = rank;
    for (int k = 0; k < nranks && nranks > 1; ++k) {
        HC_REQUIRE(h_exchange[k], "null exchange pointer");
        pe.y[k] = (double *)h_exchange[k];
        pe.flag[k] = (int *)((char *)h_exchange[k] + hc::align_up(2 * (size_t)ylen * sizeof(double), 256));
    }
    int epoch = *h_epoch;
    const dim3 cgrid((n + 1023) / 1024, nchunks);
    const int rgrid = (nrows + 7) / 8;
    auto colpass = [&](const double *rv, bool check, int guarded) {
        if (dtype == 0) {
            if (check) kb_colpass<float, true><<<cgrid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, rv, L.ypart, L.st, guarded);
            else kb_colpass<float, false><<<cgrid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, rv, L.ypart, L.st, guarded);
        } else {
            if (check) kb_colpass<double, true><<<cgrid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, rv, L.ypart, L.st, guarded);
            else kb_colpass<double, false><<<cgrid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, rv, L.ypart, L.st, guarded);
        }
    };
    auto rowpass = [&](bool raw, int guarded) {
        if (dtype == 0) {
            if (raw) kb_rowpass<float, true><<<rgrid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, L.c0, L.c1, r, eps, L.st, guarded);
            else kb_rowpass<float, false><<<rgrid, 256, 0, stream>>>((const float *)X, nrows, n, (float)vmax, L.c0, L.c1, r, eps, L.st, guarded);
        } else {
            if (raw) kb_rowpass<double, true><<<rgrid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, L.c0, L.c1, r, eps, L.st, guarded);
            else kb_rowpass<double, false><<<rgrid, 256, 0, stream>>>((const double *)X, nrows, n, vmax, L.c0, L.c1, r, eps, L.st, guarded);
        }
    };
    auto exchange = [&](int mode) {
        ++epoch;
        kb_exchange<<<nblk, 256, 0, stream>>>(L.ypart, nchunks, n, pe, epoch, nblk, L.y, L.c0, L.c1, eps, L.st, mode);
    };
    long long launches = 0, passes = 0;
    kb_init_state<<<1, 1, 0, stream>>>(L.st, iterations0);
    // r = ones; c = 1 / P^T r (:120-129); the column sums are also the first half of the first `while` test (:138-141)
    colpass(nullptr, true, 0);
    exchange(0);
    // r = 1 / (P c) (:130-134); the raw row sums (second half of the first test) come out of the same read
    rowpass(true, 0);
    launches += 4;
    passes += 2;
    SkState h;
    int queued = 0, mode = 1;
    for (;;) {
        // one loop iteration: y = P^T r -> all-reduce -> test / c = 1/y -> r = 1/(P c) -> iterations += 1
        const int batch = 3;
        for (int b = 0; b < batch; ++b) {
            colpass(r, false, 1);
            exchange(mode);
            kb_advance<<<1, 1, 0, stream>>>(L.st, mode);
            rowpass(false, 1);
            kb_post_row<<<1, 1, 0, stream>>>(L.st, max_iter);
            mode = 2;
            launches += 5;
            ++queued;
        }
        HC_CHECK_LAUNCH();
        HC_CUDA_OK(cudaMemcpyAsync(&h, L.st, sizeof h, cudaMemcpyDeviceToHost, stream));
        HC_CUDA_OK(cudaStreamSynchronize(stream));
        if (!h.cond) break;
    }
    *h_epoch = epoch;
    HC_CUDA_OK(cudaMemcpyAsync(c, h.cur ? L.c1 : L.c0, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    // matrix passes actually executed: one column pass per iteration + the converging one, one row pass per iteration
    const int it_run = h.iterations - iterations0;
    passes += (h.outside_c | h.gflags[3]) ? (2 * it_run + (h.stop ? 0 : 1)) : 1;
    h_result[0] = h.iterations;
    h_result[1] = h.stop;                 // 0 = epsilon, 1 = max_iter
    h_result[2] = h.gflags[0];            // negative / NaN entry
    h_result[3] = h.gflags[1];            // zero column sum
    h_result[4] = h.gflags[2];            // zero row sum
    h_result[5] = (int32_t)passes;
    h_result[6] = (int32_t)launches;
    h_result[7] = queued;
    return HC_OK;
}
