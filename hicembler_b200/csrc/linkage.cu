// K7: agglomerative clustering -- nearest-neighbour chain with Lance-Williams row updates on the GPU.
//
// Replaces fastcluster.linkage(y, method) (reference bam2clusters.py:385,403,411,430; third-party
// NN_chain_core + generate_SciPy_dendrogram, restated in SURVEY.md appendix A and oracle/nnchain.c)
// and scipy.cluster.hierarchy.linkage (array2scaffolds.py:163).
//
// The chain is a serial dependency (every step needs the previous argmin), so the kernel is built
// for LATENCY, not bandwidth: one persistent thread-block cluster (up to 16 CTAs on one GPC) owns
// the whole run.  The working matrix is full symmetric float64 so every nearest-neighbour search is
// one contiguous row; each CTA scans a column slice with a lexicographic (value, lowest index)
// warp-shuffle argmin, partial results are exchanged through distributed shared memory and one
// hardware cluster barrier per step, and every CTA replays the (tiny) chain logic redundantly so no
// decision has to be broadcast.  A merge updates row j contiguously and column j by strided stores.
// Tie-breaking reproduces fastcluster's scan order exactly: the previous chain element wins ties,
// otherwise the lowest index; the merged cluster lives at the larger index; after a merge the chain
// is cut back by three.  All Lance-Williams arithmetic uses explicit round-to-nearest intrinsics
// (no FMA contraction) in fastcluster's operation order so heights are bit-identical.
#include <cooperative_groups.h>
#include <math.h>
#include <stdio.h>

#include <vector>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

enum { M_SINGLE = 0, M_COMPLETE = 1, M_AVERAGE = 2, M_WEIGHTED = 3, M_WARD = 4 };
constexpr int THREADS = 1024;
constexpr int MAX_CS = 16;

// (value, index) travel as an order-preserving 64-bit key + index so that the lexicographic argmin is
// three redux.sync.min.u32 per warp instead of fifteen shuffles
struct Slot {
    uint32_t hi, lo, idx, pad;
};

struct LinkArgs {
    double *W;
    int n;
    int method;
    int *chain_g;    // [cluster size][n] private chain copies (used when the chain does not fit smem)
    int *members_g;  // [cluster size][n]
    int *rec_a, *rec_b;
    double *rec_d;
    long long *stats;  // scans, sum active over scans, merges, sum active over updates, cycle counters
    int chain_in_smem;
    double *nnv;  // [n] nearest-neighbour cache (value)
    int *nni;     // [n] nearest-neighbour cache (index, < 0 = unknown)
    int use_cache;
};

constexpr unsigned long long KEY_NONE = 0xffffffffffffffffull;

// monotone map double -> u64 (-0.0 is folded onto +0.0 first: they compare equal as doubles)
__device__ __forceinline__ unsigned long long okey(double v) {
    unsigned long long b = (unsigned long long)__double_as_longlong(__dadd_rn(v, 0.0));
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double okey_value(unsigned long long k) {
    return __longlong_as_double((long long)((k >> 63) ? (k & 0x7fffffffffffffffull) : ~k));
}

// lexicographic (key, idx) minimum over the warp; every lane gets the result
__device__ __forceinline__ void warp_argmin(unsigned long long &key, uint32_t &idx) {
    const uint32_t hi = (uint32_t)(key >> 32), lo = (uint32_t)key;
    const uint32_t mh = __reduce_min_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
    const uint32_t mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == mlo) ? idx : 0xffffffffu);
    key = ((unsigned long long)mh << 32) | mlo;
    idx = mi;
}

__device__ __forceinline__ double lw_update(int method, double a, double b, double m, double s1, double s2,
                                            double v, double avg_s, double avg_t) {
    switch (method) {
    case M_SINGLE: return a < b ? a : b;
    case M_COMPLETE: return a > b ? a : b;
    case M_AVERAGE: return __dadd_rn(__dmul_rn(avg_s, a), __dmul_rn(avg_t, b));
    case M_WEIGHTED: return __dmul_rn(__dadd_rn(a, b), 0.5);
    default: {  // ward: ((v+s1)*a - v*m + (v+s2)*b) / (s1+s2+v)
        double t = __dsub_rn(__dmul_rn(__dadd_rn(v, s1), a), __dmul_rn(v, m));
        t = __dadd_rn(t, __dmul_rn(__dadd_rn(v, s2), b));
        return __ddiv_rn(t, __dadd_rn(__dadd_rn(s1, s2), v));
    }
    }
}

__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

constexpr int EPT = 4;  // elements per thread per batch: all loads of a batch are issued before first use

// "First linkage pass": nearest neighbour (value, lowest index) of every row of the fresh working
// matrix, all SMs, one streaming read of W.  Seeds the per-row NN cache of k_nnchain.
__global__ void __launch_bounds__(256) k_row_argmin(const double *W, int n, double *nnv, int *nni) {
    __shared__ Slot ws[8];
    const int row = blockIdx.x, tid = threadIdx.x;
    const double *Wr = W + (size_t)row * n;
    unsigned long long bk = KEY_NONE;
    uint32_t bi = 0xffffffffu;
    for (int base = tid; base < n; base += EPT * 256) {
        double v[EPT];
#pragma unroll
        for (int u = 0; u < EPT; ++u) {
            const int k = base + u * 256;
            v[u] = k < n ? __ldcs(Wr + k) : __longlong_as_double(-1LL);
        }
#pragma unroll
        for (int u = 0; u < EPT; ++u) {
            if (__double_as_longlong(v[u]) != -1LL) {
                const unsigned long long kk = okey(v[u]);
                if (kk < bk) { bk = kk; bi = (uint32_t)(base + u * 256); }
            }
        }
    }
    warp_argmin(bk, bi);
    if ((tid & 31) == 0) {
        ws[tid >> 5].hi = (uint32_t)(bk >> 32);
        ws[tid >> 5].lo = (uint32_t)bk;
        ws[tid >> 5].idx = bi;
    }
    __syncthreads();
    if (tid < 32) {
        const Slot s = ws[tid & 7];
        bk = ((unsigned long long)s.hi << 32) | s.lo;
        bi = s.idx;
        warp_argmin(bk, bi);
        if (tid == 0) {
            nnv[row] = okey_value(bk);
            nni[row] = (bi == 0xffffffffu || bk == KEY_NONE) ? -1 : (int)bi;
        }
    }
}

__global__ void __launch_bounds__(THREADS, 1) k_nnchain(LinkArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const int cs = cluster.num_blocks(), rank = cluster.block_rank();
    const int n = a.n, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Slot *slots = (Slot *)smem_raw;                      // [2][MAX_CS]
    Slot *wslots = slots + 2 * MAX_CS;                   // [32] per-warp partials
    uint32_t *active = (uint32_t *)(wslots + 32);        // bitmask
    const int n_words = (n + 31) >> 5;
    int *chain = a.chain_in_smem ? (int *)(active + n_words) : a.chain_g + (size_t)rank * n;
    int *members = a.members_g + (size_t)rank * n;
    double *W = a.W;
    double *nnv = a.nnv;  // per-row nearest-neighbour cache: exact lexicographic (value, index) minimum over
    int *nni = a.nni;     // the active entries of the row, or nni < 0 = unknown (rescan when visited)
    const bool use_cache = a.use_cache != 0;

    for (int w = tid; w < n_words; w += THREADS) {
        int rem = n - (w << 5);
        active[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
    }
    for (int k = tid; k < n; k += THREADS) members[k] = 1;
    if (tid < 2 * MAX_CS) {  // unused slots stay "none" so the combine can run over all MAX_CS
        slots[tid].hi = slots[tid].lo = slots[tid].idx = 0xffffffffu;
    }
    __syncthreads();
    cluster_barrier();  // nobody publishes into a slot table that is still being initialised

    // column slice of this CTA
    const int S = (n + cs - 1) / cs;
    const int lo = min(rank * S, n), hi = min(lo + S, n);

    int parity = 0;
    int start = 0, tip = 0, idx1 = 0, idx2 = 0, nactive = n;
    double mind = 0;
    long long n_scans = 0, scan_sum = 0, upd_sum = 0, n_cached = 0, look_sum = 0;
    long long t_scan = 0, t_scan_load = 0, t_scan_csync = 0, t_upd = 0, t_upd_csync = 0, t_upd_book = 0, t_total0 = clock64();
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    int last_j = -1, last_ji = -1;  // NN of the row rewritten by the latest merge (not yet visible in nnv/nni)
    double last_jv = 0;

    // lexicographic (key, index) minimum over the whole cluster; contains one cluster barrier with
    // release/acquire semantics, so it also orders the global stores issued before it
    auto cluster_argmin = [&](unsigned long long &bk, uint32_t &bi, long long &t_csync) {
        warp_argmin(bk, bi);
        if (lane == 0) {
            wslots[warp].hi = (uint32_t)(bk >> 32);
            wslots[warp].lo = (uint32_t)bk;
            wslots[warp].idx = bi;
        }
        __syncthreads();
        if (warp == 0) {
            const Slot s = wslots[lane];
            bk = ((unsigned long long)s.hi << 32) | s.lo;
            bi = s.idx;
            warp_argmin(bk, bi);
            if (lane < cs) {  // publish this CTA's partial into every CTA's slot table
                Slot *remote = cluster.map_shared_rank(slots + parity * MAX_CS + rank, lane);
                Slot o;
                o.hi = (uint32_t)(bk >> 32);
                o.lo = (uint32_t)bk;
                o.idx = bi;
                o.pad = 0;
                *(uint4 *)remote = *(uint4 *)&o;
            }
        }
        const long long c2 = clock64();
        __syncwarp();
        cluster_barrier();
        t_csync += clock64() - c2;
        {
            const Slot s = slots[parity * MAX_CS + (lane & (MAX_CS - 1))];
            bk = ((unsigned long long)s.hi << 32) | s.lo;
            bi = s.idx;
            warp_argmin(bk, bi);
        }
        parity ^= 1;
    };

    // nearest neighbour of `row`: (value, lowest index) over its active entries
    auto nn_of = [&](int row, double &v_out, int &i_out) {
        if (use_cache) {
            if (row == last_j) {
                if (last_ji >= 0) { v_out = last_jv; i_out = last_ji; n_cached++; look_sum += nactive; return; }
            } else {
                const int ci = __ldcg(nni + row);
                const double cv = __ldcg(nnv + row);
                if (ci >= 0) { v_out = cv; i_out = ci; n_cached++; look_sum += nactive; return; }
            }
        }
        const double *Wr = W + (size_t)row * n;
        unsigned long long bk = KEY_NONE;
        uint32_t bi = 0xffffffffu;
        const long long c0 = clock64();
        for (int base = lo + tid; base < hi; base += EPT * THREADS) {
            double v[EPT];
#pragma unroll
            for (int u = 0; u < EPT; ++u) {
                const int k = base + u * THREADS;
                const bool on = k < hi && ((active[k >> 5] >> (k & 31)) & 1u);
                // diagonal holds +inf; the all-ones NaN pattern marks "skip" (inputs are NaN free)
                v[u] = on ? __ldcg(Wr + k) : __longlong_as_double(-1LL);
            }
#pragma unroll
            for (int u = 0; u < EPT; ++u) {
                if (__double_as_longlong(v[u]) != -1LL) {
                    const unsigned long long kk = okey(v[u]);
                    if (kk < bk) { bk = kk; bi = (uint32_t)(base + u * THREADS); }
                }
            }
        }
        const long long c1 = clock64();
        cluster_argmin(bk, bi, t_scan_csync);
        v_out = okey_value(bk);
        i_out = (int)bi;
        if (use_cache && rank == 0 && tid == 0) {  // read again only after a later cluster barrier
            __stcg(nnv + row, v_out);
            __stcg(nni + row, i_out);
        }
        n_scans++;
        scan_sum += nactive;
        t_scan_load += c1 - c0;
        t_scan += clock64() - c0;
    };

    for (int step = 0; step < n - 1; ++step) {
        if (tip <= 3) {
            idx1 = start;
            if (tid == 0) chain[0] = idx1;
            tip = 1;
            nn_of(idx1, mind, idx2);  // idx1 is the lowest active index: plain row minimum, lowest index on ties
        } else {
            tip -= 3;
            idx1 = chain[tip - 1];
            idx2 = chain[tip];
            mind = __ldcg(W + (size_t)idx1 * n + idx2);
        }
        int prev2;
        do {
            if (tid == 0) chain[tip] = idx2;
            double v;
            int found;
            nn_of(idx2, v, found);
            // fastcluster's scan starts from (previous element, its distance) and replaces only on a
            // strictly smaller value: the previous element wins ties, else the lowest index
            if (v < mind) mind = v;
            else found = idx1;
            idx1 = idx2;  // == chain[tip]
            idx2 = found;
            tip++;
            __syncthreads();  // chain[tip - 1] (written by thread 0 above) is visible
            prev2 = chain[tip - 2];
        } while (idx2 != prev2);

        if (rank == 0 && tid == 0) {
            a.rec_a[step] = idx1;
            a.rec_b[step] = idx2;
            a.rec_d[step] = mind;
        }
        const long long u0 = clock64();
        const int i = min(idx1, idx2), j = max(idx1, idx2);
        const double s1 = (double)members[i], s2 = (double)members[j];
        __syncthreads();  // everyone has read members[i], members[j]
        if (tid == 0) {
            members[j] = (int)(s1 + s2);
            active[i >> 5] &= ~(1u << (i & 31));
        }
        if (i == start) {  // advance to the next active index (bit i is being cleared)
            int k = i + 1;
            while (k < n) {
                uint32_t w = active[k >> 5] & (0xffffffffu << (k & 31));
                if ((k >> 5) == (i >> 5)) w &= ~(1u << (i & 31));
                if (w) { k = (k & ~31) + __ffs(w) - 1; break; }
                k = (k & ~31) + 32;
            }
            start = k;
        }
        nactive--;
        upd_sum += nactive;
        __syncthreads();
        t_upd_book += clock64() - u0;
        // Lance-Williams: new D(k, j) for every active k != j; loads of a batch first, then the stores.
        // The same pass maintains the NN cache of every row k (column i disappears, column j changes)
        // and finds the NN of the rewritten row j.
        const double avg_s = __ddiv_rn(s1, __dadd_rn(s1, s2)), avg_t = __ddiv_rn(s2, __dadd_rn(s1, s2));
        const double *Wi = W + (size_t)i * n;
        double *Wj = W + (size_t)j * n;
        const int method = a.method;
        unsigned long long jk = KEY_NONE;
        uint32_t ji = 0xffffffffu;
        for (int base = lo + tid; base < hi; base += EPT * THREADS) {
            double da[EPT], db[EPT], mv[EPT], cv[EPT];
            int ci[EPT];
            bool on[EPT];
#pragma unroll
            for (int u = 0; u < EPT; ++u) {
                const int k = base + u * THREADS;
                on[u] = k < hi && k != j && ((active[k >> 5] >> (k & 31)) & 1u);
                da[u] = db[u] = mv[u] = cv[u] = 0.0;
                ci[u] = -1;
                if (on[u]) {
                    da[u] = __ldcg(Wi + k);
                    db[u] = __ldcg(Wj + k);
                    if (method == M_WARD) mv[u] = (double)members[k];
                    if (use_cache) {
                        ci[u] = __ldcg(nni + k);
                        cv[u] = __ldcg(nnv + k);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < EPT; ++u) {
                if (on[u]) {
                    const int k = base + u * THREADS;
                    const double nv = lw_update(method, da[u], db[u], mind, s1, s2, mv[u], avg_s, avg_t);
                    __stcg(Wj + k, nv);
                    __stcg(W + (size_t)k * n + j, nv);
                    if (use_cache) {
                        const unsigned long long kk = okey(nv);
                        if (kk < jk) { jk = kk; ji = (uint32_t)k; }
                        if (ci[u] >= 0) {
                            const bool hit = ci[u] == i || ci[u] == j;
                            bool take;  // (nv, j) becomes the row's lexicographic minimum
                            if (hit) {
                                take = nv < cv[u] || (ci[u] == j && nv == cv[u]);
                                if (!take) __stcg(nni + k, -1);  // its neighbour moved away: rescan when visited
                            } else {
                                take = nv < cv[u] || (nv == cv[u] && j < ci[u]);
                            }
                            if (take) {
                                __stcg(nnv + k, nv);
                                __stcg(nni + k, j);
                            }
                        }
                    }
                }
            }
        }
        // NN of the rewritten row j (also the barrier that makes row/column j and the cache updates
        // visible to the whole cluster before the next scan)
        cluster_argmin(jk, ji, t_upd_csync);
        if (use_cache) {
            last_j = j;
            last_ji = (jk == KEY_NONE) ? -1 : (int)ji;
            last_jv = okey_value(jk);
            if (rank == 0 && tid == 0) {
                __stcg(nnv + j, last_jv);
                __stcg(nni + j, last_ji);
            }
        }
        t_upd += clock64() - u0;
    }
    if (rank == 0 && tid == 0) {
        a.stats[0] = n_scans + n_cached;   // nearest-neighbour searches of the NN-chain (what the CPU library scans)
        a.stats[1] = scan_sum + look_sum;  // sum of active sizes over them
        a.stats[12] = n_scans;             // row scans actually executed (the rest came from the NN cache)
        a.stats[13] = scan_sum;
        a.stats[2] = n - 1;
        a.stats[3] = upd_sum;
        a.stats[4] = t_scan;
        a.stats[5] = t_scan_load;
        a.stats[6] = t_scan_csync;
        a.stats[7] = t_upd;
        a.stats[8] = t_upd_csync;
        a.stats[9] = clock64() - t_total0;
        a.stats[10] = t_upd_book;
        a.stats[11] = n_cached;
    }
    cluster_barrier();  // no CTA exits while a peer may still write into its shared memory
}

size_t smem_bytes(int n, bool chain_in_smem) {
    size_t b = sizeof(Slot) * (2 * MAX_CS + 32) + (size_t)((n + 31) / 32) * 4;
    if (chain_in_smem) b += (size_t)n * 4;
    return b + 16;
}

}  // namespace

extern "C" size_t hc_linkage_workspace_bytes(int32_t n) {
    size_t b = 4096;
    b += hc::align_up((size_t)MAX_CS * n * 4, 256) * 2;  // chain + members copies
    b += hc::align_up((size_t)n * 4, 256) * 2 + hc::align_up((size_t)n * 8, 256);
    b += hc::align_up((size_t)n * 8, 256) + hc::align_up((size_t)n * 4, 256);  // NN cache
    return b;
}

namespace {
struct LinkBuffers {
    int *rec_a, *rec_b;
    double *rec_d;
    long long *stats;
};

int carve_linkage(void *workspace, size_t bytes, int n, LinkArgs &a, LinkBuffers &lb) {
    hc::Carver ws(workspace, bytes);
    lb.stats = a.stats = ws.take<long long>(16);
    a.chain_g = ws.take<int>((size_t)MAX_CS * n);
    a.members_g = ws.take<int>((size_t)MAX_CS * n);
    lb.rec_a = a.rec_a = ws.take<int>(n);
    lb.rec_b = a.rec_b = ws.take<int>(n);
    lb.rec_d = a.rec_d = ws.take<double>(n);
    a.nnv = ws.take<double>(n);
    a.nni = ws.take<int>(n);
    return ws.ok() ? 0 : -1;
}
}  // namespace

// launch only (asynchronous); several linkages may run concurrently on different streams
extern "C" int hc_linkage_launch(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes,
                                 void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(W && workspace && n >= 1 && method >= 0 && method <= 4, "bad arguments");
    if (workspace_bytes < hc_linkage_workspace_bytes(n)) {
        hc::set_error("hc_linkage: workspace too small");
        return HC_EWORKSPACE;
    }
    if (n < 2) return HC_OK;
    LinkArgs a;
    LinkBuffers lb;
    memset(&a, 0, sizeof a);
    if (carve_linkage(workspace, workspace_bytes, n, a, lb)) return HC_EWORKSPACE;
    a.W = W;
    a.n = n;
    a.method = method;
    int dev = 0, smem_max = 0;
    HC_CUDA_OK(cudaGetDevice(&dev));
    HC_CUDA_OK(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    a.chain_in_smem = smem_bytes(n, true) <= (size_t)smem_max;
    size_t smem = smem_bytes(n, a.chain_in_smem);
    HC_CUDA_OK(cudaFuncSetAttribute(k_nnchain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    HC_CUDA_OK(cudaFuncSetAttribute(k_nnchain, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));

    int want = MAX_CS;
    if (const char *e = getenv("HC_LINKAGE_CLUSTER")) want = atoi(e);
    if (n < 4096 && want > 8) want = 8;  // small problems: fewer CTAs, cheaper barrier
    if (n < 512 && want > 2) want = 2;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int cs = want;
    for (; cs >= 1; cs >>= 1) {
        attr[0].val.clusterDim.x = cs;
        attr[0].val.clusterDim.y = attr[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3(cs);
        int nclusters = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&nclusters, k_nnchain, &cfg);
        if (e == cudaSuccess && nclusters >= 1) break;
        (void)cudaGetLastError();
    }
    HC_REQUIRE(cs >= 1, "no cluster size fits");
    // measured on B200 (profiles/r01_linkage_experiments.md): the NN cache removes ~1/3 of the row scans but
    // its maintenance lengthens every Lance-Williams pass; it wins while a row slice is <= 1 element per
    // thread (n <= 16 CTAs x 1024 threads) and loses beyond
    a.use_cache = n <= MAX_CS * THREADS;
    if (const char *e = getenv("HC_LINKAGE_NNCACHE")) a.use_cache = atoi(e) != 0;
    if (a.use_cache) {  // first linkage pass: nearest neighbour of every row, all SMs
        k_row_argmin<<<n, 256, 0, stream>>>(W, n, a.nnv, a.nni);
        HC_CHECK_LAUNCH();
    }
    HC_CUDA_OK(cudaLaunchKernelEx(&cfg, k_nnchain, a));
    return HC_OK;
}

// wait for the chain, then build the scipy linkage array on the host: sqrt for ward, stable sort by
// height, union-find labels (generate_SciPy_dendrogram; O(n log n) on n-1 records)
extern "C" int hc_linkage_finish(int32_t n, int method, void *workspace, size_t workspace_bytes, double *h_Z,
                                 int64_t *h_stats, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(workspace && h_Z && n >= 1, "bad arguments");
    if (n < 2) {
        if (h_stats) memset(h_stats, 0, 8 * sizeof(int64_t));
        return HC_OK;
    }
    LinkArgs a;
    LinkBuffers lb;
    if (carve_linkage(workspace, workspace_bytes, n, a, lb)) return HC_EWORKSPACE;
    const int m = n - 1;
    std::vector<int> ra(m), rb(m);
    std::vector<double> rd(m);
    long long st[16];
    HC_CUDA_OK(cudaMemcpyAsync(ra.data(), lb.rec_a, m * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(rb.data(), lb.rec_b, m * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(rd.data(), lb.rec_d, m * sizeof(double), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(st, lb.stats, sizeof st, cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaStreamSynchronize(stream));
    hc::records_to_Z(n, method == M_WARD, ra.data(), rb.data(), rd.data(), h_Z, 0.0, nullptr);
    if (h_stats) {
        for (int k = 0; k < 4; ++k) h_stats[k] = st[k];
        h_stats[4] = st[12];
        h_stats[5] = st[13];
        h_stats[6] = st[11];
        h_stats[7] = 0;
    }
    if (getenv("HC_LINKAGE_PROFILE")) {  // SM-clock breakdown seen by thread 0 of cluster rank 0
        double tot = (double)st[9];
        fprintf(stderr, "[hc_linkage n=%d] cached NN lookups %lld | cycles total %.3g | row scans %lld: %.1f%% (load+block-reduce %.1f%%, cluster barrier %.1f%%) | "
                "updates %lld: %.1f%% (bookkeeping %.1f%%, cluster barrier %.1f%%) | cyc/scan %.0f cyc/update %.0f\n", n, st[11], tot, st[12],
                100.0 * st[4] / tot, 100.0 * st[5] / tot, 100.0 * st[6] / tot, st[2], 100.0 * st[7] / tot,
                100.0 * st[10] / tot, 100.0 * st[8] / tot, (double)st[4] / (double)(st[12] ? st[12] : 1), (double)st[7] / (double)st[2]);
    }
    return HC_OK;
}

extern "C" int hc_linkage(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes, double *h_Z,
                          int64_t *h_stats, void *stream) {
    int rc = hc_linkage_launch(W, n, method, workspace, workspace_bytes, stream);
    if (rc) return rc;
    return hc_linkage_finish(n, method, workspace, workspace_bytes, h_Z, h_stats, stream);
}
