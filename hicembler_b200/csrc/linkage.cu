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

#include <algorithm>
#include <numeric>
#include <vector>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

enum { M_SINGLE = 0, M_COMPLETE = 1, M_AVERAGE = 2, M_WEIGHTED = 3, M_WARD = 4 };
constexpr int THREADS = 1024;
constexpr int MAX_CS = 16;

struct Slot {
    double v;
    int idx;
    int pad;
};

struct LinkArgs {
    double *W;
    int n;
    int method;
    int *chain_g;    // [cluster size][n] private chain copies (used when the chain does not fit smem)
    int *members_g;  // [cluster size][n]
    int *rec_a, *rec_b;
    double *rec_d;
    long long *stats;  // scans, sum active over scans, merges, sum active over updates
    int chain_in_smem;
};

__device__ __forceinline__ bool lex_less(double v, int i, double bv, int bi) {
    return v < bv || (v == bv && i < bi);
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

    for (int w = tid; w < n_words; w += THREADS) {
        int rem = n - (w << 5);
        active[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
    }
    for (int k = tid; k < n; k += THREADS) members[k] = 1;
    __syncthreads();

    // column slice of this CTA
    const int S = (n + cs - 1) / cs;
    const int lo = min(rank * S, n), hi = min(lo + S, n);

    int parity = 0;
    int start = 0, tip = 0, idx1 = 0, idx2 = 0, nactive = n;
    double mind = 0;
    long long n_scans = 0, scan_sum = 0, upd_sum = 0;

    // argmin over the active entries of row `row`, ties -> lowest index; then the
    // fastcluster rule: keep (prev, m0) unless something is strictly smaller
    auto argmin_row = [&](int row, double m0, int prev, double &m_out, int &i_out) {
        const double *Wr = W + (size_t)row * n;
        double bv = __longlong_as_double(0x7ff0000000000000LL);
        int bi = 0x7fffffff;
        for (int k = lo + tid; k < hi; k += THREADS) {
            if ((active[k >> 5] >> (k & 31)) & 1u) {
                double v = __ldcg(Wr + k);  // diagonal holds +inf
                if (v < bv) { bv = v; bi = k; }
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (lex_less(ov, oi, bv, bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { wslots[warp].v = bv; wslots[warp].idx = bi; }
        __syncthreads();
        if (warp == 0) {
            bv = wslots[lane].v;
            bi = wslots[lane].idx;
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                double ov = __shfl_xor_sync(0xffffffffu, bv, o);
                int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (lex_less(ov, oi, bv, bi)) { bv = ov; bi = oi; }
            }
            if (lane < cs) {  // publish this CTA's partial into every CTA's slot table
                Slot *remote = cluster.map_shared_rank(slots + parity * MAX_CS + rank, lane);
                remote->v = bv;
                remote->idx = bi;
            }
        }
        cluster.sync();
        bv = __longlong_as_double(0x7ff0000000000000LL);
        bi = 0x7fffffff;
        for (int q = 0; q < cs; ++q) {
            Slot s = slots[parity * MAX_CS + q];
            if (lex_less(s.v, s.idx, bv, bi)) { bv = s.v; bi = s.idx; }
        }
        parity ^= 1;
        if (bv < m0) { m_out = bv; i_out = bi; }
        else { m_out = m0; i_out = prev; }
        n_scans++;
        scan_sum += nactive;
    };

    for (int step = 0; step < n - 1; ++step) {
        if (tip <= 3) {
            idx1 = start;
            if (tid == 0) chain[0] = idx1;
            tip = 1;
            // the diagonal is +inf and idx1 itself is active: nothing is < +inf there
            argmin_row(idx1, __longlong_as_double(0x7ff0000000000000LL), -1, mind, idx2);
        } else {
            tip -= 3;
            idx1 = chain[tip - 1];
            idx2 = chain[tip];
            mind = __ldcg(W + (size_t)idx1 * n + idx2);
        }
        int prev2;
        do {
            if (tid == 0) chain[tip] = idx2;
            int found;
            argmin_row(idx2, mind, idx1, mind, found);
            idx1 = idx2;  // == chain[tip]
            idx2 = found;
            tip++;
            prev2 = chain[tip - 2];
        } while (idx2 != prev2);

        if (rank == 0 && tid == 0) {
            a.rec_a[step] = idx1;
            a.rec_b[step] = idx2;
            a.rec_d[step] = mind;
        }
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
        // Lance-Williams: new D(k, j) for every active k != j
        const double avg_s = __ddiv_rn(s1, __dadd_rn(s1, s2)), avg_t = __ddiv_rn(s2, __dadd_rn(s1, s2));
        const double *Wi = W + (size_t)i * n;
        double *Wj = W + (size_t)j * n;
        for (int k = lo + tid; k < hi; k += THREADS) {
            if (k != j && ((active[k >> 5] >> (k & 31)) & 1u)) {
                double da = __ldcg(Wi + k), db = __ldcg(Wj + k);
                double v = a.method == M_WARD ? (double)members[k] : 0.0;
                double nv = lw_update(a.method, da, db, mind, s1, s2, v, avg_s, avg_t);
                Wj[k] = nv;
                W[(size_t)k * n + j] = nv;
            }
        }
        cluster.sync();  // row/column j visible to the whole cluster before the next scan
    }
    if (rank == 0 && tid == 0) {
        a.stats[0] = n_scans;
        a.stats[1] = scan_sum;
        a.stats[2] = n - 1;
        a.stats[3] = upd_sum;
    }
}

size_t smem_bytes(int n, bool chain_in_smem) {
    size_t b = sizeof(Slot) * (2 * MAX_CS + 32) + (size_t)((n + 31) / 32) * 4;
    if (chain_in_smem) b += (size_t)n * 4;
    return b + 16;
}

struct UF {
    std::vector<int> parent;
    explicit UF(int n) : parent(n) { std::iota(parent.begin(), parent.end(), 0); }
    int find(int x) {
        int root = x;
        while (parent[root] != root) root = parent[root];
        while (parent[x] != root) { int nx = parent[x]; parent[x] = root; x = nx; }
        return root;
    }
};

}  // namespace

extern "C" size_t hc_linkage_workspace_bytes(int32_t n) {
    size_t b = 4096;
    b += hc::align_up((size_t)MAX_CS * n * 4, 256) * 2;  // chain + members copies
    b += hc::align_up((size_t)n * 4, 256) * 2 + hc::align_up((size_t)n * 8, 256);
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
    lb.stats = a.stats = ws.take<long long>(8);
    a.chain_g = ws.take<int>((size_t)MAX_CS * n);
    a.members_g = ws.take<int>((size_t)MAX_CS * n);
    lb.rec_a = a.rec_a = ws.take<int>(n);
    lb.rec_b = a.rec_b = ws.take<int>(n);
    lb.rec_d = a.rec_d = ws.take<double>(n);
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
        if (h_stats) h_stats[0] = h_stats[1] = h_stats[2] = h_stats[3] = 0;
        return HC_OK;
    }
    LinkArgs a;
    LinkBuffers lb;
    if (carve_linkage(workspace, workspace_bytes, n, a, lb)) return HC_EWORKSPACE;
    const int m = n - 1;
    std::vector<int> ra(m), rb(m);
    std::vector<double> rd(m);
    long long st[8];
    HC_CUDA_OK(cudaMemcpyAsync(ra.data(), lb.rec_a, m * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(rb.data(), lb.rec_b, m * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(rd.data(), lb.rec_d, m * sizeof(double), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(st, lb.stats, sizeof st, cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaStreamSynchronize(stream));
    if (method == M_WARD)
        for (int k = 0; k < m; ++k) rd[k] = sqrt(rd[k]);
    std::vector<int> order(m);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return rd[x] < rd[y]; });
    UF uf(2 * n - 1);
    std::vector<int> size(2 * n - 1, 1);
    int next = n;
    for (int row = 0; row < m; ++row) {
        int k = order[row];
        int x = uf.find(ra[k]), y = uf.find(rb[k]);
        uf.parent[x] = uf.parent[y] = next;
        size[next] = size[x] + size[y];
        h_Z[4 * row + 0] = (double)std::min(x, y);
        h_Z[4 * row + 1] = (double)std::max(x, y);
        h_Z[4 * row + 2] = rd[k];
        h_Z[4 * row + 3] = (double)size[next];
        next++;
    }
    if (h_stats)
        for (int k = 0; k < 4; ++k) h_stats[k] = st[k];
    return HC_OK;
}

extern "C" int hc_linkage(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes, double *h_Z,
                          int64_t *h_stats, void *stream) {
    int rc = hc_linkage_launch(W, n, method, workspace, workspace_bytes, stream);
    if (rc) return rc;
    return hc_linkage_finish(n, method, workspace, workspace_bytes, h_Z, h_stats, stream);
}
