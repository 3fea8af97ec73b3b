// K7b: round-parallel agglomeration -- every reciprocal-nearest-neighbour pair is merged in the same
// round, on all SMs.  Second engine behind fastcluster.linkage / scipy linkage (reference
// bam2clusters.py:385,403,411,430, array2scaffolds.py:163), next to the serial NN-chain of linkage.cu.
//
// Why it is allowed.  single/complete/average/weighted/ward are REDUCIBLE Lance-Williams schemes
// (D(k, i+j) >= min(D(k,i), D(k,j))): merging a reciprocal-NN pair never destroys another reciprocal
// pair, so the NN-chain -- which also only ever merges reciprocal pairs -- and "merge them all at
// once" build the same tree whenever every decision is strict.  Where a decision is NOT strict (ties),
// fastcluster's answer depends on its scan order; this engine then refuses: it CERTIFIES each merge
// (both rows' minimum is separated from the second smallest entry by a relative margin far above
// the measured rounding difference between the engines, MARGIN) and the sorted heights (records_to_Z near_ties), and returns
// HC_EAMBIGUOUS otherwise, so the caller rebuilds the working matrix and runs the exact serial engine.
// Heights agree with the serial order to rounding (the Lance-Williams operations are the same, their
// interleaving differs); merge order and cluster sizes are identical.
//
// Why it is fast.  The serial chain is ~200k dependent steps of 2-5 us each at n = 50k (latency).
// Here a round is three launches over all SMs and the whole run is bandwidth work:
//   kr_select : row a and b = nn[a] are reciprocal -> pair list, certificate, merge record
//   kr_light  : every surviving unmerged row k updates ITS OWN entries  D(k, j_q) for all pairs q
//               (the matrix is full symmetric, so row k already holds D(k,i_q) and D(k,j_q): no
//               transposed traffic, no inter-row hazard) and keeps its NN cache (min, argmin,
//               lower bound of the rest) valid without a rescan whenever reducibility allows
//   kr_heavy  : new row of every merged cluster from its two old rows (pair x pair entries by a
//               fixed "lower j first" rule so both mirror entries get bit-identical values), exact
//               rescans of rows whose cache could not be kept; the last CTA retires the pairs.
// Round 2 structure of a run (hc_linkage_rounds):
//   initial scan (kr_light + kr_heavy, every row's candidate list)
//   bulk rounds, while a fifth of the columns merge at once: kr_select, then ONE pass that merges and compacts
//     (kb_rebuild3: old rows and per-chunk column lists staged by the TMA unit; kb_rebuild2 / kb_rebuild behind switches)
//   the rest inside kr_chain, a persistent cooperative kernel (select / build + light / rescan behind grid barriers); thin
//     rounds (1-4 pairs) take latency-optimised paths (shared-memory staged builds and rescans, eight lanes per light row)
//   the host comes back for a compaction (order preserving) whenever a quarter of the stored columns are dead.
#include <limits.h>
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"

namespace {

enum { M_SINGLE = 0, M_COMPLETE = 1, M_AVERAGE = 2, M_WEIGHTED = 3, M_WARD = 4 };
// Certificate thresholds.  The two engines differ only by the interleaving of the same Lance-Williams
// operations; the largest relative difference between their heights measured on the bench workloads is
// 3.2e-15 at n = 50 000 (ward), 4.4e-16 at n = 10 000 (average): profiles/r01_engine_discrepancy.md.
constexpr double MARGIN = 1e-13;      // relative separation required of every decision (30 x the measured maximum)
constexpr double HEIGHT_TOL = 1e-13;  // relative separation required of adjacent sorted heights (30 x)
// (HC_ROUNDS_MARGIN / HC_ROUNDS_HEIGHT_TOL override both for measurements)
constexpr int ROLE_DEAD = -2, ROLE_FREE = -1;  // >= 0: 2*pair (i side, retires) / 2*pair+1 (j side, survives)
constexpr int LIGHT_THREADS = 256, HEAVY_THREADS = 256, MAX_SEGS = 32;
constexpr int HEAVY_GRID_MAX = 2048;

struct __align__(16) Pair {  // 48 bytes; i < j, the merged cluster lives at j
    int i, j;
    double m, s1, s2, as, at;
};

// per-round counters; the persistent kernel alternates between the two sets (round parity) so that the set of the
// NEXT round can be cleared while the current one is still being read
struct Ctr {
    int P, qlen, deferred, pad;
};

struct RState {
    Ctr ctr[2];
    int n_done, nactive, ambiguous, stuck, done, round, blocks_done, last_P, last_q, exit_reason;
    unsigned int bar;  // arrival counter of the persistent kernel's grid barrier (monotone)
    unsigned long long dbgmax[4];     // latest finishing times (globaltimer) of build / light / rescan items over all warps
    unsigned long long phase_ns[16];  // persistent kernel, CTA 0: time spent in select / build+light / rescan+retire (incl. barrier)
    int amb_r, amb_b, amb_sr, amb_sb;
    double amb_v, amb_secr, amb_secb, amb_vb;
    double gaps[16];  // relative gap (second minimum / minimum - 1) of the first near-tie observations
    unsigned long long stats[10];  // 0 rescans, 1 rescan columns, 2 builds, 3 build columns, 4 light updates, 5 deferred
};

// Per-row neighbour cache: the KC nearest live columns in ascending (value, index) order and, in slot KC
// of cval, a lower bound of every live entry that is NOT listed.  A row is rescanned only when its list
// runs empty, so a row that follows a growing cluster through KC merges costs one scan, not KC.
constexpr int LC = 8, KC = LC - 1;  // LC = entries a scan keeps (power of two: bitonic merges), KC candidates + bound

struct Cols {
    int *role, *size, *orig;
    int *cidx;     // [ncols][LC]: candidate columns, -1 = none (slot KC unused)
    double *cval;  // [ncols][LC]: candidate values; slot KC = lower bound of the unlisted entries
    unsigned char *fresh, *need;  // fresh: the bound is the exact (count+1)-th smallest entry (set by a scan)
};

struct __align__(16) Part {  // partial candidate list of one column segment (values as order-preserving keys)
    unsigned long long k[LC];
    int i[LC];
    unsigned long long bound, pad;
};

struct RArgs {
    Part *parts;     // [entries][max_segs]
    int *seg_done;   // [entries] arrival counters, zero between rounds
    int max_segs;
    double *W;
    long long stride;
    int ncols, n, method, diag, bulk_segs, light8;
    double margin;
    Cols c;
    Pair *pairs;
    int *rq, *rec_a, *rec_b;
    double *rec_d;
    RState *st;
    unsigned int *bar;   // arrival counter of the persistent kernel's grid barrier, alone in its 256-byte line
};

__device__ __forceinline__ double lw(int method, double a, double b, const Pair &p, double v) {
    switch (method) {
    case M_SINGLE: return a < b ? a : b;
    case M_COMPLETE: return a > b ? a : b;
    case M_AVERAGE: return __dadd_rn(__dmul_rn(p.as, a), __dmul_rn(p.at, b));
    case M_WEIGHTED: return __dmul_rn(__dadd_rn(a, b), 0.5);
    default: {  // ward on squared distances: ((v+s1)*a - v*m + (v+s2)*b) / (s1+s2+v), fastcluster's order
        double t = __dsub_rn(__dmul_rn(__dadd_rn(v, p.s1), a), __dmul_rn(v, p.m));
        t = __dadd_rn(t, __dmul_rn(__dadd_rn(v, p.s2), b));
        return __ddiv_rn(t, __dadd_rn(__dadd_rn(p.s1, p.s2), v));
    }
    }
}

// Everything another CTA may have written earlier in the same launch (the persistent kernel runs many rounds in one
// launch) is read through L2: ld.global.cg.  In the one-kernel-per-phase launches these are ordinary L2 hits.
__device__ __forceinline__ Pair ld_pair(const Pair *p) {
    const int4 *q = (const int4 *)p;   // 48 bytes, 16-byte aligned: {i, j, m} {s1, s2} {as, at}
    const int4 a = __ldcg(q), b = __ldcg(q + 1), c = __ldcg(q + 2);
    Pair r;
    r.i = a.x; r.j = a.y;
    r.m = __hiloint2double(a.w, a.z);
    r.s1 = __hiloint2double(b.y, b.x); r.s2 = __hiloint2double(b.w, b.z);
    r.as = __hiloint2double(c.y, c.x); r.at = __hiloint2double(c.w, c.z);
    return r;
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// one instruction asks the TMA unit to bring a whole row segment into L2 (cp.async.bulk.prefetch): the register loop
// that follows then runs on L2 hits instead of paying one DRAM (+TLB) round trip per dependent iteration
__device__ __forceinline__ void bulk_prefetch_l2(const void *p, size_t bytes) {
    const size_t a = ((size_t)p + 15) & ~(size_t)15;
    const size_t e = ((size_t)p + bytes) & ~(size_t)15;
    if (e > a) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a), "r"((unsigned)(e - a)) : "memory");
}

// ---- TMA (cp.async.bulk) + mbarrier helpers ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
// (the suspend-time hint lets the hardware park the warp instead of spinning through the issue slots of the working warps)
__device__ __forceinline__ void mbar_wait(unsigned long long *b, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "W_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@!p bra W_%=;\n\t}"
        ::"r"(smem_u32(b)), "r"(parity), "r"(20000u) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool lex_lt(double x, int c, double y, int d) { return x < y || (x == y && c < d); }

// running LC smallest entries of a row in ascending (value, index) order, all in registers (static indexing only)
struct TopL {
    double v[LC];
    int i[LC];
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int s = 0; s < LC; ++s) { v[s] = __longlong_as_double(0x7ff0000000000000LL); i[s] = INT_MAX; }
    }
    __device__ __forceinline__ void insert(double x, int c) {  // sorted insert, drops the largest
#pragma unroll
        for (int s = LC - 1; s > 0; --s) {
            const bool shift = lex_lt(x, c, v[s - 1], i[s - 1]), here = lex_lt(x, c, v[s], i[s]);
            v[s] = shift ? v[s - 1] : (here ? x : v[s]);
            i[s] = shift ? i[s - 1] : (here ? c : i[s]);
        }
        if (lex_lt(x, c, v[0], i[0])) { v[0] = x; i[0] = c; }
    }
    __device__ __forceinline__ void push(double x, int c) {  // c ascending per thread
        if (x < v[LC - 1]) insert(x, c);
    }
    // this <- the LC smallest of (this U w); w ascending.  min(a[s], w[LC-1-s]) keeps exactly the LC smallest as a
    // bitonic sequence, log2(LC) half-cleaner stages sort it.
    __device__ __forceinline__ void merge_sorted(const double (&w)[LC], const int (&j)[LC]) {
#pragma unroll
        for (int s = 0; s < LC; ++s)
            if (lex_lt(w[LC - 1 - s], j[LC - 1 - s], v[s], i[s])) { v[s] = w[LC - 1 - s]; i[s] = j[LC - 1 - s]; }
#pragma unroll
        for (int d = LC / 2; d > 0; d >>= 1) {
#pragma unroll
            for (int s = 0; s < LC; ++s) {
                if ((s & d) == 0) {
                    const int t = (s + d) & (LC - 1);
                    if (lex_lt(v[t], i[t], v[s], i[s])) {
                        const double tv = v[s]; v[s] = v[t]; v[t] = tv;
                        const int ti = i[s]; i[s] = i[t]; i[t] = ti;
                    }
                }
            }
        }
    }
    __device__ __forceinline__ void xmerge(int o) {  // butterfly step: both lanes end with the same merged list
        double w[LC];
        int j[LC];
#pragma unroll
        for (int s = 0; s < LC; ++s) {
            w[s] = __shfl_xor_sync(0xffffffffu, v[s], o);
            j[s] = __shfl_xor_sync(0xffffffffu, i[s], o);
        }
        merge_sorted(w, j);
    }
    __device__ __forceinline__ void warp_reduce() {
#pragma unroll
        for (int o = 16; o; o >>= 1) xmerge(o);
    }
};

// Per-thread state of a row scan: the two smallest entries with their columns and the value of the third.
// Candidates of a row = the two smallest of every thread; everything else is >= min over threads of m3.
struct Min3 {
    double m1, m2, m3;
    int i1, i2;
    __device__ __forceinline__ void init() {
        m1 = m2 = m3 = __longlong_as_double(0x7ff0000000000000LL);
        i1 = i2 = INT_MAX;
    }
    __device__ __forceinline__ void push(double x, int c) {  // c ascending per thread
        if (x < m3) {
            if (x < m2) {
                m3 = m2;
                if (x < m1) { m2 = m1; i2 = i1; m1 = x; i1 = c; }
                else { m2 = x; i2 = c; }
            } else {
                m3 = x;
            }
        }
    }
};

__global__ void kr_init(RArgs a) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < a.n) {
        a.c.role[c] = ROLE_FREE;
        a.c.size[c] = 1;
        a.c.orig[c] = c;
        for (int k = 0; k < LC; ++k) {
            a.c.cidx[(size_t)c * LC + k] = -1;
            a.c.cval[(size_t)c * LC + k] = __longlong_as_double(0x7ff0000000000000LL);
        }
        a.c.fresh[c] = 0;
        a.c.need[c] = 1;  // the first kr_light / kr_heavy pass is the initial NN scan of every row
    }
    if (c == 0) {
        RState s;
        memset(&s, 0, sizeof s);
        s.nactive = a.n;
        *a.st = s;
    }
}

// ---- select: reciprocal nearest neighbours -> certified pair list ----------------------------
// rows r = r_first, r_first + r_stride, ...: whole warps stay together (r_first differs by lane only)
__device__ __forceinline__ void select_rows(const RArgs &a, Ctr *ct, int r_first, int r_stride) {
    RState *st = a.st;
    const int lane = threadIdx.x & 31;
    const int n_done = __ldcg(&st->n_done);
    for (int r0 = r_first - lane; r0 < a.ncols; r0 += r_stride) {
        const int r = r0 + lane;
        bool emit = false;
        int b = -1;
        double va = 0;
        // two memory levels instead of four: everything indexed by r goes out together, then everything indexed by its
        // nearest neighbour b (a round of the chain phase is a chain of dependent round trips, not bytes)
        const bool in = r < a.ncols;
        const size_t rbase = (size_t)(in ? r : 0) * LC;
        const int role_r = in ? __ldcg(a.c.role + r) : ROLE_DEAD;
        const int b0 = in ? __ldcg(a.c.cidx + rbase) : -1, c1r = in ? __ldcg(a.c.cidx + rbase + 1) : -1;
        const double v0r = __ldcg(a.c.cval + rbase), v1r = __ldcg(a.c.cval + rbase + 1), vKr = __ldcg(a.c.cval + rbase + KC);
        const int fresh_r = in ? __ldcg(a.c.fresh + r) : 0, size_r = in ? __ldcg(a.c.size + r) : 1;
        const bool cand = role_r == ROLE_FREE && b0 > r;
        const size_t bbase = (size_t)(cand ? b0 : 0) * LC;
        int c0b = -1, c1b = -1, fresh_b = 0, size_b = 1;
        double v0b = 0, v1b = 0, vKb = 0;
        if (cand) {
            c0b = __ldcg(a.c.cidx + bbase); c1b = __ldcg(a.c.cidx + bbase + 1);
            v0b = __ldcg(a.c.cval + bbase); v1b = __ldcg(a.c.cval + bbase + 1); vKb = __ldcg(a.c.cval + bbase + KC);
            fresh_b = __ldcg(a.c.fresh + b0); size_b = __ldcg(a.c.size + b0);
        }
        if (cand) {
            b = b0;
            if (c0b == r) {
                va = v0r;
                const double vb = v0b;
                // second smallest entry of either row: the next candidate (exact) or, with a single candidate left,
                // the bound of the unlisted entries (exact only straight after a scan)
                const bool two_r = c1r >= 0, two_b = c1b >= 0;
                const double sr = two_r ? v1r : vKr, sb = two_b ? v1b : vKb;
                const bool exact_r = two_r || fresh_r, exact_b = two_b || fresh_b;
                // Margin 0 where both engines compute bit-identical entries: min/max updates commute exactly
                // (single, complete), and for average/weighted every entry of a row whose cluster is still a
                // single leaf went only through the OTHER cluster's merge tree, whose operation order is fixed
                // (ward is excluded: its update uses the merge height, which is interleaving dependent).
                // Elsewhere the engines interleave two merge trees differently: demand a relative MARGIN.
                const bool exact_all = a.method == M_SINGLE || a.method == M_COMPLETE;
                const bool leaf_exact = a.method == M_AVERAGE || a.method == M_WEIGHTED;
                const double thr = __dadd_rn(va, __dmul_rn(fabs(va), a.margin));
                const bool ok_r = sr > ((exact_all || (leaf_exact && size_r == 1)) ? va : thr);
                const bool ok_b = sb > ((exact_all || (leaf_exact && size_b == 1)) ? va : thr);
                const bool ok = ok_r && ok_b && vb == va;
                if (ok) emit = true;
                else if (exact_r && exact_b) {
                    // Exact second minima and still too close.  A NEAR tie is left pending: the pair stays reciprocal
                    // (reducibility) and is merged in a later round if its rivals move away; all other certified
                    // merges go on (a run whose every merge is strict has a unique tree, whatever the order).  An
                    // EXACT tie (identical rows, e.g. empty windows) practically never resolves: stop at once.
                    st->amb_r = a.c.orig[r]; st->amb_b = a.c.orig[b];
                    st->amb_sr = size_r; st->amb_sb = size_b;
                    st->amb_v = va; st->amb_vb = vb;
                    st->amb_secr = sr; st->amb_secb = sb;
                    const unsigned long long k_obs = atomicAdd(&st->stats[6], 1ull);
                    if (k_obs < 16) st->gaps[k_obs] = (fmin(sr, sb) - va) / fabs(va);
                    if (a.diag) emit = true;  // diagnostics only: merge anyway, the result is NOT certified
                    else if (!(sr > va) || !(sb > va) || vb != va) st->ambiguous = 1;
                }
                else {  // the bound of "the rest" is stale: rescan both rows, decide next round
                    a.c.need[r] = 1;
                    a.c.need[b] = 1;
                    atomicAdd(&ct->deferred, 1);
                }
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, emit);
        if (!m) continue;
        int base = 0;
        const int leader = __ffs(m) - 1;
        if (lane == leader) base = atomicAdd(&ct->P, __popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (emit) {
            const int slot = base + __popc(m & ((1u << lane) - 1));
            Pair p;
            p.i = r;
            p.j = b;
            p.m = va;
            p.s1 = (double)__ldcg(a.c.size + r);
            p.s2 = (double)__ldcg(a.c.size + b);
            const double tot = __dadd_rn(p.s1, p.s2);
            p.as = __ddiv_rn(p.s1, tot);
            p.at = __ddiv_rn(p.s2, tot);
            a.pairs[slot] = p;
            a.c.role[r] = 2 * slot;
            a.c.role[b] = 2 * slot + 1;
            const int k = n_done + slot;
            a.rec_a[k] = a.c.orig[r];
            a.rec_b[k] = a.c.orig[b];
            a.rec_d[k] = va;
        }
    }
}

__global__ void __launch_bounds__(256) kr_select(RArgs a) {
    if (a.st->done) return;
    select_rows(a, &a.st->ctr[0], blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

// ---- light: every unmerged row updates its own D(k, j_q), keeps its NN cache --------------------
// A row is owned by a group of G = min(32, pow2ceil(P)) lanes that strides over the pairs, so a round with
// one pair handles 32 rows per warp and a round with thousands of pairs gives a row a whole warp.
__device__ __forceinline__ void light_rows(const RArgs &a, Ctr *ct, int P, long long warp0, long long nwarps) {
    RState *st = a.st;
    const int method = a.method, lane = threadIdx.x & 31;
    int G = 1;
    while (G < 32 && G < P) G <<= 1;
    const int rows_per_warp = 32 / G, g = lane / G, q0 = lane - g * G;
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (g * G));
    for (long long rbase = warp0 * rows_per_warp; rbase < a.ncols; rbase += nwarps * rows_per_warp) {
        const int row = (int)rbase + g;
        if (row >= a.ncols || __ldcg(a.c.role + row) != ROLE_FREE) continue;
        const bool need = __ldcg(a.c.need + row) != 0;
        if (P > 0) {
            double *Wr = a.W + (size_t)row * a.stride;
            const double v = (double)__ldcg(a.c.size + row);
            for (int q = q0; q < P; q += 2 * G) {  // two pairs per lane in flight
                const int q2 = q + G;
                const bool two = q2 < P;
                const Pair p = ld_pair(a.pairs + q);
                const Pair p2 = ld_pair(a.pairs + (two ? q2 : q));
                const double A = __ldcg(Wr + p.i), B = __ldcg(Wr + p.j);
                const double A2 = __ldcg(Wr + p2.i), B2 = __ldcg(Wr + p2.j);
                Wr[p.j] = lw(method, A, B, p, v);
                if (two) Wr[p2.j] = lw(method, A2, B2, p2, v);
            }
            __syncwarp(gmask);  // the new entries of this row (written by the lanes of its group) are visible below
            if (q0 == 0) {
                a.c.fresh[row] = 0;  // the bound is no longer the exact next entry
                if (!need) {
                    // Candidates that were merged this round follow their cluster to column j with the new value.
                    // Unlisted entries only move up (reducibility), so the bound stays valid; a candidate that rose
                    // above the bound leaves the list, and an empty list asks for a rescan.
                    const size_t base = (size_t)row * LC;
                    double cv[KC];
                    int ci[KC];
#pragma unroll
                    for (int k = 0; k < KC; ++k) { cv[k] = __ldcg(a.c.cval + base + k); ci[k] = __ldcg(a.c.cidx + base + k); }
                    const double lb = __ldcg(a.c.cval + base + KC);
                    bool hit = false;
#pragma unroll
                    for (int k = 0; k < KC; ++k) {
                        if (ci[k] < 0) continue;
                        const int r = __ldcg(a.c.role + ci[k]);
                        if (r < 0) continue;
                        hit = true;
                        const int j = __ldcg(&a.pairs[r >> 1].j);
                        bool dup = false;  // both sides of the pair were listed: keep one
#pragma unroll
                        for (int t = 0; t < KC; ++t) dup |= (t != k) && (ci[t] == j) && ((r & 1) == 0 || t < k);
                        if (dup && (r & 1) == 0) { ci[k] = -1; continue; }
                        ci[k] = j;
                        cv[k] = __ldcg(Wr + j);
                    }
                    if (hit) {
                        TopL out;
                        out.init();
#pragma unroll
                        for (int k = 0; k < KC; ++k)
                            if (ci[k] >= 0 && cv[k] <= lb) out.insert(cv[k], ci[k]);
#pragma unroll
                        for (int k = 0; k < KC; ++k) {
                            a.c.cval[base + k] = out.v[k];
                            a.c.cidx[base + k] = out.i[k] == INT_MAX ? -1 : out.i[k];
                        }
                        if (out.i[0] == INT_MAX) a.rq[atomicAdd(&ct->qlen, 1)] = row;
                    }
                }
            }
        }
        if (need && q0 == 0) a.rq[atomicAdd(&ct->qlen, 1)] = row;  // the rescan clears the flag
    }
}

// Thin rounds (<= 4 pairs, the chain phase): one lane per row and a dependency chain of TWO memory levels instead of
// six -- everything indexed by the row (role, size, candidate list) and the row's old entries of the merging columns
// are requested together, the roles of the candidates right behind them, and the new entries are taken from registers
// when the list is repaired.  Same arithmetic, same list, same queue as light_rows.
template <typename Tv>
__device__ __forceinline__ Tv pick4(const Tv (&v)[4], int k) {
    return k == 0 ? v[0] : (k == 1 ? v[1] : (k == 2 ? v[2] : v[3]));
}

__device__ __forceinline__ void light_rows_thin(const RArgs &a, Ctr *ct, int P, int t_first, int t_stride) {
    RState *st = a.st;
    const int method = a.method;
    Pair pr[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) pr[q] = ld_pair(a.pairs + (q < P ? q : 0));
    for (int row = t_first; row < a.ncols; row += t_stride) {
        const size_t base = (size_t)row * LC;
        double *Wr = a.W + (size_t)row * a.stride;
        const int role = __ldcg(a.c.role + row);
        const bool need = __ldcg(a.c.need + row) != 0;
        const double v = (double)__ldcg(a.c.size + row);
        double cv[KC], A[4], B[4];
        int ci[KC], rci[KC];
#pragma unroll
        for (int k = 0; k < KC; ++k) { cv[k] = __ldcg(a.c.cval + base + k); ci[k] = __ldcg(a.c.cidx + base + k); }
        const double lb = __ldcg(a.c.cval + base + KC);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            A[q] = __ldcg(Wr + pr[q].i);
            B[q] = __ldcg(Wr + pr[q].j);
        }
#pragma unroll
        for (int k = 0; k < KC; ++k) rci[k] = ci[k] >= 0 ? __ldcg(a.c.role + ci[k]) : -1;
        if (role != ROLE_FREE) continue;
        double nv[4];
        int nj[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            nv[q] = 0.0;
            nj[q] = pr[q].j;
            if (q < P) {   // (uniform: a round with one pair does one Lance-Williams step per row, not four)
                nv[q] = lw(method, A[q], B[q], pr[q], v);
                Wr[pr[q].j] = nv[q];
            }
        }
        a.c.fresh[row] = 0;  // the bound is no longer the exact next entry
        if (!need) {
            bool hit = false;
#pragma unroll
            for (int k = 0; k < KC; ++k) {
                if (ci[k] < 0) continue;
                const int r = rci[k];
                if (r < 0) continue;
                hit = true;
                const int j = pick4(nj, r >> 1);
                bool dup = false;  // both sides of the pair were listed: keep one
#pragma unroll
                for (int t = 0; t < KC; ++t) dup |= (t != k) && (ci[t] == j) && ((r & 1) == 0 || t < k);
                if (dup && (r & 1) == 0) { ci[k] = -1; continue; }
                ci[k] = j;
                cv[k] = pick4(nv, r >> 1);
            }
            if (hit) {
                TopL out;
                out.init();
#pragma unroll
                for (int k = 0; k < KC; ++k)
                    if (ci[k] >= 0 && cv[k] <= lb) out.insert(cv[k], ci[k]);
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    a.c.cval[base + k] = out.v[k];
                    a.c.cidx[base + k] = out.i[k] == INT_MAX ? -1 : out.i[k];
                }
                if (out.i[0] == INT_MAX) a.rq[atomicAdd(&ct->qlen, 1)] = row;
            }
        } else {
            a.rq[atomicAdd(&ct->qlen, 1)] = row;  // the rescan clears the flag
        }
    }
}

// The same thin-round update with EIGHT lanes per row (four rows per warp, two rows per group in flight): the one-lane
// version is a serial chain of ~700 dependent instructions per row (four Lance-Williams steps, a seven-fold insertion
// sort) with 48 registers of pair data and 220 bytes of spills.  Here lane s of a group holds candidate s of the row's
// list (lane 7 the bound) and, for s < P, pair s of the round: the steps run side by side, a merged candidate takes
// its new column and value from the pair's lane by shuffle, and the list is re-sorted by rank (seven shuffles).
// Same list as light_rows_thin: an i-side candidate is dropped when the j side of its pair is listed too, otherwise it
// follows the cluster to column j; entries above the bound leave; order = ascending (value, column).
__device__ __forceinline__ void light_rows_thin8(const RArgs &a, Ctr *ct, int P, int gwarp, int nwarps) {
    const int lane = threadIdx.x & 31, sub = lane & 7, gbase = lane & ~7, method = a.method;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int ngroups = nwarps * 4;
    Pair pr = ld_pair(a.pairs + (sub < P ? sub : 0));
    for (int r0 = gwarp * 4 + (lane >> 3); r0 - (lane >> 3) < a.ncols; r0 += 2 * ngroups) {   // warp-uniform trip count
        int row[2];
        bool act[2], need[2];
        double cv[2], A[2], B[2], v[2];
        int ci[2], rci[2], role[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            row[h] = r0 + h * ngroups;
            const bool in = row[h] < a.ncols;
            const size_t base = (size_t)(in ? row[h] : 0) * LC;
            const double *Wr = a.W + (size_t)(in ? row[h] : 0) * a.stride;
            role[h] = in ? __ldcg(a.c.role + row[h]) : ROLE_DEAD;
            need[h] = in && __ldcg(a.c.need + row[h]) != 0;
            v[h] = in ? (double)__ldcg(a.c.size + row[h]) : 1.0;
            cv[h] = in ? __ldcg(a.c.cval + base + sub) : INF;                   // lane 7: the bound of the unlisted entries
            ci[h] = (in && sub < KC) ? __ldcg(a.c.cidx + base + sub) : -1;
            A[h] = (in && sub < P) ? __ldcg(Wr + pr.i) : 0.0;
            B[h] = (in && sub < P) ? __ldcg(Wr + pr.j) : 0.0;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) rci[h] = ci[h] >= 0 ? __ldcg(a.c.role + ci[h]) : -1;
        bool push_row[2] = {false, false};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            act[h] = role[h] == ROLE_FREE;      // group-uniform
            double nv = 0.0;
            if (act[h] && sub < P) {
                nv = lw(method, A[h], B[h], pr, v[h]);
                a.W[(size_t)row[h] * a.stride + pr.j] = nv;
            }
            if (act[h] && sub == 0) a.c.fresh[row[h]] = 0;   // the bound is no longer the exact next entry
            // candidates whose cluster merges this round follow it to column j with the new value (all lanes shuffle)
            const bool merging = act[h] && !need[h] && ci[h] >= 0 && rci[h] >= 0;
            const int src = gbase | (merging ? (rci[h] >> 1) : 0);
            const int pj = __shfl_sync(0xffffffffu, pr.j, src);
            const double pv = __shfl_sync(0xffffffffu, nv, src);
            const unsigned hit_mask = __ballot_sync(0xffffffffu, merging) & (0xffu << gbase);
            bool jside_listed = false;   // is column pj (the j side) one of the ORIGINAL candidates of the row?
#pragma unroll
            for (int t = 0; t < KC; ++t) {
                const int oc = __shfl_sync(0xffffffffu, ci[h], gbase | t);
                jside_listed |= (t != sub) && oc == pj;
            }
            int nci = ci[h];
            double ncv = cv[h];
            if (merging) {
                if ((rci[h] & 1) == 0 && jside_listed) nci = -1;   // both sides were listed: keep the j side only
                else { nci = pj; ncv = pv; }
            }
            const double lb = __shfl_sync(0xffffffffu, cv[h], gbase | KC);
            const bool keep = sub < KC && nci >= 0 && ncv <= lb;
            // rank among the kept entries in (value, column) order; the others fill the tail with (inf, -1)
            int rank = 0, below_unkept = 0;
            const unsigned keep_mask = (__ballot_sync(0xffffffffu, keep) >> gbase) & 0x7fu;
#pragma unroll
            for (int t = 0; t < KC; ++t) {
                const double ov = __shfl_sync(0xffffffffu, ncv, gbase | t);
                const int oi = __shfl_sync(0xffffffffu, nci, gbase | t);
                const bool ok = (keep_mask >> t) & 1u;
                if (t != sub && ok && lex_lt(ov, oi, ncv, nci)) rank++;
                if (t < sub && !ok) below_unkept++;
            }
            const int count = __popc(keep_mask);
            if (act[h] && !need[h] && hit_mask && sub < KC) {
                const size_t base = (size_t)row[h] * LC;
                const int slot = keep ? rank : count + below_unkept;
                a.c.cval[base + slot] = keep ? ncv : INF;
                a.c.cidx[base + slot] = keep ? nci : -1;
            }
            push_row[h] = act[h] && sub == 0 && (need[h] || (hit_mask && count == 0));   // an empty list asks for a rescan
        }
        // one queue reservation per warp
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const unsigned m = __ballot_sync(0xffffffffu, push_row[h]);
            if (m) {
                const int leader = __ffs(m) - 1;
                int slot0 = 0;
                if (lane == leader) slot0 = atomicAdd(&ct->qlen, __popc(m));
                slot0 = __shfl_sync(0xffffffffu, slot0, leader);
                if (push_row[h]) a.rq[slot0 + __popc(m & ((1u << lane) - 1u))] = row[h];
            }
        }
    }
}

__global__ void __launch_bounds__(LIGHT_THREADS) kr_light(RArgs a) {
    RState *st = a.st;
    if (st->done) return;
    const int P = st->ctr[0].P;
    if (P >= 1 && P <= 4) light_rows_thin(a, &st->ctr[0], P, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
    else light_rows(a, &st->ctr[0], P, ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5,
                    ((long long)gridDim.x * blockDim.x) >> 5);
}

// ---- heavy: new rows of merged clusters, exact rescans; the last CTA retires the pairs -----------
// Work item = (entry, column segment).  With many entries a CTA takes whole rows; in the long tail of
// rounds with one or two entries a row is cut into segments over many CTAs and the last CTA to finish
// an entry combines the partial minima, so a round never waits for one CTA streaming a whole row.
template <bool VEC>
__device__ __forceinline__ void heavy_columns(const RArgs &a, int e, int P, int seg, int S, int c_lo, int c_hi,
                                              bool all_free, Min3 &best) {
    const int lane = threadIdx.x & 31, method = a.method;  // one warp per (entry, segment)
    constexpr int V = VEC ? 2 : 1, U = 2, UR = 4;  // loads in flight per thread: builds 2 x (A, B), rescans 4
    if (e < P) {
        const Pair p = ld_pair(a.pairs + e);
        const double *Wi = a.W + (size_t)p.i * a.stride;
        double *Wj = a.W + (size_t)p.j * a.stride;
        if (lane == 0 && c_hi - c_lo <= 4096) {   // (short segments = the latency-bound rounds; big rounds stream anyway)
            bulk_prefetch_l2(Wi + c_lo, (size_t)(c_hi - c_lo) * 8);
            bulk_prefetch_l2(Wj + c_lo, (size_t)(c_hi - c_lo) * 8);
        }
        // Phase A: columns of clusters that do not merge this round -- coalesced, no divergent work.
        for (int c0 = c_lo + lane * V; c0 < c_hi; c0 += U * V * 32) {
            double A[U][V], B[U][V];
            int rc[U][V], sz[U][V];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int c = c0 + u * V * 32;
#pragma unroll
                for (int k = 0; k < V; ++k) rc[u][k] = ROLE_DEAD;
                if (VEC && c + 1 < c_hi) {
                    const int2 r2 = __ldcg((const int2 *)(a.c.role + c)), s2 = __ldcg((const int2 *)(a.c.size + c));
                    const double2 a2 = __ldcg((const double2 *)(Wi + c)), b2 = __ldcg((const double2 *)(Wj + c));
                    rc[u][0] = r2.x; rc[u][V - 1] = r2.y;
                    sz[u][0] = s2.x; sz[u][V - 1] = s2.y;
                    A[u][0] = a2.x; A[u][V - 1] = a2.y;
                    B[u][0] = b2.x; B[u][V - 1] = b2.y;
                } else if (c < c_hi) {
                    rc[u][0] = __ldcg(a.c.role + c);
                    sz[u][0] = __ldcg(a.c.size + c);
                    A[u][0] = __ldcg(Wi + c);
                    B[u][0] = __ldcg(Wj + c);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    if (rc[u][k] != ROLE_FREE) continue;  // dead, or a side of some pair (phase B)
                    const int c = c0 + u * V * 32 + k;
                    const double x = lw(method, A[u][k], B[u][k], p, (double)sz[u][k]);
                    Wj[c] = x;
                    best.push(x, c);
                }
            }
        }
        // Phase B: the surviving column of every OTHER pair q: four old entries -> one new, "lower j first" so that
        // both mirror entries get bit-identical values.  The pairs of this round are dealt over the segments of
        // the entry; all four gathers of a lane are independent.  Nobody else reads or writes these entries:
        // phase A skips them and every pair owns its two columns.
        // Both merge orders are the same three Lance-Williams steps on permuted operands: select the operands,
        // not the code path (lanes of a warp hold different q, a branch would run both orders for every lane).
        const double ps = __dadd_rn(p.s1, p.s2);
        auto combine = [&](const Pair &pq, double Ai, double Bi, double Aj, double Bj) {
            const bool pf = p.j < pq.j;  // p merges first
            Pair F, Sd;
            F.m = pf ? p.m : pq.m; F.s1 = pf ? p.s1 : pq.s1; F.s2 = pf ? p.s2 : pq.s2; F.as = pf ? p.as : pq.as; F.at = pf ? p.at : pq.at;
            Sd.m = pf ? pq.m : p.m; Sd.s1 = pf ? pq.s1 : p.s1; Sd.s2 = pf ? pq.s2 : p.s2; Sd.as = pf ? pq.as : p.as; Sd.at = pf ? pq.at : p.at;
            F.i = F.j = Sd.i = Sd.j = 0;
            const double t1 = lw(method, Ai, pf ? Bi : Aj, F, pf ? pq.s1 : p.s1);
            const double t2 = lw(method, pf ? Aj : Bi, Bj, F, pf ? pq.s2 : p.s2);
            const double x = lw(method, t1, t2, Sd, pf ? ps : __dadd_rn(pq.s1, pq.s2));
            Wj[pq.j] = x;
            best.push(x, pq.j);
        };
        for (int q = seg + S * lane; q < P; q += S * 32) {  // (two pairs per lane in flight measured no faster)
            if (q == e) continue;
            const Pair pq = ld_pair(a.pairs + q);
            const double Ai = __ldcg(Wi + pq.i), Bi = __ldcg(Wj + pq.i), Aj = __ldcg(Wi + pq.j), Bj = __ldcg(Wj + pq.j);
            combine(pq, Ai, Bi, Aj, Bj);
        }
    } else if (all_free) {  // initial scan: every column is a live singleton, no role lookups
        const int row = __ldcg(a.rq + e - P);
        const double *Wr = a.W + (size_t)row * a.stride;
        constexpr int UA = 8;
        for (int c0 = c_lo + lane * V; c0 < c_hi; c0 += UA * V * 32) {
            double X[UA][V];
#pragma unroll
            for (int u = 0; u < UA; ++u) {
                const int c = c0 + u * V * 32;
#pragma unroll
                for (int k = 0; k < V; ++k) X[u][k] = __longlong_as_double(0x7ff0000000000000LL);
                if (VEC && c + 1 < c_hi) {
                    const double2 x2 = __ldcs((const double2 *)(Wr + c));
                    X[u][0] = x2.x; X[u][V - 1] = x2.y;
                } else if (c < c_hi) {
                    X[u][0] = __ldcs(Wr + c);
                }
            }
#pragma unroll
            for (int u = 0; u < UA; ++u) {
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    const int c = c0 + u * V * 32 + k;
                    if (c != row) best.push(X[u][k], c);  // (+inf padding never enters)
                }
            }
        }
    } else {
        const int row = __ldcg(a.rq + e - P);
        const double *Wr = a.W + (size_t)row * a.stride;
        if (lane == 0 && c_hi - c_lo <= 4096) bulk_prefetch_l2(Wr + c_lo, (size_t)(c_hi - c_lo) * 8);
        for (int c0 = c_lo + lane * V; c0 < c_hi; c0 += UR * V * 32) {
            double X[UR][V];
            int rc[UR][V];
#pragma unroll
            for (int u = 0; u < UR; ++u) {
                const int c = c0 + u * V * 32;
#pragma unroll
                for (int k = 0; k < V; ++k) rc[u][k] = ROLE_DEAD;
                if (VEC && c + 1 < c_hi) {
                    const int2 r2 = __ldcg((const int2 *)(a.c.role + c));
                    const double2 x2 = __ldcg((const double2 *)(Wr + c));
                    rc[u][0] = r2.x; rc[u][V - 1] = r2.y;
                    X[u][0] = x2.x; X[u][V - 1] = x2.y;
                } else if (c < c_hi) {
                    rc[u][0] = __ldcg(a.c.role + c);
                    X[u][0] = __ldcg(Wr + c);
                }
            }
#pragma unroll
            for (int u = 0; u < UR; ++u) {
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    const int c = c0 + u * V * 32 + k, r = rc[u][k];
                    if (r == ROLE_DEAD || (r >= 0 && !(r & 1)) || c == row) continue;
                    best.push(X[u][k], c);
                }
            }
        }
    }
}

// ---- warp-level selection of the LC smallest out of 64 candidates (two sorted per lane) ---------------
// Values travel as order-preserving 64-bit keys so a warp minimum is two redux.sync instructions.
constexpr unsigned long long KEY_INF = 0xfff0000000000000ull;  // dkey(+inf)
__device__ __forceinline__ unsigned long long dkey(double x) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    return b ^ ((b >> 63) ? ~0ull : 0x8000000000000000ull);
}
__device__ __forceinline__ double dunkey(unsigned long long k) {
    return __longlong_as_double((long long)((k >> 63) ? (k ^ 0x8000000000000000ull) : ~k));
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long k) {
    const unsigned hi = (unsigned)(k >> 32);
    const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? (unsigned)k : 0xffffffffu);
    return ((unsigned long long)mh << 32) | ml;
}
// in: per lane two candidates k1 <= k2 (KEY_INF / INT_MAX = none) and kb = lower bound of what the lane dropped.
// out: lane s < LC holds the s-th smallest candidate of the warp; bound (all lanes) = smallest value not selected.
__device__ __forceinline__ void warp_select(unsigned long long k1, int i1, unsigned long long k2, int i2,
                                            unsigned long long kb, unsigned long long &okey, int &oidx,
                                            unsigned long long &obound) {
    const int lane = threadIdx.x & 31;
    okey = KEY_INF;
    oidx = INT_MAX;
#pragma unroll
    for (int s = 0; s < LC; ++s) {
        const unsigned long long m = warp_min_u64(k1);
        const int wl = __ffs(__ballot_sync(0xffffffffu, k1 == m)) - 1;
        const int wi = __shfl_sync(0xffffffffu, i1, wl);
        if (lane == s) { okey = m; oidx = wi; }
        if (lane == wl) { k1 = k2; i1 = i2; k2 = KEY_INF; i2 = INT_MAX; }
    }
    const unsigned long long r1 = warp_min_u64(kb), r2 = warp_min_u64(k1);
    obound = r1 < r2 ? r1 : r2;
}

// one warp: lane l holds the sorted partial list of segment l (KEY_INF padded) and its bound kb; out: lane s < LC holds
// the s-th smallest candidate of all segments, bound (all lanes) = smallest value not selected
__device__ __forceinline__ void combine_partials(unsigned long long (&pk)[LC], int (&pi)[LC], unsigned long long kb,
                                                 unsigned long long &key, int &idx, unsigned long long &bound) {
    const int lane = threadIdx.x & 31;
    key = KEY_INF;
    idx = INT_MAX;
#pragma unroll
    for (int sel = 0; sel < LC; ++sel) {
        const unsigned long long m = warp_min_u64(pk[0]);
        const int wl = __ffs(__ballot_sync(0xffffffffu, pk[0] == m)) - 1;
        const int wi = __shfl_sync(0xffffffffu, pi[0], wl);
        if (lane == sel) { key = m; idx = wi; }
        if (lane == wl) {
#pragma unroll
            for (int t = 0; t + 1 < LC; ++t) { pk[t] = pk[t + 1]; pi[t] = pi[t + 1]; }
            pk[LC - 1] = KEY_INF;
            pi[LC - 1] = INT_MAX;
        }
    }
    const unsigned long long r1 = warp_min_u64(kb), r2 = warp_min_u64(pk[0]);
    bound = r1 < r2 ? r1 : r2;
}

// one warp: the candidate list of `row` from a finished scan (lane s < LC: s-th smallest entry; bound: smallest entry
// that was not selected)
__device__ __forceinline__ void store_list(const RArgs &a, int row, unsigned long long key, int idx, unsigned long long bound) {
    const int lane = threadIdx.x & 31;
    const size_t base = (size_t)row * LC;
    // the smallest entry outside the list: the (KC+1)-th candidate or an entry nobody kept
    const unsigned long long last = __shfl_sync(0xffffffffu, key, KC);
    const unsigned long long lb = bound < last ? bound : last;
    if (lane < KC) {
        const bool keep = idx != INT_MAX && key <= lb;
        a.c.cval[base + lane] = keep ? dunkey(key) : __longlong_as_double(0x7ff0000000000000LL);
        a.c.cidx[base + lane] = keep ? idx : -1;
    }
    if (lane == KC) a.c.cval[base + KC] = dunkey(lb);
    if (lane == 0) {
        a.c.fresh[row] = 1;
        a.c.need[row] = 0;
    }
}

// segments per entry for `total` entries on `nwarps` warps: enough items for every resident warp, at least SEG_MIN
// columns each (even length keeps the 16-byte loads aligned)
__device__ __forceinline__ int heavy_segments(const RArgs &a, int P, int total, int nwarps) {
    constexpr int SEG_MIN = 512;
    int S = 1;
    if (total > 0) {
        if (total < nwarps / 2) S = (2 * nwarps + total - 1) / total;   // up to two items per warp: short dependent chains
        // Big rounds: the pair x pair gathers of a merged row (phase B) re-read the two old rows.  With every row
        // cut into many segments only ~100 rows are being built at a time (2 x 8 x ncols bytes each), so the
        // gathers and the second write of a sector find them in the 126 MB L2 instead of going back to HBM.
        if (P >= 64 && S < a.bulk_segs) S = a.bulk_segs;
        const int smax = (a.ncols + SEG_MIN - 1) / SEG_MIN;
        if (S > smax) S = smax;
        if (S > a.max_segs) S = a.max_segs;
        if (S < 1) S = 1;
    }
    return S;
}

// One warp per (entry, column segment), entries [e_lo, e_hi) of the round's list (merged rows first, then the rows
// to rescan): warps never wait for each other, so short rows in the long tail of small rounds cost one load latency
// plus one selection.  nactive = live columns (for the byte accounting).
__device__ __forceinline__ void heavy_items(const RArgs &a, int P, int e_lo, int e_hi, int S, int gwarp, int nwarps,
                                            bool all_free, int nactive) {
    RState *st = a.st;
    const int lane = threadIdx.x & 31, ncols = a.ncols;
    int seg_len = (ncols + S - 1) / S;
    seg_len = (seg_len + 1) & ~1;
    const bool vec = (a.stride & 1) == 0;
    const long long items = (long long)(e_hi - e_lo) * S;
    for (long long w = gwarp; w < items; w += nwarps) {
        const int e = e_lo + (int)(w / S), seg = (int)(w % S);
        const int c_lo = seg * seg_len, c_hi = min(ncols, c_lo + seg_len);
        Min3 mine;
        mine.init();
        if (vec) heavy_columns<true>(a, e, P, seg, S, c_lo, c_hi, all_free, mine);
        else heavy_columns<false>(a, e, P, seg, S, c_lo, c_hi, all_free, mine);
        unsigned long long key, bound;
        int idx;
        warp_select(dkey(mine.m1), mine.i1, dkey(mine.m2), mine.i2, dkey(mine.m3), key, idx, bound);
        bool finish = S == 1;
        if (S > 1) {  // publish the partial; the last segment of the entry combines them all
            Part *pt = &a.parts[(size_t)e * a.max_segs + seg];
            if (lane < LC) { pt->k[lane] = key; pt->i[lane] = idx; }
            if (lane == 0) pt->bound = bound;
            __threadfence();
            __syncwarp();
            int last = 0;
            if (lane == 0) last = atomicAdd(&a.seg_done[e], 1) == S - 1;
            finish = __shfl_sync(0xffffffffu, last, 0) != 0;
            if (finish) {
                __threadfence();
                // lane < S takes the sorted partial list of segment `lane`
                const Part *pp = &a.parts[(size_t)e * a.max_segs + lane];
                unsigned long long pk[LC], kb = KEY_INF;
                int pi[LC];
#pragma unroll
                for (int t = 0; t < LC; ++t) {
                    pk[t] = lane < S ? __ldcg(&pp->k[t]) : KEY_INF;
                    pi[t] = lane < S ? __ldcg(&pp->i[t]) : INT_MAX;
                }
                if (lane < S) kb = __ldcg(&pp->bound);
                combine_partials(pk, pi, kb, key, idx, bound);
            }
        }
        if (finish) {
            const int row = e < P ? __ldcg(&a.pairs[e].j) : __ldcg(a.rq + e - P);
            store_list(a, row, key, idx, bound);
            if (lane == 0 && S > 1) a.seg_done[e] = 0;
        }
    }
}

// the pairs of a finished round leave the matrix: i retires, j carries the merged cluster
__device__ __forceinline__ void retire_pairs(const RArgs &a, int P, int tid, int nthreads) {
    for (int q = tid; q < P; q += nthreads) {
        const Pair p = ld_pair(a.pairs + q);
        a.c.role[p.i] = ROLE_DEAD;
        a.c.role[p.j] = ROLE_FREE;
        a.c.size[p.j] = (int)(p.s1 + p.s2);
    }
}

// one thread: close the round (counters of `cur` are read, the set `next` is cleared for the round after)
__device__ __forceinline__ void close_round(const RArgs &a, Ctr *cur, Ctr *next, int P, int total) {
    RState *st = a.st;
    // work counters of the round, derived instead of counted: one atomicAdd per updated row on ONE address (14 000
    // per round in the chain phase) serialised in L2 for ~20 us, longer than the round's real work
    const unsigned long long nact = (unsigned long long)st->nactive, np = (unsigned long long)P, nq = (unsigned long long)(total - P);
    st->stats[0] += nq;                    // rescans
    st->stats[1] += nq * nact;             // ... live columns
    st->stats[2] += np;                    // merged-row builds
    st->stats[3] += np * nact;
    st->stats[4] += (nact - 2 * np) * np;  // row x pair updates of the free rows
    st->n_done += P;
    st->nactive -= P;
    st->stats[5] += cur->deferred;
    if (st->n_done >= a.n - 1) st->done = 1;
    else if (st->ambiguous && !a.diag) st->done = 1;
    else if (P == 0 && cur->deferred == 0 && total == 0) { st->stuck = 1; st->done = 1; }  // only pending near-ties left
    st->last_P = P;
    st->last_q = total - P;
    next->P = 0;
    next->qlen = 0;
    next->deferred = 0;
    st->blocks_done = 0;
    st->round += 1;
}

template <int MINB>
__global__ void __launch_bounds__(HEAVY_THREADS, MINB) kr_heavy(RArgs a) {
    __shared__ int s_flag;
    static_assert(MAX_SEGS <= 32, "one lane per partial list in the final selection");
    RState *st = a.st;
    if (st->done) return;
    Ctr *ct = &st->ctr[0];
    const int P = ct->P, total = P + ct->qlen, tid = threadIdx.x;
    const int nwarps = (int)gridDim.x * (HEAVY_THREADS / 32), gwarp = (int)blockIdx.x * (HEAVY_THREADS / 32) + (tid >> 5);
    const int S = heavy_segments(a, P, total, nwarps);
    heavy_items(a, P, 0, total, S, gwarp, nwarps, st->round == 0, st->nactive);
    // last CTA out retires the pairs of this round and closes it
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_flag = atomicAdd(&st->blocks_done, 1) == (int)gridDim.x - 1;
    }
    __syncthreads();
    if (!s_flag) return;
    __threadfence();
    retire_pairs(a, P, tid, HEAVY_THREADS);
    __syncthreads();
    if (tid == 0) close_round(a, ct, ct, P, total);
}

// ---- the chain phase in ONE launch ---------------------------------------------------------------------------------
// After the bulk rounds a single growing cluster merges one to four pairs per round for hundreds of rounds (C3: rounds
// 12-545).  Three dependent launches per round (select, light, heavy) plus a host synchronisation every few rounds
// cost ~45 us per round for ~5 us of work.  kr_chain runs the same three phases for MANY rounds inside one
// cooperative launch, separated by grid barriers (one CTA per SM):
//   phase 1  select_rows            -> pairs of the round (counter set `par`)
//   phase 2  build of the merged rows (heavy_items over the pairs) and light_rows -- they touch disjoint rows
//   phase 3  exact rescans of the rows whose candidate list ran empty; CTA 0 retires the pairs and closes the round
// It returns to the host when the run is finished, when half of the columns are dead (the host compacts the matrix)
// or after max_rounds rounds.
struct ChainCtl {
    int max_rounds, compact, diag;
    int thin;   // > 0: thin rounds through the shared-memory staged builds / rescans (rows can be bulk-copied); rows per CTA
    int cnum, cden;   // compact when nactive * cden <= ncols * cnum (live fraction at or below cnum / cden)
};

__device__ __forceinline__ void grid_barrier(unsigned int *bar, unsigned int &target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += gridDim.x;
        __threadfence();
        atomicAdd(bar, 1u);
        // poll with a pause: 148 CTAs spinning flat out on one L2 sector slow down every access to that slice
        while (*(volatile unsigned int *)bar < target) __nanosleep(64);
        __threadfence();
    }
    __syncthreads();
}

constexpr int CHAIN_THREADS = 512, CHAIN_THREADS_C = CHAIN_THREADS;

// out-of-line copies for the persistent kernel: three inlined phases do not fit 128 registers together
__device__ __noinline__ void select_rows_ni(const RArgs &a, Ctr *ct, int r_first, int r_stride) { select_rows(a, ct, r_first, r_stride); }
__device__ __noinline__ void light_rows_thin8_ni(const RArgs &a, Ctr *ct, int P, int gwarp, int nwarps) { light_rows_thin8(a, ct, P, gwarp, nwarps); }
__device__ __noinline__ void light_rows_ni(const RArgs &a, Ctr *ct, int P, long long warp0, long long nwarps) {
    if (P >= 1 && P <= 4 && a.light8) light_rows_thin8_ni(a, ct, P, (int)warp0, (int)nwarps);
    else if (P >= 1 && P <= 4) light_rows_thin(a, ct, P, (int)warp0 * 32 + (threadIdx.x & 31), (int)nwarps * 32);
    else light_rows(a, ct, P, warp0, nwarps);
}
__device__ __noinline__ void heavy_items_ni(const RArgs &a, int P, int e_lo, int e_hi, int S, int gwarp, int nwarps, int nactive) {
    heavy_items(a, P, e_lo, e_hi, S, gwarp, nwarps, false, nactive);
}


// ---- thin rounds (1-4 pairs, the chain phase): latency-optimised builds and rescans ---------------------------------
// A thin round is a handful of rows: its time is the length of the chain of DEPENDENT memory round trips, not bytes.
// heavy_items (index load -> four sequential load iterations per segment -> selection -> publish -> fence -> atomic ->
// last arriver re-reads the partials -> list) is ~10 round trips per row, 22-26 us per phase measured.  Here:
//   * every warp takes ONE segment and brings it (values, roles, sizes) into shared memory with cp.async.bulk -- one
//     round trip whatever the segment length, no registers held -- and works it from there;
//   * a row to rescan belongs to one CTA: its 16 warps combine their partial lists through shared memory behind a
//     __syncthreads (no fence, no atomic, no second trip to L2); (a second staging buffer per warp, with the next row's
//     copy in flight during the scan, measured SLOWER: 35.8 vs 33.6 ms -- the shared memory it takes from L1 costs more)
//   * a merged row is built by two CTAs (32 segments); their partial lists go to global memory and are combined one
//     grid barrier later, at the start of the rescan phase, by a single warp (the lists are not needed before select).
// Same arithmetic, same list semantics (the KC smallest live entries + a lower bound of the rest).
#ifndef HC_THIN_SEG
#define HC_THIN_SEG 512
#endif
constexpr int THIN_RESCAN_SEG = HC_THIN_SEG, THIN_BUILD_SEG = HC_THIN_SEG / 2;
constexpr int THIN_WARP_BYTES = THIN_RESCAN_SEG * 12;   // per warp: SEG doubles + SEG ints (rescan) = 2 x SEG/2 doubles + 2 x SEG/2 ints (build)
struct __align__(16) SPart {
    unsigned long long k[LC];
    int i[LC];
    unsigned long long bound;
};
constexpr size_t CHAIN_SMEM = 16 * (size_t)THIN_WARP_BYTES + 16 * sizeof(unsigned long long) + 16 * sizeof(SPart);

// all 16 warps of the CTA: exact rescan of `row` (call with CTA-uniform arguments; ends on a __syncthreads)
__device__ __noinline__ void rescan_row_cta(const RArgs &a, int row, unsigned char *wbuf, unsigned long long *bar, unsigned &par,
                                            SPart *sp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, ncols = a.ncols;
    constexpr int NW = CHAIN_THREADS_C / 32;
    const double *Wr = a.W + (size_t)row * a.stride;
    double *vals = (double *)wbuf;
    int *roles = (int *)(wbuf + THIN_RESCAN_SEG * 8);
    Min3 best;
    best.init();
    for (int base = 0; base < ncols; base += NW * THIN_RESCAN_SEG) {
        const int span = min(NW * THIN_RESCAN_SEG, ncols - base);
        const int seg_len = ((span + NW - 1) / NW + 3) & ~3;
        const int c_lo = base + warp * seg_len, cnt = min(base + span, c_lo + seg_len) - c_lo;
        if (cnt > 0) {   // warp-uniform
            if (lane == 0) {
                const uint32_t vb = (uint32_t)((cnt + 1) & ~1) * 8u, rb = (uint32_t)((cnt + 3) & ~3) * 4u;
                mbar_expect_tx(bar, vb + rb);
                tma_load_1d(vals, Wr + c_lo, vb, bar);
                tma_load_1d(roles, a.c.role + c_lo, rb, bar);
            }
            mbar_wait(bar, par & 1u);
            par ^= 1u;
            for (int t = lane; t < cnt; t += 32) {
                const int r = roles[t], c = c_lo + t;
                if (r == ROLE_DEAD || (r >= 0 && !(r & 1)) || c == row) continue;
                best.push(vals[t], c);
            }
            __syncwarp();   // the buffer has been read: the next bulk copy may overwrite it
        }
    }
    unsigned long long key, bound;
    int idx;
    warp_select(dkey(best.m1), best.i1, dkey(best.m2), best.i2, dkey(best.m3), key, idx, bound);
    if (lane < LC) { sp[warp].k[lane] = key; sp[warp].i[lane] = idx; }
    if (lane == 0) sp[warp].bound = bound;
    __syncthreads();
    if (warp == 0) {
        unsigned long long pk[LC], kb = KEY_INF;
        int pi[LC];
#pragma unroll
        for (int t = 0; t < LC; ++t) {
            pk[t] = lane < NW ? sp[lane].k[t] : KEY_INF;
            pi[t] = lane < NW ? sp[lane].i[t] : INT_MAX;
        }
        if (lane < NW) kb = sp[lane].bound;
        combine_partials(pk, pi, kb, key, idx, bound);
        store_list(a, row, key, idx, bound);
    }
    __syncthreads();
}

// one warp: segment `seg` of 32 of the merged row of pair e (phase A of heavy_columns out of shared memory; the warp of
// segment 0 also takes the pair x pair entries, phase B); its partial list goes to parts[e][seg]
__device__ __noinline__ void build_thin_warp(const RArgs &a, int P, int e, int seg, unsigned char *wbuf, unsigned long long *bar,
                                             unsigned &par) {
    const int lane = threadIdx.x & 31, ncols = a.ncols, method = a.method;
    const Pair p = ld_pair(a.pairs + e);
    const double *Wi = a.W + (size_t)p.i * a.stride;
    double *Wj = a.W + (size_t)p.j * a.stride;
    // phase B operands (the other pairs of the round): requested first, used last
    const bool doB = seg == 0 && lane < P && lane != e;
    Pair pq = p;
    double Ai = 0, Bi = 0, Aj = 0, Bj = 0;
    if (doB) {
        pq = ld_pair(a.pairs + lane);
        Ai = __ldcg(Wi + pq.i); Bi = __ldcg(Wj + pq.i); Aj = __ldcg(Wi + pq.j); Bj = __ldcg(Wj + pq.j);
    }
    double *bA = (double *)wbuf, *bB = (double *)(wbuf + THIN_BUILD_SEG * 8);
    int *bR = (int *)(wbuf + THIN_BUILD_SEG * 16), *bS = (int *)(wbuf + THIN_BUILD_SEG * 20);
    Min3 best;
    best.init();
    for (int base = 0; base < ncols; base += 32 * THIN_BUILD_SEG) {
        const int span = min(32 * THIN_BUILD_SEG, ncols - base);
        const int seg_len = ((span + 31) / 32 + 3) & ~3;
        const int c_lo = base + seg * seg_len, cnt = min(base + span, c_lo + seg_len) - c_lo;
        if (cnt > 0) {   // warp-uniform
            if (lane == 0) {
                const uint32_t vb = (uint32_t)((cnt + 1) & ~1) * 8u, rb = (uint32_t)((cnt + 3) & ~3) * 4u;
                mbar_expect_tx(bar, 2u * vb + 2u * rb);
                tma_load_1d(bA, Wi + c_lo, vb, bar);
                tma_load_1d(bB, Wj + c_lo, vb, bar);
                tma_load_1d(bR, a.c.role + c_lo, rb, bar);
                tma_load_1d(bS, a.c.size + c_lo, rb, bar);
            }
            mbar_wait(bar, par & 1u);
            par ^= 1u;
            for (int t = lane; t < cnt; t += 32) {
                if (bR[t] != ROLE_FREE) continue;   // dead, or a side of some pair (phase B)
                const double x = lw(method, bA[t], bB[t], p, (double)bS[t]);
                Wj[c_lo + t] = x;
                best.push(x, c_lo + t);
            }
            __syncwarp();
        }
    }
    if (doB) {   // as the combine lambda of heavy_columns: "lower j first", both mirror entries bit-identical
        const double ps = __dadd_rn(p.s1, p.s2);
        const bool pf = p.j < pq.j;  // p merges first
        Pair F, Sd;
        F.m = pf ? p.m : pq.m; F.s1 = pf ? p.s1 : pq.s1; F.s2 = pf ? p.s2 : pq.s2; F.as = pf ? p.as : pq.as; F.at = pf ? p.at : pq.at;
        Sd.m = pf ? pq.m : p.m; Sd.s1 = pf ? pq.s1 : p.s1; Sd.s2 = pf ? pq.s2 : p.s2; Sd.as = pf ? pq.as : p.as; Sd.at = pf ? pq.at : p.at;
        F.i = F.j = Sd.i = Sd.j = 0;
        const double t1 = lw(method, Ai, pf ? Bi : Aj, F, pf ? pq.s1 : p.s1);
        const double t2 = lw(method, pf ? Aj : Bi, Bj, F, pf ? pq.s2 : p.s2);
        const double x = lw(method, t1, t2, Sd, pf ? ps : __dadd_rn(pq.s1, pq.s2));
        Wj[pq.j] = x;
        best.push(x, pq.j);
    }
    unsigned long long key, bound;
    int idx;
    warp_select(dkey(best.m1), best.i1, dkey(best.m2), best.i2, dkey(best.m3), key, idx, bound);
    Part *pt = &a.parts[(size_t)e * a.max_segs + seg];
    if (lane < LC) { pt->k[lane] = key; pt->i[lane] = idx; }
    if (lane == 0) pt->bound = bound;
}

// one warp, one grid barrier after build_thin_warp: the 32 partial lists of pair e -> candidate list of its merged row
__device__ __noinline__ void build_thin_finish(const RArgs &a, int e) {
    const int lane = threadIdx.x & 31;
    const Part *pp = &a.parts[(size_t)e * a.max_segs + lane];
    unsigned long long pk[LC], kb, key, bound;
    int pi[LC], idx;
#pragma unroll
    for (int t = 0; t < LC; ++t) { pk[t] = __ldcg(&pp->k[t]); pi[t] = __ldcg(&pp->i[t]); }
    kb = __ldcg(&pp->bound);
    combine_partials(pk, pi, kb, key, idx, bound);
    store_list(a, __ldcg(&a.pairs[e].j), key, idx, bound);
}

__global__ void __launch_bounds__(CHAIN_THREADS, 1) kr_chain(RArgs a, ChainCtl ctl) {
    extern __shared__ __align__(128) unsigned char chain_smem[];
    RState *st = a.st;
    const int tid = threadIdx.x, wpc = CHAIN_THREADS / 32, warp = tid >> 5;
    const int nwarps = (int)gridDim.x * wpc;
    // consecutive items go to different SMs
    const int gwarp = warp * (int)gridDim.x + (int)blockIdx.x;
    const int gthread = gwarp * 32 + (tid & 31), gthreads = nwarps * 32;
    // thin-round staging (ctl.thin): 12 KB and one mbarrier per warp, 16 partial lists per CTA
    unsigned char *wbuf = chain_smem + (size_t)warp * THIN_WARP_BYTES;
    unsigned long long *wbar = (unsigned long long *)(chain_smem + (size_t)wpc * THIN_WARP_BYTES) + warp;
    SPart *sparts = (SPart *)(chain_smem + (size_t)wpc * THIN_WARP_BYTES + wpc * sizeof(unsigned long long));
    unsigned wpar = 0;
    if (ctl.thin) {
        if ((tid & 31) == 0) {
            mbar_init(wbar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
    if (blockIdx.x == 0 && tid == 0) st->exit_reason = 0;
    unsigned int target = __ldcg(a.bar);   // all CTAs start from the same value: the previous launch ended on a barrier
    target -= target % gridDim.x;
    for (int it = 0; it < ctl.max_rounds; ++it) {
        const int round = __ldcg(&st->round), par = round & 1;
        Ctr *ct = &st->ctr[par], *nx = &st->ctr[par ^ 1];
        if (__ldcg(&st->done)) break;
        const int nactive = __ldcg(&st->nactive);
        if (ctl.compact && a.ncols >= 256 && (long long)nactive * ctl.cden <= (long long)a.ncols * ctl.cnum) {   // the host compacts and comes back
            if (blockIdx.x == 0 && tid == 0) st->exit_reason = 1;
            break;
        }
        const bool timing = blockIdx.x == 0 && tid == 0;
        unsigned long long t0 = timing ? globaltimer_ns() : 0ull;
        select_rows_ni(a, ct, gthread, gthreads);
        grid_barrier(a.bar, target);
        const int P = __ldcg(&ct->P);
        const int tk = P <= 8 ? 0 : 3;   // thin rounds (the chain phase) and thick rounds are timed apart
        if (timing) { const unsigned long long t1 = globaltimer_ns(); st->phase_ns[tk] += t1 - t0; t0 = t1; st->phase_ns[6 + (tk ? 1 : 0)] += 1; }
        // thin round: the merged rows are built by two whole CTAs each (the last ones of the grid), out of shared memory
        const bool thin_build = ctl.thin && P >= 1 && P <= 4 && (int)gridDim.x >= 4 * P;
        if (thin_build) {
            const int rb = (int)gridDim.x - 1 - (int)blockIdx.x;
            if (rb < 2 * P) build_thin_warp(a, P, rb >> 1, (rb & 1) * wpc + warp, wbuf, wbar, wpar);
            if (ctl.diag && (tid & 31) == 0 && rb < 2 * P) atomicMax(&st->dbgmax[0], globaltimer_ns());
            light_rows_ni(a, ct, P, gwarp, nwarps);
            if (ctl.diag && (tid & 31) == 0 && gwarp * 32 < a.ncols) atomicMax(&st->dbgmax[1], globaltimer_ns());
        } else {
            // the merged rows are built by the LAST warps of the grid, the light rows start at the first: in a thin
            // round no warp does both, so the two run side by side
            const int Sb = heavy_segments(a, P, P, nwarps / 2);
            const int bwarp = nwarps - 1 - gwarp;
            heavy_items_ni(a, P, 0, P, Sb, bwarp, nwarps, nactive);
            if (ctl.diag && (tid & 31) == 0 && bwarp < P * Sb) atomicMax(&st->dbgmax[0], globaltimer_ns());
            light_rows_ni(a, ct, P, gwarp, nwarps);
            if (ctl.diag && (tid & 31) == 0) atomicMax(&st->dbgmax[1], globaltimer_ns());
        }
        grid_barrier(a.bar, target);
        if (timing) {
            const unsigned long long t1 = globaltimer_ns();
            st->phase_ns[tk + 1] += t1 - t0;
            if (ctl.diag && tk == 0) {
                const unsigned long long b = __ldcg(&st->dbgmax[0]), l = __ldcg(&st->dbgmax[1]);
                if (b > t0) st->phase_ns[8] += b - t0;
                if (l > t0) st->phase_ns[9] += l - t0;
            }
            t0 = t1;
        }
        const int total = P + __ldcg(&ct->qlen);
        {
            // the partial lists of the thin builds are complete (one barrier later): one warp per pair finishes them
            if (thin_build && blockIdx.x == gridDim.x - 1 && warp < P) build_thin_finish(a, warp);
            const int R = total - P;
            if (ctl.thin && R <= ctl.thin * (int)gridDim.x) {   // few rows: one CTA per row, everything behind one memory round trip
                for (int k = blockIdx.x; k < R; k += gridDim.x) rescan_row_cta(a, __ldcg(a.rq + k), wbuf, wbar, wpar, sparts);
                if (ctl.diag && (tid & 31) == 0 && (int)blockIdx.x < R) atomicMax(&st->dbgmax[2], globaltimer_ns());
            } else {
                const int Sr = heavy_segments(a, 0, R, nwarps);
                heavy_items_ni(a, P, P, total, Sr, gwarp, nwarps, nactive);
                if (ctl.diag && (tid & 31) == 0 && gwarp < R * Sr) atomicMax(&st->dbgmax[2], globaltimer_ns());
            }
            if (blockIdx.x == 0) {
                retire_pairs(a, P, tid, CHAIN_THREADS);
                __syncthreads();
                if (tid == 0) close_round(a, ct, nx, P, total);
            }
        }
        grid_barrier(a.bar, target);
        if (timing) {
            const unsigned long long t1 = globaltimer_ns();
            st->phase_ns[tk + 2] += t1 - t0;
            if (ctl.diag && tk == 0) {
                const unsigned long long r = __ldcg(&st->dbgmax[2]);
                if (r > t0) st->phase_ns[10] += r - t0;
                st->phase_ns[11] += total - P;
            }
        }
    }
}

// ---- order-preserving compaction -----------------------------------------------------------------
// drop_i: the i side of this round's pairs is dropped as well (kb_rebuild merges and compacts in one pass)
__global__ void __launch_bounds__(1024) kc_map(const int *role, int ncols, int *newidx, int *oldidx, bool drop_i) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int c0 = 0; c0 < ncols; c0 += 1024) {
        const int c = c0 + threadIdx.x;
        const int rc = c < ncols ? role[c] : ROLE_DEAD;
        const bool live = rc != ROLE_DEAD && !(drop_i && rc >= 0 && !(rc & 1));
        const unsigned m = __ballot_sync(0xffffffffu, live);
        if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = __popc(m);
        __syncthreads();
        int before = s_base;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += s_warp[w];
        const int k = before + __popc(m & ((1u << (threadIdx.x & 31)) - 1));
        if (c < ncols) newidx[c] = live ? k : -1;
        if (live) oldidx[k] = c;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int w = 0; w < 32; ++w) t += s_warp[w];
            s_base += t;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) kc_gather(const double *Wold, long long stride_old, double *Wnew,
                                                 long long stride_new, const int *oldidx, int nnew) {
    for (int r = blockIdx.x; r < nnew; r += gridDim.x) {
        const double *src = Wold + (size_t)oldidx[r] * stride_old;
        double *dst = Wnew + (size_t)r * stride_new;
        for (int c = threadIdx.x; c < nnew; c += 256) dst[c] = __ldcs(src + oldidx[c]);
    }
}

__global__ void kc_cols(Cols o, Cols d, const int *newidx, const int *oldidx, int nnew) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnew) return;
    const int c = oldidx[k];
    d.role[k] = ROLE_FREE;
    d.size[k] = o.size[c];
    for (int t = 0; t < LC; ++t) {
        const int nn = o.cidx[(size_t)c * LC + t];
        d.cidx[(size_t)k * LC + t] = nn >= 0 ? newidx[nn] : -1;
        d.cval[(size_t)k * LC + t] = o.cval[(size_t)c * LC + t];
    }
    d.orig[k] = o.orig[c];
    d.fresh[k] = o.fresh[c];
    d.need[k] = o.need[c];
}

// ---- bulk round: merge + compact in ONE pass over the matrix ---------------------------------------
// In the first rounds a fifth or more of all clusters merge at once: kr_light then touches nearly every
// sector of every free row, kr_heavy streams two old rows per pair and scans the new one, and soon after
// a compaction copies everything again.  kb_rebuild does the same arithmetic (same Lance-Williams calls on
// the same operands: free x pair as kr_light, pair x free as phase A, pair x pair by the "lower j first"
// rule of phase B) while it writes the NEXT matrix -- dead and retiring columns dropped, order preserved --
// and takes every row's candidate list on the way: the old matrix is read once, the new one written once.
//   desc[k] = {old column, its role, its size, old column of the pair's i side} of new column k, and the pair's
//   m, s1, s2, as, at per NEW column in colp (kb_desc): all per-column data of a row pass are coalesced streams
//   out of L2, only the i-side entries are gathers (they sit next to the j side for Hi-C shaped input).
__global__ void kb_desc(Cols o, Cols d, const Pair *pairs, const int *oldidx, int nnew, int4 *desc, double *colp,
                        long long cstride) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnew) return;
    const int c = oldidx[k], rc = o.role[c], sz = o.size[c];
    int ci = c, nsz = sz;
    if (rc >= 0) {
        const Pair p = pairs[rc >> 1];
        ci = p.i;
        nsz = (int)(p.s1 + p.s2);
        colp[k] = p.m;
        colp[cstride + k] = p.s1;
        colp[2 * cstride + k] = p.s2;
        colp[3 * cstride + k] = p.as;
        colp[4 * cstride + k] = p.at;
    }
    desc[k] = make_int4(c, rc, sz, ci);
    d.role[k] = ROLE_FREE;
    d.size[k] = nsz;
    d.orig[k] = o.orig[c];
}

template <int METHOD>
__global__ void __launch_bounds__(HEAVY_THREADS, 3) kb_rebuild(RArgs a, Cols d, const int4 *__restrict__ desc,
                                                               const double *__restrict__ colp, long long cstride,
                                                               double *__restrict__ Wn, long long stride_n, int nnew) {
    const int lane = threadIdx.x & 31;
    const int nwarps = (int)gridDim.x * (HEAVY_THREADS / 32), gwarp = (int)blockIdx.x * (HEAVY_THREADS / 32) + (threadIdx.x >> 5);
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    constexpr int U = 2;
    constexpr bool NEED_S = METHOD == M_WARD, NEED_A = METHOD == M_AVERAGE;
    for (int rn = gwarp; rn < nnew; rn += nwarps) {
        const int4 dr = __ldg(desc + rn);
        const bool prow = dr.y >= 0;
        Pair p;
        p.i = dr.w;
        p.j = dr.x;
        p.m = p.s1 = p.s2 = p.as = p.at = 0.0;
        if (prow) {
            p.m = __ldg(colp + rn); p.s1 = __ldg(colp + cstride + rn); p.s2 = __ldg(colp + 2 * cstride + rn);
            p.as = __ldg(colp + 3 * cstride + rn); p.at = __ldg(colp + 4 * cstride + rn);
        }
        const double *Wj = a.W + (size_t)dr.x * a.stride;                  // the row itself / the j side of its pair
        const double *Wi = a.W + (size_t)dr.w * a.stride;                  // the i side (pair rows only)
        const double vR = (double)dr.z, ps = __dadd_rn(p.s1, p.s2);
        double *dst = Wn + (size_t)rn * stride_n;
        Min3 best;
        best.init();
        const int4 dummy = make_int4(dr.x, ROLE_FREE, 1, dr.x);
        int4 dc[U];
#pragma unroll
        for (int u = 0; u < U; ++u) dc[u] = lane + 32 * u < nnew ? __ldg(desc + lane + 32 * u) : dummy;
        for (int c0 = lane; c0 < nnew; c0 += 32 * U) {
            double A[U], B[U], Ai[U], Bi[U], qm[U], qs1[U], qs2[U], qas[U], qat[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int c = c0 + 32 * u;
                const bool pc = dc[u].y >= 0;
                B[u] = __ldg(Wj + dc[u].x);
                A[u] = prow ? __ldg(Wi + dc[u].x) : 0.0;
                Bi[u] = pc ? __ldg(Wj + dc[u].w) : 0.0;
                Ai[u] = pc && prow ? __ldg(Wi + dc[u].w) : 0.0;
                qm[u] = qs1[u] = qs2[u] = qas[u] = qat[u] = 0.0;
                if (pc) {
                    if (NEED_S) { qm[u] = __ldg(colp + c); qs1[u] = __ldg(colp + cstride + c); qs2[u] = __ldg(colp + 2 * cstride + c); }
                    if (NEED_A) { qas[u] = __ldg(colp + 3 * cstride + c); qat[u] = __ldg(colp + 4 * cstride + c); }
                }
            }
            int4 dn[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int cn = c0 + 32 * (U + u);
                dn[u] = cn < nnew ? __ldg(desc + cn) : dummy;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int c = c0 + 32 * u;
                if (c < nnew) {
                    const int rc = dc[u].y;
                    double x;
                    if (c == rn) x = INF;
                    else if (rc < 0) x = prow ? lw(METHOD, A[u], B[u], p, (double)dc[u].z) : B[u];
                    else {
                        Pair pq;
                        pq.i = dc[u].w; pq.j = dc[u].x;
                        pq.m = qm[u]; pq.s1 = qs1[u]; pq.s2 = qs2[u]; pq.as = qas[u]; pq.at = qat[u];
                        if (!prow) x = lw(METHOD, Bi[u], B[u], pq, vR);                 // as kr_light
                        else {                                                          // as phase B of kr_heavy
                            const double Aj = A[u], Bj = B[u];
                            const bool pf = p.j < pq.j;  // p merges first
                            Pair F, Sd;
                            F.m = pf ? p.m : pq.m; F.s1 = pf ? p.s1 : pq.s1; F.s2 = pf ? p.s2 : pq.s2; F.as = pf ? p.as : pq.as; F.at = pf ? p.at : pq.at;
                            Sd.m = pf ? pq.m : p.m; Sd.s1 = pf ? pq.s1 : p.s1; Sd.s2 = pf ? pq.s2 : p.s2; Sd.as = pf ? pq.as : p.as; Sd.at = pf ? pq.at : p.at;
                            F.i = F.j = Sd.i = Sd.j = 0;
                            const double t1 = lw(METHOD, Ai[u], pf ? Bi[u] : Aj, F, pf ? pq.s1 : p.s1);
                            const double t2 = lw(METHOD, pf ? Aj : Bi[u], Bj, F, pf ? pq.s2 : p.s2);
                            x = lw(METHOD, t1, t2, Sd, pf ? ps : __dadd_rn(pq.s1, pq.s2));
                        }
                    }
                    dst[c] = x;
                    if (c != rn) best.push(x, c);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) dc[u] = dn[u];
        }
        unsigned long long key, bound;
        int idx;
        warp_select(dkey(best.m1), best.i1, dkey(best.m2), best.i2, dkey(best.m3), key, idx, bound);
        const size_t base = (size_t)rn * LC;
        const unsigned long long last = __shfl_sync(0xffffffffu, key, KC);
        const unsigned long long lb = bound < last ? bound : last;
        if (lane < KC) {
            const bool keep = idx != INT_MAX && key <= lb;
            d.cval[base + lane] = keep ? dunkey(key) : INF;
            d.cidx[base + lane] = keep ? idx : -1;
        }
        if (lane == KC) {
            d.cval[base + KC] = dunkey(lb);
            d.cidx[base + KC] = -1;   // (unused slot, but kc_cols copies it)
        }
        if (lane == 0) {
            d.fresh[rn] = 1;
            d.need[rn] = 0;
        }
    }
}

// ---- kb_rebuild2: the same pass with the old rows STAGED BY THE TMA UNIT ------------------------------------------------
// kb_rebuild gathers every old entry with an 8-byte load behind an index load: ~170 instructions per new cell, two
// (rows) x two (unroll) loads in flight per lane, 2.0-2.5 TB/s.  Here a CTA builds RB_WARPS new rows in lockstep, one per
// consumer warp, chunk by chunk of RB_CH OLD columns, out of a two-stage shared-memory ring that a producer warp fills
// with cp.async.bulk copies (full / empty mbarriers): per stage the chunk of both old rows of every new row, the chunk's
// column descriptors and the per-column pair data -- metadata is read from L2 once per CTA and chunk instead of once per
// row, and no consumer instruction waits for global memory.
//   * the new columns of a chunk are contiguous (the compaction preserves order): pass 1 takes the FREE columns (copy, or
//     one Lance-Williams step for a merged row) and queues the chunk's PAIR columns, pass 2 works the queue with all lanes
//     on the same code path (one step for a free row, the three "lower j first" steps for a merged row) -- the gather
//     kernel ran both paths for every warp iteration;
//   * both sides of a column pair come out of shared memory (the i side sits a few columns before the j side for Hi-C
//     shaped input; an i side before the chunk start is a plain global load);
//   * the new row segment is assembled in shared memory and leaves in whole 256-byte warp stores.
// Merged rows come first in the row list (they cost about twice a free row), groups of RB_WARPS rows are dealt round
// robin.  Same operands and operation order as kb_rebuild, so the matrix and the merge records are bit-identical.
// (A first version with a private two-stage ring per warp and the descriptors read through L1 measured 1.6x SLOWER than
// the gather kernel: every chunk paid two exposed L2 latencies for its metadata.)
constexpr int RB_CH = 128;           // unit of the chunk table (chunk0): a kernel chunk is RB_CH or 2 * RB_CH old columns
constexpr int RB_WARPS = 8;
template <int CH>
struct __align__(16) RbStage {
    int4 sD[CH];                    // descriptors of the chunk's new columns
    double sP[3][CH + 2];           // pair data of those columns: m, s1, s2 (ward) / as, at (average)
    double rows[RB_WARPS][2][CH];   // old row chunk: [0] the row itself / j side, [1] the i side (merged rows)
};
template <int CH>
struct __align__(16) RbWarp {
    double sOut[CH];
    unsigned short q[CH];
};
template <int CH, int NST>
constexpr size_t rb_smem_bytes() {
    return NST * sizeof(RbStage<CH>) + RB_WARPS * sizeof(RbWarp<CH>) + 2 * NST * sizeof(unsigned long long);
}

// block 0: chunk0[k] = number of surviving columns whose OLD index is below k * RB_CH (oldidx is ascending)
// block 1: rowlist = the new rows of merged clusters (ascending), then the free rows (ascending); P = pairs of the round
__global__ void __launch_bounds__(1024) kb_chunks(const int *oldidx, const int4 *desc, int nnew, int nch, int P, int *chunk0,
                                                   int *rowlist) {
    if (blockIdx.x == 0) {
        for (int k = threadIdx.x; k <= nch; k += blockDim.x) {
            const int key = k * RB_CH;
            int lo = 0, hi = nnew;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (oldidx[mid] < key) lo = mid + 1; else hi = mid;
            }
            chunk0[k] = lo;
        }
        return;
    }
    __shared__ int s_warp[32];
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int c0 = 0; c0 < nnew; c0 += 1024) {
        const int c = c0 + threadIdx.x;
        const bool pr = c < nnew && desc[c].y >= 0;
        const unsigned m = __ballot_sync(0xffffffffu, pr);
        if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = __popc(m);
        __syncthreads();
        int before = s_base;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += s_warp[w];
        const int k = before + __popc(m & ((1u << (threadIdx.x & 31)) - 1));   // merged rows before c
        if (c < nnew) rowlist[pr ? k : P + (c - k)] = c;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int w = 0; w < 32; ++w) t += s_warp[w];
            s_base += t;
        }
        __syncthreads();
    }
}

template <int METHOD, int CH, int NST>
__global__ void __launch_bounds__((RB_WARPS + 1) * 32, 2) kb_rebuild2(RArgs a, Cols d, const int4 *__restrict__ desc,
                                                                       const double *__restrict__ colp, long long cstride,
                                                                       const int *__restrict__ chunk0, int nch_units,
                                                                       const int *__restrict__ rowlist,
                                                                       double *__restrict__ Wn, long long stride_n, int nnew) {
    extern __shared__ __align__(128) unsigned char rb_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    typedef RbStage<CH> Stage;
    typedef RbWarp<CH> WarpMem;
    constexpr int RB_STAGES = NST, CPU = CH / RB_CH;   // chunk-table units per kernel chunk
    Stage *stages = (Stage *)rb_smem;
    WarpMem *wsm = (WarpMem *)(rb_smem + RB_STAGES * sizeof(Stage));
    unsigned long long *full = (unsigned long long *)(rb_smem + RB_STAGES * sizeof(Stage) + RB_WARPS * sizeof(WarpMem));
    unsigned long long *empty = full + RB_STAGES;
    if (threadIdx.x == 0) {
        for (int s = 0; s < RB_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], RB_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    constexpr bool NEED_S = METHOD == M_WARD, NEED_A = METHOD == M_AVERAGE;
    constexpr int NP = NEED_S ? 3 : (NEED_A ? 2 : 0), PBASE = NEED_A ? 3 : 0;   // pair-data arrays staged per chunk
    const int ncols = a.ncols, ngroups = (nnew + RB_WARPS - 1) / RB_WARPS, nch = (ncols + CH - 1) / CH;
    unsigned it = 0;   // stage sequence number: slot = it % RB_STAGES, phase = it / RB_STAGES

    if (warp == RB_WARPS) {   // ---- producer warp: lane l < RB_WARPS feeds row l of the group, the next lanes the metadata
        for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
            const double *Wj = nullptr, *Wi = nullptr;
            bool have = false, prow = false;
            if (lane < RB_WARPS && g * RB_WARPS + lane < nnew) {
                const int4 dr = __ldg(desc + __ldg(rowlist + g * RB_WARPS + lane));
                have = true;
                prow = dr.y >= 0;
                Wj = a.W + (size_t)dr.x * a.stride;
                Wi = a.W + (size_t)dr.w * a.stride;
            }
            for (int k = 0; k < nch; ++k, ++it) {
                const int slot = it % RB_STAGES;
                if (it >= RB_STAGES) mbar_wait(&empty[slot], ((it / RB_STAGES) - 1) & 1u);
                Stage *st = stages + slot;
                const int o0 = k * CH, cnt = min(CH, ncols - o0);
                const uint32_t rbytes = (uint32_t)((cnt + 1) & ~1) * 8u;   // (stride is even: the pad element belongs to the row)
                const int n0 = __ldg(chunk0 + k * CPU), cn = __ldg(chunk0 + min((k + 1) * CPU, nch_units)) - n0;
                const uint32_t pbytes = (uint32_t)(((n0 & 1) + cn + 1) & ~1) * 8u;
                uint32_t mine = 0;
                if (have) mine = prow ? 2u * rbytes : rbytes;
                else if (lane == RB_WARPS && cn > 0) mine = (uint32_t)cn * 16u;
                else if (lane > RB_WARPS && lane <= RB_WARPS + NP && cn > 0) mine = pbytes;
                const uint32_t total = __reduce_add_sync(0xffffffffu, mine);
                if (lane == 0) mbar_expect_tx(&full[slot], total);
                __syncwarp();
                if (have) {
                    tma_load_1d(st->rows[lane][0], Wj + o0, rbytes, &full[slot]);
                    if (prow) tma_load_1d(st->rows[lane][1], Wi + o0, rbytes, &full[slot]);
                } else if (lane == RB_WARPS && cn > 0) {
                    tma_load_1d(st->sD, desc + n0, (uint32_t)cn * 16u, &full[slot]);
                } else if (lane > RB_WARPS && lane <= RB_WARPS + NP && cn > 0) {
                    const int pa = lane - RB_WARPS - 1;
                    tma_load_1d(st->sP[pa], colp + (size_t)(PBASE + pa) * cstride + (n0 & ~1), pbytes, &full[slot]);
                }
            }
        }
        return;
    }

    // ---- consumer warps: one new row each
    WarpMem *ws = wsm + warp;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const unsigned lt_mask = (1u << lane) - 1u;
    for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
        const bool have = g * RB_WARPS + warp < nnew;
        const int rn = have ? __ldg(rowlist + g * RB_WARPS + warp) : -1;
        int4 dr = make_int4(0, ROLE_FREE, 1, 0);
        if (have) dr = __ldg(desc + rn);
        const bool prow = dr.y >= 0;
        Pair p;
        p.i = dr.w;
        p.j = dr.x;
        p.m = p.s1 = p.s2 = p.as = p.at = 0.0;
        if (prow) {
            p.m = __ldg(colp + rn); p.s1 = __ldg(colp + cstride + rn); p.s2 = __ldg(colp + 2 * cstride + rn);
            p.as = __ldg(colp + 3 * cstride + rn); p.at = __ldg(colp + 4 * cstride + rn);
        }
        const double *Wj = a.W + (size_t)dr.x * a.stride, *Wi = a.W + (size_t)dr.w * a.stride;   // (i sides before a chunk)
        const double vR = (double)dr.z, ps = __dadd_rn(p.s1, p.s2);
        double *dst = Wn + (size_t)(have ? rn : 0) * stride_n;
        Min3 best;
        best.init();
        for (int k = 0; k < nch; ++k, ++it) {
            const int slot = it % RB_STAGES;
            const Stage *st = stages + slot;
            const int o0 = k * CH, n0 = __ldg(chunk0 + k * CPU), cn = __ldg(chunk0 + min((k + 1) * CPU, nch_units)) - n0;
            mbar_wait(&full[slot], (it / RB_STAGES) & 1u);
            if (have) {
                const double *sJ = st->rows[warp][0], *sI = st->rows[warp][1];
                const int podd = n0 & 1;
                // pass 1: free columns; pair columns are queued
                int nq = 0;
#pragma unroll 2
                for (int t0 = 0; t0 < cn; t0 += 32) {
                    const int t = t0 + lane, c = n0 + t;
                    const bool in = t < cn;
                    int4 dc = make_int4(0, ROLE_FREE, 1, 0);
                    if (in) dc = st->sD[t];
                    const bool pc = in && dc.y >= 0;
                    if (in && !pc) {
                        const int off = dc.x - o0;
                        const double B = sJ[off];
                        double x = B;
                        if (prow) x = lw(METHOD, sI[off], B, p, (double)dc.z);
                        if (c == rn) x = INF;
                        else best.push(x, c);
                        ws->sOut[t] = x;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, pc);
                    if (pc) ws->q[nq + __popc(m & lt_mask)] = (unsigned short)t;
                    nq += __popc(m);
                }
                __syncwarp();
                // pass 2: the surviving (j side) columns of the round's pairs
                for (int u0 = 0; u0 < nq; u0 += 32) {
                    const int u = u0 + lane;
                    if (u < nq) {
                        const int t = ws->q[u], c = n0 + t;
                        const int4 dc = st->sD[t];
                        const int off = dc.x - o0, ioff = dc.w - o0;
                        Pair pq;
                        pq.i = dc.w; pq.j = dc.x;
                        pq.m = pq.s1 = pq.s2 = pq.as = pq.at = 0.0;
                        if (NEED_S) { pq.m = st->sP[0][podd + t]; pq.s1 = st->sP[1][podd + t]; pq.s2 = st->sP[2][podd + t]; }
                        if (NEED_A) { pq.as = st->sP[0][podd + t]; pq.at = st->sP[1][podd + t]; }
                        const double Bj = sJ[off], Bi = ioff >= 0 ? sJ[ioff] : __ldg(Wj + dc.w);
                        double x;
                        if (!prow) x = lw(METHOD, Bi, Bj, pq, vR);                 // as kr_light
                        else {                                                     // as phase B of kr_heavy
                            const double Aj = sI[off], Ai = ioff >= 0 ? sI[ioff] : __ldg(Wi + dc.w);
                            const bool pf = p.j < pq.j;  // p merges first
                            Pair F, Sd;
                            F.m = pf ? p.m : pq.m; F.s1 = pf ? p.s1 : pq.s1; F.s2 = pf ? p.s2 : pq.s2; F.as = pf ? p.as : pq.as; F.at = pf ? p.at : pq.at;
                            Sd.m = pf ? pq.m : p.m; Sd.s1 = pf ? pq.s1 : p.s1; Sd.s2 = pf ? pq.s2 : p.s2; Sd.as = pf ? pq.as : p.as; Sd.at = pf ? pq.at : p.at;
                            F.i = F.j = Sd.i = Sd.j = 0;
                            const double t1 = lw(METHOD, Ai, pf ? Bi : Aj, F, pf ? pq.s1 : p.s1);
                            const double t2 = lw(METHOD, pf ? Aj : Bi, Bj, F, pf ? pq.s2 : p.s2);
                            x = lw(METHOD, t1, t2, Sd, pf ? ps : __dadd_rn(pq.s1, pq.s2));
                        }
                        if (c == rn) x = INF;
                        else best.push(x, c);
                        ws->sOut[t] = x;
                    }
                }
            }
            __syncwarp();                                   // the stage has been read by every lane
            if (lane == 0) mbar_arrive(&empty[slot]);
            if (have) {
                double *o = dst + n0 + lane;
                const double *si = ws->sOut + lane;
#pragma unroll
                for (int u = 0; u < CH / 32; ++u)
                    if (lane + 32 * u < cn) o[32 * u] = si[32 * u];
            }
            __syncwarp();                                   // the queue and sOut are free again
        }
        if (!have) continue;
        unsigned long long key, bound;
        int idx;
        warp_select(dkey(best.m1), best.i1, dkey(best.m2), best.i2, dkey(best.m3), key, idx, bound);
        const size_t base = (size_t)rn * LC;
        const unsigned long long last = __shfl_sync(0xffffffffu, key, KC);
        const unsigned long long lb = bound < last ? bound : last;
        if (lane < KC) {
            const bool keep = idx != INT_MAX && key <= lb;
            d.cval[base + lane] = keep ? dunkey(key) : INF;
            d.cidx[base + lane] = keep ? idx : -1;
        }
        if (lane == KC) {
            d.cval[base + KC] = dunkey(lb);
            d.cidx[base + KC] = -1;
        }
        if (lane == 0) {
            d.fresh[rn] = 1;
            d.need[rn] = 0;
        }
    }
}

// ---- kb_rebuild3: kb_rebuild2 driven by per-chunk column LISTS ---------------------------------------------------------
// ncu on kb_rebuild2: 5.3 warp instructions per new cell, 60 % of the issue slots -- every row re-derives, for every chunk, what
// depends on the chunk alone: the role of each column (descriptor load + test), the queue of its pair columns (ballot, popc,
// shared-memory append) and the offsets into the staged old rows.  kb_lists does that ONCE per round: for every chunk of 256
// old columns the list of its free new columns and the list of its pair new columns (offset in the new chunk | offset in the
// old chunk << 8, cluster size / i-side column, pair data), padded to 8 entries so that every slice can be bulk-copied.  The
// producer warp stages the chunk's slices next to the old rows; a consumer warp runs two dense loops over them.  Same
// operands and operation order as kb_rebuild / kb_rebuild2: bit-identical matrix and merge records.
constexpr int R3_CH = 256, R3_LIST = R3_CH + 8;
struct __align__(16) R3Stage {
    unsigned short fTO[R3_LIST], pTO[R3_LIST];   // free / pair columns: t | off << 8
    int fSize[R3_LIST], pW[R3_LIST];              // cluster size of a free column; old column of the i side of a pair column
    double pP[3][R3_LIST];                        // pair data of the pair columns: m, s1, s2 (ward) / as, at (average)
    double rows[RB_WARPS][2][R3_CH];              // old row chunk: [0] the row itself / j side, [1] the i side (merged rows)
};
struct __align__(16) R3Warp {
    double sOut[R3_CH];
};
constexpr size_t R3_SMEM = 2 * sizeof(R3Stage) + RB_WARPS * sizeof(R3Warp) + 4 * sizeof(unsigned long long);

struct R3Lists {
    int4 *ctab;                 // [nch + 1]: {first new column, first free-list entry, free columns, first pair-list entry} of a chunk
    unsigned short *fTO, *pTO;  // [cap]
    int *fSize, *pW;            // [cap]
    double *pP;                 // [5][pstride]: m, s1, s2, as, at per pair-list entry
    long long pstride;
};

// kb_lists in three small launches, a WARP per chunk (a first version did everything in one block with a thread per chunk:
// 0.5 ms per round, more than a tenth of the pass it prepares):
//   kb_lists_count: first new column of the chunk (oldidx is ascending: binary search), its free / pair column counts
//   kb_lists_scan : padded list starts (multiples of 8 entries: 16-byte aligned slices of the 2-byte lists)
//   kb_lists_fill : the lists, order preserving (ballot compaction), coalesced
__global__ void __launch_bounds__(256) kb_lists_count(const int *oldidx, const int4 *desc, int nnew, int ncols, R3Lists L) {
    const int nch = (ncols + R3_CH - 1) / R3_CH, lane = threadIdx.x & 31;
    const int k = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (k > nch) return;
    int bound[2] = {0, 0};   // lanes 0 and 1: first new column of chunk k and of chunk k + 1
    if (lane < 2) {
        const int key = (k + lane) * R3_CH;
        int lo = 0, hi = nnew;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (oldidx[mid] < key) lo = mid + 1; else hi = mid;
        }
        bound[0] = lo;
    }
    const int n0 = __shfl_sync(0xffffffffu, bound[0], 0), n1 = __shfl_sync(0xffffffffu, bound[0], 1);
    int np = 0;
    if (k < nch)
        for (int c = n0 + lane; c < n1; c += 32) np += desc[c].y >= 0;
    np = (int)hc::warp_sum_u64((unsigned long long)np);
    if (lane == 0) {
        L.ctab[k].x = n0;
        L.ctab[k].z = k < nch ? (n1 - n0) - np : 0;   // free columns
        L.ctab[k].w = k < nch ? np : 0;               // (pair columns for now; replaced by the list start in kb_lists_scan)
    }
}

__global__ void kb_lists_scan(int ncols, R3Lists L) {
    const int nch = (ncols + R3_CH - 1) / R3_CH;
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int f = 0, q = 0;
        for (int k = 0; k <= nch; ++k) {
            const int nf = L.ctab[k].z, np = L.ctab[k].w;
            L.ctab[k].y = f;
            L.ctab[k].w = q;
            f += (nf + 7) & ~7;
            q += (np + 7) & ~7;
        }
    }
}

__global__ void __launch_bounds__(256) kb_lists_fill(const int4 *desc, const double *colp, long long cstride, int ncols, R3Lists L) {
    const int nch = (ncols + R3_CH - 1) / R3_CH, lane = threadIdx.x & 31;
    const int k = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (k >= nch) return;
    const int4 ct = L.ctab[k];
    const int n0 = ct.x, n1 = L.ctab[k + 1].x, o0 = k * R3_CH;
    int f = ct.y, q = ct.w;
    for (int c0 = n0; c0 < n1; c0 += 32) {
        const int c = c0 + lane;
        const bool in = c < n1;
        int4 dsc = make_int4(0, ROLE_FREE, 0, 0);
        if (in) dsc = desc[c];
        const bool pr = in && dsc.y >= 0, fr = in && dsc.y < 0;
        const unsigned mp = __ballot_sync(0xffffffffu, pr), mf = __ballot_sync(0xffffffffu, fr);
        const unsigned below = (1u << lane) - 1u;
        const unsigned short to = (unsigned short)((c - n0) | ((dsc.x - o0) << 8));
        if (fr) {
            const int pos = f + __popc(mf & below);
            L.fTO[pos] = to;
            L.fSize[pos] = dsc.z;
        }
        if (pr) {
            const int pos = q + __popc(mp & below);
            L.pTO[pos] = to;
            L.pW[pos] = dsc.w;
#pragma unroll
            for (int a = 0; a < 5; ++a) L.pP[a * L.pstride + pos] = colp[a * cstride + c];
        }
        f += __popc(mf);
        q += __popc(mp);
    }
}

template <int METHOD>
__global__ void __launch_bounds__((RB_WARPS + 1) * 32, 2) kb_rebuild3(RArgs a, Cols d, const int4 *__restrict__ desc,
                                                                       const double *__restrict__ colp, long long cstride,
                                                                       R3Lists L, const int *__restrict__ rowlist,
                                                                       double *__restrict__ Wn, long long stride_n, int nnew) {
    extern __shared__ __align__(128) unsigned char rb_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    R3Stage *stages = (R3Stage *)rb_smem;
    R3Warp *wsm = (R3Warp *)(rb_smem + 2 * sizeof(R3Stage));
    unsigned long long *full = (unsigned long long *)(rb_smem + 2 * sizeof(R3Stage) + RB_WARPS * sizeof(R3Warp));
    unsigned long long *empty = full + 2;
    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], RB_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    constexpr bool NEED_S = METHOD == M_WARD, NEED_A = METHOD == M_AVERAGE;
    constexpr int NP = NEED_S ? 3 : (NEED_A ? 2 : 0), PBASE = NEED_A ? 3 : 0;   // pair-data arrays staged per chunk
    const int ncols = a.ncols, ngroups = (nnew + RB_WARPS - 1) / RB_WARPS, nch = (ncols + R3_CH - 1) / R3_CH;
    unsigned it = 0;   // stage sequence number: slot = it & 1, phase = it >> 1

    if (warp == RB_WARPS) {   // ---- producer warp: lane l < RB_WARPS feeds row l of the group, the next lanes the lists
        for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
            const double *Wj = nullptr, *Wi = nullptr;
            bool have = false, prow = false;
            if (lane < RB_WARPS && g * RB_WARPS + lane < nnew) {
                const int4 dr = __ldg(desc + __ldg(rowlist + g * RB_WARPS + lane));
                have = true;
                prow = dr.y >= 0;
                Wj = a.W + (size_t)dr.x * a.stride;
                Wi = a.W + (size_t)dr.w * a.stride;
            }
            const bool any_prow = __any_sync(0xffffffffu, prow);   // (free rows never read the cluster sizes)
            for (int k = 0; k < nch; ++k, ++it) {
                const int slot = it & 1;
                if (it >= 2) mbar_wait(&empty[slot], ((it >> 1) - 1) & 1u);
                R3Stage *st = stages + slot;
                const int o0 = k * R3_CH, cnt = min(R3_CH, ncols - o0);
                const uint32_t rbytes = (uint32_t)((cnt + 1) & ~1) * 8u;   // (stride is even: the pad element belongs to the row)
                const int4 ct = __ldg(L.ctab + k);
                const int nf8 = (ct.z + 7) & ~7, np8 = (((__ldg(&L.ctab[k + 1].x) - ct.x) - ct.z) + 7) & ~7;
                const int li = lane - RB_WARPS;   // list copies: 0 fTO, 1 fSize, 2 pTO, 3 pW, 4.. pP
                uint32_t mine = 0;
                if (have) mine = prow ? 2u * rbytes : rbytes;
                else if (li == 0) mine = (uint32_t)nf8 * 2u;
                else if (li == 1 && any_prow) mine = (uint32_t)nf8 * 4u;
                else if (li == 2) mine = (uint32_t)np8 * 2u;
                else if (li == 3) mine = (uint32_t)np8 * 4u;
                else if (li >= 4 && li < 4 + NP) mine = (uint32_t)np8 * 8u;
                const uint32_t total = __reduce_add_sync(0xffffffffu, mine);
                if (lane == 0) mbar_expect_tx(&full[slot], total);
                __syncwarp();
                if (have) {
                    tma_load_1d(st->rows[lane][0], Wj + o0, rbytes, &full[slot]);
                    if (prow) tma_load_1d(st->rows[lane][1], Wi + o0, rbytes, &full[slot]);
                } else if (mine) {
                    if (li == 0) tma_load_1d(st->fTO, L.fTO + ct.y, mine, &full[slot]);
                    else if (li == 1) tma_load_1d(st->fSize, L.fSize + ct.y, mine, &full[slot]);
                    else if (li == 2) tma_load_1d(st->pTO, L.pTO + ct.w, mine, &full[slot]);
                    else if (li == 3) tma_load_1d(st->pW, L.pW + ct.w, mine, &full[slot]);
                    else tma_load_1d(st->pP[li - 4], L.pP + (size_t)(PBASE + li - 4) * L.pstride + ct.w, mine, &full[slot]);
                }
            }
        }
        return;
    }

    // ---- consumer warps: one new row each
    R3Warp *ws = wsm + warp;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
        const bool have = g * RB_WARPS + warp < nnew;
        const int rn = have ? __ldg(rowlist + g * RB_WARPS + warp) : -1;
        int4 dr = make_int4(0, ROLE_FREE, 1, 0);
        if (have) dr = __ldg(desc + rn);
        const bool prow = dr.y >= 0;
        Pair p;
        p.i = dr.w;
        p.j = dr.x;
        p.m = p.s1 = p.s2 = p.as = p.at = 0.0;
        if (prow) {
            p.m = __ldg(colp + rn); p.s1 = __ldg(colp + cstride + rn); p.s2 = __ldg(colp + 2 * cstride + rn);
            p.as = __ldg(colp + 3 * cstride + rn); p.at = __ldg(colp + 4 * cstride + rn);
        }
        const double *Wj = a.W + (size_t)dr.x * a.stride, *Wi = a.W + (size_t)dr.w * a.stride;   // (i sides before a chunk)
        const double vR = (double)dr.z, ps = __dadd_rn(p.s1, p.s2);
        double *dst = Wn + (size_t)(have ? rn : 0) * stride_n;
        Min3 best;
        best.init();
        for (int k = 0; k < nch; ++k, ++it) {
            const int slot = it & 1;
            const R3Stage *st = stages + slot;
            const int4 ct = __ldg(L.ctab + k);
            const int o0 = k * R3_CH, n0 = ct.x, cn = __ldg(&L.ctab[k + 1].x) - n0, nf = ct.z, np = cn - nf;
            mbar_wait(&full[slot], (it >> 1) & 1u);
            if (have) {
                const double *sJ = st->rows[warp][0], *sI = st->rows[warp][1];
                // pass 1: free columns (copy, or one Lance-Williams step for a merged row)
                if (!prow) {
                    for (int u = lane; u < nf; u += 32) {
                        const int to = st->fTO[u], t = to & 255, c = n0 + t;
                        double x = sJ[to >> 8];
                        if (c == rn) x = INF;
                        else best.push(x, c);
                        ws->sOut[t] = x;
                    }
                } else {
                    for (int u = lane; u < nf; u += 32) {
                        const int to = st->fTO[u], t = to & 255, off = to >> 8, c = n0 + t;
                        double x = lw(METHOD, sI[off], sJ[off], p, (double)st->fSize[u]);
                        if (c == rn) x = INF;
                        else best.push(x, c);
                        ws->sOut[t] = x;
                    }
                }
                // pass 2: the surviving (j side) columns of the round's pairs
                for (int u = lane; u < np; u += 32) {
                    const int to = st->pTO[u], t = to & 255, off = to >> 8, c = n0 + t;
                    const int wcol = st->pW[u], ioff = wcol - o0;
                    Pair pq;
                    pq.i = wcol; pq.j = o0 + off;
                    pq.m = pq.s1 = pq.s2 = pq.as = pq.at = 0.0;
                    if (NEED_S) { pq.m = st->pP[0][u]; pq.s1 = st->pP[1][u]; pq.s2 = st->pP[2][u]; }
                    if (NEED_A) { pq.as = st->pP[0][u]; pq.at = st->pP[1][u]; }
                    const double Bj = sJ[off], Bi = ioff >= 0 ? sJ[ioff] : __ldg(Wj + wcol);
                    double x;
                    if (!prow) x = lw(METHOD, Bi, Bj, pq, vR);                 // as kr_light
                    else {                                                     // as phase B of kr_heavy
                        const double Aj = sI[off], Ai = ioff >= 0 ? sI[ioff] : __ldg(Wi + wcol);
                        const bool pf = p.j < pq.j;  // p merges first
                        Pair F, Sd;
                        F.m = pf ? p.m : pq.m; F.s1 = pf ? p.s1 : pq.s1; F.s2 = pf ? p.s2 : pq.s2; F.as = pf ? p.as : pq.as; F.at = pf ? p.at : pq.at;
                        Sd.m = pf ? pq.m : p.m; Sd.s1 = pf ? pq.s1 : p.s1; Sd.s2 = pf ? pq.s2 : p.s2; Sd.as = pf ? pq.as : p.as; Sd.at = pf ? pq.at : p.at;
                        F.i = F.j = Sd.i = Sd.j = 0;
                        const double t1 = lw(METHOD, Ai, pf ? Bi : Aj, F, pf ? pq.s1 : p.s1);
                        const double t2 = lw(METHOD, pf ? Aj : Bi, Bj, F, pf ? pq.s2 : p.s2);
                        x = lw(METHOD, t1, t2, Sd, pf ? ps : __dadd_rn(pq.s1, pq.s2));
                    }
                    if (c == rn) x = INF;
                    else best.push(x, c);
                    ws->sOut[t] = x;
                }
            }
            __syncwarp();                                   // the stage has been read by every lane
            if (lane == 0) mbar_arrive(&empty[slot]);
            if (have) {
                double *o = dst + n0 + lane;
                const double *si = ws->sOut + lane;
#pragma unroll
                for (int u = 0; u < R3_CH / 32; ++u)
                    if (lane + 32 * u < cn) o[32 * u] = si[32 * u];
            }
            __syncwarp();                                   // sOut is free again
        }
        if (!have) continue;
        unsigned long long key, bound;
        int idx;
        warp_select(dkey(best.m1), best.i1, dkey(best.m2), best.i2, dkey(best.m3), key, idx, bound);
        const size_t base = (size_t)rn * LC;
        const unsigned long long last = __shfl_sync(0xffffffffu, key, KC);
        const unsigned long long lb = bound < last ? bound : last;
        if (lane < KC) {
            const bool keep = idx != INT_MAX && key <= lb;
            d.cval[base + lane] = keep ? dunkey(key) : INF;
            d.cidx[base + lane] = keep ? idx : -1;
        }
        if (lane == KC) {
            d.cval[base + KC] = dunkey(lb);
            d.cidx[base + KC] = -1;
        }
        if (lane == 0) {
            d.fresh[rn] = 1;
            d.need[rn] = 0;
        }
    }
}

// closes a bulk round like the last CTA of kr_heavy does
__global__ void kb_retire(RArgs a, int nnew) {
    RState *st = a.st;
    Ctr *ct = &st->ctr[0];
    const int P = ct->P;
    const unsigned long long nfree = (unsigned long long)(nnew - P), oldcols = (unsigned long long)st->nactive;
    st->stats[0] += nfree;
    st->stats[1] += nfree * oldcols;
    st->stats[2] += (unsigned long long)P;
    st->stats[3] += (unsigned long long)P * oldcols;
    st->stats[4] += nfree * (unsigned long long)P;
    st->stats[7] += nfree * oldcols;
    st->stats[8] += (unsigned long long)P * oldcols;
    st->stats[9] += (unsigned long long)nnew * (unsigned long long)nnew;   // cells of the new matrix written by the pass
    st->n_done += P;
    st->nactive -= P;
    st->stats[5] += ct->deferred;
    if (st->n_done >= a.n - 1) st->done = 1;
    else if (st->ambiguous && !a.diag) st->done = 1;
    st->last_P = P;
    st->last_q = nnew - P;
    ct->P = 0;
    ct->qlen = 0;
    ct->deferred = 0;
    st->round += 1;
}

struct Layout {
    RState *st;
    unsigned int *bar;
    Pair *pairs;
    int *rq, *rec_a, *rec_b, *newidx, *oldidx;
    double *rec_d;
    Cols cols[2];
    double *spare;
    size_t spare_elems;
    Part *parts;
    int *seg_done;
    int4 *desc;
    double *colp;   // [5][n]: m, s1, s2, as, at of the pair behind a new column
    int *chunk0;    // [n / RB_CH + 2]: first new column of every chunk of old columns (kb_rebuild2)
    int *rowlist;   // [n]: new rows in the order kb_rebuild2 builds them (merged rows first)
    long long cstride;  // elements between the five arrays of colp (even: every array 16-byte aligned for the bulk copies)
    R3Lists r3;         // per-chunk column lists of kb_rebuild3
};

// second matrix buffer: a compaction halves the columns, a bulk round (>= 1/5 of the columns merge) keeps <= 4/5
size_t spare_elems_for(int n) {
    const size_t h = (size_t)n * 4 / 5 + 2;
    return h * h;
}

// entries of a per-chunk list array: every column once + up to 7 padding entries per chunk (multiple of 8: aligned strides)
size_t r3_list_cap(int n) { return (((size_t)n + 8 * ((size_t)n / R3_CH + 3)) + 7) & ~(size_t)7; }

bool carve(void *workspace, size_t bytes, int n, Layout &L) {
    hc::Carver ws(workspace, bytes);
    L.st = ws.take<RState>(1);
    L.bar = ws.take<unsigned int>(64);
    L.pairs = ws.take<Pair>((size_t)n / 2 + 1);
    L.rq = ws.take<int>(n);
    L.rec_a = ws.take<int>(n);
    L.rec_b = ws.take<int>(n);
    L.newidx = ws.take<int>(n);
    L.oldidx = ws.take<int>(n);
    L.rec_d = ws.take<double>(n);
    for (int k = 0; k < 2; ++k) {
        L.cols[k].role = ws.take<int>(n);
        L.cols[k].size = ws.take<int>(n);
        L.cols[k].cidx = ws.take<int>((size_t)n * LC);
        L.cols[k].orig = ws.take<int>(n);
        L.cols[k].cval = ws.take<double>((size_t)n * LC);
        L.cols[k].fresh = ws.take<unsigned char>(n);
        L.cols[k].need = ws.take<unsigned char>(n);
    }
    L.parts = ws.take<Part>((size_t)n * MAX_SEGS);
    L.seg_done = ws.take<int>(n);
    L.desc = ws.take<int4>(n);
    L.cstride = ((long long)n + 1) & ~1LL;
    L.colp = ws.take<double>((size_t)L.cstride * 5);
    L.chunk0 = ws.take<int>((size_t)n / RB_CH + 2);
    L.rowlist = ws.take<int>(n);
    {
        const size_t cap = r3_list_cap(n);
        L.r3.ctab = ws.take<int4>((size_t)n / R3_CH + 3);
        L.r3.fTO = ws.take<unsigned short>(cap);
        L.r3.pTO = ws.take<unsigned short>(cap);
        L.r3.fSize = ws.take<int>(cap);
        L.r3.pW = ws.take<int>(cap);
        L.r3.pstride = (long long)cap;
        L.r3.pP = ws.take<double>(cap * 5);
    }
    L.spare_elems = spare_elems_for(n);
    L.spare = ws.take<double>(L.spare_elems);
    return ws.ok();
}

}  // namespace

extern "C" size_t hc_linkage_rounds_workspace_bytes(int32_t n) {
    if (n < 1) n = 1;
    size_t per_col = 4 * 5 + 8 + 2 * (4 * 3 + 12 * LC + 2);
    return 8192 + (size_t)n * per_col + ((size_t)n / 2 + 1) * sizeof(Pair) + 40 * 256 + spare_elems_for(n) * 8 +
           (size_t)n * (MAX_SEGS * sizeof(Part) + 4 + 16 + 40) + 1024 + ((size_t)n / RB_CH + 2) * 4 + (size_t)n * 4 + 1024 +
           ((size_t)n / R3_CH + 3) * 16 + r3_list_cap(n) * (2 + 2 + 4 + 4 + 40) + 8 * 256;
}

// W: n x n float64 full symmetric working matrix (squared distances for ward, +inf diagonal), destroyed.
// Returns HC_OK with Z filled, or HC_EAMBIGUOUS (W is then garbage: rebuild it and call hc_linkage).
// h_stats[24]: rounds, merges, rescans, sum of live columns over rescans, sum over merged-row builds,
// row x pair light updates, compactions, near-tied sorted heights, near-tied pair observations (pairs left
// pending for a round), pairs deferred for an exact rescan.
extern "C" int hc_linkage_rounds(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes,
                                 double *h_Z, int64_t *h_stats, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(W && workspace && h_Z && n >= 1 && method >= 0 && method <= 4, "bad arguments");
    if (h_stats) memset(h_stats, 0, 24 * sizeof(int64_t));
    if (n < 2) return HC_OK;
    Layout L;
    if (workspace_bytes < hc_linkage_rounds_workspace_bytes(n) || !carve(workspace, workspace_bytes, n, L)) {
        hc::set_error("hc_linkage_rounds: workspace too small");
        return HC_EWORKSPACE;
    }
    const int sms = hc::sm_count();
    RArgs a;
    memset(&a, 0, sizeof a);
    a.W = W;
    a.stride = n;
    a.ncols = n;
    a.n = n;
    a.method = method;
    a.c = L.cols[0];
    a.pairs = L.pairs;
    a.rq = L.rq;
    a.rec_a = L.rec_a;
    a.rec_b = L.rec_b;
    a.rec_d = L.rec_d;
    a.st = L.st;
    a.bar = L.bar;
    a.parts = L.parts;
    a.seg_done = L.seg_done;
    a.max_segs = MAX_SEGS;
    a.bulk_segs = 1;  // (32 measured slower at C3: 17.8 vs 13.8 ms in the largest round; kept for experiments)
    if (const char *e = getenv("HC_ROUNDS_BULK_SEGS")) a.bulk_segs = atoi(e);
    a.light8 = 1;   // thin rounds of the persistent kernel: eight lanes per row (HC_ROUNDS_LIGHT8=0: one lane per row)
    if (const char *e = getenv("HC_ROUNDS_LIGHT8")) a.light8 = atoi(e) != 0;
    a.margin = MARGIN;
    double height_tol = HEIGHT_TOL;
    if (const char *e = getenv("HC_ROUNDS_MARGIN")) a.margin = atof(e);
    if (const char *e = getenv("HC_ROUNDS_HEIGHT_TOL")) height_tol = atof(e);
    a.diag = getenv("HC_ROUNDS_DIAG") != nullptr;  // count near-ties instead of stopping at the first (never certifies)
    int cur = 0;          // which Cols set is live
    bool in_spare = false;  // whether the live matrix is in the spare buffer
    const bool compact = getenv("HC_ROUNDS_NOCOMPACT") == nullptr;
    int cnum = 3, cden = 4;   // (measured at C3, chain phase + compactions: 1/2 44.0, 2/3 41.9, 3/4 40.5, 7/8 42.1 ms)
    // compaction threshold: live columns / stored columns <= cnum / cden (HC_ROUNDS_COMPACT=num/den)
    if (const char *e = getenv("HC_ROUNDS_COMPACT")) {
        int x = 0, y = 0;
        if (sscanf(e, "%d/%d", &x, &y) == 2 && x > 0 && y > x) { cnum = x; cden = y; }
    }
    int batch = 8;  // rounds per host synchronisation (a round after the last merge is three empty launches)
    if (const char *e = getenv("HC_ROUNDS_BATCH")) batch = atoi(e) > 0 ? atoi(e) : 8;
    long long compactions = 0;
    const bool trace = getenv("HC_ROUNDS_TRACE") != nullptr;  // diagnostics: one line per round on stderr (syncs every round)
    if (trace) batch = 1;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (trace) { cudaEventCreate(&ev0); cudaEventCreate(&ev1); }

    // HC_ROUNDS_KTIME: CUDA events around every launch (no synchronisation added), summed per kernel into
    // h_stats[10..14] (microseconds: select, light, heavy, compaction, initial scan) -- bench.py uses the shares
    const bool ktime = getenv("HC_ROUNDS_KTIME") != nullptr && !trace;
    std::vector<cudaEvent_t> kev;   // events of the current batch, in launch order
    std::vector<int> kkind;         // kind of the interval that STARTS at event k (-1: none)
    double kus[7] = {0, 0, 0, 0, 0, 0, 0};  // select, light, heavy, compaction, initial scan, kb_rebuild, kr_chain
    size_t kused = 0;
    auto kmark = [&](int kind) {
        if (!ktime) return;
        if (kused == kev.size()) { cudaEvent_t e; cudaEventCreate(&e); kev.push_back(e); kkind.push_back(-1); }
        cudaEventRecord(kev[kused], stream);
        kkind[kused++] = kind;
    };
    auto kflush = [&]() {  // after a stream synchronisation
        for (size_t k = 0; k + 1 < kused; ++k) {
            if (kkind[k] < 0) continue;
            float ms = 0;
            cudaEventElapsedTime(&ms, kev[k], kev[k + 1]);
            kus[kkind[k]] += ms * 1e3;
        }
        kused = 0;
    };
    bool first_pass = true;
    long long n_launches = 1;  // kernels launched (kr_init counted here), reported in h_stats[15]
    int heavy_launches = 0, bulk3 = 2;
    if (const char *e = getenv("HC_ROUNDS_BULK3")) bulk3 = atoi(e);
    int heavy_occ = 2;  // CTAs per SM the kernel is compiled for: 3 (80 registers, some spills) or 2 (128 registers)
    if (const char *e = getenv("HC_ROUNDS_HEAVY_OCC")) heavy_occ = atoi(e);
    auto light_heavy_on = [&](cudaStream_t sq, bool mark) {
        int lg = (a.ncols + LIGHT_THREADS / 32 - 1) / (LIGHT_THREADS / 32);
        if (lg > sms * 8) lg = sms * 8;
        if (mark) kmark(first_pass ? 4 : 1);
        kr_light<<<lg, LIGHT_THREADS, 0, sq>>>(a);
        const int hg = sms * 8 < HEAVY_GRID_MAX ? sms * 8 : HEAVY_GRID_MAX;
        if (mark) kmark(first_pass ? 4 : 2);
        // the first rounds after the initial scan are dominated by merged-row builds, where the 3-CTA build wins
        if (heavy_occ == 2 && !(heavy_launches >= 1 && heavy_launches <= bulk3)) kr_heavy<2><<<hg, HEAVY_THREADS, 0, sq>>>(a);
        else kr_heavy<3><<<hg, HEAVY_THREADS, 0, sq>>>(a);
        if (mark) kmark(-1);
        heavy_launches++;
        n_launches += 2;
        first_pass = false;
    };
    auto light_heavy = [&]() { light_heavy_on(stream, true); };
    auto rounds_on = [&](cudaStream_t sq, bool mark) {
        for (int b = 0; b < batch; ++b) {
            if (mark) kmark(0);
            kr_select<<<(a.ncols + 255) / 256, 256, 0, sq>>>(a);
            n_launches++;
            light_heavy_on(sq, mark);
        }
    };
    // A batch of rounds (3 x batch launches with fixed arguments between two compactions) is captured once into a
    // CUDA graph on a private stream and replayed on the caller's stream: the chain phase is ~1700 launches of
    // 5-20 us kernels, so the launch gaps matter.
    // Measured at C3: 75.6 vs 78.3 ms on one box, 78.6 vs 77.9 ms on another -- within run-to-run noise, so the
    // replay is opt-in (HC_ROUNDS_GRAPH=1).
    bool use_graph = !trace && !ktime && getenv("HC_ROUNDS_GRAPH") != nullptr;
    cudaStream_t cap_stream = nullptr;
    cudaGraphExec_t gexec = nullptr;
    auto drop_graph = [&]() {
        if (gexec) cudaGraphExecDestroy(gexec);
        gexec = nullptr;
    };
    auto build_graph = [&]() -> bool {
        drop_graph();
        if (!cap_stream && cudaStreamCreateWithFlags(&cap_stream, cudaStreamNonBlocking) != cudaSuccess) return false;
        cudaGraph_t g = nullptr;
        if (cudaStreamBeginCapture(cap_stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) return false;
        rounds_on(cap_stream, false);
        if (cudaStreamEndCapture(cap_stream, &g) != cudaSuccess || !g) { cudaGetLastError(); return false; }
        const cudaError_t e = cudaGraphInstantiate(&gexec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) { gexec = nullptr; cudaGetLastError(); return false; }
        return true;
    };
    HC_CUDA_OK(cudaMemsetAsync(L.seg_done, 0, (size_t)n * sizeof(int), stream));
    HC_CUDA_OK(cudaMemsetAsync(L.bar, 0, 256, stream));
    kr_init<<<(n + 255) / 256, 256, 0, stream>>>(a);
    if (trace) cudaEventRecord(ev0, stream);
    light_heavy();  // initial nearest neighbour of every row ("first linkage pass")
    HC_CHECK_LAUNCH();
    if (trace) {
        cudaEventRecord(ev1, stream);
        cudaEventSynchronize(ev1);
        float ms = 0;
        cudaEventElapsedTime(&ms, ev0, ev1);
        fprintf(stderr, "[rounds trace] initial scan us %.1f\n", ms * 1e3f);
    }
    RState h;
    // Bulk rounds: while a fifth or more of the columns merge in one round, merge and compact in one pass
    // (kb_rebuild) instead of kr_light + kr_heavy + a later compaction.  The pair count is only known after
    // kr_select, so these few rounds synchronise once each.  HC_ROUNDS_REBUILD=0 switches the path off.
    int rebuild_min = 2048;  // columns: below this a round is launch latency, not traffic
    if (const char *e = getenv("HC_ROUNDS_REBUILD")) rebuild_min = atoi(e) > 0 ? atoi(e) : INT_MAX;
    long long rebuilds = 0;
    int bulk_div = 5;   // a round is a bulk round while pairs * bulk_div >= columns (HC_ROUNDS_BULK_DIV)
    if (const char *e = getenv("HC_ROUNDS_BULK_DIV")) bulk_div = atoi(e) > 0 ? atoi(e) : 5;
    int rebuild_v = 3, rb_cfg = 0;   // 3 = list-driven TMA kernel, 2 = TMA kernel with per-row role tests, 1 = gather kernel;
    // rb_cfg (version 2 only): 0 = chunks of 256 old columns, 2 stages; 1 = 128 columns, 4 stages
    if (const char *e = getenv("HC_ROUNDS_REBUILD_V")) rebuild_v = atoi(e);
    if (const char *e = getenv("HC_ROUNDS_RB_CFG")) rb_cfg = atoi(e);
    if (rebuild_v >= 2) {
        static bool attr_set[64] = {false};   // per device, once: the kernel needs more than 48 KB of shared memory
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !attr_set[dev]) {
#define HC_RB_ATTR(M)                                                                                                              \
    HC_CUDA_OK(cudaFuncSetAttribute(kb_rebuild2<M, 256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rb_smem_bytes<256, 2>())); \
    HC_CUDA_OK(cudaFuncSetAttribute(kb_rebuild2<M, 128, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rb_smem_bytes<128, 4>())); \
    HC_CUDA_OK(cudaFuncSetAttribute(kb_rebuild3<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)R3_SMEM));
            HC_RB_ATTR(M_SINGLE) HC_RB_ATTR(M_COMPLETE) HC_RB_ATTR(M_AVERAGE) HC_RB_ATTR(M_WEIGHTED) HC_RB_ATTR(M_WARD)
#undef HC_RB_ATTR
            attr_set[dev] = true;
        }
    }
    while (a.ncols >= rebuild_min) {
        if (trace) cudaEventRecord(ev0, stream);
        kmark(0);
        kr_select<<<(a.ncols + 255) / 256, 256, 0, stream>>>(a);
        kmark(-1);
        n_launches++;
        HC_CHECK_LAUNCH();
        HC_CUDA_OK(cudaMemcpyAsync(&h, L.st, sizeof h, cudaMemcpyDeviceToHost, stream));
        HC_CUDA_OK(cudaStreamSynchronize(stream));
        const int nnew = h.nactive - h.ctr[0].P;
        const long long stride_new = (nnew + 1) & ~1;
        const bool bulk = !h.done && !h.ambiguous && (long long)h.ctr[0].P * bulk_div >= a.ncols && nnew >= 2 &&
                          (in_spare || (size_t)nnew * stride_new <= L.spare_elems);
        if (!bulk) {            // finish this round the ordinary way and go on in batches
            light_heavy();
            HC_CHECK_LAUNCH();
            break;
        }
        double *dst = in_spare ? W : L.spare;
        kmark(3);
        kc_map<<<1, 1024, 0, stream>>>(a.c.role, a.ncols, L.newidx, L.oldidx, true);
        kb_desc<<<(nnew + 255) / 256, 256, 0, stream>>>(a.c, L.cols[cur ^ 1], a.pairs, L.oldidx, nnew, L.desc, L.colp, L.cstride);
        // TMA-staged pass where the rows can be bulk-copied (16-byte aligned rows: even stride); HC_ROUNDS_REBUILD_V=1
        // keeps the gather kernel
        const bool tma_ok = (a.stride & 1) == 0 && (((uintptr_t)a.W) & 15) == 0 && rebuild_v >= 2;
        const int nch = (a.ncols + RB_CH - 1) / RB_CH;
        if (tma_ok) kb_chunks<<<2, 1024, 0, stream>>>(L.oldidx, L.desc, nnew, nch, h.ctr[0].P, L.chunk0, L.rowlist);
        if (tma_ok && rebuild_v == 3) {
            const int nch3 = (a.ncols + R3_CH - 1) / R3_CH, lg = (nch3 + 1 + 7) / 8;   // a warp per chunk (+ the sentinel), 8 warps per CTA
            kb_lists_count<<<lg, 256, 0, stream>>>(L.oldidx, L.desc, nnew, a.ncols, L.r3);
            kb_lists_scan<<<1, 32, 0, stream>>>(a.ncols, L.r3);
            kb_lists_fill<<<lg, 256, 0, stream>>>(L.desc, L.colp, L.cstride, a.ncols, L.r3);
        }
        kmark(5);
        {
            const int rg = sms * 12;  // 3 CTAs per SM (80 registers): 12.4 vs 14.2 ms in the largest round at C3
            const int rg2 = sms * 2;
            const Cols dc = L.cols[cur ^ 1];
#define HC_REBUILD_CASE(M)                                                                                                      \
    if (tma_ok && rebuild_v == 3) kb_rebuild3<M><<<rg2, (RB_WARPS + 1) * 32, R3_SMEM, stream>>>(a, dc, L.desc, L.colp, L.cstride, L.r3, L.rowlist, dst, stride_new, nnew); \
    else if (tma_ok && rb_cfg == 1) kb_rebuild2<M, 128, 4><<<rg2, (RB_WARPS + 1) * 32, rb_smem_bytes<128, 4>(), stream>>>(a, dc, L.desc, L.colp, L.cstride, L.chunk0, nch, L.rowlist, dst, stride_new, nnew); \
    else if (tma_ok) kb_rebuild2<M, 256, 2><<<rg2, (RB_WARPS + 1) * 32, rb_smem_bytes<256, 2>(), stream>>>(a, dc, L.desc, L.colp, L.cstride, L.chunk0, nch, L.rowlist, dst, stride_new, nnew); \
    else kb_rebuild<M><<<rg, HEAVY_THREADS, 0, stream>>>(a, dc, L.desc, L.colp, L.cstride, dst, stride_new, nnew);                \
    break;
            switch (method) {
            case M_SINGLE: HC_REBUILD_CASE(M_SINGLE)
            case M_COMPLETE: HC_REBUILD_CASE(M_COMPLETE)
            case M_AVERAGE: HC_REBUILD_CASE(M_AVERAGE)
            case M_WEIGHTED: HC_REBUILD_CASE(M_WEIGHTED)
            default: HC_REBUILD_CASE(M_WARD)
            }
#undef HC_REBUILD_CASE
        }
        kmark(-1);
        kb_retire<<<1, 1, 0, stream>>>(a, nnew);
        HC_CHECK_LAUNCH();
        if (trace) {
            cudaEventRecord(ev1, stream);
            cudaEventSynchronize(ev1);
            float ms = 0;
            cudaEventElapsedTime(&ms, ev0, ev1);
            fprintf(stderr, "[rounds trace] bulk round pairs %d ncols %d -> %d us %.1f\n", h.ctr[0].P, a.ncols, nnew, ms * 1e3f);
        }
        cur ^= 1;
        in_spare = !in_spare;
        a.c = L.cols[cur];
        a.W = dst;
        a.stride = stride_new;
        a.ncols = nnew;
        n_launches += tma_ok ? (rebuild_v == 3 ? 8 : 5) : 4;
        rebuilds++;
    }
    // The rest of the run in cooperative launches of the persistent kernel (one CTA per SM, many rounds per launch);
    // the one-kernel-per-phase path below stays for the per-round trace and HC_ROUNDS_PERSIST=0.
    bool persist = !trace && getenv("HC_ROUNDS_DIAG") == nullptr;
    if (const char *e = getenv("HC_ROUNDS_PERSIST")) persist = persist && atoi(e) != 0;
    int chain_grid = 0;
    int thin_on = 3;   // rescans of a round go one row per CTA while there are at most thin_on rows per CTA
    if (const char *e = getenv("HC_ROUNDS_THIN")) thin_on = atoi(e);
    if (persist) {
        int dev = 0, coop = 0, occ = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (cudaFuncSetAttribute(kr_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CHAIN_SMEM) != cudaSuccess) cudaGetLastError();
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kr_chain, CHAIN_THREADS, CHAIN_SMEM);
        chain_grid = sms;
        if (!coop || occ < 1) persist = false;
    }
    while (persist) {
        ChainCtl ctl;
        ctl.max_rounds = 1 << 20;
        ctl.compact = compact ? 1 : 0;
        ctl.cnum = cnum;
        ctl.cden = cden;
        ctl.diag = getenv("HC_ROUNDS_PHASES") != nullptr;
        // bulk copies need 16-byte aligned rows (even stride); HC_ROUNDS_THIN=0 keeps the segment-per-warp path everywhere
        ctl.thin = ((a.stride & 1) == 0 && (((uintptr_t)a.W) & 15) == 0) ? thin_on : 0;
        if (const char *e = getenv("HC_ROUNDS_CHAIN_MAX")) ctl.max_rounds = atoi(e) > 0 ? atoi(e) : 1;
        void *kargs[] = {(void *)&a, (void *)&ctl};
        kmark(6);
        HC_CUDA_OK(cudaLaunchCooperativeKernel((const void *)kr_chain, dim3(chain_grid), dim3(CHAIN_THREADS), kargs, CHAIN_SMEM, stream));
        kmark(-1);
        n_launches++;
        HC_CUDA_OK(cudaMemcpyAsync(&h, L.st, sizeof h, cudaMemcpyDeviceToHost, stream));
        HC_CUDA_OK(cudaStreamSynchronize(stream));
        kflush();
        if (h.done) break;
        if (compact && a.ncols >= 256 && (long long)h.nactive * cden <= (long long)a.ncols * cnum) {
            const int nnew = h.nactive;
            double *dst = in_spare ? W : L.spare;
            const long long stride_new = (nnew + 1) & ~1;
            HC_REQUIRE(in_spare || (size_t)nnew * stride_new <= L.spare_elems, "compaction buffer too small");
            kmark(3);
            kc_map<<<1, 1024, 0, stream>>>(a.c.role, a.ncols, L.newidx, L.oldidx, false);
            kc_gather<<<nnew < sms * 8 ? nnew : sms * 8, 256, 0, stream>>>(a.W, a.stride, dst, stride_new, L.oldidx, nnew);
            kc_cols<<<(nnew + 255) / 256, 256, 0, stream>>>(a.c, L.cols[cur ^ 1], L.newidx, L.oldidx, nnew);
            kmark(-1);
            HC_CHECK_LAUNCH();
            n_launches += 3;
            cur ^= 1;
            in_spare = !in_spare;
            a.c = L.cols[cur];
            a.W = dst;
            a.stride = stride_new;
            a.ncols = nnew;
            compactions++;
        }
    }
    for (; !persist;) {
        if (trace) cudaEventRecord(ev0, stream);
        // (instantiating a graph costs ~0.1 ms: only worth it for the long epochs of big problems)
        const bool graph_now = use_graph && a.ncols >= 16384;
        if (graph_now && !gexec && !build_graph()) use_graph = false;
        if (graph_now && use_graph) {
            HC_CUDA_OK(cudaGraphLaunch(gexec, stream));
            n_launches += 3 * batch;
        } else rounds_on(stream, true);
        if (trace) cudaEventRecord(ev1, stream);
        HC_CHECK_LAUNCH();
        HC_CUDA_OK(cudaMemcpyAsync(&h, L.st, sizeof h, cudaMemcpyDeviceToHost, stream));
        HC_CUDA_OK(cudaStreamSynchronize(stream));
        kflush();
        if (trace) {
            float ms = 0;
            cudaEventElapsedTime(&ms, ev0, ev1);
            fprintf(stderr, "[rounds trace] round %d pairs %d rescans %d ncols %d nactive %d us %.1f\n", h.round, h.last_P, h.last_q,
                    a.ncols, h.nactive, ms * 1e3f);
        }
        if (h.done) break;
        if (compact && a.ncols >= 256 && (long long)h.nactive * cden <= (long long)a.ncols * cnum) {
            const int nnew = h.nactive;
            double *dst = in_spare ? W : L.spare;
            const long long stride_new = (nnew + 1) & ~1;  // even: rows stay 16-byte aligned for the vector loads
            HC_REQUIRE(in_spare || (size_t)nnew * stride_new <= L.spare_elems, "compaction buffer too small");
            if (trace) cudaEventRecord(ev0, stream);
            kmark(3);
            kc_map<<<1, 1024, 0, stream>>>(a.c.role, a.ncols, L.newidx, L.oldidx, false);
            kc_gather<<<nnew < sms * 8 ? nnew : sms * 8, 256, 0, stream>>>(a.W, a.stride, dst, stride_new, L.oldidx, nnew);
            kc_cols<<<(nnew + 255) / 256, 256, 0, stream>>>(a.c, L.cols[cur ^ 1], L.newidx, L.oldidx, nnew);
            kmark(-1);
            HC_CHECK_LAUNCH();
            if (trace) {
                cudaEventRecord(ev1, stream);
                cudaEventSynchronize(ev1);
                float ms = 0;
                cudaEventElapsedTime(&ms, ev0, ev1);
                fprintf(stderr, "[rounds trace] compaction %d -> %d us %.1f\n", a.ncols, nnew, ms * 1e3f);
            }
            drop_graph();  // the kernel arguments change
            n_launches += 3;
            cur ^= 1;
            in_spare = !in_spare;
            a.c = L.cols[cur];
            a.W = dst;
            a.stride = stride_new;
            a.ncols = nnew;
            compactions++;
        }
    }
    if (getenv("HC_ROUNDS_PHASES"))
        fprintf(stderr, "[rounds phases] persistent kernel (CTA 0 clock, barriers included): %llu thin rounds (<= 8 pairs): select %.1f "
                        "build+light %.1f rescan+retire %.1f us; %llu thick rounds: select %.1f build+light %.1f rescan+retire %.1f us\n",
                h.phase_ns[6], h.phase_ns[0] * 1e-3, h.phase_ns[1] * 1e-3, h.phase_ns[2] * 1e-3,
                h.phase_ns[7], h.phase_ns[3] * 1e-3, h.phase_ns[4] * 1e-3, h.phase_ns[5] * 1e-3);
    if (getenv("HC_ROUNDS_PHASES"))
        fprintf(stderr, "[rounds phases] thin rounds, last warp to finish after the phase started: build items %.1f, light rows %.1f, "
                        "rescan items %.1f us (%llu rescans)\n", h.phase_ns[8] * 1e-3, h.phase_ns[9] * 1e-3, h.phase_ns[10] * 1e-3, h.phase_ns[11]);
    if (trace) { cudaEventDestroy(ev0); cudaEventDestroy(ev1); }
    for (cudaEvent_t e : kev) cudaEventDestroy(e);
    drop_graph();
    if (cap_stream) cudaStreamDestroy(cap_stream);
    const int m = n - 1;
    const bool uncertified = h.ambiguous || h.stuck || h.n_done != m || (a.diag && h.stats[6]);
    if (a.diag) {
        fprintf(stderr, "[hc_linkage_rounds diag] n=%d rounds=%d merges=%d near-tie observations=%llu, relative gaps:", n, h.round,
                h.n_done, h.stats[6]);
        for (unsigned long long k = 0; k < h.stats[6] && k < 16; ++k) fprintf(stderr, " %.3g", h.gaps[k]);
        fprintf(stderr, "\n");
    }
    if (uncertified) {
        hc::set_error("hc_linkage_rounds: %s within %.0e relative after %d merges (round %d): not certified "
                      "[leaves %d(size %d) ~ %d(size %d): d=%.17g/%.17g second minima %.17g %.17g]",
                      h.stuck ? "only near-tied pairs left" : "exact tie", a.margin, h.n_done,
                      h.round, h.amb_r, h.amb_sr, h.amb_b, h.amb_sb, h.amb_v, h.amb_vb, h.amb_secr, h.amb_secb);
        if (h_stats) { h_stats[0] = h.round; h_stats[1] = h.n_done; h_stats[8] = (int64_t)h.stats[6]; }
        if (h.n_done != m) return HC_EAMBIGUOUS;  // (diagnostics mode finishes the run: Z is filled, but NOT certified)
    }
    std::vector<int> ra(m), rb(m);
    std::vector<double> rd(m);
    HC_CUDA_OK(cudaMemcpyAsync(ra.data(), L.rec_a, m * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(rb.data(), L.rec_b, m * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaMemcpyAsync(rd.data(), L.rec_d, m * sizeof(double), cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaStreamSynchronize(stream));
    long long near = 0;
    const double t_host0 = trace ? hc::wall_seconds() : 0.0;
    hc::records_to_Z(n, method == M_WARD, ra.data(), rb.data(), rd.data(), h_Z, height_tol, &near);
    if (trace) fprintf(stderr, "[rounds trace] records_to_Z host us %.1f\n", (hc::wall_seconds() - t_host0) * 1e6);
    if (h_stats) {
        h_stats[0] = h.round;
        h_stats[1] = h.n_done;
        h_stats[2] = (int64_t)h.stats[0];
        h_stats[3] = (int64_t)h.stats[1];
        h_stats[4] = (int64_t)h.stats[3];
        h_stats[5] = (int64_t)h.stats[4];
        h_stats[6] = compactions + rebuilds;
        h_stats[7] = near;
        h_stats[8] = (int64_t)h.stats[6];
        h_stats[9] = (int64_t)h.stats[5];
        for (int k = 0; k < 5; ++k) h_stats[10 + k] = (int64_t)(kus[k] + 0.5);
        h_stats[15] = n_launches;
        h_stats[16] = (int64_t)(kus[5] + 0.5);
        h_stats[17] = (int64_t)(kus[6] + 0.5);
        h_stats[18] = (int64_t)h.stats[7];   // bulk rounds: live columns summed over the free rows of kb_rebuild
        h_stats[19] = (int64_t)h.stats[8];   // ... and over its merged rows
        h_stats[20] = rebuilds;
        h_stats[21] = persist ? 1 : 0;
        h_stats[22] = (int64_t)h.stats[9];   // bulk rounds: cells of the new matrices written
    }
    if (uncertified) return HC_EAMBIGUOUS;
    if (near) {
        hc::set_error("hc_linkage_rounds: %lld sorted heights closer than %.0e relative: merge order not certified", near,
                      height_tol);
        return HC_EAMBIGUOUS;
    }
    return HC_OK;
}
