// K1: read pair -> window binning (scatter-add into the dense contact matrix).
//
// Replaces the Python per-line loop of bam2arrays (reference bam2clusters.py:128-166), with the
// window arithmetic of get_window (:119-126) and the record filter of `samtools view -q -F`
// (:108-117).  Semantics per record: SURVEY.md appendix C.
//
// B200 design.  The matrix row a record lands in is the window of READ 1, and a coordinate-sorted
// BAM delivers read-1 records in (nearly) ascending window order.  So instead of scattering
// 64-byte read-modify-writes over a 10 GB matrix, a CTA owns a tile of R complete matrix rows in
// its 200+ KB of shared memory (one 50 000-column u32 row = 200 KB fits an sm_100 SM), streams
// exactly the slice of the pair stream that falls in those rows, accumulates with shared-memory
// atomics and writes every row ONCE with plain coalesced stores (no memset, no global atomics).
// HBM traffic = 24 B/pair + 4 B/cell = the algorithmic minimum.
//
// The slice of the stream belonging to a row tile is found from one sampled key per 256-record
// chunk (prefix-max repaired, so ANY input order stays correct): records that fall outside the
// rows their chunk advertises are "stragglers", appended to an overflow list and applied with
// global atomics afterwards; if the list overflows (input not sorted at all) the whole call is
// redone by the plain global-atomic kernel.
#include <limits.h>

#include <algorithm>

#include "common.cuh"

namespace {

constexpr int CHUNK_LOG = 8;
constexpr int CHUNK = 1 << CHUNK_LOG;  // records per chunk
constexpr int SCAN_BLOCK = 4096;       // chunks per scan block

struct BinArgs {
    const uint2 *pairs;
    long long n_pairs;
    const int32_t *offsets, *member;
    const uint8_t *region_mask;
    int32_t n_contigs;
    uint32_t window_bp;
    unsigned long long inv_bp;  // ceil(2^64 / window_bp): exact division of any u32 by one mul.hi.u64
    int32_t n_windows, mapq_min;
    uint32_t skip_flag;
    int32_t row0, row1;
    uint32_t *counts;
    int accumulate;  // add to the existing matrix instead of overwriting it
    // tiling
    int32_t R, W, n_groups, n_segs;
    long long n_chunks;
    const int32_t *chunk_lo, *chunk_hi;  // group-aligned row bounds advertised by each chunk
    const long long *tile_clo, *tile_chi;  // chunk range per row group
    uint2 *overflow;                    // stragglers: records whose row their chunk does not advertise
    unsigned long long *overflow_count;
    long long overflow_cap;
    uint2 *spill;                       // 16-bit tile cells that handed 16 384 counts over (<= n_ok / 16 384 entries)
    unsigned long long *spill_count;
    long long spill_cap;
    unsigned long long *stats;
    int stream_evict_first;  // band kernel: the record stream is read once -- keep it from evicting the rows resident in L2
    int row_evict_last;      // band kernel: zero stores and far-hit reductions of a row carry the evict_last L2 priority
};

enum { ST_FILTERED = 0, ST_OK = 1, ST_ABSENT = 2, ST_RANGE = 3, ST_KEY = 4 };

// p / window_bp for p < 2^32, exact (p * (ceil(2^64/d) - 2^64/d) < 2^32 <= 2^64/d)
__device__ __forceinline__ uint32_t div_bp(const BinArgs &a, uint32_t p) {
    return a.window_bp == 1u ? p : (uint32_t)__umul64hi((unsigned long long)p, a.inv_bp);
}

__device__ __forceinline__ int decode(const BinArgs &a, uint2 q0, uint2 q1, uint2 q2, int32_t &row,
                                      int32_t &col) {
    const int32_t ref = (int32_t)q0.x, pos = (int32_t)q0.y;
    const int32_t nref = (int32_t)q1.x, npos = (int32_t)q1.y;
    const int32_t lseq = (int32_t)q2.x;
    const uint32_t flag = q2.y & 0xffffu, mapq = (q2.y >> 16) & 0xffu;
    if ((flag & a.skip_flag) || (int32_t)mapq < a.mapq_min) return ST_FILTERED;
    const bool ref_ok = (uint32_t)ref < (uint32_t)a.n_contigs;
    if (a.region_mask && !(ref_ok && a.region_mask[ref])) return ST_FILTERED;
    const bool nref_ok = (uint32_t)nref < (uint32_t)a.n_contigs;
    if (!ref_ok || !nref_ok) return ST_ABSENT;
    if (__ldg(a.member + ref) < 0 || __ldg(a.member + nref) < 0) return ST_ABSENT;
    const int32_t o1 = __ldg(a.offsets + ref), o2 = __ldg(a.offsets + nref);
    if (o1 < 0 || o2 < 0) return ST_KEY;
    const int32_t half = (lseq > 1 ? lseq : 1) >> 1;  // len("*") == 1; Py2 integer division
    const long long p1 = (long long)pos + 1 + half, p2 = (long long)npos + 1 + half;
    if (p1 < 0 || p2 < 0) return ST_RANGE;
    const long long w1 = (long long)div_bp(a, (uint32_t)p1) + o1;
    const long long w2 = (long long)div_bp(a, (uint32_t)p2) + o2;
    if (w1 >= a.n_windows || w2 >= a.n_windows) return ST_RANGE;
    row = (int32_t)w1;
    col = (int32_t)w2;
    return ST_OK;
}

// p / window_bp for p < 2^32 with 32-bit multiplies only: hi64(p * inv) = (p * inv_hi + hi32(p * inv_lo)) >> 32
__device__ __forceinline__ uint32_t div_bp32(const BinArgs &a, uint32_t p) {
    if (a.window_bp == 1u) return p;
    const unsigned long long lo = (unsigned long long)p * (uint32_t)a.inv_bp;
    return (uint32_t)(((unsigned long long)p * (uint32_t)(a.inv_bp >> 32) + (lo >> 32)) >> 32);
}

// decode() for the tile kernel: the four table look-ups are issued together and the status is two predicates
// (filtered, ok); what kind of error a rejected record is gets sorted out by decode_reject() on the rare path.
template <bool SAME_MEMBER = false>
__device__ __forceinline__ void decode_lean(const BinArgs &a, uint2 q0, uint2 q1, uint2 q2, int32_t &row,
                                            int32_t &col, bool &filtered, bool &ok) {
    const int32_t ref = (int32_t)q0.x, pos = (int32_t)q0.y;
    const int32_t nref = (int32_t)q1.x, npos = (int32_t)q1.y;
    const int32_t lseq = (int32_t)q2.x;
    const uint32_t flag = q2.y & 0xffffu, mapq = (q2.y >> 16) & 0xffu;
    const bool ref_ok = (uint32_t)ref < (uint32_t)a.n_contigs, nref_ok = (uint32_t)nref < (uint32_t)a.n_contigs;
    const int32_t o1 = ref_ok ? __ldg(a.offsets + ref) : -1, o2 = nref_ok ? __ldg(a.offsets + nref) : -1;
    // (the membership table of bam2clusters.py:151 is the window table itself unless several window sizes are binned)
    const int32_t m1 = SAME_MEMBER ? o1 : (ref_ok ? __ldg(a.member + ref) : -1), m2 = SAME_MEMBER ? o2 : (nref_ok ? __ldg(a.member + nref) : -1);
    filtered = (flag & a.skip_flag) || (int32_t)mapq < a.mapq_min;
    if (a.region_mask) filtered = filtered || !(ref_ok && a.region_mask[ref]);
    const int32_t add = 1 + ((lseq > 1 ? lseq : 1) >> 1);  // len("*") == 1; Py2 integer division
    const long long p1 = (long long)pos + add, p2 = (long long)npos + add;  // < 2^32
    const uint32_t nw = (uint32_t)a.n_windows;
    const uint32_t d1 = div_bp32(a, (uint32_t)p1), d2 = div_bp32(a, (uint32_t)p2);
    const uint32_t w1 = d1 + (uint32_t)o1, w2 = d2 + (uint32_t)o2;  // no wrap when d < nw and o >= 0
    ok = !filtered && (m1 | m2 | o1 | o2) >= 0 && (p1 | p2) >= 0 && d1 < nw && d2 < nw && w1 < nw && w2 < nw;
    row = (int32_t)w1;
    col = (int32_t)w2;
}

// sampled key of a chunk: window of read 1 of the chunk's first record, -1 if it has none
__global__ void k_lead(BinArgs a, int32_t *lead) {
    long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= a.n_chunks) return;
    const uint2 *p = a.pairs + 3 * (c << CHUNK_LOG);
    uint2 q0 = __ldg(p), q2 = __ldg(p + 2);
    int32_t ref = (int32_t)q0.x, key = -1;
    if ((uint32_t)ref < (uint32_t)a.n_contigs) {
        int32_t o = __ldg(a.offsets + ref);
        int32_t lseq = (int32_t)q2.x;
        long long p1 = (long long)(int32_t)q0.y + 1 + ((lseq > 1 ? lseq : 1) >> 1);
        if (o >= 0 && p1 >= 0) {
            long long w = (long long)div_bp(a, (uint32_t)p1) + o;
            key = (int32_t)(w < a.n_windows ? w : a.n_windows - 1);
        }
    }
    lead[c] = key;
}

__device__ __forceinline__ int32_t block_max(int32_t v, int32_t *sh) {
    for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        int32_t w = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : INT_MIN;
        for (int o = 16; o; o >>= 1) w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
        if (threadIdx.x == 0) sh[0] = w;
    }
    __syncthreads();
    int32_t r = sh[0];
    __syncthreads();
    return r;
}

__global__ void k_scan_blockmax(const int32_t *lead, long long n, int32_t *bmax) {
    __shared__ int32_t sh[32];
    long long base = (long long)blockIdx.x * SCAN_BLOCK;
    int32_t v = INT_MIN;
    for (int i = threadIdx.x; i < SCAN_BLOCK && base + i < n; i += blockDim.x) v = max(v, lead[base + i]);
    v = block_max(v, sh);
    if (threadIdx.x == 0) bmax[blockIdx.x] = v;
}

// exclusive prefix max over the (few) block maxima, single thread block
__global__ void k_scan_carry(int32_t *bmax, int nb) {
    if (threadIdx.x == 0) {
        int32_t run = -1;
        for (int b = 0; b < nb; ++b) {
            int32_t v = bmax[b];
            bmax[b] = run;
            run = max(run, v);
        }
    }
}

// inclusive prefix max inside each scan block (1024 threads x 4 elements), then the
// group-aligned row bounds every chunk advertises
__global__ void k_scan_apply(int32_t *lead, long long n, const int32_t *bmax) {
    __shared__ int32_t sh[1024];
    long long base = (long long)blockIdx.x * SCAN_BLOCK + threadIdx.x * 4;
    int32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (base + k < n) ? lead[base + k] : INT_MIN;
    v[1] = max(v[1], v[0]);
    v[2] = max(v[2], v[1]);
    v[3] = max(v[3], v[2]);
    sh[threadIdx.x] = v[3];
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        int32_t t = threadIdx.x >= o ? sh[threadIdx.x - o] : INT_MIN;
        __syncthreads();
        sh[threadIdx.x] = max(sh[threadIdx.x], t);
        __syncthreads();
    }
    int32_t carry = max(bmax[blockIdx.x], threadIdx.x ? sh[threadIdx.x - 1] : INT_MIN);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (base + k < n) lead[base + k] = max(v[k], carry);
}

__global__ void k_chunk_bounds(BinArgs a, const int32_t *leadm, int32_t *chunk_lo, int32_t *chunk_hi) {
    long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= a.n_chunks) return;
    int32_t lo = leadm[c], hi = (c + 1 < a.n_chunks) ? leadm[c + 1] : INT_MAX;
    lo = min(max(lo, a.row0), a.row1 - 1);
    hi = min(max(hi, a.row0), a.row1 - 1);
    chunk_lo[c] = a.row0 + (lo - a.row0) / a.R * a.R;
    chunk_hi[c] = a.row0 + ((hi - a.row0) / a.R + 1) * a.R;
}

// chunk range [clo, chi) intersecting each row group: chunk_lo[c] <= r0 < chunk_hi[c]
__global__ void k_tile_ranges(BinArgs a, long long *tile_clo, long long *tile_chi) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= a.n_groups) return;
    int32_t r0 = a.row0 + g * a.R;
    long long lo = 0, hi = a.n_chunks;
    while (lo < hi) {  // first c with chunk_hi[c] > r0
        long long m = (lo + hi) >> 1;
        if (a.chunk_hi[m] > r0) hi = m; else lo = m + 1;
    }
    tile_clo[g] = lo;
    lo = 0, hi = a.n_chunks;
    while (lo < hi) {  // first c with chunk_lo[c] > r0
        long long m = (lo + hi) >> 1;
        if (a.chunk_lo[m] > r0) hi = m; else lo = m + 1;
    }
    tile_chi[g] = lo;
}

// Main kernel: persistent CTAs, each walks over (row group, column segment) tiles with stride
// gridDim.x.  Per tile: stream the pair slice, accumulate into the shared-memory tile with
// WARP-AGGREGATED atomics (Hi-C pairs pile up on the diagonal cell: without aggregation a third of
// a row's records serialise on one shared-memory address), then one fused flush+zero pass that
// writes every matrix cell of the tile exactly once with 16-byte coalesced stores.  The stores of
// tile t drain while the records of tile t+1 are being loaded.
// CELL16: tile cells are 16 bits, two per shared-memory word, so a 50 000-column row takes 100 KB and TWO CTAs
// (512 threads each) share an SM: one streams its records while the other flushes.  A cell never overflows its
// half word: the add that carries the stored value across a multiple of 16 384 takes 16 384 back out and files it
// in the overflow list (weight bit), so the stored value stays below 16 384 + what is in flight (< 65 536).
template <int UNROLL, bool CELL16>
__global__ void __launch_bounds__(CELL16 ? 512 : 1024, CELL16 ? 2 : 1) k_bin_tiles(BinArgs a) {
    extern __shared__ __align__(16) uint32_t tile[];
    const int T = blockDim.x, lane_id = threadIdx.x & 31;
    const int n_tiles = a.n_groups * a.n_segs;
    const int max_cells = CELL16 ? (a.R * a.W + 1) / 2 : a.R * a.W;  // shared-memory words
    for (int i = threadIdx.x; i < max_cells; i += T) tile[i] = 0;
    __syncthreads();
    uint32_t n_lines = 0, n_ok = 0, n_absent = 0, n_range = 0, n_key = 0, n_over = 0, n_outside = 0;
    const bool vec_ok = ((a.n_windows & 3) == 0) && ((a.W & 3) == 0) && ((((uintptr_t)a.counts) & 15) == 0);

    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const int g = t / a.n_segs, s = t - g * a.n_segs;
        const int32_t r0 = a.row0 + g * a.R, r1 = min(r0 + a.R, a.row1);
        const int32_t c0 = s * a.W, c1 = min(c0 + a.W, a.n_windows);
        const int W = c1 - c0;
        const long long clo = a.tile_clo[g], chi = a.tile_chi[g];
        const long long beg = clo << CHUNK_LOG;
        const long long end = min(chi << CHUNK_LOG, a.n_pairs);
        // Software-pipelined record stream: each thread keeps two half batches (HB records each) in registers and
        // loads the next half batch BEFORE it works on the current one, so the record loads of a warp are always
        // in flight while it decodes / adds (warp-uniform trip counts: the ballots / matches need the whole warp).
        constexpr int HB = UNROLL == 3 ? 1 : UNROLL / 2;  // UNROLL 3: three one-record stages
        auto fetch = [&](long long wb, uint2 (&q)[HB][3], int32_t (&lo_)[HB], int32_t (&hi_)[HB]) {
#pragma unroll
            for (int u = 0; u < HB; ++u) {
                const long long idx = wb + lane_id + (long long)u * T;
                if (idx < end) {
                    const uint2 *p = a.pairs + 3 * idx;
                    q[u][0] = __ldg(p);
                    q[u][1] = __ldg(p + 1);
                    q[u][2] = __ldg(p + 2);
                    const long long c = idx >> CHUNK_LOG;
                    lo_[u] = __ldg(a.chunk_lo + c);
                    hi_[u] = __ldg(a.chunk_hi + c);
                } else {
                    q[u][0] = q[u][1] = q[u][2] = make_uint2(0, 0);
                    lo_[u] = hi_[u] = 0;
                }
            }
        };
        auto work = [&](long long idx, uint2 q0, uint2 q1, uint2 q2, int32_t lo_in, int32_t hi_in) {
                const bool live = idx < end;
                int32_t row = 0, col = 0;
                bool filtered, ok;
                decode_lean(a, q0, q1, q2, row, col, filtered, ok);
                filtered = filtered || !live;
                ok = ok && live;
                const int32_t lo = lo_in, hi = hi_in;
                const bool owner = live && (lo == r0) && (s == 0);  // this tile accounts for the record
                if (owner) {
                    n_lines += !filtered;
                    n_ok += ok;
                    if (!filtered && !ok) {  // rare: which error the reference would raise / skip
                        int32_t rr, cc;
                        const int st = decode(a, q0, q1, q2, rr, cc);
                        n_absent += st == ST_ABSENT;
                        n_range += st == ST_RANGE;
                        n_key += st == ST_KEY;
                    }
                }
                const bool mine = ok && row >= r0 && row < r1 && col >= c0 && col < c1;
                // warp-aggregated shared-memory atomics: one add per distinct cell per warp
                const uint32_t cell = mine ? (uint32_t)((row - r0) * W + (col - c0)) : 0xffffffffu;
                const unsigned peers = __match_any_sync(0xffffffffu, cell);
                bool spill = false;
                if (mine && (__ffs(peers) - 1) == lane_id) {
                    const uint32_t inc = (uint32_t)__popc(peers);
                    if (CELL16) {
                        const uint32_t sh = (cell & 1u) * 16u;
                        const uint32_t h = (atomicAdd(&tile[cell >> 1], inc << sh) >> sh) & 0xffffu;
                        if (((h + inc) >> 14) != (h >> 14)) {
                            atomicSub(&tile[cell >> 1], 0x4000u << sh);
                            spill = true;
                        }
                    } else {
                        atomicAdd(&tile[cell], inc);
                    }
                }
                if (CELL16) {
                    const unsigned ms = __ballot_sync(0xffffffffu, spill);
                    if (ms) {
                        const int leader = __ffs(ms) - 1;
                        unsigned long long slot = 0;
                        if (lane_id == leader) slot = atomicAdd(a.spill_count, (unsigned long long)__popc(ms));
                        slot = __shfl_sync(0xffffffffu, slot, leader);
                        if (spill) {
                            slot += __popc(ms & ((1u << lane_id) - 1));
                            if ((long long)slot < a.spill_cap) a.spill[slot] = make_uint2((uint32_t)row, col);
                        }
                    }
                }
                // straggler: row not advertised by its chunk -> nobody's tile sees it
                const bool in_block = row >= a.row0 && row < a.row1;
                n_outside += owner && ok && !in_block;  // binnable, but not a row of this call's block
                const bool strag = owner && ok && (row < lo || row >= hi) && in_block;
                const unsigned m = __ballot_sync(0xffffffffu, strag);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    unsigned long long slot = 0;
                    if (lane_id == leader) slot = atomicAdd(a.overflow_count, (unsigned long long)__popc(m));
                    slot = __shfl_sync(0xffffffffu, slot, leader);
                    if (strag) {
                        slot += __popc(m & ((1u << lane_id) - 1));
                        if ((long long)slot < a.overflow_cap) a.overflow[slot] = make_uint2(row, col);
                        n_over++;
                    }
                }
        };
        auto work_half = [&](long long wb, uint2 (&q)[HB][3], int32_t (&lo_)[HB], int32_t (&hi_)[HB]) {
#pragma unroll
            for (int u = 0; u < HB; ++u) work(wb + lane_id + (long long)u * T, q[u][0], q[u][1], q[u][2], lo_[u], hi_[u]);
        };
        {
            const long long step = (long long)T * HB;
            long long wb = beg + (threadIdx.x - lane_id);
            uint2 qa[HB][3], qb[HB][3];
            int32_t la[HB], ha[HB], lb[HB], hb[HB];
            fetch(wb, qa, la, ha);
            if (UNROLL == 3) {  // loads run two records ahead of the decode
                uint2 qc[HB][3];
                int32_t lc[HB], hc_[HB];
                fetch(wb + step, qb, lb, hb);
                for (; wb < end; wb += 3 * step) {
                    fetch(wb + 2 * step, qc, lc, hc_);
                    work_half(wb, qa, la, ha);
                    fetch(wb + 3 * step, qa, la, ha);
                    if (wb + step < end) work_half(wb + step, qb, lb, hb);
                    fetch(wb + 4 * step, qb, lb, hb);
                    if (wb + 2 * step < end) work_half(wb + 2 * step, qc, lc, hc_);
                }
            } else
            for (; wb < end; wb += 2 * step) {
                fetch(wb + step, qb, lb, hb);
                work_half(wb, qa, la, ha);
                fetch(wb + 2 * step, qa, la, ha);
                if (wb + step < end) work_half(wb + step, qb, lb, hb);
            }
        }
        __syncthreads();
        // fused flush + zero: every tile cell is written to the matrix exactly once, coalesced
        const int rows = r1 - r0;
        if (CELL16) {  // (the host picks this kernel only when vec_ok, c0 % 4 == 0 and W % 4 == 0)
            const int W4 = W >> 2;
            if (rows == 1 && !a.accumulate) {  // the N > ~29 000 case: one matrix row per tile
                uint2 *src = (uint2 *)tile;
                uint4 *dst = (uint4 *)(a.counts + (size_t)(r0 - a.row0) * a.n_windows + c0);
#pragma unroll 4
                for (int i = threadIdx.x; i < W4; i += T) {
                    const uint2 w = src[i];
                    src[i] = make_uint2(0, 0);
                    __stcs(dst + i, make_uint4(w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16));
                }
            } else
            for (int i = threadIdx.x; i < rows * W4; i += T) {
                const int rr = rows == 1 ? 0 : i / W4, cc = i - rr * W4;
                uint2 *src = (uint2 *)(tile + rr * (W >> 1)) + cc;  // four 16-bit cells
                uint4 *dst = (uint4 *)(a.counts + (size_t)(r0 - a.row0 + rr) * a.n_windows + c0) + cc;
                const uint2 w = *src;
                uint4 v = make_uint4(w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16);
                if (a.accumulate) {
                    const uint4 o = *dst;
                    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                }
                __stcs(dst, v);
                *src = make_uint2(0, 0);
            }
        } else if (vec_ok && ((c0 & 3) == 0) && ((W & 3) == 0) && rows == 1) {
            // one matrix row per tile (N > ~29 000): no index arithmetic in the loop
            const int W4 = W >> 2;
            uint4 *src = (uint4 *)tile;
            uint4 *dst = (uint4 *)(a.counts + (size_t)(r0 - a.row0) * a.n_windows + c0);
            if (a.accumulate) {
                for (int i = threadIdx.x; i < W4; i += T) {
                    uint4 v = src[i];
                    const uint4 o = dst[i];
                    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                    __stcs(dst + i, v);
                    src[i] = make_uint4(0, 0, 0, 0);
                }
            } else {
#pragma unroll 4
                for (int i = threadIdx.x; i < W4; i += T) {
                    __stcs(dst + i, src[i]);
                    src[i] = make_uint4(0, 0, 0, 0);
                }
            }
        } else if (vec_ok && ((c0 & 3) == 0) && ((W & 3) == 0)) {
            const int W4 = W >> 2;
            for (int i = threadIdx.x; i < rows * W4; i += T) {
                const int rr = i / W4, cc = i - rr * W4;
                uint4 *src = (uint4 *)(tile + rr * W) + cc;
                uint4 *dst = (uint4 *)(a.counts + (size_t)(r0 - a.row0 + rr) * a.n_windows + c0) + cc;
                uint4 v = *src;
                if (a.accumulate) {
                    const uint4 o = *dst;
                    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                }
                __stcs(dst, v);
                *src = make_uint4(0, 0, 0, 0);
            }
        } else {
            for (int i = threadIdx.x; i < rows * W; i += T) {
                const int rr = i / W, cc = i - rr * W;
                uint32_t *dst = a.counts + (size_t)(r0 - a.row0 + rr) * a.n_windows + c0 + cc;
                uint32_t v = tile[i];
                if (a.accumulate) v += *dst;
                *dst = v;
                tile[i] = 0;
            }
        }
        __syncthreads();
    }
    unsigned long long v[7] = {n_lines, n_ok, n_absent, n_range, n_key, n_over, n_outside};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        unsigned long long t = hc::warp_sum_u64(v[k]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(a.stats + (k == 6 ? 10 : k), t);
    }
}

// ---- band kernel: one matrix row per tile, records staged by TMA --------------------------------------------------
// For matrices too wide for two u32 rows per SM (n_windows > ~29 000) the row tile above needs 16-bit cells and
// leaves no shared memory for staging the record stream: its loads live in registers and the kernel stalls on them
// (ncu: 29 % of the stall samples on the first use of the record loads, 46 % issue slots busy).
// Hi-C pairs are near-diagonal: at C3 ~80 % of the records of a row fall within +-1024 windows of it.  This kernel
// keeps only that BAND of the row in shared memory (u32 cells, 8 KB) and
//   * zeroes the rest of the row in global memory itself (plain 16-byte stores: the row's 200 KB stay in the 126 MB
//     L2 while the ~300 resident CTAs work on ~300 rows = 60 MB),
//   * adds the far hits (trans contacts, long-range cis) with global reductions (RED.ADD), resolved in L2 on the
//     lines it has just written, so DRAM still sees every matrix cell once,
//   * streams the records through a ring of shared-memory stages filled by the TMA unit (cp.async.bulk + mbarrier):
//     1024 records per stage, BAND_STAGES stages, one elected thread issues the copies -- across tile boundaries, so the
//     first records of the next tile arrive while the band of the current one is flushed.
// Counts, statistics, straggler handling and accumulate mode are those of k_bin_tiles.
constexpr int BAND_RPT = 2;         // records per thread and stage

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
__device__ __forceinline__ void mbar_wait(unsigned long long *b, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "W_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra W_%=;\n\t}"
        ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// the same copy with an L2 eviction-priority hint (createpolicy): data read exactly once should leave L2 first
__device__ __forceinline__ unsigned long long l2_policy_evict_first() {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ unsigned long long l2_policy_evict_last() {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void tma_load_1d_hint(void *dst, const void *src, uint32_t bytes, unsigned long long *bar,
                                                 unsigned long long pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}

// the CTA's tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... as one stream of stages; every thread walks it alike
template <int BAND_STAGE>
struct BandCursor {
    int t;               // tile (row group) of the stage
    long long idx, end;  // first record of the stage, end of the tile's slice
    bool valid;
    __device__ __forceinline__ void open(const BinArgs &a, int tile) {
        t = tile;
        valid = t < a.n_groups;
        if (valid) {
            idx = a.tile_clo[t] << CHUNK_LOG;
            end = min(a.tile_chi[t] << CHUNK_LOG, a.n_pairs);
            if (idx > end) idx = end;                          // (an empty slice still gets one, empty, stage)
        } else idx = end = 0;
    }
    __device__ __forceinline__ void next(const BinArgs &a) {   // after a stage [idx, idx + BAND_STAGE)
        idx += BAND_STAGE;
        if (idx >= end) open(a, t + (int)gridDim.x);
    }
    __device__ __forceinline__ bool last_of_tile() const { return idx + BAND_STAGE >= end; }
};

template <bool SAME_MEMBER, int NT, int NSTAGE>
__global__ void __launch_bounds__(NT, 1024 / NT) k_bin_band(BinArgs a, int half_band) {
    constexpr int BAND_THREADS = NT, BAND_STAGE = NT * BAND_RPT, BAND_STAGES = NSTAGE, CPS = NT / CHUNK;  // chunks per record slot
    static_assert(NT == 256 || NT == 512, "a record slot is one or two 256-record chunks");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint2 *ring = (uint2 *)smem_raw;                                      // [BAND_STAGES][BAND_STAGE * 3]
    uint32_t *band = (uint32_t *)(smem_raw + (size_t)BAND_STAGES * BAND_STAGE * 24);
    __shared__ unsigned long long full[BAND_STAGES], empty[BAND_STAGES];
    __shared__ int4 sbound[BAND_STAGES][BAND_RPT];   // row bounds advertised by the 256-record chunks of a stage
    const int tid = threadIdx.x, lane_id = tid & 31;
    const int band_cap = 2 * half_band + 8;
    for (int i = tid; i < band_cap; i += BAND_THREADS) band[i] = 0;
    if (tid == 0) {
        for (int s = 0; s < BAND_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], BAND_THREADS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t n_lines = 0, n_ok = 0, n_absent = 0, n_range = 0, n_key = 0, n_over = 0, n_outside = 0;
    const int nw = a.n_windows;

    // producer state (thread 0): the stage stream runs BAND_STAGES - 1 ahead of the consumers
    BandCursor<BAND_STAGE> prod, cons;
    prod.open(a, blockIdx.x);
    cons.open(a, blockIdx.x);
    int issued = 0;
    // chunk bounds of the stage that will be issued next: requested one issue ahead, so that the producer never waits
    // for them (the consumers read them from shared memory with their records)
    int4 nb[BAND_RPT];
    auto fetch_bounds = [&]() {
        if (!prod.valid) return;
        const long long c = prod.idx >> CHUNK_LOG;
#pragma unroll
        for (int k = 0; k < BAND_RPT; ++k) {
            const long long c0 = min(c + CPS * k, a.n_chunks - 1), c1 = min(c + CPS * k + (CPS - 1), a.n_chunks - 1);
            nb[k] = make_int4(__ldg(a.chunk_lo + c0), __ldg(a.chunk_hi + c0), __ldg(a.chunk_lo + c1), __ldg(a.chunk_hi + c1));
        }
    };
    if (tid == 0) fetch_bounds();
    const unsigned long long pol_stream = l2_policy_evict_first(), pol_row = l2_policy_evict_last();
    auto issue = [&]() {      // thread 0 only
        if (!prod.valid) return;
        const int slot = issued % BAND_STAGES;
        if (issued >= BAND_STAGES) mbar_wait(&empty[slot], (uint32_t)(((issued / BAND_STAGES) - 1) & 1));
#pragma unroll
        for (int k = 0; k < BAND_RPT; ++k) sbound[slot][k] = nb[k];
        const long long cnt = min((long long)BAND_STAGE, prod.end - prod.idx);
        if (cnt > 0) {
            // multiple of 8, TMA wants 16: round up inside the record array, down at its very end (the thread that
            // holds the array's last record then fetches its last 8 bytes itself)
            const uint32_t bytes = (uint32_t)(cnt * 24);
            const uint32_t b16 = (prod.idx + cnt == a.n_pairs) ? (bytes & ~15u) : ((bytes + 15u) & ~15u);
            mbar_expect_tx(&full[slot], b16);
            if (a.stream_evict_first) tma_load_1d_hint(ring + (size_t)slot * BAND_STAGE * 3, a.pairs + 3 * prod.idx, b16, &full[slot], pol_stream);
            else tma_load_1d(ring + (size_t)slot * BAND_STAGE * 3, a.pairs + 3 * prod.idx, b16, &full[slot]);
        } else {
            mbar_arrive(&full[slot]);                       // empty slice: complete the phase without a copy
        }
        ++issued;
        prod.next(a);
        fetch_bounds();
    };
    if (tid == 0)
        for (int k = 0; k < BAND_STAGES - 1; ++k) issue();

    int32_t r0 = 0, b0 = 0, b1 = 0;
    // the part of a tile's row outside the band is zeroed before its first reduction can arrive: by the CTA itself, at
    // the end of the previous tile (the stores drain under the flush), with one fence + barrier for both
    auto zero_row = [&](int t) {
        const int32_t r = a.row0 + t;
        const int32_t z0 = max(0, (r - half_band) & ~3), z1 = min(nw, (r + half_band + 4) & ~3);
        uint4 *dst = (uint4 *)(a.counts + (size_t)(r - a.row0) * nw);
        const int n4 = nw >> 2, s0 = z0 >> 2, s1 = z1 >> 2;
        const uint4 z = make_uint4(0, 0, 0, 0);
        if (a.row_evict_last) {
            for (int i = tid; i < n4; i += BAND_THREADS)
                if (i < s0 || i >= s1)
                    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %1, %1, %1}, %2;" ::"l"(dst + i), "r"(0u), "l"(pol_row) : "memory");
        } else {
            for (int i = tid; i < n4; i += BAND_THREADS)
                if (i < s0 || i >= s1) dst[i] = z;
        }
    };
    auto open_tile = [&](int t) {
        r0 = a.row0 + t;
        b0 = max(0, (r0 - half_band) & ~3);
        b1 = min(nw, (r0 + half_band + 4) & ~3);
    };
    if (cons.valid) {
        if (!a.accumulate) { zero_row(cons.t); __threadfence(); }
        __syncthreads();
        open_tile(cons.t);
    }
    int consumed = 0;
    while (cons.valid) {
        if (tid == 0) issue();
        const int slot = consumed % BAND_STAGES;
        mbar_wait(&full[slot], (uint32_t)((consumed / BAND_STAGES) & 1));
        const uint2 *rp = ring + (size_t)slot * BAND_STAGE * 3;
        uint2 q[BAND_RPT][3];
        int4 bd[BAND_RPT];
#pragma unroll
        for (int k = 0; k < BAND_RPT; ++k) {
            const uint2 *p = rp + 3 * (tid + k * BAND_THREADS);
            q[k][0] = p[0]; q[k][1] = p[1]; q[k][2] = p[2];
            bd[k] = sbound[slot][k];
        }
        __syncwarp();
        if (lane_id == 0) mbar_arrive(&empty[slot]);         // this warp holds its records in registers
        // (32-bit positions inside the stage: how many records it holds, and where the array's last record sits)
        const int remain = (int)min((long long)BAND_STAGE, cons.end - cons.idx);
        const int tail = (cons.idx + remain == a.n_pairs) ? remain - 1 : -1;
        // decode all records of the thread first (their table look-ups go out together), count afterwards: the
        // reduction below is a volatile asm with a memory clobber, no load of the next record moves above it
        if (tail >= 0) {   // (only the stage that holds the array's very last record)
#pragma unroll
            for (int k = 0; k < BAND_RPT; ++k)
                if (tid + k * BAND_THREADS == tail) q[k][2] = __ldg(a.pairs + 3 * (cons.idx + tail) + 2);
        }
        int32_t rows[BAND_RPT], cols[BAND_RPT];
        bool filt[BAND_RPT], oks[BAND_RPT];
#pragma unroll
        for (int k = 0; k < BAND_RPT; ++k) {
            const bool live = tid + k * BAND_THREADS < remain;
            if (!live) q[k][0] = q[k][1] = q[k][2] = make_uint2(0, 0);
            rows[k] = cols[k] = 0;
            decode_lean<SAME_MEMBER>(a, q[k][0], q[k][1], q[k][2], rows[k], cols[k], filt[k], oks[k]);
            filt[k] = filt[k] || !live;
            oks[k] = oks[k] && live;
        }
#pragma unroll
        for (int k = 0; k < BAND_RPT; ++k) {
            const int rel = tid + k * BAND_THREADS;
            const bool live = rel < remain;
            const int32_t lo = (CPS == 1 || tid < CHUNK) ? bd[k].x : bd[k].z, hi = (CPS == 1 || tid < CHUNK) ? bd[k].y : bd[k].w;
            const int32_t row = rows[k], col = cols[k];
            const bool filtered = filt[k], ok = oks[k];
            const bool owner = live && lo == r0;             // this tile accounts for the record
            if (owner) {
                n_lines += !filtered;
                n_ok += ok;
                if (!filtered && !ok) {
                    int32_t rr, cc;
                    const int st = decode(a, q[k][0], q[k][1], q[k][2], rr, cc);
                    n_absent += st == ST_ABSENT;
                    n_range += st == ST_RANGE;
                    n_key += st == ST_KEY;
                }
            }
            const bool mine = ok && row == r0;
            const bool near = mine && col >= b0 && col < b1;
            // a third of a row's pairs land on its diagonal cell: those are counted with one ballot per warp; the other
            // band cells rarely coincide inside a warp and take plain shared-memory atomics
            const bool diag = mine && col == r0;
            const unsigned dm = __ballot_sync(0xffffffffu, diag);
            if (dm && lane_id == (__ffs(dm) - 1)) atomicAdd(&band[r0 - b0], (uint32_t)__popc(dm));
            if (near && !diag) atomicAdd(&band[col - b0], 1u);
            if (mine && !near) {  // far hit: a reduction resolved in L2, nothing comes back (the warp does not wait for it)
                if (a.row_evict_last)
                    asm volatile("red.relaxed.gpu.global.add.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(a.counts + (size_t)(r0 - a.row0) * nw + col), "r"(1u), "l"(pol_row) : "memory");
                else
                    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(a.counts + (size_t)(r0 - a.row0) * nw + col), "r"(1u) : "memory");
            }
            const bool in_block = row >= a.row0 && row < a.row1;
            n_outside += owner && ok && !in_block;
            const bool strag = owner && ok && (row < lo || row >= hi) && in_block;
            const unsigned m = __ballot_sync(0xffffffffu, strag);
            if (m) {
                const int leader = __ffs(m) - 1;
                unsigned long long sl = 0;
                if (lane_id == leader) sl = atomicAdd(a.overflow_count, (unsigned long long)__popc(m));
                sl = __shfl_sync(0xffffffffu, sl, leader);
                if (strag) {
                    sl += __popc(m & ((1u << lane_id) - 1));
                    if ((long long)sl < a.overflow_cap) a.overflow[sl] = make_uint2(row, col);
                    n_over++;
                }
            }
        }
        const bool last = cons.last_of_tile();
        ++consumed;
        cons.next(a);
        if (last) {           // flush the band (every band cell is written exactly once) and prepare the next row
            __syncthreads();
            uint32_t *dst = a.counts + (size_t)(r0 - a.row0) * nw + b0;
            for (int i = tid * 4; i < b1 - b0; i += BAND_THREADS * 4) {
                uint4 v = *(uint4 *)(band + i);
                if (a.accumulate) {
                    const uint4 o = *(uint4 *)(dst + i);
                    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                }
                *(uint4 *)(dst + i) = v;
                *(uint4 *)(band + i) = make_uint4(0, 0, 0, 0);
            }
            if (cons.valid) {
                if (!a.accumulate) { zero_row(cons.t); __threadfence(); }
                open_tile(cons.t);
            }
            __syncthreads();
        }
    }
    unsigned long long v[7] = {n_lines, n_ok, n_absent, n_range, n_key, n_over, n_outside};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        unsigned long long t = hc::warp_sum_u64(v[k]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(a.stats + (k == 6 ? 10 : k), t);
    }
}

// Spilled 16 384-count blocks are always applied (their list cannot overflow); the straggler list only when it
// held every straggler -- otherwise k_bin_stragglers re-scans the records.  Both decisions are taken on the device,
// BEFORE anything a second pass would add again is in `counts` (accumulate mode stays exact).
__global__ void k_apply_overflow(BinArgs a) {
    const long long ns = min((long long)*a.spill_count, a.spill_cap);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < ns; i += (long long)gridDim.x * blockDim.x) {
        const uint2 e = a.spill[i];
        atomicAdd(a.counts + (size_t)((int32_t)e.x - a.row0) * a.n_windows + e.y, 0x4000u);
    }
    const long long n = (long long)*a.overflow_count;
    if (n > a.overflow_cap) return;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uint2 e = a.overflow[i];
        atomicAdd(a.counts + (size_t)((int32_t)e.x - a.row0) * a.n_windows + e.y, 1u);
    }
}

// The straggler list overflowed (input not sorted by read-1 window): every straggler -- a record whose row lies
// outside the rows [chunk_lo, chunk_hi) its chunk advertises, so that no tile saw it -- is added with a global
// atomic instead.  The tiles have already been flushed with everything else, each record is counted exactly once.
__global__ void __launch_bounds__(256) k_bin_stragglers(BinArgs a) {
    if ((long long)*a.overflow_count <= a.overflow_cap) return;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < a.n_pairs;
         idx += (long long)gridDim.x * blockDim.x) {
        const uint2 *p = a.pairs + 3 * idx;
        int32_t row, col;
        if (decode(a, __ldg(p), __ldg(p + 1), __ldg(p + 2), row, col) != ST_OK) continue;
        const long long c = idx >> CHUNK_LOG;
        const int32_t lo = __ldg(a.chunk_lo + c), hi = __ldg(a.chunk_hi + c);
        if ((row < lo || row >= hi) && row >= a.row0 && row < a.row1)
            atomicAdd(a.counts + (size_t)(row - a.row0) * a.n_windows + col, 1u);
    }
}

// Fallback: plain global atomics on a zeroed matrix, any input order.
__global__ void __launch_bounds__(256) k_bin_atomic(BinArgs a) {
    unsigned long long n_lines = 0, n_ok = 0, n_absent = 0, n_range = 0, n_key = 0, n_outside = 0;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < a.n_pairs;
         idx += (long long)gridDim.x * blockDim.x) {
        const uint2 *p = a.pairs + 3 * idx;
        int32_t row, col;
        int st = decode(a, __ldg(p), __ldg(p + 1), __ldg(p + 2), row, col);
        n_lines += st != ST_FILTERED;
        n_ok += st == ST_OK;
        n_absent += st == ST_ABSENT;
        n_range += st == ST_RANGE;
        n_key += st == ST_KEY;
        if (st == ST_OK && row >= a.row0 && row < a.row1)
            atomicAdd(a.counts + (size_t)(row - a.row0) * a.n_windows + col, 1u);
        n_outside += st == ST_OK && !(row >= a.row0 && row < a.row1);
    }
    unsigned long long v[6] = {n_lines, n_ok, n_absent, n_range, n_key, n_outside};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        unsigned long long t = hc::warp_sum_u64(v[k]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(a.stats + (k == 5 ? 10 : k), t);
    }
}

// ---- several window sizes from ONE read of the records ---------------------------------------------------------
// bam2arrays bins every kept line into all window sizes of the call (inner loop bam2clusters.py:154-161).  The
// per-cluster matrices of the scaffolding stage (5 / 10 / 2 kb, bam2scaffolds.py:379) are small enough to live in
// the 126 MB L2 together, so each record is read and filtered once and added to every matrix with a warp-aggregated
// global atomic (resolved in L2); the privatised-row kernel above stays the path for one big matrix.
constexpr int MAX_WS = 4;
struct MultiArgs {
    const uint2 *pairs;
    long long n_pairs;
    const int32_t *offsets[MAX_WS];
    const uint8_t *region_mask;
    int32_t n_contigs, n_ws;
    uint32_t window_bp[MAX_WS];
    unsigned long long inv_bp[MAX_WS];
    int32_t n_windows[MAX_WS];
    int32_t mapq_min;
    uint32_t skip_flag;
    uint32_t *counts[MAX_WS];
    unsigned long long *stats;
};

__global__ void __launch_bounds__(256) k_bin_multi(MultiArgs a) {
    unsigned long long n_lines = 0, n_ok = 0, n_absent = 0, n_range = 0, n_key = 0;
    const int lane = threadIdx.x & 31;
    const long long total = (a.n_pairs + 31) & ~31ll;   // whole warps take part in the matches
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        bool live = idx < a.n_pairs;
        uint2 q0 = make_uint2(0, 0), q1 = q0, q2 = q0;
        if (live) {
            const uint2 *p = a.pairs + 3 * idx;
            q0 = __ldg(p); q1 = __ldg(p + 1); q2 = __ldg(p + 2);
        }
        const int32_t ref = (int32_t)q0.x, pos = (int32_t)q0.y, nref = (int32_t)q1.x, npos = (int32_t)q1.y;
        const int32_t lseq = (int32_t)q2.x;
        const uint32_t flag = q2.y & 0xffffu, mapq = (q2.y >> 16) & 0xffu;
        const bool ref_ok = (uint32_t)ref < (uint32_t)a.n_contigs, nref_ok = (uint32_t)nref < (uint32_t)a.n_contigs;
        live = live && !(flag & a.skip_flag) && (int32_t)mapq >= a.mapq_min;
        if (a.region_mask) live = live && ref_ok && a.region_mask[ref];
        n_lines += live;
        // membership is tested against the first table only (bam2clusters.py:151)
        const bool member = live && ref_ok && nref_ok && __ldg(a.offsets[0] + ref) >= 0 && __ldg(a.offsets[0] + nref) >= 0;
        n_absent += live && !member;
        const int32_t half = (lseq > 1 ? lseq : 1) >> 1;
        const long long p1 = (long long)pos + 1 + half, p2 = (long long)npos + 1 + half;
        bool ok_all = member;
#pragma unroll
        for (int ii = 0; ii < MAX_WS; ++ii) {
            if (ii >= a.n_ws) break;
            uint32_t cell = 0xffffffffu;
            if (member && ok_all) {
                const int32_t o1 = __ldg(a.offsets[ii] + ref), o2 = __ldg(a.offsets[ii] + nref);
                if (o1 < 0 || o2 < 0) { n_key++; ok_all = false; }
                else if (p1 < 0 || p2 < 0) { n_range++; ok_all = false; }
                else {
                    const uint32_t wb = a.window_bp[ii];
                    const long long w1 = (long long)(wb == 1u ? (uint32_t)p1 : (uint32_t)__umul64hi((unsigned long long)(uint32_t)p1, a.inv_bp[ii])) + o1;
                    const long long w2 = (long long)(wb == 1u ? (uint32_t)p2 : (uint32_t)__umul64hi((unsigned long long)(uint32_t)p2, a.inv_bp[ii])) + o2;
                    if (w1 >= a.n_windows[ii] || w2 >= a.n_windows[ii]) { n_range++; ok_all = false; }
                    else cell = (uint32_t)(w1 * a.n_windows[ii] + w2);   // full orientation, like hc_bin_pairs
                }
            }
            const unsigned peers = __match_any_sync(0xffffffffu, cell);
            if (cell != 0xffffffffu && (__ffs(peers) - 1) == lane) atomicAdd(a.counts[ii] + cell, (uint32_t)__popc(peers));
        }
        n_ok += member && ok_all;
    }
    unsigned long long v[5] = {n_lines, n_ok, n_absent, n_range, n_key};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        unsigned long long t = hc::warp_sum_u64(v[k]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(a.stats + k, t);
    }
}

// ---- upto -------------------------------------------------------------------------------
__device__ __forceinline__ bool is_line(const uint2 *pairs, long long idx, const uint8_t *region_mask,
                                        int32_t n_contigs, int32_t mapq_min, uint32_t skip_flag) {
    uint2 q2 = __ldg(pairs + 3 * idx + 2);
    uint32_t flag = q2.y & 0xffffu, mapq = (q2.y >> 16) & 0xffu;
    if ((flag & skip_flag) || (int32_t)mapq < mapq_min) return false;
    if (region_mask) {
        int32_t ref = (int32_t)__ldg(pairs + 3 * idx).x;
        if (!((uint32_t)ref < (uint32_t)n_contigs && region_mask[ref])) return false;
    }
    return true;
}

__global__ void k_count_lines(const uint2 *pairs, long long n, const uint8_t *region_mask,
                              int32_t n_contigs, int32_t mapq_min, uint32_t skip_flag,
                              unsigned long long *block_lines) {
    __shared__ unsigned long long sh[32];
    long long base = (long long)blockIdx.x * 4096;
    unsigned long long cnt = 0;
    for (int i = threadIdx.x; i < 4096 && base + i < n; i += blockDim.x)
        cnt += is_line(pairs, base + i, region_mask, n_contigs, mapq_min, skip_flag);
    cnt = hc::warp_sum_u64(cnt);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < (blockDim.x >> 5); ++w) t += sh[w];
        block_lines[blockIdx.x] = t;
    }
}

// single block: locate the block holding the upto-th line, then the record inside it
__global__ void k_find_cutoff(const uint2 *pairs, long long n, const uint8_t *region_mask,
                              int32_t n_contigs, int32_t mapq_min, uint32_t skip_flag,
                              const unsigned long long *block_lines, long long n_blocks,
                              long long upto, long long *result) {
    __shared__ long long s_block, s_before;
    __shared__ int s_warp_cnt[32];
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        long long b = 0;
        for (; b < n_blocks; ++b) {
            if (run + block_lines[b] >= (unsigned long long)upto) break;
            run += block_lines[b];
        }
        s_block = b;
        s_before = (long long)run;
    }
    __syncthreads();
    if (s_block >= n_blocks) {  // fewer than upto lines in total
        if (threadIdx.x == 0) { result[0] = n; result[1] = s_before; }
        return;
    }
    long long base = s_block * 4096, need = upto - s_before;  // need >= 1
    // 1024 threads x 4 consecutive records, ordered scan
    int flags[4], mine = 0;
    for (int k = 0; k < 4; ++k) {
        long long idx = base + threadIdx.x * 4 + k;
        flags[k] = idx < n && is_line(pairs, idx, region_mask, n_contigs, mapq_min, skip_flag);
        mine += flags[k];
    }
    int incl = mine;
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) s_warp_cnt[threadIdx.x >> 5] = incl;
    __syncthreads();
    int before = incl - mine;
    for (int w = 0; w < (threadIdx.x >> 5); ++w) before += s_warp_cnt[w];
    for (int k = 0; k < 4; ++k) {
        if (flags[k]) {
            before++;
            if (before == need) { result[0] = base + threadIdx.x * 4 + k + 1; result[1] = upto; }
        }
    }
}

__global__ void k_region_filter(const uint2 *pairs, long long n, const int32_t *tlen,
                                const int32_t *ref_span, const int32_t *contig_size,
                                const int32_t *member, int32_t n_contigs, int32_t min_size,
                                int32_t nth, int32_t mapq_min, uint32_t skip_flag, uint8_t *keep,
                                int32_t *dist) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n;
         k += (long long)gridDim.x * blockDim.x) {
        uint2 q0 = __ldg(pairs + 3 * k), q1 = __ldg(pairs + 3 * k + 1), q2 = __ldg(pairs + 3 * k + 2);
        int32_t ref = (int32_t)q0.x, pos = (int32_t)q0.y, nref = (int32_t)q1.x;
        uint32_t flag = q2.y & 0xffffu, mapq = (q2.y >> 16) & 0xffu;
        uint8_t kp = 0;
        int32_t d = -1;
        if (!(flag & skip_flag) && (int32_t)mapq >= mapq_min && (uint32_t)ref < (uint32_t)n_contigs &&
            member[ref] >= 0 && pos >= 0) {
            // regions (s+1, s+minSize), s = 0, step, 2*step..., s <= size - minSize (1-based, inclusive)
            long long step = (long long)nth * min_size;
            long long first = (long long)pos + 1, last = (long long)pos + max(ref_span ? ref_span[k] : 1, 1);
            long long smax = (long long)contig_size[ref] - min_size;  // largest admissible s
            if (smax >= 0) {
                // region index range overlapping [first, last]: s+1 <= last and s+minSize >= first
                long long i_hi = (last - 1) / step;                   // s <= last-1
                long long lo_s = first - min_size;                   // s >= first - minSize
                long long i_lo = lo_s <= 0 ? 0 : (lo_s + step - 1) / step;
                long long i_max = smax / step;
                if (i_hi > i_max) i_hi = i_max;
                if (i_lo <= i_hi) {
                    kp = (uint8_t)min((long long)255, i_hi - i_lo + 1);  // printed once per region hit
                    if (nref == ref) {
                        int32_t t = tlen ? tlen[k] : 0;
                        d = t < 0 ? -t : t;
                    }
                }
            }
        }
        keep[k] = kp;
        dist[k] = d;
    }
}

// stats[5] = stragglers, stats[6] = 1 when the sorted-input path handled everything (0: forced atomic kernel, or the
// straggler list overflowed and the records were re-scanned), stats[7] = tile shape
__global__ void k_bin_finish(unsigned long long *stats, long long overflow_cap, long long shape) {
    const long long n_over = (long long)stats[8];
    stats[5] = overflow_cap < 0 ? 0ull : (unsigned long long)n_over;
    stats[6] = (overflow_cap >= 0 && n_over <= overflow_cap) ? 1ull : 0ull;
    stats[7] = (unsigned long long)shape;
}

}  // namespace

extern "C" size_t hc_bin_pairs_workspace_bytes(int64_t n_pairs, int64_t n_rows) {
    size_t n_chunks = (size_t)((n_pairs + CHUNK - 1) / CHUNK) + 1;
    size_t nb = (n_chunks + SCAN_BLOCK - 1) / SCAN_BLOCK + 1;
    size_t bytes = 4096;
    bytes += hc::align_up(n_chunks * 4, 256) * 3;        // lead/chunk_lo/chunk_hi
    bytes += hc::align_up(nb * 4, 256);                  // block maxima
    bytes += hc::align_up((size_t)n_rows * 8, 256) * 2;  // tile ranges (<= one group per row)
    bytes += hc::align_up((size_t)(n_pairs / 16 + 65536) * 8, 256);  // straggler list
    bytes += hc::align_up((size_t)(n_pairs / 16384 + 1024) * 8, 256);  // spill list
    return bytes;
}

extern "C" int hc_bin_pairs(const hc_pair_t *pairs, int64_t n_pairs, const int32_t *offsets,
                            const int32_t *member, const uint8_t *region_mask, int32_t n_contigs,
                            uint32_t window_bp, int32_t n_windows, int32_t mapq_min,
                            uint32_t skip_flag, int32_t row0, int32_t row1, uint32_t *counts,
                            void *workspace, size_t workspace_bytes, int mode, int64_t *h_stats,
                            void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(n_pairs >= 0 && n_windows > 0 && window_bp > 0, "bad sizes");
    HC_REQUIRE(row0 >= 0 && row0 < row1 && row1 <= n_windows, "bad row block");
    HC_REQUIRE(offsets && counts && workspace, "null pointer");
    HC_REQUIRE(((uintptr_t)pairs & 7) == 0, "pairs must be 8-byte aligned");
    if (workspace_bytes < hc_bin_pairs_workspace_bytes(n_pairs, row1 - row0)) {
        hc::set_error("hc_bin_pairs: workspace too small");
        return HC_EWORKSPACE;
    }
    const int32_t rows = row1 - row0;
    BinArgs a;
    memset(&a, 0, sizeof a);
    a.pairs = (const uint2 *)pairs;
    a.n_pairs = n_pairs;
    a.offsets = offsets;
    a.member = member ? member : offsets;
    a.region_mask = region_mask;
    a.n_contigs = n_contigs;
    a.window_bp = window_bp;
    a.inv_bp = window_bp == 1u ? 0ull : (~0ull) / window_bp + 1ull;
    a.n_windows = n_windows;
    a.mapq_min = mapq_min;
    a.skip_flag = skip_flag;
    a.row0 = row0;
    a.row1 = row1;
    a.counts = counts;
    a.n_chunks = (n_pairs + CHUNK - 1) / CHUNK;

    // tile shape: R rows x W columns of u32 in shared memory
    int dev = 0, smem_max = 0;
    HC_CUDA_OK(cudaGetDevice(&dev));
    HC_CUDA_OK(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const int sms = hc::sm_count();
    // 16-bit tile cells when a u32 row no longer leaves room for two CTAs per SM (n_windows > ~29 000) and the
    // vector flush applies (n_windows % 4 == 0, 16-byte aligned matrix); HC_BIN_CELL16=0/1 overrides
    const bool vec_host = (n_windows & 3) == 0 && (((uintptr_t)counts) & 15) == 0;
    bool cell16 = vec_host && (long long)n_windows * 4 > (smem_max - 1024) / 2;
    if (const char *e = getenv("HC_BIN_CELL16")) cell16 = vec_host && atoi(e) != 0;
    // one-row tiles (n_windows > ~29 000): the band kernel (TMA-staged record ring, far hits as L2 reductions);
    // HC_BIN_BAND=0/1 overrides, HC_BIN_BAND_B sets the half width of the band in windows
    bool band = cell16 && (((uintptr_t)pairs) & 15) == 0 && n_windows >= 4096;
    if (const char *e = getenv("HC_BIN_BAND")) band = atoi(e) != 0 && vec_host && (((uintptr_t)pairs) & 15) == 0;
    int half_band = 1024;
    if (const char *e = getenv("HC_BIN_BAND_B")) half_band = atoi(e);
    if (half_band < 4) half_band = 4;
    if (half_band > 8192) half_band = 8192;
    a.stream_evict_first = 1;
    if (const char *e = getenv("HC_BIN_EVICT_FIRST")) a.stream_evict_first = atoi(e) != 0;
    a.row_evict_last = 1;   // measured at C3: 5.83 -> 5.75 ms, 0.5 GB less DRAM write traffic
    if (const char *e = getenv("HC_BIN_ROW_EVICT_LAST")) a.row_evict_last = atoi(e) != 0;
    half_band &= ~3;
    long long max_cells = (smem_max - 1024) / (cell16 ? 2 : 4);
    if (const char *e = getenv("HC_BIN_MAX_CELLS")) max_cells = atoll(e);
    a.n_segs = (int)((n_windows + max_cells - 1) / max_cells);
    a.W = (n_windows + a.n_segs - 1) / a.n_segs;
    if (a.n_segs > 1) a.W = (a.W + 3) & ~3;  // 16-byte aligned column segments
    int Rmax = (int)(max_cells / a.W);
    if (Rmax < 1) Rmax = 1;
    // prefer tiles of at most half an SM's shared memory: two resident CTAs overlap flush and stream
    if ((max_cells / 2) / a.W >= 1 && !getenv("HC_BIN_MAX_CELLS")) Rmax = (int)((max_cells / 2) / a.W);
    int Rwant = (rows + 4 * sms - 1) / (4 * sms);  // aim for >= 4 tiles per SM
    a.R = Rwant < 1 ? 1 : (Rwant > Rmax ? Rmax : Rwant);
    if (const char *e = getenv("HC_BIN_ROWS")) a.R = atoi(e) < 1 ? 1 : (atoi(e) > Rmax ? Rmax : atoi(e));
    if (band) { a.R = 1; a.n_segs = 1; a.W = n_windows; }
    if (cell16 && (a.W & 3) != 0) cell16 = false;  // (cannot happen: W = n_windows or a multiple of 4)
    a.n_groups = (rows + a.R - 1) / a.R;

    hc::Carver ws(workspace, workspace_bytes);
    unsigned long long *stats = ws.take<unsigned long long>(16);
    int32_t *lead = ws.take<int32_t>(a.n_chunks + 1);
    int32_t *chunk_lo = ws.take<int32_t>(a.n_chunks + 1);
    int32_t *chunk_hi = ws.take<int32_t>(a.n_chunks + 1);
    long long nb = (a.n_chunks + SCAN_BLOCK - 1) / SCAN_BLOCK;
    int32_t *bmax = ws.take<int32_t>(nb + 1);
    long long *tile_clo = ws.take<long long>(a.n_groups);
    long long *tile_chi = ws.take<long long>(a.n_groups);
    a.overflow_cap = n_pairs / 16 + 65536;
    a.overflow = ws.take<uint2>(a.overflow_cap);
    a.spill_cap = n_pairs / 16384 + 1024;
    a.spill = ws.take<uint2>(a.spill_cap);
    if (!ws.ok()) {
        hc::set_error("hc_bin_pairs: workspace carve overflow");
        return HC_EWORKSPACE;
    }
    a.stats = stats;
    a.overflow_count = stats + 8;
    a.spill_count = stats + 9;
    a.chunk_lo = chunk_lo;
    a.chunk_hi = chunk_hi;
    a.tile_clo = tile_clo;
    a.tile_chi = tile_chi;

    a.accumulate = (mode & 2) != 0;
    bool fast = ((mode & 1) == 0) && n_pairs > 0;
    if (fast) {
        HC_CUDA_OK(cudaMemsetAsync(stats, 0, 16 * sizeof(unsigned long long), stream));
        int cb = (int)((a.n_chunks + 255) / 256);
        k_lead<<<cb, 256, 0, stream>>>(a, lead);
        k_scan_blockmax<<<(int)nb, 256, 0, stream>>>(lead, a.n_chunks, bmax);
        k_scan_carry<<<1, 32, 0, stream>>>(bmax, (int)nb);
        k_scan_apply<<<(int)nb, 1024, 0, stream>>>(lead, a.n_chunks, bmax);
        k_chunk_bounds<<<cb, 256, 0, stream>>>(a, lead, chunk_lo, chunk_hi);
        k_tile_ranges<<<(a.n_groups + 127) / 128, 128, 0, stream>>>(a, tile_clo, tile_chi);
        size_t smem = hc::align_up((size_t)a.R * a.W * (cell16 ? 2 : 4), 16);
        int threads = cell16 ? 512 : 1024;
        if (const char *e = getenv("HC_BIN_THREADS")) threads = atoi(e);
        if (cell16 && threads > 512) threads = 512;
        if (cell16)
            HC_CUDA_OK(cudaFuncSetAttribute(k_bin_tiles<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
        else
            HC_CUDA_OK(cudaFuncSetAttribute(k_bin_tiles<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
        // persistent: as many CTAs as fit (tiles <= ~110 KB let two CTAs share an SM and overlap
        // one CTA's flush with the other's record stream)
        int per_sm = (int)((size_t)(smem_max + 1024) / (smem + 1024));
        if (per_sm > 2048 / threads) per_sm = 2048 / threads;
        if (cell16 && per_sm > 2) per_sm = 2;
        if (per_sm < 1) per_sm = 1;
        long long n_tiles = (long long)a.n_groups * a.n_segs;
        int grid = (int)std::min<long long>(n_tiles, (long long)sms * per_sm);
        if (band) {
            int nt = 256, nst = 3;   // four CTAs per SM: C3 5.9 ms (512 x 4: 6.2 ms, 256 x 4: 8.0 ms)
            if (const char *e = getenv("HC_BIN_BAND_THREADS")) nt = atoi(e) == 256 ? 256 : 512;
            if (const char *e = getenv("HC_BIN_BAND_STAGES")) nst = atoi(e) == 3 ? 3 : 4;
            const size_t bsmem2 = (size_t)nst * nt * BAND_RPT * 24 + (size_t)(2 * half_band + 8) * 4;
            const int bgrid = (int)std::min<long long>(a.n_groups, (long long)sms * (1024 / nt));
            const bool same = a.member == a.offsets;
#define HC_BAND_LAUNCH(SM_, NT_, NS_)                                                                                       \
    do {                                                                                                                    \
        HC_CUDA_OK(cudaFuncSetAttribute(k_bin_band<SM_, NT_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bsmem2)); \
        k_bin_band<SM_, NT_, NS_><<<bgrid, NT_, bsmem2, stream>>>(a, half_band);                                             \
    } while (0)
            if (nt == 512 && nst == 4) { if (same) HC_BAND_LAUNCH(true, 512, 4); else HC_BAND_LAUNCH(false, 512, 4); }
            else if (nt == 512) { if (same) HC_BAND_LAUNCH(true, 512, 3); else HC_BAND_LAUNCH(false, 512, 3); }
            else if (nst == 4) { if (same) HC_BAND_LAUNCH(true, 256, 4); else HC_BAND_LAUNCH(false, 256, 4); }
            else { if (same) HC_BAND_LAUNCH(true, 256, 3); else HC_BAND_LAUNCH(false, 256, 3); }
#undef HC_BAND_LAUNCH
        } else if (cell16) k_bin_tiles<3, true><<<grid, threads, smem, stream>>>(a);
        else k_bin_tiles<2, false><<<grid, threads, smem, stream>>>(a);
        HC_CHECK_LAUNCH();
        k_apply_overflow<<<4 * sms, 256, 0, stream>>>(a);
        k_bin_stragglers<<<16 * sms, 256, 0, stream>>>(a);   // returns at once unless the straggler list overflowed
        HC_CHECK_LAUNCH();
    } else {
        HC_CUDA_OK(cudaMemsetAsync(stats, 0, 16 * sizeof(unsigned long long), stream));
        if (!a.accumulate) HC_CUDA_OK(cudaMemsetAsync(counts, 0, (size_t)rows * n_windows * 4, stream));
        if (n_pairs > 0) {
            k_bin_atomic<<<16 * sms, 256, 0, stream>>>(a);
            HC_CHECK_LAUNCH();
        }
    }
    k_bin_finish<<<1, 1, 0, stream>>>(stats, fast ? a.overflow_cap : -1, ((long long)a.R << 32) | (uint32_t)a.n_segs);
    HC_CHECK_LAUNCH();
    if (!h_stats) return HC_OK;   // asynchronous call: hc_bin_pairs_stats() reads the device stats later
    return hc_bin_pairs_stats(workspace, h_stats, stream_);
}

extern "C" int hc_bin_pairs_stats(const void *workspace, int64_t *h_stats, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(workspace && h_stats, "null pointer");
    unsigned long long h[16];
    HC_CUDA_OK(cudaMemcpyAsync(h, workspace, sizeof h, cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaStreamSynchronize(stream));
    for (int k = 0; k < 8; ++k) h_stats[k] = (int64_t)h[k];
    h_stats[8] = (int64_t)h[10];
    if (h[4]) {
        hc::set_error("contig present in chr2window[0] but absent for this window size (%llu records)", h[4]);
        return HC_EKEY;
    }
    if (h[3]) {
        hc::set_error("window index out of range for %llu records", h[3]);
        return HC_ERANGE;
    }
    return HC_OK;
}

extern "C" int hc_bin_pairs_multi(const hc_pair_t *pairs, int64_t n_pairs, int32_t n_ws, const int32_t *const *h_offsets,
                                  const uint8_t *region_mask, int32_t n_contigs, const uint32_t *h_window_bp,
                                  const int32_t *h_n_windows, int32_t mapq_min, uint32_t skip_flag,
                                  uint32_t *const *h_counts, int accumulate, void *workspace, size_t workspace_bytes,
                                  int64_t *h_stats, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(n_ws >= 1 && n_ws <= MAX_WS && n_pairs >= 0, "1..4 window sizes per call");
    HC_REQUIRE(h_offsets && h_window_bp && h_n_windows && h_counts && workspace && workspace_bytes >= 256, "null pointer");
    HC_REQUIRE(((uintptr_t)pairs & 7) == 0, "pairs must be 8-byte aligned");
    MultiArgs a;
    memset(&a, 0, sizeof a);
    a.pairs = (const uint2 *)pairs;
    a.n_pairs = n_pairs;
    a.region_mask = region_mask;
    a.n_contigs = n_contigs;
    a.n_ws = n_ws;
    a.mapq_min = mapq_min;
    a.skip_flag = skip_flag;
    a.stats = (unsigned long long *)workspace;
    for (int ii = 0; ii < n_ws; ++ii) {
        HC_REQUIRE(h_offsets[ii] && h_counts[ii] && h_window_bp[ii] > 0 && h_n_windows[ii] > 0, "bad window table");
        HC_REQUIRE((long long)h_n_windows[ii] * h_n_windows[ii] < 0xffffffffll, "matrix too large for the one-pass kernel");
        a.offsets[ii] = h_offsets[ii];
        a.window_bp[ii] = h_window_bp[ii];
        a.inv_bp[ii] = h_window_bp[ii] == 1u ? 0ull : (~0ull) / h_window_bp[ii] + 1ull;
        a.n_windows[ii] = h_n_windows[ii];
        a.counts[ii] = h_counts[ii];
        if (!accumulate)
            HC_CUDA_OK(cudaMemsetAsync(h_counts[ii], 0, (size_t)h_n_windows[ii] * h_n_windows[ii] * 4, stream));
    }
    HC_CUDA_OK(cudaMemsetAsync(workspace, 0, 16 * sizeof(unsigned long long), stream));
    if (n_pairs > 0) {
        k_bin_multi<<<8 * hc::sm_count(), 256, 0, stream>>>(a);
        HC_CHECK_LAUNCH();
    }
    if (!h_stats) return HC_OK;
    return hc_bin_pairs_stats(workspace, h_stats, stream_);
}

extern "C" int hc_upto_cutoff(const hc_pair_t *pairs, int64_t n_pairs, const uint8_t *region_mask,
                              int32_t n_contigs, int32_t mapq_min, uint32_t skip_flag, int64_t upto,
                              void *workspace, size_t workspace_bytes, int64_t *h_n_effective,
                              int64_t *h_lines, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(h_n_effective && h_lines && upto > 0, "bad arguments");
    long long nb = (n_pairs + 4095) / 4096;
    if (workspace_bytes < (size_t)(nb + 2) * 8 + 256) {
        hc::set_error("hc_upto_cutoff: workspace too small");
        return HC_EWORKSPACE;
    }
    if (n_pairs == 0) { *h_n_effective = 0; *h_lines = 0; return HC_OK; }
    long long *result = (long long *)workspace;
    unsigned long long *block_lines = (unsigned long long *)((char *)workspace + 256);
    k_count_lines<<<(int)nb, 256, 0, stream>>>((const uint2 *)pairs, n_pairs, region_mask, n_contigs,
                                               mapq_min, skip_flag, block_lines);
    k_find_cutoff<<<1, 1024, 0, stream>>>((const uint2 *)pairs, n_pairs, region_mask, n_contigs, mapq_min,
                                          skip_flag, block_lines, nb, upto, result);
    HC_CHECK_LAUNCH();
    long long h[2];
    HC_CUDA_OK(cudaMemcpyAsync(h, result, sizeof h, cudaMemcpyDeviceToHost, stream));
    HC_CUDA_OK(cudaStreamSynchronize(stream));
    *h_n_effective = h[0];
    *h_lines = h[1];
    return HC_OK;
}

extern "C" int hc_contig_region_filter(const hc_pair_t *pairs, int64_t n_pairs, const int32_t *tlen,
                                       const int32_t *ref_span, const int32_t *contig_size,
                                       const int32_t *member, int32_t n_contigs, int32_t min_size,
                                       int32_t nth, int32_t mapq_min, uint32_t skip_flag,
                                       uint8_t *keep, int32_t *dist, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(min_size > 0 && nth > 0 && keep && dist && contig_size && member, "bad arguments");
    if (n_pairs == 0) return HC_OK;
    k_region_filter<<<8 * hc::sm_count(), 256, 0, stream>>>((const uint2 *)pairs, n_pairs, tlen, ref_span,
                                                            contig_size, member, n_contigs, min_size, nth,
                                                            mapq_min, skip_flag, keep, dist);
    HC_CHECK_LAUNCH();
    return HC_OK;
}

// ---- compact wire format (16 bytes per record) -> hc_pair_t -------------------------------------------
// Host records cross PCIe once per job; at 55 GB/s the 24-byte form costs 0.22 s for 5e8 records, more than
// the whole device path.  hc_pair16_t carries exactly what get_window (bam2clusters.py:119-126) and the
// samtools filter (:108-111) consume: the two read MIDPOINTS (pos + len(SEQ)//2, the only way POS and SEQ
// enter, :123 and :155-156), both reference ids, FLAG and MAPQ.  Expanded records have pos = midpoint and
// l_seq = 0 (SEQ "*": len//2 == 0), so every window, filter decision and count is unchanged.
__global__ void __launch_bounds__(256) k_unpack_pairs16(const uint4 *packed, long long n, uint2 *out) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n;
         k += (long long)gridDim.x * blockDim.x) {
        const uint4 q = __ldcs(packed + k);
        const uint32_t r1 = q.z & 0xfffffu, r2 = q.w & 0xfffffu;
        const int32_t ref = r1 == 0xfffffu ? -1 : (int32_t)r1, nref = r2 == 0xfffffu ? -1 : (int32_t)r2;
        const uint32_t flag = q.z >> 20, mapq = (q.w >> 20) & 0xffu;
        uint2 *o = out + 3 * k;
        o[0] = make_uint2((uint32_t)ref, q.x);
        o[1] = make_uint2((uint32_t)nref, q.y);
        o[2] = make_uint2(0u, flag | (mapq << 16));
    }
}

extern "C" int hc_unpack_pairs16(const hc_pair16_t *packed, int64_t n_pairs, hc_pair_t *pairs, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    HC_REQUIRE(n_pairs >= 0 && (n_pairs == 0 || (packed && pairs)), "bad arguments");
    HC_REQUIRE((((uintptr_t)packed) & 15) == 0 && (((uintptr_t)pairs) & 7) == 0, "misaligned buffers");
    if (n_pairs == 0) return HC_OK;
    k_unpack_pairs16<<<16 * hc::sm_count(), 256, 0, stream>>>((const uint4 *)packed, n_pairs, (uint2 *)pairs);
    HC_CHECK_LAUNCH();
    return HC_OK;
}
