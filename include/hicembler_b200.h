/* hicembler_b200 -- C ABI of the B200 (sm_100a) hot path of HiCembler.
 *
 * Every entry point takes plain pointers and sizes.  Pointers are DEVICE pointers unless the
 * parameter name starts with ``h_`` (host).  ``stream`` is a cudaStream_t passed as void*
 * (NULL = default stream).  Nothing here allocates caller-visible memory: outputs and
 * workspaces are caller-allocated (the Python layer uses torch tensors as buffers).
 * All functions return 0 on success or a negative hc_status; hc_last_error() gives the text.
 *
 * The reference (lpryszcz/HiCembler) has no FFI of its own: the boundary it exposes is a set of
 * Python callables plus two library calls (fastcluster.linkage, SinkhornKnopp.fit).  Each entry
 * point below cites the reference lines it replaces; INTEGRATION.md shows the ctypes binding a
 * maintainer adds on the reference side.
 */
#ifndef HICEMBLER_B200_H
#define HICEMBLER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HC_VERSION 200
#define HC_MAX_RANKS 8   /* GPUs of one NVLink / NVSwitch node that share a row-sharded matrix */

typedef enum {
    HC_OK = 0,
    HC_EINVAL = -1,   /* bad argument */
    HC_ERANGE = -2,   /* window index outside the matrix: reference raises IndexError (bam2clusters.py:161) */
    HC_EKEY = -3,     /* contig in chr2window[0] but not in chr2window[ii]: reference raises KeyError (:151,155) */
    HC_ENAN = -4,     /* NaN dissimilarity: fastcluster raises */
    HC_ENEG = -5,     /* negative entry in P: SinkhornKnopp.fit asserts (sinkhorn_knopp.py:110) */
    HC_ECUDA = -6,    /* CUDA runtime error */
    HC_EWORKSPACE = -7, /* workspace too small */
    HC_ENOTRI = -8,   /* condensed length is not n*(n-1)/2: fastcluster raises ValueError */
    HC_EAMBIGUOUS = 1 /* hc_linkage_rounds only, not an error: a tie keeps the round-parallel engine from
                       * certifying fastcluster's merge order; rebuild W and call hc_linkage */
} hc_status;

/* One alignment record, 24 bytes, little endian (BAM core fields of
 * bin/snap/SNAPLib/Bam.h:91-105 consumed at bam2clusters.py:145-146). pos/next_pos are 0-based. */
typedef struct {
    int32_t refID, pos, next_refID, next_pos, l_seq;
    uint16_t flag;
    uint8_t mapq, pad;
} hc_pair_t;

/* Compact wire form of the same record, 16 bytes, for records that cross PCIe (host decoder -> device):
 *   mid      = pos + len(SEQ)//2       (0-based; the only way POS and SEQ enter get_window, bam2clusters.py:123)
 *   next_mid = next_pos + len(SEQ)//2  (the mate's window uses read 1's length, :155-156)
 *   ref_flag = refID (20 bits, 0xFFFFF = "*") | FLAG << 20;  nref_mapq = next_refID (20 bits) | MAPQ << 20.
 * hc_unpack_pairs16 expands it on the device to hc_pair_t with pos = mid, next_pos = next_mid, l_seq = 0
 * (SEQ "*": len//2 == 0): windows, -q/-F decisions and counts are those of the original records.
 * Needs n_contigs < 2^20 - 1; TLEN / CIGAR side arrays of the contig-level path are not carried. */
typedef struct {
    uint32_t mid, next_mid, ref_flag, nref_mapq;
} hc_pair16_t;

int hc_version(void);
int hc_last_error(char *h_buf, size_t n);
/* number of SMs / bytes of HBM of the current device (for grid sizing by the caller) */
int hc_set_device(int device);
int hc_device_info(int *h_sm_count, size_t *h_total_mem, int *h_cc_major, int *h_cc_minor);

/* ---------------------------------------------------------------------------------------
 * K1  read pair -> window binning.   Replaces the per-line loop of bam2arrays
 * (bam2clusters.py:128-166, get_window :119-126, samtools filter :108-117) and, with
 * window_bp = UINT32_MAX (one window per contig), the Counter loop of _bam2array (:197-202).
 *
 * counts is the FULL (not triangular) u32 matrix block rows [row0,row1) x n_windows with
 * row = window of read 1, column = window of the mate; hc_fold_counts() turns it into the
 * reference's upper-triangular float32 array or into the symmetrised matrix.
 *
 * offsets[c]  = chr2window[ii][contig c]  (first window of contig c, -1 = absent)
 * member[c]   = chr2window[0][contig c]   (membership test of :151; NULL = same as offsets)
 * region_mask = per-contig 0/1 for ``samtools view bam <contig>...`` (NULL = no regions)
 * n_pairs     = records to read (apply hc_upto_cutoff first for ``upto``)
 * h_stats[9]  = {lines kept by -q/-F/regions, pairs binned, skipped (contig absent),
 *               out-of-range, key errors, stragglers (records whose row their chunk does not advertise),
 *               1 = the sorted-input path handled everything / 0 = forced atomic kernel or the straggler
 *               list overflowed and the stragglers were re-scanned, tile shape,
 *               binnable records whose row lies outside [row0,row1) (left to the call that owns that block)}
 * mode: bit 0 clear = auto (row-privatised shared-memory tiles; records of unsorted input that no tile sees are
 *       added with global atomics afterwards, decided on the device), bit 0 set = force the global-atomic kernel;
 *       bit 1 set = accumulate into counts (several BAM files, bam2clusters.py:133) instead of
 *       overwriting it.
 */
size_t hc_bin_pairs_workspace_bytes(int64_t n_pairs, int64_t n_rows);
int hc_bin_pairs(const hc_pair_t *pairs, int64_t n_pairs,
                 const int32_t *offsets, const int32_t *member, const uint8_t *region_mask,
                 int32_t n_contigs, uint32_t window_bp, int32_t n_windows,
                 int32_t mapq_min, uint32_t skip_flag, int32_t row0, int32_t row1,
                 uint32_t *counts, void *workspace, size_t workspace_bytes,
                 int mode, int64_t *h_stats, void *stream);

/* h_stats may be NULL: the call is then asynchronous (no device-to-host copy, no stream synchronisation; every
 * fast-path / fallback decision is taken on the device) and the stats, which live in the first 128 bytes of
 * `workspace`, are read later with hc_bin_pairs_stats, which also returns HC_EKEY / HC_ERANGE for the call. */
int hc_bin_pairs_stats(const void *workspace, int64_t *h_stats, void *stream);
/* All window sizes of one bam2arrays call from ONE read of the records (inner loop bam2clusters.py:154-161): each
 * kept line is added to counts[ii] for every ii < n_ws (<= 4) with warp-aggregated global atomics -- meant for the
 * small per-cluster matrices of the scaffolding stage (windowSize 5 / 10 / 2 kb, bam2scaffolds.py:379), which stay
 * resident in L2 together.  h_offsets / h_counts: HOST arrays of n_ws DEVICE pointers; h_window_bp / h_n_windows:
 * host arrays.  Membership is tested against h_offsets[0] (:151); a contig absent from a later table is HC_EKEY
 * (:155), a window past the matrix HC_ERANGE (:161).  workspace >= 256 bytes; h_stats[9] as hc_bin_pairs. */
int hc_bin_pairs_multi(const hc_pair_t *pairs, int64_t n_pairs, int32_t n_ws, const int32_t *const *h_offsets,
                       const uint8_t *region_mask, int32_t n_contigs, const uint32_t *h_window_bp,
                       const int32_t *h_n_windows, int32_t mapq_min, uint32_t skip_flag, uint32_t *const *h_counts,
                       int accumulate, void *workspace, size_t workspace_bytes, int64_t *h_stats, void *stream);

int hc_unpack_pairs16(const hc_pair16_t *packed, int64_t n_pairs, hc_pair_t *pairs, void *stream);

/* Host-only BAM decoder (csrc/bam_decode.cu): replaces the `samtools view` text pipe of
 * bam2clusters.py:108-117,139-146.  BGZF blocks (bin/snap/SNAPLib/Bam.h:322-350) are inflated and CRC-checked
 * by n_threads zlib workers (0 = all cores); alignment records (Bam.h:91-105) are handed over in file order,
 * unfiltered (-q / -F run in hc_bin_pairs), as hc_pair_t (wire 24) or hc_pair16_t (wire 16), straight into the
 * caller's (pinned) buffer.  h_tlen / h_ref_span (may be NULL): TLEN and the CIGAR reference span (>= 1) for
 * the contig-level path (:183-189).  hc_bam_next returns up to `capacity` records per call, *h_n = 0 at end. */
int hc_bam_open(const char *path, int32_t n_threads, void **h_handle, int32_t *h_n_ref,
                int64_t *h_uncompressed_bytes);
int hc_bam_next(void *handle, int32_t wire, void *h_out, int64_t capacity, int32_t *h_tlen,
                int32_t *h_ref_span, int64_t *h_n);
int hc_bam_close(void *handle);

/* ``upto`` of bam2arrays (:139-141): number of leading records that contain exactly ``upto``
 * lines surviving -q/-F/regions (n_pairs if fewer survive).  workspace >= 8*(n_pairs/4096+2). */
int hc_upto_cutoff(const hc_pair_t *pairs, int64_t n_pairs, const uint8_t *region_mask,
                   int32_t n_contigs, int32_t mapq_min, uint32_t skip_flag, int64_t upto,
                   void *workspace, size_t workspace_bytes, int64_t *h_n_effective,
                   int64_t *h_lines, void *stream);

/* Contig-level side outputs of _bam2array (:183-189, region queries :179,222): for every record
 * kept by -q/-F whose alignment [pos, pos+ref_span) overlaps a sampled region
 * (s+1, s+minSize), s = 0, nth*minSize, ... of its contig, write keep[k]=1 and
 * dist[k] = |tlen| (same contig) or -1 (other contig); else keep[k]=0. */
int hc_contig_region_filter(const hc_pair_t *pairs, int64_t n_pairs, const int32_t *tlen,
                            const int32_t *ref_span, const int32_t *contig_size,
                            const int32_t *member, int32_t n_contigs, int32_t min_size,
                            int32_t nth, int32_t mapq_min, uint32_t skip_flag, uint8_t *keep,
                            int32_t *dist, void *stream);

/* ---------------------------------------------------------------------------------------
 * K2  fold / symmetrise.
 * hc_fold_counts: counts (u32, n x n, full orientation) -> float32 n x n
 *   mode 0: reference layout of bam2arrays: upper triangle incl. diagonal, zeros below
 *   mode 1: symmetrised  d += d.T - diag(d)  (bam2clusters.py:42,511; bam2scaffolds.py:262),
 *           plus ``add`` (the +1 pseudo-count of normalize(), :43); d_vmax receives max(out)
 * out may alias counts (in place).  d_vmax: device float (may be NULL for mode 0). */
int hc_fold_counts(const uint32_t *counts, int32_t n, int mode, float add, float *out,
                   float *d_vmax, void *stream);
/* in-place ``d += d.T - np.diag(d.diagonal())`` (+add) for float32 (dtype 0) / float64 (1);
 * d_vmax (same dtype) receives max. Equivalent to array2scaffolds.py:117-118. */
int hc_symmetrize(void *d, int32_t n, int dtype, double add, void *d_vmax, void *stream);
/* tail of bam2array (bam2clusters.py:244): (array / sizes).T / sizes on the contig-level counts
 * (u32, full orientation) -> float64 n x n, transposed (lower triangular) like the reference */
int hc_contig_normalize(const uint32_t *counts, int32_t n, const int32_t *sizes, double *out,
                        void *stream);

/* ---------------------------------------------------------------------------------------
 * K3-K5  Sinkhorn-Knopp (bin/sinkhorn_knopp/sinkhorn_knopp/sinkhorn_knopp.py:93-163 as called
 * by normalize(), bam2clusters.py:36-49).  X is the row block [row0,row0+nrows) x n of the
 * matrix, dtype 0 = float32 / 1 = float64; P = X / vmax elementwise in X's dtype (vmax = 1 for a
 * plain fit).  r (nrows) and c (n) are float64.
 *   hc_sk_colpass : ypart[b][j] = sum over rows of row-chunk b of P[i][j] * r[i]   (r NULL = ones)
 *                   also flags negative / NaN entries on the first call (d_flags[0]/[1])
 *   hc_sk_colfinish: y[j] = sum_b ypart[b][j]  (fixed order => deterministic)
 *   hc_sk_update_c : c = 1/y ; d_flags[2] |= any(y == 0); if c_prev != NULL also evaluates the
 *                   stopping test of :138-141 on column sums c_prev[j]*y[j] -> d_flags[3] = 1
 *                   when some column is outside 1 +- eps
 *   hc_sk_rowpass : x[i] = sum_j P[i][j]*c[j]  (c NULL = ones); r[i] = 1/x[i] if r != NULL
 *   hc_sk_apply   : out[i][j] = (r[i]*P[i][j])*c[j]   (out dtype 0 = f32, 1 = f64)
 * Row-sharded use: every rank runs colpass+colfinish on its block, all-reduces y (NCCL sum), then
 * update_c / rowpass locally.
 * Symmetric P (the folded contact matrix d + d.T of bam2clusters.py:505-509 always is): both matvecs
 * are y = P v and read the upper triangle of 256 x 256 tiles only (2 n^2 bytes instead of 4 n^2):
 *   hc_sk_sympass  : part[b][k] = contribution of tile (min(K,b), max(K,b)) to y[k], k in block K
 *                    (v NULL = ones; part holds hc_sk_sym_num_blocks(n) x n float64; d_flags != NULL
 *                    also flags negative / NaN entries in d_flags[0])
 *   hc_sk_symfinish: y[k] = sum_b part[b][k] in fixed order; if r != NULL: r = 1/y and
 *                    d_flags[4] |= any(y == 0) */
int32_t hc_sk_num_row_chunks(int32_t nrows);
int hc_sk_colpass(const void *X, int dtype, int32_t nrows, int32_t n, double vmax,
                  const double *r, double *ypart, int32_t *d_flags, void *stream);
int hc_sk_colfinish(const double *ypart, int32_t n_chunks, int32_t n, double *y, void *stream);
int hc_sk_update_c(const double *y, const double *c_prev, int32_t n, double eps, double *c,
                   int32_t *d_flags, void *stream);
int hc_sk_rowpass(const void *X, int dtype, int32_t nrows, int32_t n, double vmax,
                  const double *c, double *x, double *r, int32_t *d_flags, void *stream);
int32_t hc_sk_sym_num_blocks(int32_t n);
int hc_sk_sympass(const void *X, int dtype, int32_t n, double vmax, const double *v, double *part,
                  int32_t *d_flags, void *stream);
int hc_sk_symfinish(const double *part, int32_t n, double *y, double *r, int32_t *d_flags, void *stream);
/* Self-check of the hoisted-reciprocal form of P = X / vmax used by the float32 passes: every float of
 * the binades [1,4) divided by each of the nb divisors d_b must equal the IEEE division bit for bit;
 * *d_mismatches (device u64) receives the number of differences (expected 0). */
int hc_selftest_scaled_division(const float *d_b, int32_t nb, unsigned long long *d_mismatches, void *stream);
/* *d_flag |= any(v outside 1 +- eps): the first evaluation of the while test (:138-141) on raw P */
int hc_sk_outside(const double *v, int32_t n, double eps, int32_t *d_flag, void *stream);
int hc_sk_apply(const void *X, int dtype, int32_t nrows, int32_t n, double vmax, const double *r,
                const double *c, void *out, int out_dtype, void *stream);

/* ---------------------------------------------------------------------------------------
 * Row-sharded build of one matrix over the GPUs of a node (csrc/sharded.cu).  The reference's analogue is
 * the split of the contigs over Pool workers and the reduction in the parent (bam2clusters.py:224-230,
 * bam2plot.py:138-161).  Rank k owns rows [k*per, (k+1)*per), per % 64 == 0.
 *
 * hc_shared_alloc / _open / _close / _free: device memory that peers can map (cudaMalloc + cudaIpc handle, zeroed).
 *   The 64-byte handle travels through the caller's process group; hc_shared_open maps a PEER's buffer (a process
 *   cannot open its own handle).
 * hc_fold_block: d += d.T - diag(d) (+add) for the row block of `rank` (bam2clusters.py:42,511):
 *   out[i-row0][j] = C[i][j] + C[j][i]; C[j][i] is read from the counts block of the rank that owns row j
 *   (h_counts_blocks: HOST array of nranks device pointers, peer-mapped), so the transposition is the kernel's own
 *   NVLink read and not a separate collective.  out: float32 rows x n, must not alias a counts block; d_vmax
 *   (device float, zeroed by the caller) receives the block maximum -- all-reduce(max) it over the ranks.
 * hc_sk_fit_block: SinkhornKnopp.fit (sinkhorn_knopp.py:114-163) on the row block X of P = X / vmax, whole fit in one
 *   call.  The column-sum vector of every pass is all-reduced by a one-shot exchange through peer memory fused with
 *   the column-chunk reduction and the c = 1/y update / stopping test; the loop condition lives on the device
 *   (kernels of iterations past convergence return at once), the host reads it once per 3 queued iterations.
 *   h_exchange: HOST array of nranks device pointers to the ranks' exchange regions (hc_sk_exchange_bytes(n) bytes
 *   each, from hc_shared_alloc, peer-mapped; NULL when nranks == 1).  *h_epoch: exchange counter, starts at 0 and is
 *   carried from call to call (all ranks make the same calls).  r (nrows) and c (n): float64 device outputs.
 *   h_result[8] = {iterations (iterations0 carried in, :150-152), stop (0 epsilon / 1 max_iter), negative-or-NaN
 *   entry, zero column sum, zero row sum -- the three flags all-reduced so every rank raises / warns alike --,
 *   matrix passes executed, kernels launched, loop iterations queued}. */
int hc_sharded_preload(void);   /* load the kernels of the row-sharded path now (no lazy load inside an exchange) */
int hc_shared_alloc(size_t bytes, void **h_dptr, void *h_handle64);
int hc_shared_free(void *dptr);
int hc_shared_open(const void *h_handle64, void **h_dptr);
int hc_shared_close(void *dptr);
int hc_fold_block(const uint32_t *const *h_counts_blocks, int32_t nranks, int32_t rank, int32_t per, int32_t n,
                  float add, float *out, float *d_vmax, void *stream);
size_t hc_sk_exchange_bytes(int32_t n);
size_t hc_sk_fit_block_workspace_bytes(int32_t nrows, int32_t n);
int hc_sk_fit_block(const void *X, int dtype, int32_t nrows, int32_t n, double vmax, double eps, int32_t max_iter,
                    int32_t iterations0, void *const *h_exchange, int32_t nranks, int32_t rank, int32_t *h_epoch,
                    void *workspace, size_t workspace_bytes, double *r, double *c, int32_t *h_result, void *stream);

/* ---------------------------------------------------------------------------------------
 * K6  contacts -> distances, written straight into the linkage working matrix.
 * Reads the UPPER triangle of X (n x n, dtype 0/1) -- the part d[np.triu_indices(n,1)]
 * selects (bam2clusters.py:385,411; array2scaffolds.py:163) -- and writes the full symmetric
 * float64 matrix W (n x n).  v = X[i][j], optionally balanced on the fly as
 * (r[i]*(v/vmax))*c[j] when r != NULL.
 *   kind 0: identity                       kind 1: (1/(a*(v+1)))**(1/b)   (bam2clusters.py:248-251,516)
 *   kind 2: log(vmax2) - log(v+1) computed in X's dtype (array2scaffolds.py:31), p0 = vmax2
 *   kind 3: (1/(a*v))**(1/b)  -- distance_func itself, no pseudo-count (bam2clusters.py:248-251)
 * square != 0 stores the squared distance (fastcluster works on squares for ward).
 * d_flags[0] is set if a NaN is produced. */
int hc_transform_to_working(const void *X, int dtype, int32_t n, int kind, double p0, double p1,
                            const double *r, const double *c, double vmax, int square, double *W,
                            int32_t *d_flags, void *stream);
/* The same from a ROW-SHARDED matrix: block k = rows [k*per, (k+1)*per) lives in h_blocks[k] (a HOST array of device
 * pointers -- on a multi-GPU node the peer-mapped row blocks of the other ranks, read over NVLink while the tile is
 * transformed: gather and transform are one kernel).  per % 64 == 0; r has all n rows. */
int hc_transform_gather(const void *const *h_blocks, int32_t nblocks, int32_t per, int dtype, int32_t n, int kind,
                        double p0, double p1, const double *r, const double *c, double vmax, int square,
                        double *W, int32_t *d_flags, void *stream);
/* elementwise transform of a full matrix (same kinds), out dtype = in dtype (kind 2) or f64 */
int hc_transform(const void *X, int dtype, int64_t count, int kind, double p0, double p1,
                 void *out, int out_dtype, void *stream);
/* y = d[np.triu_indices(n,1)] as float64 (condensed) and the inverse expansion */
int hc_condense(const void *X, int dtype, int32_t n, double *y, void *stream);
int hc_expand_condensed(const double *y, int32_t n, int square, double *W, int32_t *d_flags,
                        void *stream);

/* ---------------------------------------------------------------------------------------
 * K7  agglomerative clustering: GPU nearest-neighbour chain + Lance-Williams updates.
 * Replaces fastcluster.linkage(y, method) (bam2clusters.py:385,403,411,430) and
 * sch.linkage (array2scaffolds.py:163).  W is the full symmetric float64 working matrix from
 * hc_transform_to_working / hc_expand_condensed (squared for ward) and is destroyed.
 * method: 0 single, 1 complete, 2 average, 3 weighted, 4 ward.
 * h_Z receives the scipy linkage array (n-1) x 4 float64 (host).  h_stats[8] = {NN searches
 * of the chain, sum of active sizes over them, merges, sum of active sizes over updates, row scans
 * actually executed (the others were answered by the per-row NN cache), sum of active sizes over
 * executed scans, cached lookups, 0}.  The first four equal what the CPU library does step for step. */
size_t hc_linkage_workspace_bytes(int32_t n);
int hc_linkage(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes,
               double *h_Z, int64_t *h_stats, void *stream);
/* the same in two halves, so that independent linkages (the 30 bootstrap runs of cluster_contigs,
 * bam2clusters.py:427-431, or the per-cluster trees of bam2scaffolds.py:294-298) overlap on
 * different streams: launch is asynchronous, finish waits on the stream and labels the tree. */
int hc_linkage_launch(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes,
                      void *stream);
int hc_linkage_finish(int32_t n, int method, void *workspace, size_t workspace_bytes, double *h_Z,
                      int64_t *h_stats, void *stream);

/* K7b  round-parallel engine for the same call sites (csrc/linkage_rounds.cu): merges every reciprocal
 * nearest-neighbour pair per round on all SMs.  Same W contract as hc_linkage.  Every merge and the
 * sorted heights are certified against ties (1e-13 relative, 30 x the largest height difference
 * measured between the two engines); when certification fails the
 * call returns HC_EAMBIGUOUS with W destroyed and the caller falls back to hc_linkage, whose scan
 * order reproduces fastcluster's tie-breaking.  When it returns HC_OK, Z[:,0:2] and Z[:,3] equal
 * hc_linkage's and Z[:,2] agrees to rounding (same Lance-Williams operations, different interleaving).
 * A NEAR tie only leaves its pair pending (it is merged in a later round once it is strict); the call
 * gives up on an EXACT tie or when only near-tied pairs are left.
 * h_stats[24]: rounds, merges, rescans, sum live columns over rescans, sum live columns over merged-row
 * builds, row x pair updates, compactions, near-tied heights, pending near-tie observations, pairs
 * deferred for an exact rescan; [10..14] microseconds per kernel (select, light, heavy, compaction, initial
 * scan) when HC_ROUNDS_KTIME is set, else 0; [15] kernels launched; [16] microseconds in kb_rebuild (bulk rounds:
 * a fifth or more of the columns merge, done as one merge+compact pass over the matrix; HC_ROUNDS_REBUILD=0 switches
 * that path off), [17] microseconds in kr_chain (the persistent kernel that runs all later rounds, many per
 * cooperative launch; HC_ROUNDS_PERSIST=0 switches it off), [18] / [19] live columns summed over the free / merged
 * rows of the bulk rounds (the algorithmic bytes of kb_rebuild are 8 x [18] + 24 x [19]), [20] bulk rounds,
 * [21] 1 when the persistent kernel ran, [22] cells of the new matrices written by the bulk rounds. */
size_t hc_linkage_rounds_workspace_bytes(int32_t n);
int hc_linkage_rounds(double *W, int32_t n, int method, void *workspace, size_t workspace_bytes,
                      double *h_Z, int64_t *h_stats, void *stream);

/* Host-only tail shared by both engines (csrc/dendrogram.cu): n-1 merge records (a leaf of either cluster,
 * merge height; squared heights when ward != 0) -> scipy linkage array Z[(n-1)*4], as fastcluster's
 * generate_SciPy_dendrogram does after NN_chain_core (behind bam2clusters.py:385, array2scaffolds.py:163):
 * STABLE sort by height, union-find relabelling, row = (min label, max label, height, size).  h_d is
 * overwritten (sqrt for ward).  h_near_ties (may be NULL) receives the number of adjacent sorted heights
 * closer than rel_tol (relative).  No device work. */
int hc_merges_to_linkage(int32_t n, int ward, const int32_t *h_a, const int32_t *h_b, double *h_d, double *h_Z,
                         double rel_tol, int64_t *h_near_ties);

/* ---------------------------------------------------------------------------------------
 * K8  block reductions on the resident matrix.
 * hc_segment_mean_2d: _average_reduce_2d (bam2scaffolds.py:53-71) for window->contig keys whose
 *   windows form one contiguous run each: group k (in np.unique(keys) order) covers windows
 *   [seg_begin[k], seg_end[k]); float32 in, float32 out (n_seg x n_seg); tmp = n_seg x n floats.
 * hc_normalize_rows: normalize_rows (array2scaffolds.py:33-41), float32 in place;
 *   workspace >= 20*n + 1024 bytes. */
int hc_segment_mean_2d(const float *A, int32_t n, const int32_t *seg_begin, const int32_t *seg_end,
                       int32_t n_seg, float *tmp, float *out, void *stream);
int hc_normalize_rows(float *A, int32_t n, void *workspace, size_t workspace_bytes, void *stream);
/* hc_segment_mean_2d_static: hc_segment_mean_2d for callers that cannot allocate (CUDA graph capture, tight serving
 *   loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 *   Same inputs, same output bits; the n_seg x n intermediate lives in a file-scope `static __device__`
 *   array of HC_SEGMEAN_STATIC_MAX_ELEMS floats (8 MiB, held by every GPU that loads the library until exit), so the
 *   call performs no cudaMalloc / cudaFree and needs no tmp.  n * n_seg > HC_SEGMEAN_STATIC_MAX_ELEMS -> HC_EWORKSPACE
 *   with a message in hc_last_error().  All calls on a device share the one array (it keeps the last call's contents;
 *   stage 1 overwrites every element stage 2 reads, so nothing needs clearing): calls that overlap on the GPU --
 *   different streams without ordering -- race; order them on one stream or with events. */
#define HC_SEGMEAN_STATIC_MAX_ELEMS (1 << 21)
int hc_segment_mean_2d_static(const float *A, int32_t n, const int32_t *seg_begin, const int32_t *seg_end,
                              int32_t n_seg, float *out, void *stream);
/* hc_block_sums4: the corner-block sums of join_scaffolds (bam2scaffolds.py:155-161 = array2scaffolds.py:287-293),
 *   batched over the independent joins of one tree level.  Job j: scaffold-1 windows idx[job_off[j] ..
 *   +job_n1[j]), scaffold-2 windows right behind them (job_n2[j]); out[4*j + q], q = 0..3 in the reference's
 *   order (start-start, start-end, end-start, end-end), float64.  A = symmetric float32 window matrix (n x n). */
int hc_block_sums4(const float *A, int32_t n, const int32_t *idx, const int64_t *job_off, const int32_t *job_n1,
                   const int32_t *job_n2, int32_t n_jobs, double *out, void *stream);

/* hc_matrix_max: np.max over `count` elements (dtype 0 float32 / 1 float64; NaN propagates) into *d_max (same dtype);
 *   d_partial: n_partial scratch elements of the same dtype (array2scaffolds.py:31 ``np.max(x+1)``). */
int hc_matrix_max(const void *X, int dtype, int64_t count, void *d_partial, int32_t n_partial, void *d_max, void *stream);
/* hc_gather_elements: out[k] = A[rows[k]][cols[k]] -- the (anti)diagonal entries the other join_scaffolds
 *   (array2scaffolds.py:274-309: np.diag of the rectangle between two scaffolds, :288-291) compares, batched over the
 *   independent joins of one tree level; the four short float32 sums are then added up in numpy's order on the host. */
int hc_gather_elements(const float *A, int32_t n, const int32_t *rows, const int32_t *cols, int64_t m, float *out,
                       void *stream);

/* Alternative normalisations of the symmetric matrix, dtype 0 = float32, 1 = float64.
 * hc_gather_square: out (m x m) = X[keep][:, keep], the zero-diagonal filter of normalize_diagional
 *   (bam2clusters.py:55-57); also the sub-matrix selection of cluster_contigs (:428).
 * hc_diag_normalize: out = d * ((d*n2).T*n2), n2 = mean/diag(d) in the matrix dtype (bam2clusters.py:64-65);
 *   `mean` = np.mean(d.diagonal()) supplied by the caller (numpy's pairwise float32 sum stays numpy's);
 *   out must not alias D.
 * hc_window_size_normalize: d *= 1.*windowSize**2 / (sizes*sizes.T) in place (bam2scaffolds.py:40);
 *   sizes = int64 window lengths, ws2 = windowSize**2; float64 factor, product cast back to the dtype. */
int hc_gather_square(const void *X, int dtype, int32_t n, const int32_t *keep, int32_t m, void *out, void *stream);
int hc_diag_normalize(const void *D, int dtype, int32_t n, double mean, void *out, void *stream);
int hc_window_size_normalize(void *D, int dtype, int32_t n, const int64_t *sizes, double ws2, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* HICEMBLER_B200_H */
