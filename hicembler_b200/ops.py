"""Device-resident operators: torch CUDA tensors in, torch CUDA tensors out.

torch is used for device memory and streams only; every computation is a call into
libhicembler_b200.so (hand-written sm_100a kernels).  These are the building blocks the
reference-named drop-ins in ``bam2clusters.py`` / ``bam2scaffolds.py`` / ``array2scaffolds.py`` /
``sinkhorn_knopp.py`` / ``fastcluster.py`` are made of, and the "fast path" that keeps N x N data in
HBM from pair records to the linkage array.
"""
from __future__ import annotations

import ctypes as C
import os
import warnings

import numpy as np
import torch

from . import _lib
from .pairs import PAIR_DTYPE, SKIP_FLAG

METHODS = {"single": 0, "complete": 1, "average": 2, "weighted": 3, "ward": 4}
KIND_IDENTITY, KIND_POWER, KIND_LOGMAX, KIND_POWER_RAW = 0, 1, 2, 3
_UINT32_MAX = 0xFFFFFFFF


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev(t: torch.Tensor):
    if not t.is_cuda:
        raise RuntimeError("hicembler_b200 operators need CUDA tensors (there is no CPU fallback)")
    _lib.check(_lib.lib().hc_set_device(t.device.index))


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _dtype_code(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.float64:
        return 1
    raise TypeError("matrix must be float32 or float64, got %s" % t.dtype)


def device_info():
    sm, mem, major, minor = C.c_int(), C.c_size_t(), C.c_int(), C.c_int()
    _lib.check(_lib.lib().hc_device_info(C.byref(sm), C.byref(mem), C.byref(major), C.byref(minor)))
    return dict(sm_count=sm.value, total_mem=mem.value, cc=(major.value, minor.value))


def to_device_records(rec, device="cuda"):
    """numpy PAIR_DTYPE array / int32 [n,6] tensor -> int32 [n,6] CUDA tensor (24-byte records)."""
    if isinstance(rec, torch.Tensor):
        t = rec
    else:
        rec = np.ascontiguousarray(rec)
        if rec.dtype != PAIR_DTYPE:
            raise TypeError("records must have pairs.PAIR_DTYPE")
        t = torch.from_numpy(rec.view(np.int32).reshape(-1, 6))
    if t.dtype != torch.int32 or t.dim() != 2 or t.shape[1] != 6:
        raise TypeError("record tensor must be int32 [n, 6]")
    return t.to(device, non_blocking=True).contiguous()


# ---------------------------------------------------------------------------------------------
# K1
# ---------------------------------------------------------------------------------------------
def upto_cutoff(records, mapq, upto, region_mask=None, skip_flag=SKIP_FLAG):
    """(n_effective, lines): leading records containing ``upto`` lines that survive -q/-F."""
    _dev(records)
    n = records.shape[0]
    ws = torch.empty(8 * (n // 4096 + 2) + 512, dtype=torch.uint8, device=records.device)
    n_eff, lines = C.c_int64(), C.c_int64()
    n_contigs = 0 if region_mask is None else region_mask.numel()
    _lib.check(_lib.lib().hc_upto_cutoff(_ptr(records), n, _ptr(region_mask), n_contigs, int(mapq),
                                         skip_flag, int(upto), _ptr(ws), ws.numel(),
                                         C.byref(n_eff), C.byref(lines), _stream()))
    return n_eff.value, lines.value


def _bin_stats_dict(stats):
    return dict(lines=stats[0], binned=stats[1], skipped_absent=stats[2], out_of_range=stats[3],
                key_errors=stats[4], overflow=stats[5], fast_path=bool(stats[6]),
                tile_rows=stats[7] >> 32, col_segments=stats[7] & 0xFFFFFFFF, outside_rows=stats[8])


class DeferredBinStats:
    """Stats of an asynchronous hc_bin_pairs call: nothing was copied to the host and the stream was not synchronised;
    ``read()`` does both (and raises IndexError / KeyError if the call met the reference's error cases)."""

    def __init__(self, workspace, stream):
        self.workspace, self.stream, self._st = workspace, stream, None

    def snapshot(self):
        """keep the 128 stats bytes in a buffer of their own, so that the workspace can serve the next call"""
        keep = torch.empty(128, dtype=torch.uint8, device=self.workspace.device)
        keep.copy_(self.workspace[:128], non_blocking=True)
        self.workspace = keep
        return self

    def read(self):
        if self._st is None:
            stats = (C.c_int64 * 16)()
            rc = _lib.lib().hc_bin_pairs_stats(_ptr(self.workspace), stats, self.stream)
            self._st = _bin_stats_dict(stats)
            _lib.check(rc)
        return self._st


def bin_pairs(records, offsets, window_bp, n_windows, mapq, member=None, region_mask=None,
              n_pairs=None, rows=None, skip_flag=SKIP_FLAG, mode=0, out=None, workspace=None,
              return_stats=False, defer_stats=False):
    """Scatter-add read pairs into the full-orientation u32 count matrix (K1).

    offsets/member: int32 CUDA tensors indexed by refID (-1 = absent).  ``window_bp=None`` bins
    one window per contig (contig-level matrix).  Returns int32-typed tensor holding u32 counts of
    shape (rows, n_windows).  ``defer_stats=True``: asynchronous call (no host synchronisation); returns
    (out, DeferredBinStats) -- the workspace must stay untouched until the stats have been read."""
    _dev(records)
    dev = records.device
    n = records.shape[0] if n_pairs is None else int(n_pairs)
    row0, row1 = (0, n_windows) if rows is None else rows
    if out is None:
        out = torch.empty((row1 - row0, n_windows), dtype=torch.int32, device=dev)
    need = _lib.lib().hc_bin_pairs_workspace_bytes(n, row1 - row0)
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty(need, dtype=torch.uint8, device=dev)
    stats = (C.c_int64 * 16)()
    wbp = _UINT32_MAX if window_bp is None else int(window_bp)
    stream = _stream()
    rc = _lib.lib().hc_bin_pairs(_ptr(records), n, _ptr(offsets), _ptr(member), _ptr(region_mask),
                                 offsets.numel(), wbp, int(n_windows), int(mapq), skip_flag,
                                 row0, row1, _ptr(out), _ptr(workspace), workspace.numel(), mode,
                                 None if defer_stats else stats, stream)
    _lib.check(rc)
    if defer_stats:
        return out, DeferredBinStats(workspace, stream)
    st = _bin_stats_dict(stats)
    return (out, st) if return_stats else out


def bin_pairs_multi(records, offsets_list, window_bps, n_windows_list, mapq, region_mask=None, n_pairs=None,
                    skip_flag=SKIP_FLAG, outs=None, accumulate=False):
    """All window sizes of one bam2arrays call from ONE read of the records (hc_bin_pairs_multi; <= 4 sizes).
    offsets_list[0] is also the membership table (bam2clusters.py:151).  Returns (list of count tensors, stats)."""
    _dev(records)
    dev = records.device
    k = len(window_bps)
    n = records.shape[0] if n_pairs is None else int(n_pairs)
    if outs is None:
        outs = [None] * k
    outs = [o if o is not None else torch.empty((nw, nw), dtype=torch.int32, device=dev) for o, nw in zip(outs, n_windows_list)]
    ws = torch.empty(256, dtype=torch.uint8, device=dev)
    stats = (C.c_int64 * 16)()
    offs = (C.c_void_p * k)(*[o.data_ptr() for o in offsets_list])
    cnts = (C.c_void_p * k)(*[o.data_ptr() for o in outs])
    wbp = (C.c_uint32 * k)(*[int(w) for w in window_bps])
    nws = (C.c_int32 * k)(*[int(w) for w in n_windows_list])
    rc = _lib.lib().hc_bin_pairs_multi(_ptr(records), n, k, offs, _ptr(region_mask), offsets_list[0].numel(), wbp, nws,
                                       int(mapq), skip_flag, cnts, int(bool(accumulate)), _ptr(ws), ws.numel(), stats,
                                       _stream())
    _lib.check(rc)
    return outs, _bin_stats_dict(stats)


def contig_region_filter(records, tlen, ref_span, contig_size, member, min_size, nth, mapq,
                         skip_flag=SKIP_FLAG):
    _dev(records)
    n = records.shape[0]
    keep = torch.empty(n, dtype=torch.uint8, device=records.device)
    dist = torch.empty(n, dtype=torch.int32, device=records.device)
    _lib.check(_lib.lib().hc_contig_region_filter(
        _ptr(records), n, _ptr(tlen), _ptr(ref_span), _ptr(contig_size), _ptr(member),
        contig_size.numel(), int(min_size), int(nth), int(mapq), skip_flag, _ptr(keep), _ptr(dist),
        _stream()))
    return keep, dist


# ---------------------------------------------------------------------------------------------
# K2
# ---------------------------------------------------------------------------------------------
def fold_counts(counts, mode="upper", add=0.0, inplace=False):
    """counts (u32 full orientation) -> float32: 'upper' (what bam2arrays returns) or 'sym'
    (after d += d.T - diag(d), plus ``add``).  Returns (matrix, vmax tensor or None)."""
    _dev(counts)
    n = counts.shape[0]
    assert counts.shape == (n, n) and counts.dtype == torch.int32
    out = counts.view(torch.float32) if inplace else torch.empty((n, n), dtype=torch.float32,
                                                                 device=counts.device)
    vmax = torch.zeros(1, dtype=torch.float32, device=counts.device)
    _lib.check(_lib.lib().hc_fold_counts(_ptr(counts), n, 0 if mode == "upper" else 1, float(add),
                                         _ptr(out), _ptr(vmax), _stream()))
    return out, vmax


def symmetrize_(d, add=0.0):
    """In place ``d += d.T - np.diag(d.diagonal())`` (+ add).  Returns (d, max tensor)."""
    _dev(d)
    n = d.shape[0]
    assert d.shape == (n, n) and d.is_contiguous()
    vmax = torch.zeros(1, dtype=d.dtype, device=d.device)
    _lib.check(_lib.lib().hc_symmetrize(_ptr(d), n, _dtype_code(d), float(add), _ptr(vmax), _stream()))
    return d, vmax


# ---------------------------------------------------------------------------------------------
# K3-K5
# ---------------------------------------------------------------------------------------------
class SinkhornResult:
    __slots__ = ("r", "c", "iterations", "stopping_condition", "warned", "matrix_passes", "launches")


class _CudaSinkhornPasses:
    """The per-rank kernels behind sk_driver.sinkhorn_driver: one row block of P = X / vmax."""

    def __init__(self, X, vmax, epsilon, symmetric=False):
        _dev(X)
        self.L = _lib.lib()
        self.X, self.vmax, self.eps = X, float(vmax), float(epsilon)
        self.nrows, self.n = X.shape
        self.dt = _dtype_code(X)
        f64 = dict(dtype=torch.float64, device=X.device)
        # symmetric P, whole matrix on this rank: both matvecs read the upper triangle only
        self.sym = bool(symmetric) and self.nrows == self.n
        self.nchunks = self.L.hc_sk_sym_num_blocks(self.n) if self.sym else self.L.hc_sk_num_row_chunks(self.nrows)
        self.ypart = torch.empty((self.nchunks, self.n), **f64)
        self.y, self.c, self.c_new = (torch.empty(self.n, **f64) for _ in range(3))
        self.x, self.r = torch.empty(self.nrows, **f64), torch.empty(self.nrows, **f64)
        self.flags = torch.zeros(8, dtype=torch.int32, device=X.device)
        self.outside = torch.zeros(1, dtype=torch.int32, device=X.device)
        self.s = _stream()
        self.matrix_passes = 0   # full reads of the row block (a symmetric pass counts (nb+1)/(2 nb))
        self.launches = 0

    def _sympass(self, v, y, r=None, check=False):
        L = self.L
        self.matrix_passes += (self.nchunks + 1) / (2.0 * self.nchunks)
        self.launches += 2
        _lib.check(L.hc_sk_sympass(_ptr(self.X), self.dt, self.n, self.vmax, _ptr(v) if v is not None else None,
                                   _ptr(self.ypart), _ptr(self.flags) if check else None, self.s))
        _lib.check(L.hc_sk_symfinish(_ptr(self.ypart), self.n, _ptr(y), _ptr(r) if r is not None else None,
                                     _ptr(self.flags), self.s))

    def colpass(self, use_r, check=False):
        """y_local = P_block^T r (r = ones when use_r is False); returns the tensor to all-reduce"""
        if self.sym:
            self._sympass(self.r if use_r else None, self.y, check=check)
            return self.y
        L = self.L
        self.matrix_passes += 1
        self.launches += 2
        _lib.check(L.hc_sk_colpass(_ptr(self.X), self.dt, self.nrows, self.n, self.vmax,
                                   _ptr(self.r) if use_r else None, _ptr(self.ypart),
                                   _ptr(self.flags) if check else None, self.s))
        _lib.check(L.hc_sk_colfinish(_ptr(self.ypart), self.nchunks, self.n, _ptr(self.y), self.s))
        return self.y

    def update_c(self, test):
        """c = 1/y; with test=True first evaluates the column-sum test on c*y -> True if outside"""
        self.launches += 1
        if test:
            self.flags[3] = 0
            _lib.check(self.L.hc_sk_update_c(_ptr(self.y), _ptr(self.c), self.n, self.eps, _ptr(self.c_new),
                                             _ptr(self.flags), self.s))
            if not bool(self.flags[3].item()):
                return False
            self.c, self.c_new = self.c_new, self.c
            return True
        _lib.check(self.L.hc_sk_update_c(_ptr(self.y), None, self.n, self.eps, _ptr(self.c), _ptr(self.flags), self.s))
        return True

    def raw_outside(self):
        """first evaluation of the while test on P itself (row sums, column sums); local flag tensor.
        The test is an OR over all row and column sums: the column sums are already there (first column pass,
        all-reduced when sharded, so every rank decides alike), and only if none of them is outside 1 +- eps
        does the row-sum pass over the matrix have to run (never for an unnormalised contact matrix)."""
        L = self.L
        self.launches += 1
        _lib.check(L.hc_sk_outside(_ptr(self.y), self.n, self.eps, _ptr(self.outside), self.s))
        if self.sym or bool(self.outside.item()):     # symmetric: the row sums ARE the column sums
            return self.outside
        self.matrix_passes += 1
        self.launches += 2
        _lib.check(L.hc_sk_rowpass(_ptr(self.X), self.dt, self.nrows, self.n, self.vmax, None, _ptr(self.x), None,
                                   _ptr(self.flags), self.s))
        self.flags[4] = 0
        _lib.check(L.hc_sk_outside(_ptr(self.x), self.nrows, self.eps, _ptr(self.outside), self.s))
        return self.outside

    def rowpass(self):
        """r = 1/(P_block c)"""
        if self.sym:
            self._sympass(self.c, self.x, self.r)
            return
        self.matrix_passes += 1
        self.launches += 1
        _lib.check(self.L.hc_sk_rowpass(_ptr(self.X), self.dt, self.nrows, self.n, self.vmax, _ptr(self.c),
                                        _ptr(self.x), _ptr(self.r), _ptr(self.flags), self.s))

    def read_flags(self):
        h = self.flags.tolist()
        return dict(negative=bool(h[0]), zero_col=bool(h[2]), zero_row=bool(h[4]))

    def flag_value(self, t):
        return bool(t.item())


def sinkhorn_fit(X, vmax=1.0, max_iter=1000, epsilon=1e-3, iterations0=0, allreduce=None, symmetric=False):
    """Sinkhorn-Knopp scaling vectors for P = X / vmax (sinkhorn_knopp.py:93-163).

    X: (nrows, n) CUDA float32/float64 row block (the whole matrix when not sharded).
    ``allreduce(t)`` sums a tensor over ranks in place (row-sharded use: the float64 column-sum
    vector once per column pass).  ``iterations0`` is SinkhornKnopp._iterations carried over from
    earlier fits (:150-152).  ``symmetric=True`` is the caller's promise that X == X.T bitwise (the folded
    contact matrix): the whole-matrix fit then reads the upper triangle only."""
    from .sk_driver import sinkhorn_driver
    passes = _CudaSinkhornPasses(X, vmax, epsilon, symmetric)
    iterations, stop, warned = sinkhorn_driver(passes, max_iter, iterations0, allreduce)
    res = SinkhornResult()
    res.r, res.c, res.iterations, res.stopping_condition, res.warned = passes.r, passes.c, iterations, stop, warned
    res.matrix_passes, res.launches = passes.matrix_passes, passes.launches
    return res


def sinkhorn_apply(X, r, c, vmax=1.0, out_dtype=torch.float64, out=None):
    """P_eps = D1.dot(P).dot(D2) (:159-161) as an elementwise scaling."""
    _dev(X)
    nrows, n = X.shape
    if out is None:
        out = torch.empty((nrows, n), dtype=out_dtype, device=X.device)
    _lib.check(_lib.lib().hc_sk_apply(_ptr(X), _dtype_code(X), nrows, n, float(vmax), _ptr(r), _ptr(c),
                                      _ptr(out), _dtype_code(out), _stream()))
    return out


# ---------------------------------------------------------------------------------------------
# K6 / K7
# ---------------------------------------------------------------------------------------------
def transform_to_working(X, kind=KIND_IDENTITY, p0=0.0, p1=1.0, r=None, c=None, vmax=1.0,
                         square=False, out=None):
    """Upper triangle of X -> full symmetric float64 linkage working matrix (K6)."""
    _dev(X)
    n = X.shape[0]
    assert X.shape == (n, n) and X.is_contiguous()
    if out is None:
        out = torch.empty((n, n), dtype=torch.float64, device=X.device)
    flags = torch.zeros(2, dtype=torch.int32, device=X.device)
    _lib.check(_lib.lib().hc_transform_to_working(_ptr(X), _dtype_code(X), n, kind, float(p0), float(p1),
                                                  _ptr(r), _ptr(c), float(vmax), int(bool(square)),
                                                  _ptr(out), _ptr(flags), _stream()))
    return out, flags


def transform(X, kind, p0=0.0, p1=1.0, out_dtype=None):
    _dev(X)
    X = X.contiguous()
    if out_dtype is None:      # numpy keeps a float32 array float32 through log / pow with scalar parameters
        out_dtype = X.dtype if kind in (KIND_LOGMAX, KIND_POWER, KIND_POWER_RAW) else torch.float64
    out = torch.empty(X.shape, dtype=out_dtype, device=X.device)
    _lib.check(_lib.lib().hc_transform(_ptr(X), _dtype_code(X), X.numel(), kind, float(p0), float(p1),
                                       _ptr(out), _dtype_code(out), _stream()))
    return out


def condense(X):
    _dev(X)
    n = X.shape[0]
    y = torch.empty(n * (n - 1) // 2, dtype=torch.float64, device=X.device)
    _lib.check(_lib.lib().hc_condense(_ptr(X.contiguous()), _dtype_code(X), n, _ptr(y), _stream()))
    return y


def expand_condensed(y, n, square=False):
    _dev(y)
    W = torch.empty((n, n), dtype=torch.float64, device=y.device)
    flags = torch.zeros(2, dtype=torch.int32, device=y.device)
    _lib.check(_lib.lib().hc_expand_condensed(_ptr(y), n, int(bool(square)), _ptr(W), _ptr(flags), _stream()))
    return W, flags


def linkage_working(W, method="ward", nan_flags=None, return_stats=False):
    """NN-chain linkage of a working matrix (destroyed).  Returns Z (numpy (n-1, 4) float64)."""
    _dev(W)
    n = W.shape[0]
    if nan_flags is not None and int(nan_flags[0].item()):
        raise FloatingPointError("NaN dissimilarity value.")
    L = _lib.lib()
    ws = torch.empty(L.hc_linkage_workspace_bytes(n), dtype=torch.uint8, device=W.device)
    Z = np.empty((max(n - 1, 0), 4), dtype=np.float64)
    stats = (C.c_int64 * 8)()
    _lib.check(L.hc_linkage(_ptr(W), n, METHODS[method], _ptr(ws), ws.numel(),
                            Z.ctypes.data_as(C.c_void_p), stats, _stream()))
    if return_stats:
        return Z, dict(scans=stats[0], scan_active=stats[1], merges=stats[2], update_active=stats[3],
                       row_scans=stats[4], row_scan_active=stats[5], cached_lookups=stats[6])
    return Z


def linkage_rounds(W, method="ward", nan_flags=None, workspace=None):
    """Round-parallel linkage of a working matrix (destroyed), csrc/linkage_rounds.cu.  Returns
    (Z, stats) when every merge and the sorted heights could be certified tie-free, else (None, stats):
    the caller rebuilds W and uses linkage_working, whose scan order reproduces fastcluster's ties."""
    _dev(W)
    n = W.shape[0]
    if nan_flags is not None and int(nan_flags[0].item()):
        raise FloatingPointError("NaN dissimilarity value.")
    L = _lib.lib()
    need = L.hc_linkage_rounds_workspace_bytes(n)
    ws = workspace if workspace is not None and workspace.numel() >= need else \
        torch.empty(need, dtype=torch.uint8, device=W.device)
    Z = np.empty((max(n - 1, 0), 4), dtype=np.float64)
    stats = (C.c_int64 * 24)()
    rc = L.hc_linkage_rounds(_ptr(W), n, METHODS[method], _ptr(ws), ws.numel(),
                             Z.ctypes.data_as(C.c_void_p), stats, _stream())
    st = dict(engine="rounds", rounds=stats[0], merges=stats[1], rescans=stats[2], rescan_cols=stats[3],
              build_cols=stats[4], light_updates=stats[5], compactions=stats[6], near_tied_heights=stats[7],
              pending_near_ties=stats[8], deferred_rescans=stats[9],
              kernel_us=dict(select=stats[10], light=stats[11], heavy=stats[12], compaction=stats[13],
                             initial_scan=stats[14], rebuild=stats[16], chain=stats[17]), launches=stats[15],
              bulk_rescan_cols=stats[18], bulk_build_cols=stats[19], bulk_rounds=stats[20], persistent=bool(stats[21]), bulk_new_cells=stats[22],
              certified=(rc == _lib.HC_OK))
    if rc == _lib.HC_EAMBIGUOUS:
        st["reason"] = _lib.last_error()
        if os.environ.get("HC_ROUNDS_DIAG") and st["merges"] == n - 1:
            st["Z_uncertified"] = Z     # diagnostics only (tests / measurements), never returned as a result
        return None, st
    _lib.check(rc)
    return Z, st


def linkage_auto(make_working, method="ward", engine="auto", return_stats=False):
    """make_working() -> (W, nan_flags) builds a fresh working matrix.  engine: "exact" = serial NN-chain
    (fastcluster's tie order), "rounds" = round-parallel (raises if it cannot certify), "auto" = rounds,
    falling back to exact on a fresh matrix when a tie is met."""
    if engine not in ("auto", "exact", "rounds"):
        raise ValueError("engine must be auto, exact or rounds")
    W, flags = make_working()
    if engine != "exact" and W.shape[0] >= 2:
        Z, st = linkage_rounds(W, method, nan_flags=flags)
        if Z is not None:
            return (Z, st) if return_stats else Z
        if engine == "rounds":
            raise ArithmeticError("round-parallel linkage could not certify the merge order: " + st.get("reason", ""))
        del W
        W, flags = make_working()
        Z, st2 = linkage_working(W, method, nan_flags=flags, return_stats=True)
        st2.update(engine="exact", fallback_from_rounds=st)
        return (Z, st2) if return_stats else Z
    Z, st = linkage_working(W, method, nan_flags=flags, return_stats=True)
    st["engine"] = "exact"
    return (Z, st) if return_stats else Z


def linkage_batch(Ws, method="ward"):
    """Independent linkages overlapped on separate streams (each is one thread-block cluster)."""
    if not Ws:
        return []
    _dev(Ws[0])
    L = _lib.lib()
    cur = torch.cuda.current_stream()
    streams = [torch.cuda.Stream() for _ in range(min(len(Ws), 8))]
    jobs = []
    for k, W in enumerate(Ws):
        n = W.shape[0]
        st = streams[k % len(streams)]
        st.wait_stream(cur)
        ws = torch.empty(L.hc_linkage_workspace_bytes(n), dtype=torch.uint8, device=W.device)
        _lib.check(L.hc_linkage_launch(_ptr(W), n, METHODS[method], _ptr(ws), ws.numel(),
                                       C.c_void_p(st.cuda_stream)))
        jobs.append((n, ws, st))
    out = []
    for n, ws, st in jobs:
        Z = np.empty((max(n - 1, 0), 4), dtype=np.float64)
        _lib.check(L.hc_linkage_finish(n, METHODS[method], _ptr(ws), ws.numel(),
                                       Z.ctypes.data_as(C.c_void_p), None, C.c_void_p(st.cuda_stream)))
        out.append(Z)
    return out


# ---------------------------------------------------------------------------------------------
# K8
# ---------------------------------------------------------------------------------------------
def segment_mean_2d(A, seg_begin, seg_end):
    _dev(A)
    n, k = A.shape[0], seg_begin.numel()
    tmp = torch.empty((k, n), dtype=torch.float32, device=A.device)
    out = torch.empty((k, k), dtype=torch.float32, device=A.device)
    _lib.check(_lib.lib().hc_segment_mean_2d(_ptr(A), n, _ptr(seg_begin), _ptr(seg_end), k, _ptr(tmp),
                                             _ptr(out), _stream()))
    return out


def block_sums4(A, jobs):
    """Corner-block sums of join_scaffolds for a batch of independent joins: ``jobs`` = list of
    (indices1, indices2) window lists into the symmetric float32 matrix ``A``.  Returns a (len(jobs), 4)
    float64 numpy array (one launch, one D2H read)."""
    _dev(A)
    n = A.shape[0]
    assert A.dtype == torch.float32 and A.shape == (n, n) and A.is_contiguous()
    if not jobs:
        return np.empty((0, 4))
    n1 = np.array([len(a) for a, b in jobs], dtype=np.int32)
    n2 = np.array([len(b) for a, b in jobs], dtype=np.int32)
    tot = n1.astype(np.int64) + n2
    off = np.concatenate([[0], np.cumsum(tot)[:-1]]).astype(np.int64)
    flat = np.fromiter((i for a, b in jobs for part in (a, b) for i in part), dtype=np.int32, count=int(tot.sum()))
    if flat.size and (flat.min() < 0 or flat.max() >= n):
        raise IndexError("window index out of range")
    dev = A.device
    up = lambda a: torch.from_numpy(a).to(dev)
    out = torch.empty((len(jobs), 4), dtype=torch.float64, device=dev)
    idx = up(flat) if flat.size else torch.zeros(1, dtype=torch.int32, device=dev)
    d_off, d_n1, d_n2 = up(off), up(n1), up(n2)  # kept alive (and distinct) until the launch is queued
    _lib.check(_lib.lib().hc_block_sums4(_ptr(A), n, _ptr(idx), _ptr(d_off), _ptr(d_n1), _ptr(d_n2), len(jobs),
                                         _ptr(out), _stream()))
    return out.cpu().numpy()


def matrix_max(X):
    """np.max(X) for a float32 / float64 CUDA tensor (hc_matrix_max); returns a Python float."""
    _dev(X)
    X = X.contiguous()
    part = torch.empty(1024, dtype=X.dtype, device=X.device)
    out = torch.empty(1, dtype=X.dtype, device=X.device)
    _lib.check(_lib.lib().hc_matrix_max(_ptr(X), _dtype_code(X), X.numel(), _ptr(part), 1024, _ptr(out), _stream()))
    return float(out.item())


def gather_elements(A, rows, cols):
    """A[rows[k], cols[k]] for int32 index arrays (numpy) -> float32 numpy array (one launch, one read-back)."""
    _dev(A)
    n = A.shape[0]
    assert A.dtype == torch.float32 and A.shape == (n, n) and A.is_contiguous()
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    m = len(rows)
    if m == 0:
        return np.empty(0, dtype=np.float32)
    if rows.min() < 0 or rows.max() >= n or cols.min() < 0 or cols.max() >= n:
        raise IndexError("window index out of range")
    d_r, d_c = torch.from_numpy(rows).to(A.device), torch.from_numpy(cols).to(A.device)
    out = torch.empty(m, dtype=torch.float32, device=A.device)
    _lib.check(_lib.lib().hc_gather_elements(_ptr(A), n, _ptr(d_r), _ptr(d_c), m, _ptr(out), _stream()))
    return out.cpu().numpy()


def unpack_pairs16(packed, out=None):
    """int32 [n, 4] CUDA tensor of hc_pair16_t -> int32 [n, 6] hc_pair_t records (csrc/bin_pairs.cu)."""
    _dev(packed)
    n = packed.shape[0]
    assert packed.dtype == torch.int32 and packed.shape[1] == 4 and packed.is_contiguous()
    if out is None:
        out = torch.empty((n, 6), dtype=torch.int32, device=packed.device)
    _lib.check(_lib.lib().hc_unpack_pairs16(_ptr(packed), n, _ptr(out), _stream()))
    return out


def gather_square(X, keep):
    """X[keep][:, keep] for an int32 index tensor (bam2clusters.py:55-57, :428)."""
    _dev(X)
    n, m = X.shape[0], int(keep.numel())
    out = torch.empty((m, m), dtype=X.dtype, device=X.device)
    _lib.check(_lib.lib().hc_gather_square(_ptr(X), _dtype_code(X), n, _ptr(keep), m, _ptr(out), _stream()))
    return out


def diag_normalize(D, mean):
    """D * ((D*n2).T*n2), n2 = mean/diag(D) (bam2clusters.py:64-65); returns a new tensor."""
    _dev(D)
    out = torch.empty_like(D)
    _lib.check(_lib.lib().hc_diag_normalize(_ptr(D), _dtype_code(D), D.shape[0], float(mean), _ptr(out), _stream()))
    return out


def window_size_normalize_(D, sizes, window_size):
    """In place D *= 1.*windowSize**2 / (sizes*sizes.T) (bam2scaffolds.py:40); sizes int64 tensor [n]."""
    _dev(D)
    assert sizes.dtype == torch.int64 and sizes.numel() == D.shape[0]
    _lib.check(_lib.lib().hc_window_size_normalize(_ptr(D), _dtype_code(D), D.shape[0], _ptr(sizes),
                                                   1. * window_size ** 2, _stream()))
    return D


def normalize_rows_(A):
    _dev(A)
    n = A.shape[0]
    ws = torch.empty(20 * n + 4096, dtype=torch.uint8, device=A.device)
    _lib.check(_lib.lib().hc_normalize_rows(_ptr(A), n, _ptr(ws), ws.numel(), _stream()))
    return A
