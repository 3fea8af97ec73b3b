"""Row-sharded build of ONE contact matrix over the GPUs of an NVLink / NVSwitch node: one process per GPU
(torch.distributed for rendezvous, barriers and the few scalar reductions), rank k owns rows [k*per, (k+1)*per).

The data-path exchanges are not collectives but peer-memory reads inside the kernels (csrc/sharded.cu): the fold reads
the transposed tiles out of the owning rank's HBM, the Sinkhorn column sums are all-reduced by a one-shot exchange
fused into the update kernel, the linkage rank transforms the row blocks of all ranks straight from their memory.
Peer memory = hc_shared_alloc (cudaMalloc + cudaIpc handle); the 64-byte handles travel through the process group.

The reference's analogue: the contigs of a BAM are split over Pool workers and the partial matrices summed in the
parent (bam2clusters.py:224-230, bam2plot.py:138-161).  Linkage of a single dendrogram does not shard ("replicas
only"): independent linkages (bootstrap runs, per-cluster trees) are dealt round-robin to the ranks.

``LocalGroup`` runs the same code for several emulated ranks inside ONE process on ONE device (a thread and a stream
per rank, plain local pointers in the peer tables): that is how the single-GPU parity tests cover 2..8 row blocks."""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np
import torch

from . import _lib

MAX_RANKS = 8


def row_block(n, rank, world, align=64):
    """Contiguous, balanced row block [r0, r1) of rank ``rank``; boundaries are multiples of ``align`` (the tile
    kernels work on 64 x 64 tiles) except the last one."""
    per = rows_per_rank(n, world, align)
    r0 = min(rank * per, n)
    return r0, min(r0 + per, n)


def rows_per_rank(n, world, align=64):
    per = -(-n // world)
    return -(-per // align) * align


def effective_world(n, world, align=64):
    """Ranks that actually own rows (a small matrix on many GPUs leaves the last ones empty)."""
    per = rows_per_rank(n, world, align)
    return -(-n // per)


def assign_round_robin(n_items, rank, world):
    """Independent units (bootstrap linkages, clusters) of this rank: items rank, rank+world, ..."""
    return list(range(rank, n_items, world))


def make_allreduce(group=None):
    import torch.distributed as dist

    def allreduce(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return t
    return allreduce


class SharedBuffer:
    """Device memory other ranks can map: cudaMalloc'ed by the library (zeroed), exported as a 64-byte IPC handle."""

    def __init__(self, nbytes, device):
        self.nbytes, self.device = int(nbytes), torch.device(device)
        _lib.check(_lib.lib().hc_set_device(self.device.index))
        p, h = C.c_void_p(), C.create_string_buffer(64)
        _lib.check(_lib.lib().hc_shared_alloc(self.nbytes, C.byref(p), h))
        self.ptr, self.handle = p.value, h.raw
        self._freed = False

    def tensor(self, dtype, shape, offset=0):
        """torch view of the buffer (``__cuda_array_interface__``; the buffer must outlive the tensor)."""
        shape = tuple(int(s) for s in shape)
        ts = {torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.uint8: "|u1", torch.int64: "<i8"}[dtype]

        class _View:
            pass
        v = _View()
        v.__cuda_array_interface__ = {"shape": shape, "typestr": ts, "data": (self.ptr + offset, False), "version": 2}
        t = torch.as_tensor(v, device=self.device)
        t._hc_owner = self
        return t

    def free(self):
        if not self._freed:
            self._freed = True
            _lib.lib().hc_shared_free(C.c_void_p(self.ptr))

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class _PeerMap:
    """pointer table of one shared buffer per rank; the peers' entries are IPC mappings that are closed with it"""

    def __init__(self, ptrs, opened):
        self.ptrs, self._opened = ptrs, opened

    def table(self):
        return (C.c_void_p * len(self.ptrs))(*self.ptrs)

    def close(self):
        for p in self._opened:
            _lib.lib().hc_shared_close(C.c_void_p(p))
        self._opened = []


class ProcessGroup:
    """one process per GPU (torchrun): torch.distributed is the control plane"""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def peer_map(self, buf: SharedBuffer):
        handles = [None] * self.world
        self.dist.all_gather_object(handles, buf.handle, group=self.group)
        ptrs, opened = [], []
        for k, h in enumerate(handles):
            if k == self.rank:
                ptrs.append(buf.ptr)
                continue
            p = C.c_void_p()
            _lib.check(_lib.lib().hc_shared_open(h, C.byref(p)))
            ptrs.append(p.value)
            opened.append(p.value)
        return _PeerMap(ptrs, opened)

    def barrier(self):
        """all ranks' work queued so far is complete (device work included)"""
        torch.cuda.current_stream().synchronize()
        self.dist.barrier(group=self.group)

    def allreduce_max_float(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return t

    def gather_rows(self, local, counts):
        """concatenate the ranks' float64 row vectors (r of the Sinkhorn fit: 8 bytes per matrix row) on every rank"""
        parts = [torch.empty(c, dtype=local.dtype, device=local.device) for c in counts]
        self.dist.all_gather(parts, local.contiguous(), group=self.group) if len(set(counts)) == 1 else \
            self._all_gather_uneven(parts, local)
        return torch.cat(parts)

    def _all_gather_uneven(self, parts, local):
        for k, p in enumerate(parts):
            if k == self.rank:
                p.copy_(local)
            self.dist.broadcast(p, src=self.dist.get_global_rank(self.group, k) if self.group is not None else k,
                                group=self.group)

    def broadcast_object(self, obj, src=0):
        box = [obj]
        self.dist.broadcast_object_list(box, src=src, group=self.group)
        return box[0]

    def broadcast_tensor(self, t, src=0):
        self.dist.broadcast(t, src=self.dist.get_global_rank(self.group, src) if self.group is not None else src,
                            group=self.group)
        return t


class LocalGroup:
    """several emulated ranks in one process on one device (tests): ``run(fn)`` calls ``fn(member)`` on one thread and
    one CUDA stream per rank; shared buffers are ordinary local allocations and the peer tables hold their pointers."""

    class Member:
        emulated = True

        def __init__(self, parent, rank):
            self.parent, self.rank, self.world = parent, rank, parent.world

        def peer_map(self, buf):
            reg = self.parent._exchange(self.rank, buf.ptr)
            return _PeerMap(list(reg), [])

        def barrier(self):
            torch.cuda.current_stream().synchronize()
            self.parent._barrier.wait()

        def allreduce_max_float(self, t):
            vals = self.parent._exchange(self.rank, float(t.item()))
            t.fill_(max(vals))
            return t

        def gather_rows(self, local, counts):
            torch.cuda.current_stream().synchronize()
            parts = self.parent._exchange(self.rank, local)
            return torch.cat([p.clone() for p in parts])

        def broadcast_object(self, obj, src=0):
            return self.parent._exchange(self.rank, obj)[src]

        def broadcast_tensor(self, t, src=0):
            torch.cuda.current_stream().synchronize()
            t.copy_(self.parent._exchange(self.rank, t)[src])
            torch.cuda.current_stream().synchronize()
            self.parent._barrier.wait()
            return t

    def __init__(self, world, device="cuda:0"):
        self.world, self.device = world, torch.device(device)
        self._barrier = threading.Barrier(world)
        self._slots = [None] * world

    def _exchange(self, rank, value):
        self._slots[rank] = value
        self._barrier.wait()
        out = list(self._slots)
        self._barrier.wait()
        return out

    def run(self, fn):
        results, errors = [None] * self.world, [None] * self.world

        def work(k):
            try:
                torch.cuda.set_device(self.device)
                with torch.cuda.stream(torch.cuda.Stream(device=self.device)):
                    results[k] = fn(LocalGroup.Member(self, k))
                    torch.cuda.current_stream().synchronize()
            except BaseException as e:      # noqa: BLE001 -- re-raised in the caller's thread
                errors[k] = e
                self._barrier.abort()
        threads = [threading.Thread(target=work, args=(k,)) for k in range(self.world)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for e in errors:
            if e is not None and not isinstance(e, threading.BrokenBarrierError):
                raise e
        for e in errors:
            if e is not None:
                raise e
        return results


class RowShard:
    """Per-rank state of a row-sharded N x N build: the rank's row block and the peer-visible buffers."""

    def __init__(self, n, group, device=None, with_counts=True):
        self.n, self.g = int(n), group
        self.rank = group.rank
        self.world = effective_world(self.n, group.world)     # ranks that own rows
        self.per = rows_per_rank(self.n, group.world)
        self.row0, self.row1 = row_block(self.n, self.rank, group.world)
        self.nrows = max(0, self.row1 - self.row0)
        self.active = self.nrows > 0
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        if self.world > MAX_RANKS:
            raise ValueError("at most %d ranks share a matrix" % MAX_RANKS)
        L = _lib.lib()
        _lib.check(L.hc_set_device(self.device.index))
        _lib.check(L.hc_sharded_preload())                    # before the collective calls below: see csrc/sharded.cu
        nb = max(self.nrows, 1) * self.n * 4
        # u32 counts block, read by peers in the fold (with_counts=False: a Sinkhorn-only job on a given matrix)
        self.counts_buf = SharedBuffer(nb if with_counts else 256, self.device)
        self.sym_buf = SharedBuffer(nb, self.device)          # float32 symmetrised block, read by the linkage rank
        self.exch_buf = SharedBuffer(L.hc_sk_exchange_bytes(self.n), self.device)
        self.counts = self.counts_buf.tensor(torch.int32, (max(self.nrows, 1), self.n))[:self.nrows] if with_counts else None
        self.sym = self.sym_buf.tensor(torch.float32, (max(self.nrows, 1), self.n))[:self.nrows]
        self.counts_map = group.peer_map(self.counts_buf)
        self.sym_map = group.peer_map(self.sym_buf)
        self.exch_map = group.peer_map(self.exch_buf)
        self.epoch = C.c_int32(0)
        # everything the Sinkhorn fit touches is allocated up front: between the first exchange kernel of a fit and
        # the last, no rank may enter a device-synchronising call (cudaMalloc / cudaFree) -- with emulated ranks on
        # one device that would wait for a kernel that is itself waiting for this rank's next launch
        f64 = dict(dtype=torch.float64, device=self.device)
        self.r, self.c = torch.empty(max(self.nrows, 1), **f64)[:self.nrows], torch.empty(self.n, **f64)
        self.fit_ws = torch.empty(L.hc_sk_fit_block_workspace_bytes(max(self.nrows, 1), self.n), dtype=torch.uint8,
                                  device=self.device)
        self.row_counts = [max(0, min(self.n, (k + 1) * self.per) - k * self.per) for k in range(group.world)]

    def close(self):
        self.g.barrier()
        for m in (self.counts_map, self.sym_map, self.exch_map):
            m.close()
        self.counts = self.sym = None
        for b in (self.counts_buf, self.sym_buf, self.exch_buf):
            b.free()

    # ---- the three exchange kernels ----------------------------------------------------------------------------
    def fold(self, add=1.0):
        """S_k = rows of d + d.T - diag(d) + add (bam2clusters.py:42-43,511); returns the GLOBAL maximum (float)."""
        from .ops import _stream
        vmax = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.g.barrier()                                     # every rank's counts block is complete
        if self.active:
            _lib.check(_lib.lib().hc_fold_block(self.counts_map.table(), self.world, self.rank, self.per, self.n,
                                                float(add), C.c_void_p(self.sym.data_ptr()), C.c_void_p(vmax.data_ptr()),
                                                _stream()))
        self.g.allreduce_max_float(vmax)
        return float(vmax.item())

    def sinkhorn_fit(self, vmax, max_iter=1000, epsilon=1e-5, iterations0=0):
        """SinkhornKnopp.fit on the row blocks (sinkhorn_knopp.py:114-163): returns an ops.SinkhornResult with r
        (this rank's rows) and c (all columns)."""
        import warnings
        from .ops import SinkhornResult, _stream
        from .sk_driver import _SUPPORT
        L = _lib.lib()
        res = SinkhornResult()
        res.r, res.c = self.r, self.c
        out = (C.c_int32 * 8)()
        if getattr(self.g, "emulated", False):
            self.g.barrier()                                 # (see __init__: no allocation while a peer may be waiting)
        if self.active:
            ws = self.fit_ws
            _lib.check(L.hc_sk_fit_block(C.c_void_p(self.sym.data_ptr()), 0, self.nrows, self.n, float(vmax), float(epsilon),
                                         int(max_iter), int(iterations0), self.exch_map.table(), self.world, self.rank,
                                         C.byref(self.epoch), C.c_void_p(ws.data_ptr()), ws.numel(),
                                         C.c_void_p(res.r.data_ptr()), C.c_void_p(res.c.data_ptr()), out, _stream()))
        out = list(out)
        if self.world < self.g.world:                        # ranks without rows learn the outcome from rank 0
            out = self.g.broadcast_object(out if self.active else None, src=0)
            self.g.broadcast_tensor(res.c, src=0)
        res.iterations, res.stopping_condition = out[0], ("max_iter" if out[1] else "epsilon")
        res.matrix_passes, res.launches = out[5], out[6]
        res.warned = False
        if out[2]:
            raise AssertionError("P must be non-negative (sinkhorn_knopp.py:110)")
        for flag in (out[3], out[4]):
            if flag:
                warnings.warn(_SUPPORT, UserWarning)
                res.warned = True
        return res

    def prepare_transform(self, r_local):
        """collective: every rank's S block and Sinkhorn row vector are final; returns the full-length r (or None)"""
        r_full = self.g.gather_rows(r_local, self.row_counts) if r_local is not None else None
        self.g.barrier()
        return r_full

    def build_working(self, kind, p0, p1, r_full, c, vmax, square, out=None):
        """linkage working matrix on THIS rank from the row blocks of all ranks, read over NVLink inside the transform
        kernel (no collective: callable again when the linkage has to be repeated on a fresh matrix)."""
        from .ops import _stream
        if out is None:
            out = torch.empty((self.n, self.n), dtype=torch.float64, device=self.device)
        flags = torch.zeros(2, dtype=torch.int32, device=self.device)
        _lib.check(_lib.lib().hc_transform_gather(
            self.sym_map.table(), self.world, self.per, 0, self.n, int(kind), float(p0), float(p1),
            C.c_void_p(r_full.data_ptr()) if r_full is not None else None,
            C.c_void_p(c.data_ptr()) if c is not None else None, float(vmax), int(bool(square)),
            C.c_void_p(out.data_ptr()), C.c_void_p(flags.data_ptr()), _stream()))
        return out, flags


def record_slice_for_rows(window_of_record, n_records, row0, row1):
    """[a, b): the records of a coordinate-sorted stream whose read-1 window lies in [row0, row1).
    ``window_of_record(i)`` -> window of record i (monotone for sorted input); binary search, O(log n) probes."""
    def first_at_least(w):
        lo, hi = 0, n_records
        while lo < hi:
            mid = (lo + hi) // 2
            if window_of_record(mid) >= w:
                hi = mid
            else:
                lo = mid + 1
        return lo
    return first_at_least(row0), first_at_least(row1)
