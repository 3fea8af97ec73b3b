"""Device-resident hot path: pair records -> windowed contact matrix -> (Sinkhorn-Knopp) ->
distances -> linkage array, the sequence bam2clusters() / contigs2scaffold() run in the reference
(bam2clusters.py:493-522, bam2scaffolds.py:243-266) with all N x N data staying in HBM."""
from __future__ import annotations

import torch

from . import ops


class StageTimer:
    """CUDA-event timer per stage (events on the current stream)."""

    def __init__(self, enabled=True):
        self.enabled, self.marks = enabled, []

    def mark(self, name):
        if self.enabled:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.marks.append((name, e))

    def result_ms(self):
        if not self.marks:
            return {}
        self.marks[-1][1].synchronize()
        out = {}
        for (n0, e0), (n1, e1) in zip(self.marks[:-1], self.marks[1:]):
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        return out


def pairs_to_dendrogram(records, offsets, window_bp, n_windows, mapq=10, params=(0.5, 1.1), method="ward",
                        balance=False, max_iter=1000, epsilon=0.00001, keep=False, timer=None,
                        member=None, region_mask=None, workspace=None, engine="auto", counts_stats=None):
    """records: int32 [P,6] CUDA tensor; offsets: int32 CUDA table refID -> first window.

    balance=False is the reference at HEAD (normalize() commented out, bam2clusters.py:512);
    balance=True applies normalize() (:36-49) before distance_func(x+1, a, b) (:516).
    engine: "auto" (round-parallel linkage, exact serial NN-chain when a tie is met), "exact", "rounds".
    Returns dict(Z, stats, [r, c, sk_iterations, sk_stop, working, sym])."""
    t = timer or StageTimer(False)
    t.mark("start")
    if counts_stats is not None:     # binned already (host_pairs_to_dendrogram bins chunk by chunk under the copy)
        counts, st = counts_stats
    else:
        counts, st = ops.bin_pairs(records, offsets, window_bp, n_windows, mapq, member=member,
                                   region_mask=region_mask, workspace=workspace, return_stats=True)
    t.mark("bin_pairs")
    out = {"stats": st}
    a, b = params
    if balance:
        S1, vmax = ops.fold_counts(counts, "sym", add=1.0, inplace=True)
        t.mark("fold")
        vm = float(vmax.item())
        res = ops.sinkhorn_fit(S1, vm, max_iter, epsilon, symmetric=True)   # S1 = d + d.T
        t.mark("sinkhorn")
        out.update(r=res.r, c=res.c, sk_iterations=res.iterations, sk_stop=res.stopping_condition, vmax=vm,
                   sk_passes=res.matrix_passes, sk_launches=res.launches)
        def make_working():
            return ops.transform_to_working(S1, ops.KIND_POWER, a, b, res.r, res.c, vm, square=(method == "ward"))
        sym = S1
    else:
        S, vmax = ops.fold_counts(counts, "sym", inplace=True)
        t.mark("fold")
        def make_working():
            return ops.transform_to_working(S, ops.KIND_POWER, a, b, square=(method == "ward"))
        sym = S
    if keep:
        out["working"] = make_working()[0]
        out["sym"] = sym
    built = []

    def timed_make_working():      # the first build is the "transform" stage; a rebuild after an uncertified
        r_ = make_working()        # round-parallel run is charged to "linkage"
        if not built:
            t.mark("transform")
        built.append(1)
        return r_
    Z, lst = ops.linkage_auto(timed_make_working, method, engine=engine, return_stats=True)
    t.mark("linkage")
    out["Z"] = Z
    out["linkage_stats"] = lst
    return out


_NO_REF = 1 << 30   # RNAME "*" (refID -1; 0xFFFFF in the wire form) sorts last in a coordinate-sorted BAM


def _chunk_plan(host_records, offsets_cpu, n_windows, n_chunks):
    """Cut coordinate-sorted host records at contig boundaries into ~n_chunks pieces and give each piece the row
    range its read-1 windows can fall into: [first window of its first contig, first window after its last
    contig + 1) -- the +1 covers a midpoint that spills one window past the end of a contig.
    None when the window table is not in contig order (then the caller bins in one piece)."""
    import numpy as np
    P, width = host_records.shape
    col = host_records[:, 0] if width == 6 else host_records[:, 2]

    def ref_at(i):
        v = int(col[i])
        if width == 4:
            v &= 0xfffff
            return _NO_REF if v == 0xfffff else v
        return _NO_REF if v < 0 else v
    off = np.asarray(offsets_cpu, dtype=np.int64)
    pres = off >= 0
    if pres.sum() == 0 or not np.all(np.diff(off[pres]) > 0):
        return None
    nxt = np.full(len(off) + 1, n_windows, dtype=np.int64)      # first window of the first present contig >= c
    nxt[:-1] = np.minimum.accumulate(np.where(pres, off, n_windows)[::-1])[::-1]
    bounds = [0]
    for k in range(1, n_chunks):
        t = k * P // n_chunks
        if t <= bounds[-1]:
            continue
        r = ref_at(t - 1)
        lo, hi = t, P
        while lo < hi:                                          # first record of the next contig
            mid = (lo + hi) // 2
            if ref_at(mid) > r:
                hi = mid
            else:
                lo = mid + 1
        if bounds[-1] < lo < P:
            bounds.append(lo)
    bounds.append(P)
    plan = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        c_lo, c_hi = ref_at(a), ref_at(b - 1)
        if c_hi < c_lo or (c_lo != _NO_REF and c_lo >= len(off)):
            return None
        if c_lo == _NO_REF:
            r0 = r1 = 0                                         # only unmapped records: nothing to bin
        else:
            r0 = int(nxt[c_lo])
            r1 = n_windows if c_hi >= len(off) - 1 else min(n_windows, int(nxt[c_hi + 1]) + 1)
        plan.append((a, b, c_lo, c_hi, r0, r1))
    return plan


def host_pairs_to_dendrogram(host_records, offsets, window_bp, n_windows, staging=None, **kw):
    """The same path from HOST records (what a BAM decoder hands over): ``host_records`` is a (pinned) CPU int32
    tensor, [P, 6] = hc_pair_t or [P, 4] = the 16-byte wire form hc_pair16_t (pairs.pack_pairs16), which is
    expanded on the device (hc_unpack_pairs16).  ``staging``: dict reused between calls for the device
    buffers.  Everything after the copy is pairs_to_dendrogram."""
    dev = offsets.device
    staging = {} if staging is None else staging
    P, width = host_records.shape
    if width not in (4, 6):
        raise ValueError("host records must be int32 [P, 6] (hc_pair_t) or [P, 4] (hc_pair16_t)")

    def buf(name, w):
        t = staging.get(name)
        if t is None or t.shape[0] < P:
            t = staging[name] = torch.empty((P, w), dtype=torch.int32, device=dev)
        return t[:P]
    chunks = staging.get("chunks", 8)
    plan = None
    if chunks > 1 and P >= staging.get("min_chunk_records", 1 << 22) and kw.get("member") is None and kw.get("region_mask") is None:
        key = ("offsets_cpu", offsets.data_ptr(), offsets.numel())
        if staging.get("offsets_key") != key:
            staging["offsets_key"], staging["offsets_cpu"] = key, offsets.cpu().numpy()
        plan = _chunk_plan(host_records, staging["offsets_cpu"], n_windows, chunks)
    rec = buf("rec24", 6)
    wire = buf("rec16", 4) if width == 4 else None
    if plan is None or len(plan) < 2:
        if width == 6:
            rec.copy_(host_records, non_blocking=True)
        else:
            wire.copy_(host_records, non_blocking=True)
            ops.unpack_pairs16(wire, out=rec)
        return pairs_to_dendrogram(rec, offsets, window_bp, n_windows, **kw)
    # Chunked: all copies are queued on a copy stream; while chunk k+1 crosses PCIe, chunk k is expanded and
    # binned into its row range (accumulate mode on a zeroed matrix: a boundary row may get records of two chunks).
    cur = torch.cuda.current_stream()
    cs = staging.get("copy_stream")
    if cs is None:
        cs = staging["copy_stream"] = torch.cuda.Stream(device=dev)
    cs.wait_stream(cur)                       # the staging buffers may still be read by the previous call
    events = []
    dst = rec if width == 6 else wire
    with torch.cuda.stream(cs):
        for a, b, *_ in plan:
            dst[a:b].copy_(host_records[a:b], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
            events.append(ev)
    counts = torch.zeros((n_windows, n_windows), dtype=torch.int32, device=dev)
    mapq = kw.get("mapq", 10)
    pending = []
    for (a, b, c_lo, c_hi, r0, r1), ev in zip(plan, events):
        cur.wait_event(ev)
        if width == 4:
            ops.unpack_pairs16(wire[a:b], out=rec[a:b])
        if r1 <= r0:
            continue
        # asynchronous calls: no host synchronisation between the pieces.  A record of the piece that can be binned
        # but whose row lies outside the piece's row range is counted by the kernel (outside_rows): the guard that
        # the input really is sorted by contig
        _, later = ops.bin_pairs(rec[a:b], offsets, window_bp, n_windows, mapq, rows=(r0, r1), mode=2,
                                 out=counts[r0:r1], workspace=kw.get("workspace"), defer_stats=True)
        pending.append(later.snapshot())
    total = None
    for later in pending:
        st = later.read()
        if total is None:
            total = dict(st)
        else:
            for k_ in ("lines", "binned", "skipped_absent", "out_of_range", "key_errors", "overflow", "outside_rows"):
                total[k_] += st[k_]
            total["fast_path"] = total["fast_path"] and st["fast_path"]
    if total is None or total["outside_rows"]:    # not sorted by contig after all: bin in one piece
        del counts
        return pairs_to_dendrogram(rec, offsets, window_bp, n_windows, **kw)
    total["chunks"] = len(plan)
    return pairs_to_dendrogram(rec, offsets, window_bp, n_windows, counts_stats=(counts, total), **kw)


# ------------------------------------------------------------------------------------------------------------
# one job over the GPUs of a node: row-sharded matrix build + Sinkhorn, linkage on one rank
# ------------------------------------------------------------------------------------------------------------
def window_of_records(rec_np, offsets_np, window_bp):
    """read-1 window of host records (numpy int32 [k, 6] hc_pair_t or [k, 4] hc_pair16_t rows), -1 when the contig is
    absent: the key the coordinate-sorted stream is cut by (get_window, bam2clusters.py:119-126)."""
    import numpy as np
    rec_np = np.atleast_2d(rec_np)
    if rec_np.shape[1] == 6:
        ref = rec_np[:, 0].astype(np.int64)
        lseq = np.maximum(rec_np[:, 4].astype(np.int64), 1)
        p1 = rec_np[:, 1].astype(np.int64) + 1 + lseq // 2
    else:
        v = rec_np[:, 2].astype(np.int64) & 0xfffff
        ref = np.where(v == 0xfffff, -1, v)
        p1 = (rec_np[:, 0].astype(np.int64) & 0xffffffff) + 1
    ok = (ref >= 0) & (ref < len(offsets_np))
    off = np.where(ok, np.asarray(offsets_np, dtype=np.int64)[np.clip(ref, 0, len(offsets_np) - 1)], -1)
    return np.where(off >= 0, p1 // int(window_bp) + off, -1)


def shard_plan(records, offsets_np, window_bp, n_windows, world):
    """Record slice [a, b) of every rank's row block -- the job the BAM index (.bai) does for ``samtools view bam
    <regions>`` (bam2clusters.py:108-117): a property of the input, computed once per record buffer.  ``records``:
    host or device int32 tensor [P, 4|6] sorted by (contig, position).  Unmapped / absent contigs sort by their
    neighbours (a monotone fill), and the result is verified by the binning kernels themselves (outside_rows)."""
    import numpy as np
    from .parallel import row_block
    P = records.shape[0]
    step = max(1, P // 65536)
    idx = torch.arange(0, P, step, device=records.device)
    samp = records[idx].cpu().numpy()
    w = window_of_records(samp, offsets_np, window_bp)
    w = np.maximum.accumulate(np.where(w < 0, 0, w))          # monotone fill
    idx = idx.cpu().numpy()
    plan = []
    for k in range(world):
        r0, r1 = row_block(n_windows, k, world)
        cut = []
        for row in (r0, r1):
            if row >= n_windows and row == r1:
                cut.append(P)
                continue
            s = int(np.searchsorted(w, row, side="left"))      # first sample with window >= row
            lo = int(idx[s - 1]) if s > 0 else 0
            hi = int(idx[s]) if s < len(idx) else P
            if hi > lo:
                blk = window_of_records(records[lo:hi].cpu().numpy(), offsets_np, window_bp)
                blk = np.maximum.accumulate(np.where(blk < 0, 0, blk))
                lo += int(np.searchsorted(blk, row, side="left"))
            cut.append(lo)
        plan.append((cut[0], max(cut[0], cut[1])))
    plan[0] = (0, plan[0][1])
    plan[-1] = (plan[-1][0], P)
    for k in range(1, world):                                  # slices partition the stream
        plan[k] = (plan[k - 1][1], max(plan[k - 1][1], plan[k][1]))
    return plan


def sharded_pairs_to_dendrogram(shard, records, offsets, window_bp, mapq=10, params=(0.5, 1.1), method="ward",
                                balance=True, max_iter=1000, epsilon=0.00001, plan=None, timer=None, engine="auto",
                                staging=None, root=0, workspace=None, local_slice=False):
    """pairs -> Z as ONE job over the ranks of ``shard`` (a parallel.RowShard): every rank bins the slice of the
    record stream that falls into its row block (host records: only that slice crosses the rank's own PCIe link),
    the fold / Sinkhorn / transform exchanges read peer memory over NVLink, the dendrogram is built on ``root``.
    Same result as pairs_to_dendrogram on one GPU: counts bit-exact, r / c to rounding (the column sums are added in a
    different order), identical merge order.  Returns the pairs_to_dendrogram dict on ``root``, {"stats": ...}
    elsewhere.  ``records``: the FULL stream, int32 [P, 6] on the device or [P, 4|6] (pinned) on the host -- or, with
    ``local_slice=True`` and a ``plan``, only this rank's slice of it (what a rank keeps in its own pinned buffer)."""
    import numpy as np
    n = shard.n
    g = shard.g
    t = timer or StageTimer(False)
    t.mark("start")
    staging = {} if staging is None else staging
    if plan is None:
        if local_slice:
            raise ValueError("local_slice=True needs the plan the slices were cut by")
        key = ("plan", records.data_ptr(), records.shape[0], g.world)
        if staging.get("plan_key") != key:
            staging["plan_key"] = key
            staging["plan"] = shard_plan(records, offsets.cpu().numpy(), window_bp, n, g.world)
        plan = staging["plan"]
    a, b = plan[shard.rank]
    if local_slice:                                            # ``records`` already is this rank's slice [a, b)
        lo, hi = 0, b - a
    else:
        lo, hi = a, b
    dev = shard.device
    if records.is_cuda:
        rec = records[lo:hi]
    else:                                                      # this rank's slice over this rank's PCIe link
        width = records.shape[1]
        need = max(b - a, 1)

        def buf(name, wdt):
            x = staging.get(name)
            if x is None or x.shape[0] < need:
                x = staging[name] = torch.empty((need, wdt), dtype=torch.int32, device=dev)
            return x[:b - a]
        rec = buf("rec24", 6)
        if width == 6:
            rec.copy_(records[lo:hi], non_blocking=True)
        else:
            wire = buf("rec16", 4)
            wire.copy_(records[lo:hi], non_blocking=True)
            ops.unpack_pairs16(wire, out=rec)
    t.mark("h2d")
    st = None
    if shard.active:
        _, later = ops.bin_pairs(rec, offsets, window_bp, n, mapq, rows=(shard.row0, shard.row1), out=shard.counts,
                                 workspace=workspace, defer_stats=True)
        st = later.read()
    t.mark("bin_pairs")
    # certificate of the slicing: no binnable record of any slice belongs to another rank's rows
    flag = torch.tensor([float(st["outside_rows"]) if st else 0.0], dtype=torch.float32, device=dev)
    if float(g.allreduce_max_float(flag).item()) > 0:
        raise ValueError("records are not sorted by read-1 window: a row-sharded build needs coordinate-sorted input "
                         "(or bin every rank's row block from the full stream with ops.bin_pairs(rows=...))")
    out = {"stats": st, "slice": (a, b)}
    vmax = shard.fold(add=1.0 if balance else 0.0)
    t.mark("fold")
    a_, b_ = params
    r = c = None
    if balance:
        res = shard.sinkhorn_fit(vmax, max_iter, epsilon)
        r, c = res.r, res.c
        out.update(r=r, c=c, sk_iterations=res.iterations, sk_stop=res.stopping_condition, vmax=vmax,
                   sk_passes=res.matrix_passes, sk_launches=res.launches)
    t.mark("sinkhorn")
    r_full = shard.prepare_transform(r)
    if shard.rank == root:
        made = []

        def make_working():                                    # local: repeatable after an uncertified rounds run
            Wf = shard.build_working(ops.KIND_POWER, a_, b_, r_full, c, vmax if balance else 1.0,
                                     square=(method == "ward"))
            if not made:
                t.mark("transform")
            made.append(1)
            return Wf
        Z, lst = ops.linkage_auto(make_working, method, engine=engine, return_stats=True)
        out["Z"], out["linkage_stats"] = Z, lst
    t.mark("linkage")
    g.barrier()                                                # peers' blocks stay alive until the root is done
    return out
