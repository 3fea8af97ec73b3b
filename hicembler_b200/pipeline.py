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
    total, ok = None, True
    for (a, b, c_lo, c_hi, r0, r1), ev in zip(plan, events):
        cur.wait_event(ev)
        if width == 4:
            ops.unpack_pairs16(wire[a:b], out=rec[a:b])
        if not ok or r1 <= r0:
            continue
        # guard: every record of the piece belongs to its contig range (the input really is sorted by contig)
        lo_, hi_ = torch.aminmax(rec[a:b, 0].to(torch.int64) & 0xffffffff)
        if int(lo_) < c_lo or (c_hi != _NO_REF and int(hi_) > c_hi):
            ok = False
            continue
        _, st = ops.bin_pairs(rec[a:b], offsets, window_bp, n_windows, mapq, rows=(r0, r1), mode=2,
                              out=counts[r0:r1], workspace=kw.get("workspace"), return_stats=True)
        if total is None:
            total = dict(st)
        else:
            for k_ in ("lines", "binned", "skipped_absent", "out_of_range", "key_errors", "overflow"):
                total[k_] += st[k_]
            total["fast_path"] = total["fast_path"] and st["fast_path"]
    if not ok or total is None:               # not sorted by contig after all: bin in one piece
        del counts
        return pairs_to_dendrogram(rec, offsets, window_bp, n_windows, **kw)
    total["chunks"] = len(plan)
    return pairs_to_dendrogram(rec, offsets, window_bp, n_windows, counts_stats=(counts, total), **kw)
