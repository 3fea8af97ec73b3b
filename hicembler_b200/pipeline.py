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
                        member=None, region_mask=None, workspace=None, engine="auto"):
    """records: int32 [P,6] CUDA tensor; offsets: int32 CUDA table refID -> first window.

    balance=False is the reference at HEAD (normalize() commented out, bam2clusters.py:512);
    balance=True applies normalize() (:36-49) before distance_func(x+1, a, b) (:516).
    engine: "auto" (round-parallel linkage, exact serial NN-chain when a tie is met), "exact", "rounds".
    Returns dict(Z, stats, [r, c, sk_iterations, sk_stop, working, sym])."""
    t = timer or StageTimer(False)
    t.mark("start")
    counts, st = ops.bin_pairs(records, offsets, window_bp, n_windows, mapq, member=member,
                               region_mask=region_mask, workspace=workspace, return_stats=True)
    t.mark("bin_pairs")
    out = {"stats": st}
    a, b = params
    if balance:
        S1, vmax = ops.fold_counts(counts, "sym", add=1.0, inplace=True)
        t.mark("fold")
        vm = float(vmax.item())
        res = ops.sinkhorn_fit(S1, vm, max_iter, epsilon)
        t.mark("sinkhorn")
        out.update(r=res.r, c=res.c, sk_iterations=res.iterations, sk_stop=res.stopping_condition, vmax=vm)
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
    if width == 6:
        rec = buf("rec24", 6)
        rec.copy_(host_records, non_blocking=True)
    else:
        wire = buf("rec16", 4)
        wire.copy_(host_records, non_blocking=True)
        rec = ops.unpack_pairs16(wire, out=buf("rec24", 6))
    return pairs_to_dendrogram(rec, offsets, window_bp, n_windows, **kw)
