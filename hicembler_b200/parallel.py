"""Multi-GPU plumbing: one process per GPU (torch.distributed, NCCL over NVLink / NVSwitch).

Only the Sinkhorn column-sum vector has a real exchange step on this path: the matrix
is sharded by row block, every rank runs the column pass on its block and the N float64 partial
column sums are all-reduced (0.4 MB at N = 50k -- latency bound).  Linkage of a single dendrogram is
"replicas only"; independent linkages (bootstrap runs, per-cluster trees) go to different GPUs."""
from __future__ import annotations


def row_block(n, rank, world, align=8):
    """Contiguous, balanced row block [r0, r1) of rank ``rank``; boundaries are multiples of
    ``align`` (the row-pass kernel handles 8 rows per CTA) except the last one."""
    per = -(-n // world)
    per = -(-per // align) * align
    r0 = min(rank * per, n)
    return r0, min(r0 + per, n)


def make_allreduce(group=None):
    import torch.distributed as dist

    def allreduce(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return t
    return allreduce


def assign_round_robin(n_items, rank, world):
    """Independent units (bootstrap linkages, clusters) of this rank: items rank, rank+world, ..."""
    return list(range(rank, n_items, world))
