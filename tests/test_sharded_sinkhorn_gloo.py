"""N > 1 host logic on CPU: world-size-2 `gloo` run of the row-sharded Sinkhorn driver.

The control flow in hicembler_b200/sk_driver.py is what every rank executes around the CUDA passes;
here the per-rank passes are a numpy stand-in (test infrastructure) and the all-reduce of the column
sums goes over gloo.  Both ranks must stop at the same iteration with the vectors the unsharded
oracle (vendored SinkhornKnopp restatement, sinkhorn_knopp.py:93-163) produces."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from util import hic_like_counts

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class NumpyPasses:
    """Row block of P with the same interface as ops._CudaSinkhornPasses."""

    def __init__(self, P_block, eps):
        self.P = np.asarray(P_block, dtype=np.float64)
        self.eps = eps
        self.nrows, self.n = self.P.shape
        self.r = np.ones(self.nrows)
        self.c = np.ones(self.n)
        self.y = torch.zeros(self.n, dtype=torch.float64)
        self.x = np.zeros(self.nrows)
        self.flags = dict(negative=bool((self.P < 0).any()), zero_col=False, zero_row=False)

    def colpass(self, use_r, check=False):
        r = self.r if use_r else np.ones(self.nrows)
        self.y.copy_(torch.from_numpy(self.P.T @ r))
        return self.y

    def update_c(self, test):
        y = self.y.numpy()
        if test:
            s = self.c * y
            if not ((s < 1 - self.eps) | (s > 1 + self.eps)).any():
                return False
        self.flags["zero_col"] |= bool((y == 0).any())
        self.c = 1.0 / y
        return True

    def raw_outside(self):
        rs, cs = self.P.sum(1), self.y.numpy()
        out = ((rs < 1 - self.eps) | (rs > 1 + self.eps)).any() or ((cs < 1 - self.eps) | (cs > 1 + self.eps)).any()
        return torch.tensor([int(out)], dtype=torch.int32)

    def rowpass(self):
        self.x = self.P @ self.c
        self.flags["zero_row"] |= bool((self.x == 0).any())
        self.r = 1.0 / self.x

    def read_flags(self):
        return dict(self.flags)

    def flag_value(self, t):
        return bool(t.item())


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, P, eps, max_iter, out_dir):
    for p in (ROOT, os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from hicembler_b200.parallel import make_allreduce, row_block
    from hicembler_b200.sk_driver import sinkhorn_driver
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r0, r1 = row_block(P.shape[0], rank, world, align=8)
        passes = NumpyPasses(P[r0:r1], eps)
        try:
            it, stop, warned = sinkhorn_driver(passes, max_iter, 0, make_allreduce())
        except AssertionError:
            it, stop = -1, "assert"
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), r=passes.r, c=passes.c, it=it, stop=stop, r0=r0, r1=r1)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,eps,max_iter", [(96, 1e-5, 1000), (61, 1e-8, 1000), (40, 1e-12, 3)])
def test_row_sharded_driver_world2_matches_unsharded_oracle(tmp_path, n, eps, max_iter):
    import hicembler_oracle as O
    d = hic_like_counts(n, seed=n)
    d = d + d.T - np.diag(d.diagonal()) + 1
    P = (d / d.max()).astype(np.float64)
    sk = O.SinkhornKnopp(max_iter=max_iter, epsilon=eps)
    want = sk.fit(P.copy())
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), P, eps, max_iter, str(tmp_path)), nprocs=world, join=True)
    res = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % k)) for k in range(world)]
    assert int(res[0]["it"]) == int(res[1]["it"]) == sk._iterations
    assert str(res[0]["stop"]) == str(res[1]["stop"]) == sk._stopping_condition
    assert np.array_equal(res[0]["c"], res[1]["c"])            # every rank holds the same column vector
    r = np.concatenate([res[0]["r"], res[1]["r"]])
    assert int(res[0]["r1"]) == int(res[1]["r0"]) and int(res[1]["r1"]) == n
    got = (r[:, None] * P) * res[0]["c"][None, :]
    assert np.max(np.abs(got / want - 1)) <= 1e-12


def test_round_robin_assignment_covers_all_units():
    from hicembler_b200.parallel import assign_round_robin, row_block
    for world in (1, 2, 4, 8):
        units = sorted(u for r in range(world) for u in assign_round_robin(30, r, world))
        assert units == list(range(30))
        rows = [row_block(50_000, r, world) for r in range(world)]
        assert rows[0][0] == 0 and rows[-1][1] == 50_000
        assert all(a[1] == b[0] for a, b in zip(rows[:-1], rows[1:]))


def test_negative_entry_on_one_rank_raises_on_every_rank(tmp_path):
    """the AssertionError of sinkhorn_knopp.py:110 must reach all ranks (no rank may go on to the next
    all-reduce alone)"""
    n = 64
    d = hic_like_counts(n, seed=3)
    d = d + d.T - np.diag(d.diagonal()) + 1
    P = (d / d.max()).astype(np.float64)
    P[50, 3] = -0.5                                  # second rank's block only
    mp.spawn(_worker, args=(2, _free_port(), P, 1e-5, 100, str(tmp_path)), nprocs=2, join=True)
    res = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % k)) for k in range(2)]
    assert str(res[0]["stop"]) == str(res[1]["stop"]) == "assert"
