"""Row-sharded build of one matrix (csrc/sharded.cu, hicembler_b200/parallel.py, pipeline.sharded_pairs_to_dendrogram).

On a one-GPU box the ranks are EMULATED: parallel.LocalGroup runs the very same per-rank code on one thread and one
stream per rank, with local pointers in the peer tables -- the exchange kernels (peer reads of the fold, the one-shot
all-reduce inside the Sinkhorn update, the gather inside the transform) run exactly as they do over NVLink.  With two
or more devices test_two_process_nccl_job additionally runs the real thing under torchrun (cudaIpc peer mappings)."""
import os
import subprocess
import sys
import warnings

import numpy as np
import pytest

import hicembler_oracle as O
from util import hic_like_counts, small_case

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _full_counts(n, seed):
    """full-orientation u32 count matrix (not symmetric) with a Hi-C like band"""
    rng = np.random.default_rng(seed)
    i, j = np.indices((n, n))
    return rng.poisson(30.0 / (1.0 + np.abs(i - j)) ** 1.08 + 0.05).astype(np.int32)


@pytest.mark.parametrize("n,world", [(256, 2), (1000, 3), (1472, 8), (200, 8)])
def test_fold_block_equals_single_device_fold(n, world):
    """d += d.T - diag(d) + 1 assembled from the ranks' blocks (transposed tiles read from the owner's buffer) ==
    hc_fold_counts on the whole matrix, bit for bit; the maximum too"""
    from hicembler_b200 import ops
    from hicembler_b200.parallel import LocalGroup, RowShard
    C = _full_counts(n, n)
    want, vm = ops.fold_counts(torch.from_numpy(C).cuda(), "sym", add=1.0)
    want, vm = want.cpu().numpy(), float(vm.item())

    def rank_fn(member):
        sh = RowShard(n, member)
        if sh.active:
            sh.counts.copy_(torch.from_numpy(C[sh.row0:sh.row1]).cuda())
        v = sh.fold(add=1.0)
        out = sh.sym.cpu().numpy().copy()
        sh.close()
        return sh.row0, out, v
    for row0, blk, v in LocalGroup(world).run(rank_fn):
        assert v == vm
        assert np.array_equal(blk, want[row0:row0 + blk.shape[0]])


@pytest.mark.parametrize("n,world,dt", [(300, 2, np.float32), (1024, 4, np.float32), (777, 8, np.float32)])
def test_sharded_sinkhorn_matches_vendored_library_semantics(n, world, dt):
    """r, c, iterations and stopping condition of the row-sharded fit (fused exchange, loop condition on the device)
    == O.SinkhornKnopp (pinned against the vendored library) on normalize()'s input"""
    from hicembler_b200.parallel import LocalGroup, RowShard
    d = hic_like_counts(n, 3, dtype=dt)
    d += d.T - np.diag(d.diagonal())
    d += 1
    vmax = float(d.max())
    sk = O.SinkhornKnopp(max_iter=1000, epsilon=1e-5)
    want = sk.fit(d / d.max())
    r_want, c_want = sk._r, sk._c

    def rank_fn(member):
        sh = RowShard(n, member)
        if sh.active:
            sh.sym.copy_(torch.from_numpy(d[sh.row0:sh.row1]).cuda())
        member.barrier()
        res = sh.sinkhorn_fit(vmax, 1000, 1e-5)
        out = (sh.row0, res.r.cpu().numpy().copy(), res.c.cpu().numpy().copy(), res.iterations, res.stopping_condition,
               res.matrix_passes)
        sh.close()
        return out
    results = LocalGroup(world).run(rank_fn)
    for row0, r, c, it, stop, passes in results:
        assert it == sk._iterations and stop == sk._stopping_condition
        assert np.allclose(c, c_want, rtol=1e-12, atol=0)
        assert np.allclose(r, r_want[row0:row0 + len(r)], rtol=1e-12, atol=0)
        assert np.array_equal(c, results[0][2])                 # every rank holds the same bits
        assert passes == 2 * it + 3
    bal = (np.concatenate([x[1] for x in results])[:, None] * (d / d.max()).astype(np.float64)) * results[0][2][None, :]
    assert np.max(np.abs(bal / want - 1)) <= 1e-12


def test_sharded_sinkhorn_max_iter_and_already_balanced():
    from hicembler_b200.parallel import LocalGroup, RowShard
    P = np.array([[.011, .15], [1.71, .1]])                     # sinkhorn_knopp.py:61-69 / tests:33-39
    n = 128
    big = np.kron(np.eye(n // 2), P).astype(np.float32) + np.float32(1e-3)
    for max_iter, eps in ((5, 1e-3), (1000, 1e-3)):
        sk = O.SinkhornKnopp(max_iter=max_iter, epsilon=eps)
        sk.fit(big.astype(np.float64))

        def rank_fn(member):
            sh = RowShard(n, member)
            sh.sym.copy_(torch.from_numpy(big[sh.row0:sh.row1]).cuda())
            member.barrier()
            res = sh.sinkhorn_fit(1.0, max_iter, eps)
            out = (res.iterations, res.stopping_condition, res.c.cpu().numpy().copy())
            sh.close()
            return out
        for it, stop, c in LocalGroup(2).run(rank_fn):
            assert (it, stop) == (sk._iterations, sk._stopping_condition)
            assert np.allclose(c, sk._c, rtol=1e-10)
    # doubly stochastic input: the first `while` test fails at once -> 0 iterations, 'epsilon'
    ds = np.full((128, 128), 1.0 / 128, dtype=np.float32)

    def rank_fn2(member):
        sh = RowShard(128, member)
        sh.sym.copy_(torch.from_numpy(ds[sh.row0:sh.row1]).cuda())
        member.barrier()
        res = sh.sinkhorn_fit(1.0, 1000, 1e-3)
        sh.close()
        return res.iterations, res.stopping_condition
    assert LocalGroup(2).run(rank_fn2) == [(0, "epsilon")] * 2


def test_sharded_sinkhorn_flags_reach_every_rank():
    """a negative entry (AssertionError) or an empty row (UserWarning) in ONE rank's block must make every
    rank raise / warn -- the flags travel with the all-reduced vector"""
    from hicembler_b200.parallel import LocalGroup, RowShard
    n = 192
    d = hic_like_counts(n, 5) + 1
    d += d.T
    neg = d.copy()
    neg[150, 7] = -1.0                                          # rank 2's block only

    def run(mat):
        def rank_fn(member):
            sh = RowShard(n, member)
            sh.sym.copy_(torch.from_numpy(mat[sh.row0:sh.row1]).cuda())
            member.barrier()
            try:
                res = sh.sinkhorn_fit(float(mat.max()), 50, 1e-5)
                return "ok", int(res.warned)
            except AssertionError:
                return "assert", 0
            finally:
                sh.close()
        # warnings.catch_warnings swaps process-global state: it is entered ONCE here, around the rank threads
        with warnings.catch_warnings(record=True) as w:
            warnings.simplefilter("always")
            out = LocalGroup(3).run(rank_fn)
        return out, [x for x in w if issubclass(x.category, UserWarning)]
    out, w = run(neg)
    assert out == [("assert", 0)] * 3
    zero = d.copy()
    zero[10, :] = 0
    zero[:, 10] = 0
    out, w = run(zero)
    assert out == [("ok", 1)] * 3 and len(w) == 3


@pytest.mark.parametrize("n,world,balance", [(512, 2, True), (1000, 4, True), (640, 8, False)])
def test_transform_gather_equals_single_device_transform(n, world, balance):
    from hicembler_b200 import ops
    from hicembler_b200.parallel import LocalGroup, RowShard
    d = hic_like_counts(n, 9)
    d += d.T - np.diag(d.diagonal())
    d += 1
    rng = np.random.default_rng(1)
    r, c = (rng.random(n) + 0.5, rng.random(n) + 0.5) if balance else (None, None)
    X = torch.from_numpy(d).cuda()
    want, _ = ops.transform_to_working(X, ops.KIND_POWER, 0.5, 1.1, None if r is None else torch.from_numpy(r).cuda(),
                                       None if c is None else torch.from_numpy(c).cuda(), float(d.max()), square=True)
    want = want.cpu().numpy()

    def rank_fn(member):
        sh = RowShard(n, member)
        if sh.active:
            sh.sym.copy_(torch.from_numpy(d[sh.row0:sh.row1]).cuda())
        rl = None if r is None else torch.from_numpy(r[sh.row0:sh.row1]).cuda()
        r_full = sh.prepare_transform(rl)
        out = None
        if member.rank == 0:
            W, flags = sh.build_working(ops.KIND_POWER, 0.5, 1.1, r_full, None if c is None else torch.from_numpy(c).cuda(),
                                        float(d.max()), True)
            out = W.cpu().numpy()
        member.barrier()
        sh.close()
        return out
    got = LocalGroup(world).run(rank_fn)[0]
    assert np.array_equal(got, want)


@pytest.mark.parametrize("world,host,balance", [(2, False, True), (4, True, True), (3, True, False), (8, False, True)])
def test_sharded_job_equals_single_gpu_job(world, host, balance):
    """pairs -> Z as one row-sharded job == the single-GPU pipeline: counts bit-exact, same Sinkhorn iterations,
    identical merge order, heights to rounding"""
    from hicembler_b200 import ops, synth
    from hicembler_b200.pairs import pack_pairs16
    from hicembler_b200.parallel import LocalGroup, RowShard
    from hicembler_b200.pipeline import pairs_to_dendrogram, sharded_pairs_to_dendrogram
    n = 1200
    g = synth.make_genome(60, n, 5000, 5, seed=4, min_tail=2000)
    rec_t = synth.make_pairs(g, 400000, seed=4, device="cuda")
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    single = pairs_to_dendrogram(rec_t, off, g.window_bp, n, balance=balance, keep=True, method="ward")
    records = rec_t
    if host:
        records = torch.empty((rec_t.shape[0], 4), dtype=torch.int32).pin_memory()
        records.copy_(pack_pairs16(rec_t))

    def rank_fn(member):
        sh = RowShard(n, member)
        out = sharded_pairs_to_dendrogram(sh, records, off, g.window_bp, balance=balance, method="ward")
        res = dict(row0=sh.row0, sym=sh.sym.cpu().numpy().copy(), Z=out.get("Z"), it=out.get("sk_iterations"),
                   binned=out["stats"]["binned"] - out["stats"]["outside_rows"] if out["stats"] else 0,
                   engine=out.get("linkage_stats", {}).get("engine"))
        sh.close()
        return res
    results = LocalGroup(world).run(rank_fn)
    sym = single["sym"].cpu().numpy()
    for res in results:
        assert np.array_equal(res["sym"], sym[res["row0"]:res["row0"] + res["sym"].shape[0]])
        if balance:
            assert res["it"] == single["sk_iterations"]
    assert sum(r["binned"] for r in results) == single["stats"]["binned"]
    Z, want = results[0]["Z"], single["Z"]
    assert np.array_equal(Z[:, [0, 1, 3]], want[:, [0, 1, 3]])
    assert np.allclose(Z[:, 2], want[:, 2], rtol=1e-9, atol=0)


def test_sharded_job_refuses_unsorted_records():
    from hicembler_b200 import synth
    from hicembler_b200.parallel import LocalGroup, RowShard
    from hicembler_b200.pipeline import sharded_pairs_to_dendrogram
    n = 600
    g = synth.make_genome(30, n, 5000, 3, seed=2, min_tail=2000)
    rec_t = synth.make_pairs(g, 100000, seed=2, device="cuda")
    rec_t = rec_t[torch.randperm(rec_t.shape[0], device="cuda")].contiguous()
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()

    def rank_fn(member):
        sh = RowShard(n, member)
        try:
            sharded_pairs_to_dendrogram(sh, rec_t, off, g.window_bp)
            return "ok"
        except ValueError:
            return "refused"
        finally:
            sh.close()
    assert LocalGroup(2).run(rank_fn) == ["refused"] * 2


def test_c5_shaped_sharded_sinkhorn_eight_blocks():
    """BASELINE configs[4] (150k windows, dense float32, 8 row blocks, Sinkhorn only) scaled to one device: N = 4096 dense
    counts generated like bench.py --mode sinkhorn, 8 emulated ranks, against O.SinkhornKnopp"""
    from hicembler_b200.parallel import LocalGroup, RowShard
    n = 4096
    i = np.arange(n, dtype=np.float32)[:, None]
    j = np.arange(n, dtype=np.float32)[None, :]
    lam = 40.0 / (1.0 + np.abs(i - j)) ** 1.08 + 0.05
    h = np.abs(np.modf(np.sin((i * j + i + j) * 12.9898) * 43758.5453)[0])
    X = (np.floor(lam * (0.5 + h)) + 1.0).astype(np.float32)
    vmax = float(X.max())
    sk = O.SinkhornKnopp(max_iter=1000, epsilon=1e-5)
    sk.fit(X / np.float32(vmax), materialize=False)

    def rank_fn(member):
        sh = RowShard(n, member)
        sh.sym.copy_(torch.from_numpy(X[sh.row0:sh.row1]).cuda())
        member.barrier()
        res = sh.sinkhorn_fit(vmax, 1000, 1e-5)
        out = (sh.row0, res.r.cpu().numpy().copy(), res.c.cpu().numpy().copy(), res.iterations, res.stopping_condition)
        sh.close()
        return out
    for row0, r, c, it, stop in LocalGroup(8).run(rank_fn):
        assert (it, stop) == (sk._iterations, sk._stopping_condition)
        assert np.allclose(c, sk._c, rtol=1e-11)
        assert np.allclose(r, sk._r[row0:row0 + len(r)], rtol=1e-11)


def test_two_process_nccl_job():
    """the real thing: two processes, two GPUs, cudaIpc peer mappings over NVLink (skipped on a one-GPU box)"""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29561", os.path.join(ROOT, "scripts", "sharded_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "sharded check ok" in out.stdout
