"""BASELINE.json's full sizes (C2: 50 M pairs / 10 000 windows, C3: 500 M pairs / 50 000 windows) through
size-independent properties: a checksum of the counts, symmetry, unit row/column sums of the balanced matrix,
a well-formed dendrogram, and the two linkage engines agreeing with each other (the serial engine is the one
pinned bit-for-bit against the oracle at small sizes in test_gpu_parity.py)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _well_formed(Z, n):
    assert Z.shape == (n - 1, 4)
    assert np.all(np.diff(Z[:, 2]) >= 0)                       # sorted by height (generate_SciPy_dendrogram)
    a, b = Z[:, 0].astype(np.int64), Z[:, 1].astype(np.int64)
    assert np.all(a < b) and np.all(b < n + np.arange(n - 1))  # (min, max) labels, children exist already
    used = np.concatenate([a, b])
    assert len(np.unique(used)) == 2 * (n - 1)                 # every node merged exactly once
    size = np.ones(2 * n - 1)
    for k in range(n - 1):
        size[n + k] = size[a[k]] + size[b[k]]
    assert np.array_equal(size[n:], Z[:, 3]) and Z[-1, 3] == n


@pytest.mark.parametrize("config", ["C2", "C3"])
def test_full_size_properties(cuda, config):
    from hicembler_b200 import ops, synth
    from hicembler_b200.pipeline import pairs_to_dendrogram
    free, _ = torch.cuda.mem_get_info()
    cfg = synth.CONFIGS[config]
    N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
    if free < 24 * P * 3 + 40 * N * N:
        pytest.skip("not enough device memory for %s" % config)
    g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=3, min_tail=2000)
    rec = synth.make_pairs(g, P, seed=3, device=cuda)
    off = torch.as_tensor(g.window_offsets(), dtype=torch.int32, device=cuda)
    # --- counts: checksum = records surviving -q 10 -F 3980, every one of them inside the matrix
    counts, st = ops.bin_pairs(rec, off, ws, N, 10, return_stats=True)
    word = rec[:, 5].to(torch.int64)
    kept = int((((word & 0xffff) & 3980) == 0).logical_and((word >> 16) >= 10).sum())
    assert st["binned"] == kept == int(counts.sum(dtype=torch.int64)) and st["out_of_range"] == 0
    rows = torch.randint(0, N, (64,), device=cuda)
    S, vmax = ops.fold_counts(counts, "sym", add=1.0, inplace=True)       # counts is consumed
    assert torch.equal(S[rows][:, rows], S[rows][:, rows].T)             # symmetric
    assert float(S.max()) == float(vmax.item())
    # --- Sinkhorn-Knopp: diag(r) (S/vmax) diag(c) has unit column sums within epsilon and unit row sums
    vm = float(vmax.item())
    res = ops.sinkhorn_fit(S, vm, 1000, 1e-5)
    assert res.stopping_condition == "epsilon"
    Pn = (S[rows] / vm).double()                                       # d/vmax in float32 like the reference (:46)
    row_sums = (res.r[rows, None] * Pn * res.c[None, :]).sum(1)
    assert torch.allclose(row_sums, torch.ones_like(row_sums), rtol=0, atol=1e-6)   # (float32 d/vmax rounding)
    col = torch.zeros(N, dtype=torch.float64, device=cuda)
    for s0 in range(0, N, 2048):
        col += ((res.r[s0:s0 + 2048, None] * (S[s0:s0 + 2048] / vm).double()).sum(0)) * res.c
    assert float((col - 1).abs().max()) <= 1e-5
    del S, counts, Pn, col
    # --- both linkage engines on the same records: identical merges, heights to rounding
    kw = dict(mapq=10, params=(0.5, 1.1), method=cfg["method"], balance=True)
    fast = pairs_to_dendrogram(rec, off, ws, N, engine="auto", **kw)
    exact = pairs_to_dendrogram(rec, off, ws, N, engine="exact", **kw)
    _well_formed(exact["Z"], N)
    assert exact["linkage_stats"]["engine"] == "exact"
    Zf, Ze = fast["Z"], exact["Z"]
    assert np.array_equal(Zf[:, [0, 1, 3]], Ze[:, [0, 1, 3]])
    assert np.max(np.abs(Zf[:, 2] - Ze[:, 2]) / Ze[:, 2]) <= 1e-12      # BASELINE.json asks for 1e-5
    if fast["linkage_stats"]["engine"] == "rounds":                      # certified: no fallback happened
        assert fast["linkage_stats"]["certified"] and fast["linkage_stats"]["merges"] == N - 1
