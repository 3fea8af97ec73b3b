"""The .npz-driven variant of the pipeline (reference array2scaffolds.py): transform, the diagonal-sum join_scaffolds,
tree2scaffold on the distance_matrix2tree tree, clusters2scaffolds and the array2scaffolds() driver, against the oracle
restatements (pinned to the reference's own code by tests/test_oracle_pinned.py)."""
import gzip
import os

import numpy as np
import pytest

import hicembler_oracle as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _case(seed=3, n_contigs=9, lo=1, hi=7):
    rng = np.random.default_rng(seed)
    nwin = rng.integers(lo, hi, n_contigs)
    bin_chr = np.array([("ctg%02d" % c) for c, k in enumerate(nwin) for _ in range(k)])
    n = len(bin_chr)
    i, j = np.indices((n, n))
    d = np.triu(rng.poisson(30.0 / (1.0 + np.abs(i - j)) + 0.2).astype(np.float32))
    d += d.T - np.diag(d.diagonal())
    return d, bin_chr


@pytest.mark.parametrize("minWindows", [1, 2, 3])
def test_diag_join_scaffolds_and_tree2scaffold(minWindows):
    from hicembler_b200.array2scaffolds import distance_matrix2tree, join_scaffolds, tree2scaffold
    d, bin_chr = _case()
    c2i = O.get_contig2indices(bin_chr)
    names = list(np.unique(bin_chr))
    rng = np.random.default_rng(0)
    for _ in range(20):
        a, b = rng.choice(len(names), 2, replace=False)
        s1, s2 = [(names[a], bool(rng.integers(2)))], [(names[b], 0)]
        assert join_scaffolds(s1, s2, d, c2i, minWindows) == O.join_scaffolds_diag(s1, s2, d, c2i, minWindows)
    red = O._average_reduce_2d(d, bin_chr)
    Z = O.linkage_scipy(O.condense(O.transform_logmax(red).astype(np.float64) + rng.random(red.shape) * 1e-9), "ward")
    want = O.tree2scaffold(Z, names, d, bin_chr, minWindows, join=O.join_scaffolds_diag, children="dm2t")
    assert tree2scaffold(distance_matrix2tree(Z, names), d, bin_chr, None, minWindows) == want


def test_transform_and_matrix_max():
    from hicembler_b200 import ops
    from hicembler_b200.array2scaffolds import transform
    d, _ = _case(seed=5, n_contigs=30)
    for dt in (np.float32, np.float64):
        x = d.astype(dt)
        assert ops.matrix_max(torch.from_numpy(x).cuda()) == float(x.max())
        got, want = transform(x), O.transform_logmax(x)
        assert got.dtype == dt
        tol = 4 * np.spacing(dt(want.max()))
        assert np.max(np.abs(got - want)) <= tol
    x = d.copy()
    x[3, 4] = np.nan
    assert np.isnan(ops.matrix_max(torch.from_numpy(x).cuda()))


def _write_cache(tmp_path, sizes, ws, seed):
    windows = [(c, s, min(s + ws, size)) for c, size in sizes for s in range(0, size, ws)]
    n = len(windows)
    rng = np.random.default_rng(seed)
    i, j = np.indices((n, n))
    chrom = np.array([int(w[0][3:]) // 3 for w in windows])            # three contigs per "chromosome"
    lam = 60.0 / (1.0 + np.abs(i - j)) ** 1.1 + 0.3 + 25.0 * (chrom[:, None] == chrom[None, :])
    d = np.triu(rng.poisson(lam).astype(np.float32))
    fn = str(tmp_path / "m.npz")
    np.savez_compressed(fn, d)
    with gzip.open(fn[:-4] + ".windows.tab.gz", "wt") as out:
        out.write("\n".join("\t".join(map(str, w)) for w in windows) + "\n")
    return fn, d, windows


def _oracle_cluster_scaffold(d, windows, contigs, minWindows):
    _d, _bc, _bp, _ = O.load_matrix(d.copy(), windows, chrs=contigs)
    red = O.transform_logmax(O._average_reduce_2d(_d, _bc))
    Z = O.linkage_scipy(O.condense(red), "ward")
    return O.tree2scaffold(Z, list(np.unique(_bc)), _d, _bc, minWindows, join=O.join_scaffolds_diag, children="dm2t")


def test_clusters2scaffolds_matches_oracle(tmp_path):
    from hicembler_b200.array2scaffolds import clusters2scaffolds
    sizes = [("ctg%02d" % k, s) for k, s in enumerate([40000, 25000, 30000, 45000, 20000, 35000, 30000, 25000, 40000])]
    fn, d, windows = _write_cache(tmp_path, sizes, 5000, seed=7)
    clusters = [["ctg00", "ctg01", "ctg02"], ["ctg03", "ctg04", "ctg05"], ["ctg06", "ctg07", "ctg08"]]
    got = clusters2scaffolds(clusters, fn, 2, [])
    for contigs, sc in zip(clusters, got):
        assert sc == _oracle_cluster_scaffold(d, windows, contigs, 2)
    # dealt over two ranks: every cluster is built by exactly one of them
    parts = [clusters2scaffolds(clusters, fn, 2, [], rank=r, world=2) for r in range(2)]
    assert [a if a is not None else b for a, b in zip(*parts)] == got
    assert all((a is None) != (b is None) for a, b in zip(*parts))


def test_array2scaffolds_driver(tmp_path):
    """array2scaffolds() end to end from the .npz cache and a FASTA: clusters == the oracle's get_clusters on the same
    tree, scaffolds == the oracle's per-cluster scaffolds, files written like the reference's"""
    from hicembler_b200.array2scaffolds import array2scaffolds
    sizes = [("ctg%02d" % k, s) for k, s in enumerate([40000, 25000, 30000, 45000, 20000, 35000, 30000, 25000, 40000])]
    fn, d, windows = _write_cache(tmp_path, sizes, 5000, seed=11)
    fa = str(tmp_path / "g.fa")
    rng = np.random.default_rng(1)
    seqs = {}
    with open(fa, "w") as out:
        for name, size in sizes:
            seqs[name] = "".join(rng.choice(list("ACGT"), size))
            out.write(">%s\n" % name)
            for s in range(0, size, 60):
                out.write(seqs[name][s:s + 60] + "\n")
    outbase = str(tmp_path / "out")
    scaffolds, fastafn = array2scaffolds(outbase, fn, fn, fa, 1, 2)
    # oracle: the same stages
    _d, bc, bp, c2s = O.load_matrix(d.copy(), windows)
    names = ["%s %7sk" % (c.split()[0], s // 1000) for c, (s, e) in zip(bc, bp)]
    Z = O.linkage_scipy(O.condense(O.transform_logmax(_d)), "ward")
    want_clusters = O.get_clusters(Z, names, c2s, bc)
    # All leaves of the (ultrametric) tree are equally far from the root up to rounding, so WHICH subtree the
    # reference cuts first -- the order of the clusters -- hangs on the last bits of the heights; the GPU heights
    # differ from scipy's in those bits (float32 logf).  The clusters themselves and their scaffolds must agree.
    got_clusters = [l.rstrip("\n").split("\t") for l in open(outbase + ".clusters.tab")]
    assert sorted(got_clusters) == sorted(want_clusters)
    want = [_oracle_cluster_scaffold(d, windows, contigs, 2) for contigs in got_clusters]
    assert scaffolds == want
    tab = open(outbase + ".scaffolds.tab").read()
    assert tab == "".join("\t".join("%s%s" % (c, "-" if o else "+") for c, o in sc) + "\n" for sc in want if sc)
    fa_txt, _ = O.report_scaffolds([[(c, o, 0) for c, o in sc] for sc in want], seqs)
    assert open(fastafn).read() == fa_txt
    # the tree file: same topology, branch lengths to float32 precision (CUDA logf vs glibc logf differ by ulps)
    import re
    num = re.compile(r":([0-9.eE+-]+)")
    got_nw, want_nw = open(outbase + ".nw").read(), O.distance_matrix2tree_newick(Z, names)
    assert num.sub(":", got_nw) == num.sub(":", want_nw)
    a, b = np.array(num.findall(got_nw), dtype=float), np.array(num.findall(want_nw), dtype=float)
    assert np.allclose(a, b, rtol=1e-4, atol=1e-5)
