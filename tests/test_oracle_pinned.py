"""The CPU oracle (oracle/hicembler_oracle.py) against the REFERENCE'S OWN CODE, executed from /root/reference by
oracle/ref_exec.py (documented Py2->Py3 shims, SAM text instead of the samtools pipe).  These are the checks that turn
"restatement only" into "pinned": every case runs the reference function and the restatement on the same seeded input.

Live part: needs /root/reference (this container); skipped elsewhere.  The same reference outputs are committed as
tests/golden/reference_pinned.npz by oracle/make_golden.py and re-checked by test_*_golden below, which run anywhere."""
import gzip
import os
import pickle
import sys

import numpy as np
import pytest

import hicembler_oracle as O  # conftest.py puts oracle/ on sys.path
import ref_exec as R

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

live = pytest.mark.skipif(not R.available(), reason="/root/reference not present (GPU box): live pinning runs in the build container")


def _genome(n_contigs=12, n_windows=60, ws=5000, seed=5, n_pairs=4000, **kw):
    from hicembler_b200 import synth
    from hicembler_b200.pairs import records_to_sam
    g = synth.make_genome(n_contigs, n_windows, ws, 3, seed=seed)
    rec = synth.records_numpy(synth.make_pairs(g, n_pairs, seed=seed, **kw))
    return g, rec, records_to_sam(rec, g.names)


def _hic(n, seed, dtype=np.float32, scale=40.0):
    rng = np.random.default_rng(seed)
    i, j = np.indices((n, n))
    return np.triu(rng.poisson(scale / (1.0 + np.abs(i - j)) ** 1.08 + 0.05).astype(dtype))


# ------------------------------------------------------------------------------------------------------------
# binning
# ------------------------------------------------------------------------------------------------------------
@live
def test_get_window_matches_reference():
    ref = R.load("bam2clusters.py")
    c2w = {"a": 0, "b": 17}
    for chrom in c2w:
        for start in (1, 2, 99, 4999, 5000, 5001, 123456):
            for length in (1, 2, 35, 100, 101, 151):
                for ws in (1000, 2000, 5000, 60000):
                    assert ref.get_window(chrom, start, length, ws, c2w) == O.get_window(chrom, start, length, ws, c2w)


@live
@pytest.mark.parametrize("case", ["multi_ws", "upto", "two_bams_upto", "regions", "mapq_shadow", "seq_star"])
def test_bam2arrays_matches_reference(case):
    g, rec, sam = _genome(with_mates=True, bad_flag_frac=0.02, vary_len=True)
    sam2 = sam[::-1]
    if case == "mapq_shadow":       # :145 overwrites the mapq argument with the last line's MAPQ string
        rec = rec.copy()
        hi = np.flatnonzero(rec["mapq"] >= 10)
        rec["mapq"][hi[::3]] = 30
        rec["mapq"][hi[-1]] = 45
        from hicembler_b200.pairs import records_to_sam
        sam = sam2 = records_to_sam(rec, g.names)
    if case == "seq_star":          # SEQ "*": len == 1
        sam = ["\t".join(l.split("\t")[:9] + ["*", "*"]) for l in sam]
    ws_kb = [5, 10, 2] if case == "multi_ws" else [5]
    wsz, windows, chr2window, _, _ = O.fasta2windows(g.contig_sizes, ws_kb, minSize=2000, filterwindows=0)
    ref = R.load("bam2clusters.py", sam_sources={"a.bam": sam, "b.bam": sam2})
    kw, bams, sams = {}, ["a.bam"], [sam]
    if case == "upto":
        kw = dict(upto=1500)
    if case in ("two_bams_upto", "mapq_shadow"):
        kw, bams, sams = dict(upto=5000), ["a.bam", "b.bam"], [sam, sam2]
    if case == "regions":
        kw = dict(regions=[g.names[1], g.names[4]])
    want = ref.bam2arrays(windows, wsz, chr2window, bams, 10, verbose=0, **kw)
    got = O.bam2arrays(windows, wsz, chr2window, sams, 10, **kw)
    assert len(want) == len(got)
    for a, b in zip(want, got):
        assert a.dtype == b.dtype == np.float32 and a.sum() > 0
        assert np.array_equal(a, b)


@live
def test_bam2arrays_errors_match_reference():
    g, rec, sam = _genome()
    wsz, windows, chr2window, _, _ = O.fasta2windows(g.contig_sizes, [5, 10], minSize=2000, filterwindows=0)
    ref = R.load("bam2clusters.py", sam_sources={"a.bam": sam})
    c2w = [dict(chr2window[0]), {k: v for k, v in list(chr2window[1].items())[:-2]}]     # KeyError (:155)
    with pytest.raises(KeyError):
        ref.bam2arrays(windows, wsz, c2w, ["a.bam"], 10, verbose=0)
    with pytest.raises(KeyError):
        O.bam2arrays(windows, wsz, c2w, [sam], 10)
    short = [windows[0][:-3], windows[1]]                                                  # IndexError (:161)
    with pytest.raises(IndexError):
        ref.bam2arrays(short, wsz, chr2window, ["a.bam"], 10, verbose=0)
    with pytest.raises(IndexError):
        O.bam2arrays(short, wsz, chr2window, [sam], 10)


@live
def test_bam2array_contig_level_matches_reference():
    g, rec, sam = _genome(n_contigs=10, n_windows=80, ws=5000, n_pairs=6000, with_mates=True)
    from hicembler_b200.pairs import records_to_sam
    tlen = (rec["next_pos"].astype(np.int64) - rec["pos"]).astype(np.int32)
    sam = records_to_sam(rec, g.names, tlen=tlen)
    minSize, windows, chr2window, contig2size = O.contigs2windows(g.contig_sizes, 2000)
    ref = R.load("bam2clusters.py", sam_sources={"a.bam": sam})
    want, want_d = ref.bam2array(windows, contig2size, chr2window, ["a.bam"], mapq=10, minSize=minSize, threads=3, verbose=0)
    got, got_d = O.bam2array(windows, contig2size, chr2window, [sam], mapq=10, minSize=minSize, threads=3)
    assert want.dtype == got.dtype and want.sum() > 0
    assert np.array_equal(want, got)
    assert {c: sorted(v) for c, v in want_d.items()} == {c: sorted(v) for c, v in got_d.items()}


# ------------------------------------------------------------------------------------------------------------
# normalisations and transforms
# ------------------------------------------------------------------------------------------------------------
@live
@pytest.mark.parametrize("n,seed,dt", [(150, 3, np.float32), (120, 5, np.float64)])
def test_normalize_matches_reference(n, seed, dt):
    ref = R.load("bam2clusters.py")
    d1, d2 = _hic(n, seed, dt), _hic(n, seed, dt)
    want, got = ref.normalize(d1), O.normalize(d2)
    assert np.array_equal(want, got) and np.array_equal(d1, d2)      # result and the in-place mutation of the argument


@live
def test_normalize_diagional_matches_reference():
    ref = R.load("bam2clusters.py")
    d = _hic(90, 8)
    d[[5, 40]] = 0
    d[:, [5, 40]] = 0
    bc = np.array(["c%d" % (k // 9) for k in range(90)])
    bp = np.array([(k * 10, k * 10 + 10) for k in range(90)])
    w = ref.normalize_diagional(d.copy(), bc, bp)
    g = O.normalize_diagional(d.copy(), bc, bp)
    assert np.array_equal(w[0], g[0]) and np.array_equal(w[1], g[1]) and np.array_equal(w[2], g[2])


@live
def test_distance_and_contact_func_match_reference():
    ref = R.load("bam2clusters.py")
    a2s = R.load("array2scaffolds.py")
    x = np.abs(np.random.default_rng(3).normal(3, 2, (50, 50))) + 0.01
    for a, b in ((0.5, 1.1), (2.3, 0.7), (1e-3, 0.05)):
        assert np.array_equal(ref.distance_func(x, a, b), O.distance_func(x, a, b))
        assert np.array_equal(ref.contact_func(x, a, b), O.contact_func(x, a, b))
    for dt in (np.float32, np.float64):
        d = _hic(64, 2, dt)
        w, g = a2s.transform(d), O.transform_logmax(d)
        assert w.dtype == g.dtype == dt and np.array_equal(w, g)


@live
def test_row_and_window_normalisations_match_reference():
    a2s, b2s = R.load("array2scaffolds.py"), R.load("bam2scaffolds.py")
    d = _hic(70, 4)
    d += d.T
    d[13] = 0
    assert np.array_equal(a2s.normalize_rows(d.copy()), O.normalize_rows(d.copy()))
    assert np.array_equal(b2s.normalize_rows(d.copy()), O.normalize_rows(d.copy()))
    bc = np.array(["c%d" % (k // 10) for k in range(70)])
    bp = np.array([(0, 5000) if k % 10 else (0, 3100) for k in range(70)])
    u = _hic(70, 6)
    w = b2s.normalize_window_size(u.copy(), bc, bp)
    g = O.normalize_window_size(u.copy(), bc, bp)
    assert np.array_equal(w[0], g[0])
    keys = np.array(["c%02d" % (k // 7) for k in range(70)])
    for mod in (a2s, b2s):
        assert np.array_equal(mod._average_reduce_2d(d, keys), O._average_reduce_2d(d, keys))


# ------------------------------------------------------------------------------------------------------------
# window tables (FastaIndex + contigs2windows + fasta2windows)
# ------------------------------------------------------------------------------------------------------------
def _write_fasta(path, contig_sizes):
    rng = np.random.default_rng(1)
    with open(path, "w") as out:
        for name, size in contig_sizes:
            seq = "".join(rng.choice(list("ACGT"), size))
            out.write(">%s\n" % name)
            for s in range(0, size, 60):
                out.write(seq[s:s + 60] + "\n")


@live
def test_window_tables_match_reference(tmp_path):
    """Python 2 iterates FastaIndex in dict-hash order; under Python 3 the same code iterates in file order, which is
    the order the restatement takes as explicit input."""
    sizes = [("ctg%03d" % k, s) for k, s in enumerate([41000, 3500, 12000, 1500, 27000, 9000, 2600, 60000])]
    fa = str(tmp_path / "g.fa")
    _write_fasta(fa, sizes)
    b2c, b2s = R.load("bam2clusters.py"), R.load("bam2scaffolds.py")
    want = b2c.contigs2windows(open(fa), minSize=2000, verbose=0)
    got = O.contigs2windows(sizes, 2000)
    assert want[0] == got[0] and want[1] == got[1] and want[2] == got[2] and want[3] == got[3]
    for ws_kb, kw in (([5, 10, 2], dict(skipShorter=1, filterwindows=0)), ([5, 10, 2], dict(skipShorter=0, filterwindows=0)),
                      ([1, 5, 3], dict(skipShorter=1, filterwindows=1))):      # 5 kb fails the 1 % of genome filter
        w = b2s.fasta2windows(open(fa), ws_kb, 0, minSize=2000, **kw)
        g = O.fasta2windows(sizes, ws_kb, minSize=2000, **kw)
        assert list(w[0]) == list(g[0]) and w[1] == g[1] and w[2] == g[2] and w[3] == g[3] and w[4] == g[4]


# ------------------------------------------------------------------------------------------------------------
# trees: getNewick / array2tree / cluster_contigs (fastcluster = oracle/nnchain.c stand-in on both sides)
# ------------------------------------------------------------------------------------------------------------
@live
def test_array2tree_newick_matches_reference(tmp_path):
    ref = R.load("bam2clusters.py")
    d = _hic(40, 11, np.float64)
    d += d.T - np.diag(d.diagonal())
    dist = O.transform_distance(d, 0.5, 1.1) + np.random.default_rng(1).random((40, 40)) * 1e-3
    dist = np.triu(dist, 1) + np.triu(dist, 1).T
    names = ["w%02d" % k for k in range(40)]
    t = ref.array2tree(dist, names, method="ward")
    import scipy.cluster.hierarchy as sch
    Z = O.linkage_fastcluster(O.condense(dist), "ward")
    from hicembler_b200.bam2clusters import newick_from_linkage
    nw = newick_from_linkage(Z, names)                           # the product's iterative Newick writer (host code)
    import ete3_min
    assert ete3_min.Tree(nw).write() == t.write()
    assert nw == ref.getNewick(sch.to_tree(Z, False), "", sch.to_tree(Z, False).dist, names)


@live
def test_cluster_contigs_matches_reference(tmp_path):
    ref = R.load("bam2clusters.py")
    d = _hic(60, 21, np.float64, scale=20.0)
    for b in range(0, 60, 20):
        d[b:b + 20, b:b + 20] += np.triu(np.random.default_rng(b).poisson(30, (20, 20)))
    d += d.T - np.diag(d.diagonal())
    dist = O.transform_distance(d, 0.5, 1.1) + np.random.default_rng(2).random((60, 60)) * 1e-6
    dist = np.triu(dist, 1) + np.triu(dist, 1).T
    bin_chr = np.array(["c%02d" % k for k in range(60)])
    np.random.seed(1234)
    want = ref.cluster_contigs(str(tmp_path / "o"), dist, bin_chr, None, iterations=5, minchr=2)
    np.random.seed(1234)
    got, nchr, Z, msl = O.cluster_contigs(dist, bin_chr, iterations=5, minchr=2)
    assert want == got and len(got) == nchr
    want2 = ref.cluster_contigs(str(tmp_path / "o"), dist, bin_chr, None, nchr=4)
    assert want2 == O.cluster_contigs(dist, bin_chr, nchr=4)[0]


# ------------------------------------------------------------------------------------------------------------
# distance-decay fit
# ------------------------------------------------------------------------------------------------------------
@live
def test_distance_parameters_match_reference(tmp_path):
    g, rec, sam = _genome(n_contigs=6, n_windows=120, ws=2000, n_pairs=20000, seed=9)
    ref = R.load("bam2clusters.py", sam_sources={"a.bam": sam})
    c2s = dict(g.contig_sizes)
    wd, wc = ref.get_distances_contacts(["a.bam"], 10, c2s, 2000, icontigs=3)
    gd, gc = O.get_distances_contacts([sam], 10, c2s, 2000, icontigs=3)
    assert np.array_equal(wd, gd) and len(wc) == len(gc)
    for a, b in zip(wc, gc):
        assert np.array_equal(np.asarray(a), np.asarray(b))
    wp = ref.estimate_distance_parameters(str(tmp_path / "w"), ["a.bam"], 10, c2s, 2000, icontigs=3)
    gp = O.estimate_distance_parameters([sam], 10, c2s, 2000, icontigs=3)
    assert np.allclose(wp, gp, rtol=1e-12)
    with open(str(tmp_path / "w.distance.params"), "rb") as fh:
        assert np.array_equal(pickle.load(fh), wp)


# ------------------------------------------------------------------------------------------------------------
# scaffolding: join_scaffolds (both variants), tree2scaffold, scaffold2gaps, load_matrix, get_clusters
# ------------------------------------------------------------------------------------------------------------
def _scaffold_case(seed=3, n_contigs=7):
    rng = np.random.default_rng(seed)
    nwin = rng.integers(1, 7, n_contigs)
    bin_chr = np.array([("ctg%02d" % c) for c, k in enumerate(nwin) for _ in range(k)])
    n = len(bin_chr)
    i, j = np.indices((n, n))
    d = rng.poisson(30.0 / (1.0 + np.abs(i - j)) ** 1.0 + 0.2).astype(np.float32)
    d = np.triu(d)
    d += d.T - np.diag(d.diagonal())
    return d, bin_chr


@live
@pytest.mark.parametrize("minWindows", [1, 2, 3])
def test_join_scaffolds_and_tree2scaffold_match_reference(minWindows):
    b2s, a2s = R.load("bam2scaffolds.py"), R.load("array2scaffolds.py")
    import ete3_min
    d, bin_chr = _scaffold_case()
    c2i = O.get_contig2indices(bin_chr)
    names = list(np.unique(bin_chr))
    rng = np.random.default_rng(0)
    for _ in range(30):
        a, b = rng.choice(len(names), 2, replace=False)
        s1, s2 = [(names[a], bool(rng.integers(2)))], [(names[b], 0)]
        assert b2s.join_scaffolds(s1, s2, d, c2i, minWindows) == O.join_scaffolds(s1, s2, d, c2i, minWindows)
        assert a2s.join_scaffolds(s1, s2, d, c2i, minWindows) == O.join_scaffolds_diag(s1, s2, d, c2i, minWindows)
    red = O._average_reduce_2d(d, bin_chr)
    Z = O.linkage_scipy(O.condense(O.transform_logmax(red).astype(np.float64) + rng.random(red.shape) * 1e-9), "ward")
    for mod, join in ((b2s, O.join_scaffolds), (a2s, O.join_scaffolds_diag)):
        t = a2s.distance_matrix2tree(Z, names)
        want = mod.tree2scaffold(t, d, bin_chr, None, minWindows)
        got = O.tree2scaffold(Z, names, d, bin_chr, minWindows, join=join)
        assert want == got


@live
def test_scaffold2gaps_matches_reference(tmp_path):
    b2s = R.load("bam2scaffolds.py")
    sizes = [("ctg%02d" % k, s) for k, s in enumerate([41000, 23500, 12000, 31500, 27000])]
    fa = str(tmp_path / "g.fa")
    _write_fasta(fa, sizes)
    ws = 5000
    bin_chr = np.array([c for c, s in sizes for _ in range(0, s, ws)])
    n = len(bin_chr)
    rng = np.random.default_rng(4)
    i, j = np.indices((n, n))
    d = rng.poisson(200.0 / (1.0 + np.abs(i - j)) ** 1.2 + 0.5).astype(np.float32)
    d = np.triu(d)
    d += d.T - np.diag(d.diagonal())
    scaffold = [("ctg02", 0), ("ctg00", 1), ("ctg04", 0), ("ctg01", 0), ("ctg03", 1)]
    params = (0.01, 1.2)
    want = b2s.scaffold2gaps(d, bin_chr, scaffold, params, open(fa))
    got = O.scaffold2gaps(d, bin_chr, scaffold, params, dict(sizes))
    assert want == got


@live
def test_load_matrix_matches_reference(tmp_path):
    a2s = R.load("array2scaffolds.py")
    sizes = [("ctg%02d" % k, s) for k, s in enumerate([41000, 23500, 12000, 31500, 27000])]
    ws = 5000
    windows = [(c, s, min(s + ws, size)) for c, size in sizes for s in range(0, size, ws)]
    n = len(windows)
    d = _hic(n, 12)
    fn = str(tmp_path / "m.npz")
    np.savez_compressed(fn, d)
    with gzip.open(fn[:-4] + ".windows.tab.gz", "wt") as out:
        out.write("\n".join("\t".join(map(str, w)) for w in windows) + "\n")
    cases = [dict(), dict(remove_shorter=False), dict(chrs=["ctg01", "ctg03"]),
             dict(scaffolds=[[("ctg02", 0), ("ctg00", 1)]]), dict(scaffolds=[[("ctg02", 0)], [("ctg04", 1), ("ctg01", 0)]])]
    for kw in cases:
        w = a2s.load_matrix(fn, **kw)
        g = O.load_matrix(d.copy(), windows, **kw)
        assert np.array_equal(w[0], g[0]), kw
        assert list(w[1]) == list(g[1]) and np.array_equal(np.asarray(w[2]), np.asarray(g[2])) and w[3] == g[3], kw


@live
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_get_clusters_and_tree_match_reference(tmp_path, seed):
    a2s = R.load("array2scaffolds.py")
    rng = np.random.default_rng(seed)
    n_contigs, per = 9, 4
    bin_chr = np.array(["ctg%02d" % c for c in range(n_contigs) for _ in range(per)])
    n = len(bin_chr)
    d = rng.poisson(2.0, (n, n)).astype(np.float32)
    for b in range(0, n, 3 * per):
        d[b:b + 3 * per, b:b + 3 * per] += rng.poisson(40, (3 * per, 3 * per))
    d = np.triu(d)
    d += d.T - np.diag(d.diagonal())
    names = ["%s %7sk" % (c, k) for k, c in enumerate(bin_chr)]
    dist = O.transform_logmax(d).astype(np.float64) + rng.random((n, n)) * 1e-9
    Z = O.linkage_scipy(O.condense(dist), "ward")
    contig2size = {c: per * 5000 for c in np.unique(bin_chr)}
    t = a2s.distance_matrix2tree(Z, names)
    assert t.write() == O.distance_matrix2tree_newick(Z, names)
    try:
        want = a2s.get_clusters(str(tmp_path / "w"), t, contig2size, bin_chr)
    except IndexError:
        with pytest.raises(IndexError):
            O.get_clusters(Z, names, contig2size, bin_chr)
        return
    got = O.get_clusters(Z, names, contig2size, bin_chr)
    assert want == got
    assert open(str(tmp_path / "w.clusters.tab")).read() == "".join("\t".join(c) + "\n" for c in got)


# ------------------------------------------------------------------------------------------------------------
# the product's HOST-side tree code (no GPU involved) against the reference executed here
# ------------------------------------------------------------------------------------------------------------
@live
@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_product_linkage_tree_and_get_clusters_match_reference(tmp_path, seed):
    from hicembler_b200.array2scaffolds import distance_matrix2tree, get_clusters
    a2s = R.load("array2scaffolds.py")
    rng = np.random.default_rng(seed)
    n_contigs, per = 9, 4
    bin_chr = np.array(["ctg%02d" % c for c in range(n_contigs) for _ in range(per)])
    n = len(bin_chr)
    d = rng.poisson(2.0, (n, n)).astype(np.float32)
    for b in range(0, n, 3 * per):
        d[b:b + 3 * per, b:b + 3 * per] += rng.poisson(40, (3 * per, 3 * per))
    d = np.triu(d)
    d += d.T - np.diag(d.diagonal())
    names = ["%s %7sk" % (c, k) for k, c in enumerate(bin_chr)]
    Z = O.linkage_scipy(O.condense(O.transform_logmax(d).astype(np.float64) + rng.random((n, n)) * 1e-9), "ward")
    contig2size = {c: per * 5000 for c in np.unique(bin_chr)}
    t_ref, t = a2s.distance_matrix2tree(Z, names), distance_matrix2tree(Z, names)
    assert t.write() == t_ref.write() and len(t) == len(t_ref)
    assert [t.name[k] for k in t.leaves()] == [x.name for x in t_ref]
    try:
        want = a2s.get_clusters(str(tmp_path / "w"), t_ref, contig2size, bin_chr)
    except IndexError:
        with pytest.raises(IndexError):
            get_clusters(str(tmp_path / "g"), t, contig2size, bin_chr)
        return
    got = get_clusters(str(tmp_path / "g"), t, contig2size, bin_chr)
    assert want == got
    assert open(str(tmp_path / "w.clusters.tab")).read() == open(str(tmp_path / "g.clusters.tab")).read()


@live
def test_product_diag_cells_match_reference_slices():
    """the window pairs the product gathers for array2scaffolds.join_scaffolds are exactly the np.diag slices of the
    reference (:286-291), including i == 0 (``[-0:]`` = everything) and n1 == n2"""
    from hicembler_b200.array2scaffolds import _diag_cells
    rng = np.random.default_rng(0)
    for n1, n2 in [(5, 3), (4, 4), (7, 1), (2, 2), (9, 6), (3, 2)]:
        idx = list(rng.permutation(40)[:n1 + n2])
        d = rng.random((40, 40)).astype(np.float32)
        _d = d[:, idx][idx, :]
        i = n2 // 2
        dflip = np.fliplr(_d[:n1, n2:])
        want = (np.diag(_d, k=n1)[:i], np.diag(dflip)[:i], np.diag(dflip, k=n2 - n1)[-i:], np.diag(_d, k=n2)[-i:])
        got = _diag_cells(idx[:n1], idx[n1:])
        for w, cells in zip(want, got):
            assert np.array_equal(w, np.array([d[r, c] for r, c in cells], dtype=np.float32))


# ------------------------------------------------------------------------------------------------------------
# committed reference outputs (run anywhere): oracle/make_golden.py wrote them by EXECUTING the reference
# ------------------------------------------------------------------------------------------------------------
def test_reference_pinned_golden():
    z = np.load(os.path.join(GOLDEN, "reference_pinned.npz"), allow_pickle=False)
    d = z["hic_f32"]
    assert np.array_equal(O.transform_logmax(d.copy()), z["transform_logmax"])
    assert np.array_equal(O.normalize_rows(z["sym_f32"].copy()), z["normalize_rows"])
    keys = np.array([s for s in z["keys"]])
    assert np.array_equal(O._average_reduce_2d(z["sym_f32"], keys), z["average_reduce_2d"])
    x = z["dist_x"]
    assert np.array_equal(O.distance_func(x, 0.5, 1.1), z["distance_func"])
    w = O.normalize_diagional(d.copy(), keys, z["bin_position"])
    assert np.array_equal(w[0], z["normalize_diagional"])
    assert np.array_equal(O.normalize_window_size(d.copy(), keys, z["bin_position"])[0], z["normalize_window_size"])
