"""Distance-decay fit (bam2clusters.py:257-358) and the bam2clusters() driver (:459-537): host logic
against the oracle restatement on CPU, the device-backed parts under ``-m gpu``."""
import gzip
import os
import pickle

import numpy as np
import pytest

import hicembler_oracle as O
from util import sam_lines, small_case


def _c2dists(seed, n_contigs=30, ws=2000, empty_trans=False):
    rng = np.random.default_rng(seed)
    contig2size, c2d = {}, {}
    for i in range(n_contigs):
        c = "ctg%03d" % i
        contig2size[c] = int(rng.integers(20_000, 900_000))
        m = int(rng.integers(200, 3000))
        # power-law insert sizes, some beyond the last bin, zeros, and -1 for trans mates
        v = (ws * 0.2 * (rng.pareto(1.1, m) + 0.05)).astype(np.int64)
        v[rng.random(m) < 0.15] = -1
        v[rng.random(m) < 0.02] = 0
        c2d[c] = v.tolist()
    if empty_trans:
        c2d["ctg001"] = [-1, -1, 0, 0, 1]  # sums to -1: negative normalisation like the reference
        c2d["ctg002"] = [0, 0, -1, -1]     # max == 0 -> skipped (:309)
    return contig2size, c2d


@pytest.mark.parametrize("seed,ws,limit,contigs,odd", [(1, 2000, 30, (), False), (2, 5000, 12, (), True),
                                                      (3, 1500, 30, tuple("ctg%03d" % i for i in range(0, 30, 2)), True)])
def test_insert_size_profile_matches_oracle(seed, ws, limit, contigs, odd):
    from hicembler_b200.bam2clusters import _insert_size_profile
    contig2size, c2d = _c2dists(seed, ws=ws, empty_trans=odd)
    want = O.distance_fit_inputs(contig2size, ws, c2d, limit, contigs)
    got = _insert_size_profile(contig2size, ws, c2d, limit, list(contigs))
    assert len(want[0]) > 50
    for w, g in zip(want, got):
        assert len(w) == len(g)
    assert np.array_equal(np.array(got[0]), np.array(want[0])) and np.array_equal(np.array(got[1]), np.array(want[1]))
    assert got[2] == want[2]
    assert all(a == b for a, b in zip(got[3], want[3]))


def test_insert_size_profile_empty_list_raises_like_max():
    from hicembler_b200.bam2clusters import _insert_size_profile
    with pytest.raises(ValueError):
        _insert_size_profile({"a": 10}, 2000, {"a": []}, 30, [])
    with pytest.raises(ValueError):
        O.distance_fit_inputs({"a": 10}, 2000, {"a": []}, 30, ())


def test_estimate_distance_parameters_from_insert_sizes(tmp_path):
    """same curve_fit inputs -> same parameters; the pickle holds them (:346-348)"""
    import warnings
    from hicembler_b200.bam2clusters import contact_func, estimate_distance_parameters
    rng = np.random.default_rng(7)
    ws = 2000
    # insert sizes drawn so that bin counts follow 1/(a x^b): strictly positive distances only
    contig2size, c2d = {}, {}
    for i in range(8):
        c = "c%d" % i
        contig2size[c] = 500_000 + i
        edges = np.arange(0, 45 * ws, ws)
        p = contact_func((edges[1:] / 1000.), 0.02, 1.1)
        p = p / p.sum()
        k = rng.choice(len(p), size=20000, p=p)
        c2d[c] = (edges[k] + rng.integers(1, ws, len(k))).tolist()
    outbase = str(tmp_path / "auto")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        want = O.estimate_distance_parameters(contig2size=contig2size, windowSize=ws, c2dists=c2d)
        got = estimate_distance_parameters(outbase, contig2size=contig2size, windowSize=ws, c2dists=c2d)
    assert np.array_equal(got, want)
    assert 0.9 < got[1] < 1.3
    with open(outbase + ".distance.params", "rb") as fh:
        assert np.array_equal(pickle.load(fh), got)


# ------------------------------------------------------------------------------------------
# device-backed
# ------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("ws", [5000, 700])
def test_get_distances_contacts_matches_oracle(cuda, ws):
    from hicembler_b200.bam2clusters import get_distances_contacts
    g, rec = small_case(n_contigs=8, n_windows=120, window_bp=5000, n_pairs=30000, seed=3)
    contig2size = dict(zip(g.names, (int(s) for s in g.sizes)))
    lines = sam_lines(g, rec)
    wd, wc = O.get_distances_contacts([lines], 10, contig2size, ws, icontigs=3)
    gd, gc = get_distances_contacts([{"records": rec, "names": g.names}], 10, contig2size, ws, icontigs=3)
    assert np.array_equal(gd, wd) and len(gd) > 3
    for a, b in zip(gc, wc):
        assert np.array_equal(np.sort(np.asarray(a, dtype=np.float64)), np.sort(np.asarray(b, dtype=np.float64)))
    assert sum(float(np.sum(c)) for c in gc) > 0


@pytest.mark.gpu
def test_bam2clusters_driver_matches_oracle_flow(cuda, tmp_path):
    """FASTA + alignments in, clusters / cache files out: same clusters as the oracle's
    bam2array -> fit -> symmetrise -> distance_func -> cluster_contigs with the same sub-samples"""
    import warnings
    from hicembler_b200 import pairs
    from hicembler_b200.bam2clusters import bam2clusters
    # >= 50 contigs above N95: the fit reads c2dists of the 50 longest contigs of the WHOLE assembly (:305-308)
    g, rec = small_case(n_contigs=90, n_windows=900, window_bp=5000, n_chrom=4, n_pairs=80000, seed=11)
    fasta = str(tmp_path / "contigs.fa")
    with open(fasta, "w") as out:
        for c, s in zip(g.names, g.sizes):
            out.write(">%s\n" % c)
            seq = "ACGT" * (int(s) // 4 + 1)
            for k in range(0, int(s), 60):
                out.write(seq[k:min(k + 60, int(s))] + "\n")
    lines = sam_lines(g, rec)
    _, tlen, span = pairs.sam_to_records(lines, {n: i for i, n in enumerate(g.names)})
    bam = [dict(records=rec, names=g.names, tlen=tlen, ref_span=span)]
    # oracle flow
    sizes = [(c, int(s)) for c, s in zip(g.names, g.sizes)]
    minSize, windows, chr2window, contig2size = O.contigs2windows(sizes, minSize=2000)
    K = len(windows)
    assert K >= 50
    d, c2d = O.bam2array(windows, contig2size, chr2window, [lines], mapq=10, minSize=minSize, threads=3)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        params = O.estimate_distance_parameters(contig2size=contig2size, windowSize=minSize, c2dists=c2d)
    d = O.symmetrize(d)
    dist = O.distance_func(d + 1, *params)
    rng = np.random.default_rng(0)
    sel = [rng.permutation(K)[:int(round(K * 0.66))] for _ in range(30)]
    want, nchr, _, _ = O.cluster_contigs(dist, [w[0] for w in windows], selections=sel)
    want = [c for c in want if c]
    outdir = str(tmp_path / "out")
    os.makedirs(outdir)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        got = bam2clusters(bam, fasta, outdir, minSize=2000, mapq=10, verbose=0, selections=sel)
    assert got == want and len(got) >= 3
    base = os.path.join(outdir, "auto")
    with open(base + ".distance.params", "rb") as fh:
        assert np.allclose(pickle.load(fh), params, rtol=1e-9)
    with gzip.open(base + ".windows.tab.gz", "rt") as fh:
        assert [l.split("\t")[0] for l in fh.read().splitlines()] == [w[0] for w in windows]
    assert [l[:-1].split("\t") for l in open(base + ".clusters.tab")] == got
    # second call: clusters come from the cache file (:472-476); a fixed nchr reuses auto.* (:466-469)
    assert bam2clusters(bam, fasta, outdir, verbose=0) == got
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        got4 = bam2clusters(bam, fasta, outdir, nchr=4, verbose=0)
    assert os.path.islink(os.path.join(outdir, "chr_4.npz")) and len(got4) == 4
    want4, _, _, _ = O.cluster_contigs(dist, [w[0] for w in windows], nchr=4)
    assert got4 == [c for c in want4 if c]
