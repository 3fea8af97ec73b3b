"""Scaffolding stage (bam2scaffolds.py:126-366): host logic against the oracle restatement on CPU, the
corner-block sums / tree walk / drivers on the GPU (``-m gpu``)."""
import os
import warnings

import numpy as np
import pytest

import hicembler_oracle as O
from util import hic_like_counts, sam_lines, small_case


class _FI:
    """stand-in built like tests/test_gpu_parity.py's: sizes only"""

    def __new__(cls, sizes):
        from hicembler_b200.FastaIndex import FastaIndex
        fi = FastaIndex.__new__(FastaIndex)
        fi.id2stats = {c: [int(s)] for c, s in sizes}
        fi.genomeSize = sum(int(s) for _, s in sizes)
        return fi


def _write_fasta(path, g, seed=0):
    rng = np.random.default_rng(seed)
    seqs = {}
    with open(path, "w") as out:
        for c, s in zip(g.names, g.sizes):
            seq = "".join(rng.choice(list("ACGTacgtN"), int(s)))
            seqs[c] = seq
            out.write(">%s some description\n" % c)
            for k in range(0, len(seq), 70):
                out.write(seq[k:k + 70] + "\n")
    return seqs


def test_get_reversed_and_indices():
    from hicembler_b200.bam2scaffolds import get_contig2indices, get_indices, get_reversed
    bin_chr = np.array(["b", "b", "a", "a", "a", "c"])
    c2i = get_contig2indices(bin_chr)
    assert c2i == O.get_contig2indices(bin_chr)
    sc = [("a", 0), ("zz", 1), ("b", 1), ("c", 0)]
    assert get_reversed(sc) == O.get_reversed(sc) == [("c", True), ("b", False), ("zz", False), ("a", True)]
    assert get_indices(sc, c2i) == O.get_indices(sc, c2i) == [2, 3, 4, 1, 0, 5]


def test_scaffold2gaps_matches_oracle():
    from hicembler_b200.bam2scaffolds import scaffold2gaps
    rng = np.random.default_rng(3)
    names = np.array(sorted("c%02d" % i for i in range(12)))
    sizes = [(c, int(rng.integers(3000, 3_000_000))) for c in names]
    d = (rng.random((12, 12)) * 30).astype(np.float32)
    d[2, 5] = d[5, 2] = 0.0          # no contacts: infinite distance -> capped gap
    d[7, 8] = d[8, 7] = 1e-4
    scaffold = [(names[i], bool(i % 3 == 0)) for i in rng.permutation(12)]
    scaffold = [(names[2], False), (names[5], True)] + [s for s in scaffold if s[0] not in (names[2], names[5])]
    for params in ((0.02, 1.1), (np.float64(1.0), np.float64(1.0)), (5e-5, 0.9)):
        want = O.scaffold2gaps(d, names, scaffold, params, dict(sizes))
        got = scaffold2gaps(d, names, scaffold, params, _FI(sizes))
        assert got == want
        assert got[0][-1][2] == 0 and all(isinstance(x[2], int) and x[2] >= 100 for x in got[0][:-1])
    with pytest.raises(IndexError):
        scaffold2gaps(d, names, [], (1.0, 1.0), _FI(sizes))


def test_fasta_get_sequence_and_report(tmp_path):
    from hicembler_b200.FastaIndex import FastaIndex
    from hicembler_b200.bam2scaffolds import report_scaffolds
    g, _ = small_case(n_contigs=6, n_windows=12, window_bp=500, n_pairs=10)
    fasta = str(tmp_path / "c.fa")
    seqs = _write_fasta(fasta, g)
    fi = FastaIndex(fasta)
    for c in g.names:
        assert fi.get_sequence(c) == seqs[c]
        assert fi.get_sequence(c, reverse=True) == O.get_reverse_complement(seqs[c])
    with pytest.raises(KeyError):
        fi.get_sequence("nope")
    n = g.names
    scaffolds = [[(n[0], 0, 120), (n[3], True, 100), (n[1], False, 0)], [], [(n[2], 1, 0)]]
    outbase = str(tmp_path / "auto.10k")
    fn = report_scaffolds(outbase, scaffolds, fi)
    fa, tab = O.report_scaffolds(scaffolds, seqs)
    assert open(fn).read() == fa and open(outbase + ".scaffolds.tab").read() == tab
    assert tab == "%s+\t%s-\t%s+\n%s-\n" % (n[0], n[3], n[1], n[2])


def _random_tree(n, rng):
    """a valid linkage array with random topology (heights irrelevant for the walk)"""
    alive = list(range(n))
    Z = np.zeros((n - 1, 4))
    for k in range(n - 1):
        i, j = rng.choice(len(alive), 2, replace=False)
        a, b = alive[i], alive[j]
        Z[k] = (min(a, b), max(a, b), k + 1.0, 0)
        alive = [x for x in alive if x not in (a, b)] + [n + k]
    return Z


@pytest.mark.parametrize("seed,minWindows", [(5, 2), (6, 3), (7, 1)])
def test_tree2scaffold_host_logic_with_numpy_corner_sums(monkeypatch, seed, minWindows):
    """the level-batched walk (children order, swap, minWindows rules, orientation table) against the oracle's
    node-by-node walk; the device corner sums are replaced by a numpy stand-in of hc_block_sums4's contract"""
    from hicembler_b200 import bam2scaffolds as B
    rng = np.random.default_rng(seed)
    K = 25
    names = np.array(["k%02d" % i for i in range(K)])
    runs = rng.integers(1, 7, K)
    bin_chr = np.concatenate([[names[c]] * runs[c] for c in rng.permutation(K)])
    d = hic_like_counts(len(bin_chr), seed)
    batches = []

    def sums(A, jobs):
        batches.append(len(jobs))
        out = np.zeros((len(jobs), 4))
        for j, (i1, i2) in enumerate(jobs):
            h, w = len(i1) // 2, len(i2) // 2
            w_end = w if w else len(i2)
            r = (i1[:h], i1[len(i1) - h:])
            c = (i2[:w], i2[len(i2) - w_end:])
            for q in range(4):
                out[j, q] = d[np.ix_(r[q >> 1], c[q & 1])].sum(dtype=np.float64)
        return out

    monkeypatch.setattr(B.ops, "block_sums4", sums)
    monkeypatch.setattr(B, "_device_matrix", lambda x: x)
    Z = _random_tree(K, rng)
    want = O.tree2scaffold(Z, names, d, bin_chr, minWindows)
    assert B.tree2scaffold((Z, list(names)), d, bin_chr, None, minWindows) == want
    assert len(batches) < K - 1 and sum(batches) <= K - 1      # one launch per tree level, not per node


# ------------------------------------------------------------------------------------------
# device-backed
# ------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_block_sums4_against_numpy(cuda):
    import torch
    from hicembler_b200 import ops
    rng = np.random.default_rng(1)
    n = 300
    d = hic_like_counts(n, 5)
    A = torch.from_numpy(d).cuda()
    jobs = []
    for n1, n2 in ((2, 2), (3, 3), (7, 4), (150, 149), (1, 1), (64, 3), (5, 0)):
        perm = rng.permutation(n)
        jobs.append((perm[:n1].tolist(), perm[n1:n1 + n2].tolist()))
    got = ops.block_sums4(A, jobs)
    for (i1, i2), row in zip(jobs, got):
        _d = d[:, i1 + i2][i1 + i2, :]
        n1, i = len(i1), len(i2) // 2
        d1, d2 = _d[:n1 // 2, n1:], _d[n1 - n1 // 2:n1, n1:]
        # the reference's slices, including [-0:] = the whole axis when n2 == 1
        want = [np.sum(x, dtype=np.float64) for x in (d1[:, :i], d1[:, -i:], d2[:, :i], d2[:, -i:])]
        assert np.array_equal(row, np.array(want))
    assert ops.block_sums4(A, []).shape == (0, 4)
    with pytest.raises(IndexError):
        ops.block_sums4(A, [([0, n], [1, 2])])


@pytest.mark.gpu
@pytest.mark.parametrize("seed,minWindows", [(1, 2), (2, 3), (3, 5), (4, 1)])
def test_tree2scaffold_matches_oracle_walk(cuda, seed, minWindows):
    from hicembler_b200.bam2clusters import HCTree
    from hicembler_b200.bam2scaffolds import join_scaffolds, tree2scaffold
    rng = np.random.default_rng(seed)
    K = 40
    names = np.array(["k%02d" % i for i in range(K)])
    runs = rng.integers(1, 9, K)
    order = rng.permutation(K)                      # window table order differs from np.unique order
    bin_chr = np.concatenate([[names[c]] * runs[c] for c in order])
    n = len(bin_chr)
    d = hic_like_counts(n, seed + 10)
    Z = _random_tree(K, rng)
    want = O.tree2scaffold(Z, names, d, bin_chr, minWindows)
    got = tree2scaffold(HCTree(Z, names), d, bin_chr, None, minWindows)
    assert got == want and len(got) > K // 3
    assert tree2scaffold((Z, list(names)), d, bin_chr, None, minWindows) == want
    c2i = O.get_contig2indices(bin_chr)
    s1, s2 = [(names[0], 0), (names[3], 1)], [(names[5], 1)]
    assert join_scaffolds(s1, s2, d, c2i, 1) == O.join_scaffolds(s1, s2, d, c2i, 1)
    assert join_scaffolds(s2, s1, d, c2i, 100) == O.join_scaffolds(s2, s1, d, c2i, 100) == []


@pytest.mark.gpu
def test_contigs2scaffold_matches_oracle(cuda):
    from hicembler_b200.bam2scaffolds import contigs2scaffold
    g, rec = small_case(n_contigs=30, n_windows=500, window_bp=5000, n_chrom=3, n_pairs=120000, seed=21)
    bam = [{"records": rec, "names": g.names}]
    lines = sam_lines(g, rec)
    params = (np.float64(0.05), np.float64(1.05))
    chrom = np.asarray(g.chrom_of)
    n_multi = 0
    for ci in range(3):
        cluster = [g.names[i] for i in np.flatnonzero(chrom == ci)]
        for ws, mw in ((5, 2), (10, 3)):
            want = O.contigs2scaffold([lines], g.contig_sizes, [ws], cluster, 10, mw, params)
            got = contigs2scaffold([ci + 1, bam, _FI(g.contig_sizes), [ws], cluster, 10, mw, params])
            assert got == want
            n_multi += len(got) > 3
    assert n_multi >= 4
    # one contig / no contig
    one = [g.names[0]]
    assert contigs2scaffold([1, bam, _FI(g.contig_sizes), [5], one, 10, 2, params]) == [(g.names[0], 0, 0)]
    assert contigs2scaffold([1, bam, _FI(g.contig_sizes), [5], ["absent"], 10, 2, params]) == []


@pytest.mark.gpu
def test_bam2scaffolds_driver_files(cuda, tmp_path):
    """FASTA + alignments in -> clusters, per-window-size scaffolds.fa / scaffolds.tab out; the scaffolds are
    the oracle's for the clusters and fitted parameters the driver used"""
    import pickle
    from hicembler_b200 import pairs
    from hicembler_b200.FastaIndex import FastaIndex
    from hicembler_b200.bam2scaffolds import bam2scaffolds
    g, rec = small_case(n_contigs=90, n_windows=900, window_bp=5000, n_chrom=4, n_pairs=80000, seed=11)
    fasta = str(tmp_path / "contigs.fa")
    seqs = _write_fasta(fasta, g)
    lines = sam_lines(g, rec)
    _, tlen, span = pairs.sam_to_records(lines, {n: i for i, n in enumerate(g.names)})
    bam = [dict(records=rec, names=g.names, tlen=tlen, ref_span=span)]
    outdir = str(tmp_path / "out")
    os.makedirs(outdir)
    with pytest.raises(SystemExit):
        bam2scaffolds(bam, fasta, outdir, 2000, [10 ** 6], 10, 1, 100, 0, 2, 0, 0)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert bam2scaffolds(bam, open(fasta), outdir, 2000, [5, 10 ** 6], 10, 1, 100, 0, 2, 4, 0) is None
        clusters = bam2scaffolds(bam, fasta, outdir, 2000, [5], 10, 1, 100, 0, 2, 4, 0, metagenomics=1)
    assert len(clusters) == 4
    base = os.path.join(outdir, "chr_4.5k")
    with open(base + ".distance.params", "rb") as fh:
        params = pickle.load(fh)
    contig2size = dict(g.contig_sizes)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        want_params = O.estimate_distance_parameters([lines], 10, contig2size, 5000)
    assert np.allclose(params, want_params, rtol=1e-9)
    want = [O.contigs2scaffold([lines], g.contig_sizes, [5], cl, 10, 2, want_params) for cl in clusters]
    fa, tab = O.report_scaffolds(want, seqs)
    assert open(base + ".scaffolds.tab").read() == tab
    assert open(base + ".scaffolds.fa").read() == fa
    assert FastaIndex(base + ".scaffolds.fa").genomeSize > sum(contig2size[c] for cl in want for c, o, gp in cl)


@pytest.mark.gpu
def test_load_matrix_matches_oracle(cuda, tmp_path):
    """array2scaffolds.py:51-121: cache files in, filtered / reordered / symmetrised / row-normalised matrix out"""
    import gzip
    from hicembler_b200.array2scaffolds import load_matrix
    g, rec = small_case(n_contigs=14, n_windows=150, window_bp=5000, n_pairs=40000, seed=8)
    wsz, windows, chr2window, _, _ = O.fasta2windows(g.contig_sizes, [5], skipShorter=0, minSize=2000, filterwindows=0)
    d = O.bam2arrays(windows, wsz, chr2window, [sam_lines(g, rec)], 10)[0]
    wins = windows[0]
    assert len({e - s for c, s, e in wins}) > 2          # contig tails: windows shorter than 5 kb get dropped
    fname = str(tmp_path / "x.5k.npz")
    np.savez_compressed(fname, d)
    with gzip.open(fname[:-4] + ".windows.tab.gz", "wt") as out:
        out.write("\n".join("\t".join(map(str, w)) for w in wins) + "\n")
    names = g.names
    sc1 = [(names[3], 0), (names[1], 1), ("absent", 0), (names[7], 0)]
    sc2 = [(names[0], 1), (names[5], 0)]
    for kw in (dict(), dict(remove_shorter=False), dict(chrs=[names[2], names[9], names[4]]),
               dict(scaffolds=[sc1]), dict(scaffolds=[sc1, sc2]), dict(chrs=[names[1], names[3]], scaffolds=[sc1])):
        wd, wc, wp, ws = O.load_matrix(d.copy(), wins, **kw)
        gd, gc, gp, gs = load_matrix(fname, **kw)
        assert gd.dtype == np.float32 and gd.shape == wd.shape and gd.shape[0] > 0
        assert np.allclose(gd, wd, rtol=3e-7, atol=0)
        assert list(gc) == list(wc) and np.array_equal(gp, wp) and gs == ws
