"""Parity of the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): contact counts bit-exact; balanced matrix <= 1e-5 relative with
identical iteration count / stopping condition; identical linkage merge order (Z[:, :2], Z[:, 3])
with heights <= 1e-5 relative (we hold them to 1e-12)."""
import os
import warnings

import numpy as np
import pytest

import hicembler_oracle as O
from util import hic_like_counts, sam_lines, small_case

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module", autouse=True)
def _need_cuda(cuda):
    # the product path must be the CUDA library: fail loudly if it is missing
    from hicembler_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "libhicembler_b200.so not built"
    _lib.lib()


def _tables(g, ws_kb=(5,), **kw):
    return O.fasta2windows(g.contig_sizes, list(ws_kb), minSize=2000, filterwindows=0, **kw)


def _bam(g, rec):
    return {"records": rec, "names": g.names}


# ------------------------------------------------------------------------------------------
# K1 binning
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kw", [dict(), dict(with_mates=True, bad_flag_frac=0.02, vary_len=True),
                                dict(sort=False)])
def test_bam2arrays_multi_window_bit_exact(kw):
    from hicembler_b200.bam2clusters import bam2arrays
    g, rec = small_case(n_pairs=20000, **kw)
    wsz, windows, chr2window, _, _ = _tables(g, (5, 10, 2))
    want = O.bam2arrays(windows, wsz, chr2window, [sam_lines(g, rec)], 10)
    got = bam2arrays(windows, wsz, chr2window, [_bam(g, rec)], 10, verbose=0)
    assert len(got) == 3
    for a, b in zip(got, want):
        assert a.dtype == np.float32 and np.array_equal(a, b)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("env", [dict(), dict(HC_BIN_MAX_CELLS="700", HC_BIN_ROWS="3"), dict(HC_BIN_CELL16="1"),
                                 dict(HC_BIN_CELL16="1", HC_BIN_MAX_CELLS="1200", HC_BIN_ROWS="3"),
                                 dict(HC_BIN_CELL16="0"), dict(HC_BIN_BAND="1"), dict(HC_BIN_BAND="1", HC_BIN_BAND_B="8"),
                                 dict(HC_BIN_BAND="1", HC_BIN_BAND_B="200")])
def test_bin_pairs_paths_agree(mode, env, monkeypatch):
    """shared-memory tile path (also with several column segments / odd tile heights) and the
    global-atomic path give the same full-orientation counts"""
    from hicembler_b200 import ops
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    g, rec = small_case(n_contigs=40, n_windows=1000, window_bp=5000, n_chrom=4, n_pairs=300000, seed=9)
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    cnt, st = ops.bin_pairs(ops.to_device_records(rec), off, g.window_bp, 1000, 10, mode=mode, return_stats=True)
    full = cnt.cpu().numpy().astype(np.int64)
    want = O.bin_records(rec, g.window_bp, g.window_offsets(), 1000, 10)
    got = np.triu(full + full.T - np.diag(full.diagonal()))
    assert np.array_equal(got, want.astype(np.int64))
    assert st["binned"] == int(want.sum()) and st["fast_path"] == (mode == 0)
    kept = ((rec["flag"] & 3980) == 0) & (rec["mapq"] >= 10)
    assert st["lines"] == int(kept.sum())
    up, _ = ops.fold_counts(cnt, "upper")
    assert np.array_equal(up.cpu().numpy(), want)


@pytest.mark.parametrize("cell16", ["0", "1"])
def test_bin_hot_cells_beyond_16_bits(cell16, monkeypatch):
    """16-bit tile cells hand 16 384 counts at a time to the overflow list: cells far beyond 65 535 stay exact
    (both halves of a shared-memory word, accumulate mode included)."""
    from hicembler_b200 import ops
    from hicembler_b200.pairs import records_from_columns
    monkeypatch.setenv("HC_BIN_CELL16", cell16)
    n_win, ws = 64, 1000
    off = torch.arange(0, n_win, 16, dtype=torch.int32).cuda()            # 4 contigs of 16 windows
    hot = [(0, 2500, 0, 2500, 200001), (0, 2500, 0, 3500, 70000), (1, 500, 2, 9500, 131073), (1, 700, 1, 1700, 16384),
           (3, 15100, 3, 15200, 65536)]
    cols = [[], [], [], []]
    for r1, p1, r2, p2, k in hot:
        for c, v in zip(cols, (r1, p1, r2, p2)):
            c.append(np.full(k, v, dtype=np.int64))
    ref, pos, nref, npos = (np.concatenate(c) for c in cols)
    order = np.lexsort((pos, ref))
    n = len(ref)
    rec = records_from_columns(ref[order], pos[order], nref[order], npos[order], np.full(n, 100), np.full(n, 99),
                               np.full(n, 60))
    dev = ops.to_device_records(rec)
    cnt, st = ops.bin_pairs(dev, off, ws, n_win, 10, return_stats=True)
    want = np.zeros((n_win, n_win), dtype=np.int64)
    for r1, p1, r2, p2, k in hot:
        want[r1 * 16 + (p1 + 51) // ws, r2 * 16 + (p2 + 51) // ws] += k
    assert st["fast_path"] and st["binned"] == n
    assert np.array_equal(cnt.cpu().numpy().view(np.uint32).astype(np.int64), want)
    cnt2 = ops.bin_pairs(dev, off, ws, n_win, 10, mode=2, out=cnt)         # accumulate a second "BAM"
    assert np.array_equal(cnt2.cpu().numpy().view(np.uint32).astype(np.int64), 2 * want)


def test_unsorted_input_falls_back_and_stays_exact():
    from hicembler_b200 import ops
    g, rec = small_case(n_contigs=40, n_windows=1000, window_bp=5000, n_chrom=4, n_pairs=400000, seed=3, sort=False)
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    cnt, st = ops.bin_pairs(ops.to_device_records(rec), off, g.window_bp, 1000, 10, return_stats=True)
    up, _ = ops.fold_counts(cnt, "upper")
    assert np.array_equal(up.cpu().numpy(), O.bin_records(rec, g.window_bp, g.window_offsets(), 1000, 10))


def test_unsorted_input_accumulate_mode_counts_once():
    """when the straggler list overflows in accumulate mode nothing may be added twice.  Unsorted
    records are accumulated onto a non-zero matrix: the result must be exactly matrix + counts."""
    from hicembler_b200 import ops
    g, rec = small_case(n_contigs=40, n_windows=1000, window_bp=5000, n_chrom=4, n_pairs=400000, seed=3, sort=False)
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    dev = ops.to_device_records(rec)
    cnt, st = ops.bin_pairs(dev, off, g.window_bp, 1000, 10, return_stats=True)
    assert not st["fast_path"] and st["overflow"] > len(rec) // 16 + 65536      # the list did overflow
    once = cnt.cpu().numpy().astype(np.int64)
    cnt2, st2 = ops.bin_pairs(dev, off, g.window_bp, 1000, 10, mode=2, out=cnt, return_stats=True)
    assert not st2["fast_path"] and st2["binned"] == st["binned"]
    assert np.array_equal(cnt2.cpu().numpy().astype(np.int64), 2 * once)
    up, _ = ops.fold_counts(cnt2, "upper")
    assert np.array_equal(up.cpu().numpy(), 2 * O.bin_records(rec, g.window_bp, g.window_offsets(), 1000, 10))


def test_bam2arrays_two_unsorted_bams_and_mapq_shadowing():
    """two name-sorted (here: shuffled) BAMs through the drop-in; and the reference's re-use of the last line's MAPQ
    as -q for the second BAM (bam2clusters.py:145), opt-in on the product side, default in the oracle"""
    from hicembler_b200.bam2clusters import bam2arrays
    g, rec = small_case(n_contigs=30, n_windows=400, window_bp=5000, n_chrom=3, n_pairs=150000, seed=6, sort=False)
    rec = rec.copy()
    hi = np.flatnonzero(rec["mapq"] >= 10)
    rec["mapq"][hi[::3]] = 30
    rec["mapq"][hi[-1]] = 45
    wsz, windows, chr2window, _, _ = _tables(g, (5,))
    lines = sam_lines(g, rec)
    want_intent = O.bam2arrays(windows, wsz, chr2window, [lines, lines], 10, shadow_mapq=False)[0]
    want_ref = O.bam2arrays(windows, wsz, chr2window, [lines, lines], 10)[0]
    assert want_ref.sum() < want_intent.sum()
    got = bam2arrays(windows, wsz, chr2window, [_bam(g, rec), _bam(g, rec)], 10, verbose=0)[0]
    assert np.array_equal(got, want_intent)
    got = bam2arrays(windows, wsz, chr2window, [_bam(g, rec), _bam(g, rec)], 10, verbose=0, mapq_shadowing=True)[0]
    assert np.array_equal(got, want_ref)


def test_bin_pairs_multi_one_pass_equals_per_size_calls(golden_dir):
    """all window sizes of a call from one read of the records (bam2clusters.py:154-161) == one call per size ==
    the reference-executed golden arrays"""
    from hicembler_b200 import ops
    gld = np.load(os.path.join(golden_dir, "binning.npz"))
    dev = torch.from_numpy(gld["records"]).cuda()
    offs = [torch.from_numpy(gld["offsets%d" % ii]).cuda() for ii in range(3)]
    wbp = [int(w) for w in gld["window_bp"]]
    nws = [gld["array%d" % ii].shape[0] for ii in range(3)]
    outs, st = ops.bin_pairs_multi(dev, offs, wbp, nws, 10)
    for ii in range(3):
        one = ops.bin_pairs(dev, offs[ii], wbp[ii], nws[ii], 10, member=offs[0])
        assert torch.equal(outs[ii], one)
        up, _ = ops.fold_counts(outs[ii], "upper")
        assert np.array_equal(up.cpu().numpy(), gld["array%d" % ii])
    assert st["binned"] == int(gld["array0"].sum())
    outs2, _ = ops.bin_pairs_multi(dev, offs, wbp, nws, 10, outs=[o.clone() for o in outs], accumulate=True)
    assert all(torch.equal(b, 2 * a) for a, b in zip(outs, outs2))
    # a contig of table 0 missing from table 2 -> KeyError like the reference (:155); a short matrix -> IndexError (:161)
    bad = offs[2].clone()
    bad[int(torch.nonzero(offs[2] >= 0)[-1])] = -1
    with pytest.raises(KeyError):
        ops.bin_pairs_multi(dev, [offs[0], offs[1], bad], wbp, nws, 10)
    with pytest.raises(IndexError):
        ops.bin_pairs_multi(dev, offs, wbp, [nws[0], nws[1] - 3, nws[2]], 10)


def test_bin_pairs_deferred_stats():
    from hicembler_b200 import ops
    g, rec = small_case(n_contigs=40, n_windows=1000, window_bp=5000, n_chrom=4, n_pairs=100000, seed=2)
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    dev = ops.to_device_records(rec)
    cnt, st = ops.bin_pairs(dev, off, g.window_bp, 1000, 10, return_stats=True)
    cnt2, later = ops.bin_pairs(dev, off, g.window_bp, 1000, 10, defer_stats=True)
    assert later.read() == st and torch.equal(cnt, cnt2)
    _, later = ops.bin_pairs(dev, off, g.window_bp, 990, 10, defer_stats=True)      # windows past the matrix
    with pytest.raises(IndexError):
        later.read()


@pytest.mark.parametrize("half_band", ["4", "64", "1024"])
def test_bin_band_kernel_edge_cases(half_band, monkeypatch):
    """the TMA-staged band kernel (one row per tile, far hits as global reductions): odd record counts (the 8-byte tail
    of the last TMA copy), rows without records, accumulate mode, unsorted input, a row block of a sharded build"""
    from hicembler_b200 import ops
    monkeypatch.setenv("HC_BIN_BAND", "1")
    monkeypatch.setenv("HC_BIN_BAND_B", half_band)
    g, rec = small_case(n_contigs=40, n_windows=1200, window_bp=5000, n_chrom=4, n_pairs=200001, seed=13)
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    want = O.bin_records(rec, g.window_bp, g.window_offsets(), 1200, 10).astype(np.int64)

    def upper(c):
        full = c.cpu().numpy().astype(np.int64)
        return np.triu(full + full.T - np.diag(full.diagonal()))
    for n in (200001, 200000, 511, 513, 1):
        cnt, st = ops.bin_pairs(ops.to_device_records(rec[:n]), off, g.window_bp, 1200, 10, return_stats=True)
        assert np.array_equal(upper(cnt), O.bin_records(rec[:n], g.window_bp, g.window_offsets(), 1200, 10).astype(np.int64)), n
        assert st["fast_path"]
    dev = ops.to_device_records(rec)
    cnt = ops.bin_pairs(dev, off, g.window_bp, 1200, 10)
    cnt2 = ops.bin_pairs(dev, off, g.window_bp, 1200, 10, mode=2, out=cnt.clone())
    assert torch.equal(cnt2, 2 * cnt)
    blk = ops.bin_pairs(dev, off, g.window_bp, 1200, 10, rows=(256, 640))          # a rank's row block
    assert torch.equal(blk, cnt[256:640])
    g2, rec2 = small_case(n_contigs=40, n_windows=1200, window_bp=5000, n_chrom=4, n_pairs=300000, seed=3, sort=False)
    cnt3, st3 = ops.bin_pairs(ops.to_device_records(rec2), torch.from_numpy(g2.window_offsets().astype(np.int32)).cuda(),
                              g2.window_bp, 1200, 10, return_stats=True)
    assert np.array_equal(upper(cnt3), O.bin_records(rec2, g2.window_bp, g2.window_offsets(), 1200, 10).astype(np.int64))
    assert not st3["fast_path"]


def test_nearly_sorted_input_uses_overflow_list():
    from hicembler_b200 import ops
    g, rec = small_case(n_contigs=40, n_windows=1000, window_bp=5000, n_chrom=4, n_pairs=300000, seed=4)
    rec = rec.copy()
    rng = np.random.default_rng(0)
    idx = rng.choice(len(rec), 500, replace=False)
    rec[idx] = rec[rng.permutation(idx)]       # 500 records out of place
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    cnt, st = ops.bin_pairs(ops.to_device_records(rec), off, g.window_bp, 1000, 10, return_stats=True)
    assert st["fast_path"] and st["overflow"] > 0
    up, _ = ops.fold_counts(cnt, "upper")
    assert np.array_equal(up.cpu().numpy(), O.bin_records(rec, g.window_bp, g.window_offsets(), 1000, 10))


def test_upto_regions_absent_contigs_two_bams():
    from hicembler_b200.bam2clusters import bam2arrays
    g, rec = small_case(n_pairs=3000)
    sub = g.names[2:9]
    wsz, windows, chr2window, _, _ = _tables(g, (5,), contigs=sub)
    lines = sam_lines(g, rec)
    for upto in (0, 1, 700, 10 ** 6):
        want = O.bam2arrays(windows, wsz, chr2window, [lines, lines], 10, upto=upto, regions=sub)[0]
        got = bam2arrays(windows, wsz, chr2window, [_bam(g, rec), _bam(g, rec)], 10, upto=upto, regions=sub, verbose=0)[0]
        assert np.array_equal(got, want), upto


def test_pair16_unpack_and_binning_bit_exact():
    """16-byte wire records -> hc_unpack_pairs16 -> hc_bin_pairs == the 24-byte path == the oracle."""
    from hicembler_b200 import ops, synth
    from hicembler_b200.pairs import pack_pairs16, unpack_pairs16, PAIR_DTYPE
    g, rec = small_case(n_contigs=40, n_windows=400, n_pairs=200000, with_mates=True, bad_flag_frac=0.03,
                        vary_len=True)
    rec = rec.copy()
    rec["l_seq"][::11] = 0
    packed = pack_pairs16(rec)
    dev = torch.from_numpy(packed.view(np.int32).reshape(-1, 4)).cuda()
    got = ops.unpack_pairs16(dev).cpu().numpy().view(PAIR_DTYPE).reshape(-1)
    want = unpack_pairs16(packed)
    assert got.tobytes() == want.tobytes()
    ws = g.window_bp
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    c24 = ops.bin_pairs(synth.records_tensor(rec, "cuda"), off, ws, g.n_windows(), 10)
    c16 = ops.bin_pairs(ops.unpack_pairs16(dev), off, ws, g.n_windows(), 10)
    assert torch.equal(c24, c16)
    assert ops.unpack_pairs16(dev[:0]).shape == (0, 6)


def test_empty_input():
    from hicembler_b200.bam2clusters import bam2arrays
    g, rec = small_case(n_pairs=100)
    wsz, windows, chr2window, _, _ = _tables(g, (5,))
    got = bam2arrays(windows, wsz, chr2window, [_bam(g, rec[:0])], 10, verbose=0)[0]
    assert got.shape == (len(windows[0]),) * 2 and not got.any()


def test_index_and_key_errors():
    from hicembler_b200.bam2clusters import bam2arrays
    g, rec = small_case(n_pairs=200)
    wsz, windows, chr2window, _, _ = _tables(g, (5,))
    bad = rec.copy()
    last = len(g.names) - 1
    bad["refID"][:] = last
    bad["next_refID"][:] = last
    bad["pos"][:] = g.sizes[last] + 10000
    bad["mapq"][:] = 60
    bad["flag"][:] = 99
    with pytest.raises(IndexError):
        bam2arrays(windows, wsz, chr2window, [_bam(g, bad)], 10, verbose=0)
    # contig in chr2window[0] but not in chr2window[1] -> KeyError like bam2clusters.py:155
    wsz2, windows2, chr2window2, _, _ = _tables(g, (5, 10))
    victim = g.names[int(rec["refID"][0])]
    c2w = [dict(chr2window2[0]), {k: v for k, v in chr2window2[1].items() if k != victim}]
    ok = rec.copy()
    ok["mapq"][:] = 60
    ok["flag"][:] = 99
    with pytest.raises(KeyError):
        bam2arrays(windows2, wsz2, c2w, [_bam(g, ok)], 10, verbose=0)


def test_golden_binning(golden_dir):
    from hicembler_b200 import ops, pairs
    gld = np.load(os.path.join(golden_dir, "binning.npz"))
    rec = np.ascontiguousarray(gld["records"]).view(pairs.PAIR_DTYPE).reshape(-1)
    dev = ops.to_device_records(rec)
    member = torch.from_numpy(gld["offsets0"]).cuda()
    for ii, ws in enumerate(gld["window_bp"]):
        a = gld["array%d" % ii]
        off = torch.from_numpy(gld["offsets%d" % ii]).cuda()
        cnt = ops.bin_pairs(dev, off, int(ws), a.shape[0], 10, member=member)
        assert np.array_equal(ops.fold_counts(cnt, "upper")[0].cpu().numpy(), a)
    n_eff, lines = ops.upto_cutoff(dev, 10, 1500)
    a = gld["array0_upto1500"]
    cnt = ops.bin_pairs(dev, member, int(gld["window_bp"][0]), a.shape[0], 10, n_pairs=n_eff)
    assert lines == 1500 and np.array_equal(ops.fold_counts(cnt, "upper")[0].cpu().numpy(), a)


def test_c1_config_counts_bit_exact():
    """BASELINE.json configs[0]: 200 contigs, 20 Mb, 1 M pairs, 10 kb windows"""
    from hicembler_b200 import ops, synth
    cfg = synth.CONFIGS["C1"]
    g = synth.make_genome(cfg["n_contigs"], cfg["n_windows"], cfg["window_bp"], cfg["n_chrom"], seed=1)
    rec_t = synth.make_pairs(g, cfg["n_pairs"], seed=1, device="cuda")
    rec = synth.records_numpy(rec_t)
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    cnt, st = ops.bin_pairs(rec_t, off, g.window_bp, cfg["n_windows"], 10, return_stats=True)
    want = O.bin_records(rec, g.window_bp, g.window_offsets(), cfg["n_windows"], 10)
    assert st["fast_path"]
    assert np.array_equal(ops.fold_counts(cnt, "upper")[0].cpu().numpy(), want)


def test_contig_level_bam2array():
    from hicembler_b200.bam2clusters import bam2array
    from hicembler_b200 import pairs
    g, rec = small_case(n_contigs=10, n_windows=80, window_bp=5000, n_pairs=6000)
    minSize, windows, chr2window, contig2size = O.contigs2windows(g.contig_sizes, minSize=2000)
    lines = sam_lines(g, rec)
    want, want_d = O.bam2array(windows, contig2size, chr2window, [lines], mapq=10, minSize=minSize, threads=3)
    _, tlen, span = pairs.sam_to_records(lines, {n: i for i, n in enumerate(g.names)})
    got, got_d = bam2array(windows, contig2size, chr2window, [dict(records=rec, names=g.names, tlen=tlen, ref_span=span)],
                           mapq=10, minSize=minSize, verbose=0)
    assert got.dtype == np.float64 and np.array_equal(got, want)
    assert set(got_d) == set(want_d)
    for c in want_d:
        assert sorted(got_d[c]) == sorted(want_d[c])


# ------------------------------------------------------------------------------------------
# K2 symmetrise
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 31, 64, 257])
@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_symmetrize_in_place(n, dt):
    from hicembler_b200 import ops
    d = np.random.default_rng(n).random((n, n)).astype(dt)
    want = O.symmetrize(d.copy())
    X = torch.from_numpy(d.copy()).cuda()
    _, vmax = ops.symmetrize_(X)
    assert np.array_equal(X.cpu().numpy(), want) and vmax.item() == want.max()
    half = O.symmetrize_halfdiag(np.triu(d).copy())
    Y = torch.from_numpy(np.triu(d).copy()).cuda()
    ops.symmetrize_(Y)
    assert np.array_equal(Y.cpu().numpy(), half)


def test_fold_counts_sym_plus_one():
    from hicembler_b200 import ops
    rng = np.random.default_rng(2)
    C = rng.integers(0, 50, (130, 130)).astype(np.int32)
    X = torch.from_numpy(C).cuda()
    up = np.triu(C + C.T - np.diag(C.diagonal())).astype(np.float32)
    want = O.symmetrize(up.copy()) + 1
    got, vmax = ops.fold_counts(X, "sym", add=1.0)
    assert np.array_equal(got.cpu().numpy(), want) and vmax.item() == want.max()
    got2, _ = ops.fold_counts(X.clone(), "sym", add=1.0, inplace=True)
    assert np.array_equal(got2.cpu().numpy(), want)


# ------------------------------------------------------------------------------------------
# K3-K5 Sinkhorn-Knopp
# ------------------------------------------------------------------------------------------
def test_sinkhorn_reference_unit_tests():
    """bin/sinkhorn_knopp/tests/test_sinkhorn_knopp.py:9-99 against the CUDA implementation"""
    from hicembler_b200.sinkhorn_knopp import SinkhornKnopp
    sk = SinkhornKnopp()
    assert sk._max_iter == 1000 and sk._epsilon == 1e-3
    for kw in (dict(max_iter='1'), dict(epsilon='1'), dict(max_iter=-1), dict(epsilon=1), dict(epsilon=0)):
        with pytest.raises(AssertionError):
            SinkhornKnopp(**kw)
    for bad in (1, [[1, 2], [1, 2], [1, 2]], [[1, 2], [1, -1]]):
        with pytest.raises(AssertionError):
            SinkhornKnopp().fit(bad)
    sk = SinkhornKnopp(max_iter=5)
    assert sk._stopping_condition is None
    sk.fit(np.eye(2))
    assert sk._stopping_condition == 'epsilon'
    sk.fit(np.array([[.011, .15], [1.71, .1]]))
    assert sk._stopping_condition == 'max_iter' and sk._iterations == 5
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        SinkhornKnopp().fit(np.array([[0., 0.], [1., 0.]]))
        assert w[0].category is UserWarning
    for P, eps in (([[1, 2], [3, 4]], 1e-3), ([[1.4, .2, 4], [3, 4, .7], [.4, 6, 1]], 1e-8)):
        sk = SinkhornKnopp(epsilon=eps)
        Pt = sk.fit(np.asarray(P))
        assert sk._stopping_condition == 'epsilon'
        for s in (Pt.sum(0), Pt.sum(1)):
            assert np.all(s < 1 + eps) and np.all(s > 1 - eps)
    sk = SinkhornKnopp()
    P = np.eye(2)
    assert np.all(sk.fit(P) == P)
    assert sk._D1.ndim == 2 and np.all(sk._D1 == P) and np.all(sk._D2 == P)
    # docstring known answer (sinkhorn_knopp.py:61-69)
    P_ds = SinkhornKnopp().fit([[.011, .15], [1.71, .1]])
    assert np.allclose(P_ds, [[0.06102561, 0.93897439], [0.93809928, 0.06190072]], atol=1e-8)


@pytest.mark.parametrize("name", ["doc", "t2x2", "t3x3", "eye2", "rand64", "rand97_f32"])
def test_sinkhorn_golden(golden_dir, name):
    from hicembler_b200.sinkhorn_knopp import SinkhornKnopp
    gold = np.load(os.path.join(golden_dir, "sinkhorn.npz"))
    sk = SinkhornKnopp(epsilon=float(gold[name + "_eps"]))
    out = sk.fit(gold[name + "_P"])
    assert out.dtype == np.float64
    assert np.allclose(out, gold[name + "_out"], rtol=1e-10, atol=0)
    assert sk._iterations == int(gold[name + "_iter"])
    assert sk._stopping_condition == str(gold[name + "_stop"])


@pytest.mark.parametrize("key", ["norm150_float32", "norm257_float32", "norm120_float64"])
def test_normalize_golden(golden_dir, key):
    from hicembler_b200.bam2clusters import normalize
    gold = np.load(os.path.join(golden_dir, "sinkhorn.npz"))
    d = gold[key + "_in"].copy()
    res = normalize(d)
    assert np.array_equal(d, gold[key + "_mutated"])
    want = gold[key + "_out"]
    assert res.dtype == np.float64 and np.max(np.abs(res / want - 1)) <= 1e-5


def test_normalize_medium_vs_oracle():
    from hicembler_b200.bam2clusters import normalize
    d = hic_like_counts(1203, 21)
    want_sk = O.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
    want = O.normalize(d.copy(), sk=want_sk, materialize=False)
    from hicembler_b200 import ops
    from hicembler_b200.sinkhorn_knopp import SinkhornKnopp
    X = torch.from_numpy(d.copy()).cuda()
    _, vmax = ops.symmetrize_(X, add=1.0)
    sk = SinkhornKnopp(max_iter=1000, epsilon=0.00001)
    got = sk.fit_device(X, float(vmax.item())).cpu().numpy()
    assert sk._iterations == want_sk._iterations and sk._stopping_condition == want_sk._stopping_condition
    assert np.max(np.abs(got / want - 1)) <= 1e-5
    res = ops.sinkhorn_fit(X, float(vmax.item()), 1000, 0.00001)
    assert np.allclose(res.r.cpu().numpy(), want_sk._r, rtol=1e-9) and np.allclose(res.c.cpu().numpy(), want_sk._c, rtol=1e-9)


@pytest.mark.parametrize("n,dtype", [(1203, np.float32), (256, np.float32), (257, np.float32), (1022, np.float32),
                                     (37, np.float64), (771, np.float64)])
def test_sinkhorn_symmetric_pass(n, dtype):
    """upper-triangle passes (hc_sk_sympass) against the general column / row passes and the oracle:
    same iteration count and stopping condition, scaling vectors to rounding (different summation order)"""
    from hicembler_b200 import ops
    d = hic_like_counts(n, 21 + n).astype(dtype)
    X = torch.from_numpy(d.copy()).cuda()
    _, vmax = ops.symmetrize_(X, add=1.0)
    assert torch.equal(X, X.T)
    vm = float(vmax.item())
    gen = ops.sinkhorn_fit(X, vm, 1000, 0.00001)
    sym = ops.sinkhorn_fit(X, vm, 1000, 0.00001, symmetric=True)
    assert sym.iterations == gen.iterations and sym.stopping_condition == gen.stopping_condition
    nb = (n + 255) // 256
    assert abs(sym.matrix_passes - gen.matrix_passes * (nb + 1) / (2.0 * nb)) < 1e-9
    for a, b in ((sym.r, gen.r), (sym.c, gen.c)):
        assert torch.allclose(a, b, rtol=1e-12, atol=0)
    want_sk = O.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
    O.normalize(d.copy(), sk=want_sk, materialize=False)
    assert sym.iterations == want_sk._iterations
    assert np.allclose(sym.r.cpu().numpy(), want_sk._r, rtol=1e-9) and np.allclose(sym.c.cpu().numpy(), want_sk._c, rtol=1e-9)
    # deterministic: the partial sums are combined in a fixed order
    again = ops.sinkhorn_fit(X, vm, 1000, 0.00001, symmetric=True)
    assert torch.equal(again.r, sym.r) and torch.equal(again.c, sym.c)


def test_scaled_division_is_ieee():
    """P = X / vmax with the reciprocal taken once (common.cuh ScaledDivF) must be the IEEE division bit for bit:
    all 2^24 floats of two binades against 3000 divisors, then odd operands through the symmetric pass itself"""
    from hicembler_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(11)
    b = np.concatenate([
        rng.integers(0x3F800000, 0x40000000, 2900, dtype=np.uint32).view(np.float32) * np.float32(2.0) ** rng.integers(-40, 40, 2900),
        np.array([1, 2, 3, 7, 1.9999999, 1.0000001, 41, 50001, 2 ** 32, 4294967040.0, 1e-30, 1e30, 3e-39, 5e-18, 7e17,
                  0.3333333, 16777215, 16777216, 0.99999994, 123456.79], dtype=np.float32)]).astype(np.float32)
    d_b = torch.from_numpy(b).cuda()
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    _lib.check(L.hc_selftest_scaled_division(d_b.data_ptr(), b.size, bad.data_ptr(), None))
    torch.cuda.synchronize()
    assert int(bad.item()) == 0
    # through hc_sk_sympass: y = P e_j is column j of X / vmax exactly
    n = 300
    A = (rng.integers(0, 0x5F000000, (n, n), dtype=np.uint32).view(np.float32)).copy()   # subnormals up to 2^63
    # (finite quotients only: 0 * inf would turn the unit-vector matvec into NaN)
    A[rng.random((n, n)) < 0.1] = 0.0
    A = np.triu(A) + np.triu(A, 1).T
    X = torch.from_numpy(A).cuda()
    nb = L.hc_sk_sym_num_blocks(n)
    part = torch.empty((nb, n), dtype=torch.float64, device="cuda")
    y = torch.empty(n, dtype=torch.float64, device="cuda")
    for vmax in (3.0, 41.0, 1.0e-15, 7.0e29, 123456.79):
        for j in (0, 17, 255, 256, 299):
            v = torch.zeros(n, dtype=torch.float64, device="cuda")
            v[j] = 1.0
            _lib.check(L.hc_sk_sympass(X.data_ptr(), 0, n, vmax, v.data_ptr(), part.data_ptr(), None, None))
            _lib.check(L.hc_sk_symfinish(part.data_ptr(), n, y.data_ptr(), None, None, None))
            with np.errstate(over="ignore", under="ignore"):
                want = (A[:, j] / np.float32(vmax)).astype(np.float64)
            assert np.array_equal(y.cpu().numpy(), want), (vmax, j)


def test_sinkhorn_symmetric_pass_flags():
    from hicembler_b200 import ops
    X = torch.ones((300, 300), dtype=torch.float32, device="cuda")
    X[5, :] = 0
    X[:, 5] = 0
    with pytest.warns(UserWarning):
        ops.sinkhorn_fit(X, 1.0, 10, 1e-3, symmetric=True)
    X = torch.ones((300, 300), dtype=torch.float32, device="cuda")
    X[299, 7] = X[7, 299] = -1.0
    with pytest.raises(AssertionError):
        ops.sinkhorn_fit(X, 1.0, 10, 1e-3, symmetric=True)
    # an already doubly stochastic symmetric matrix: zero iterations
    X = torch.full((512, 512), 1.0 / 512, dtype=torch.float64, device="cuda")
    res = ops.sinkhorn_fit(X, 1.0, 10, 1e-3, symmetric=True)
    assert res.iterations == 0 and res.stopping_condition == "epsilon"


# ------------------------------------------------------------------------------------------
# K6 transform
# ------------------------------------------------------------------------------------------
def _ulp_diff(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return np.max(np.abs(a - b) / np.spacing(np.maximum(np.abs(a), np.abs(b))))


def test_distance_func_and_logmax():
    from hicembler_b200.bam2clusters import distance_func
    from hicembler_b200.array2scaffolds import transform
    d = O.symmetrize(hic_like_counts(300, 5).astype(np.float64))
    want = O.distance_func(d + 1, 0.37, 1.21)
    got = distance_func(d + 1, 0.37, 1.21)
    # CUDA pow() is documented at 2 ulp; the exp(e*log(u)) short form used for |e*log(u)| < 1 stays within 3
    assert got.dtype == np.float64 and _ulp_diff(got, want) <= 3
    assert distance_func(3.0, 0.5, 1.1) == O.distance_func(3.0, 0.5, 1.1)
    d32 = d.astype(np.float32)
    want32 = O.transform_logmax(d32)
    got32 = transform(d32)
    assert got32.dtype == np.float32 and np.max(np.abs(got32 - want32)) <= 4 * np.spacing(np.float32(want32.max()))
    assert _ulp_diff(transform(d), O.transform_logmax(d)) <= 4


@pytest.mark.parametrize("b", [0.05, 0.1, 0.4, 1.21, 7.0])
def test_distance_func_small_and_large_b(b):
    """the series fast path must not be taken when |z| = |log1p(y)/b| is not small (b from curve_fit can
    be anything); checked on near-1 arguments (the fast-path domain) and on ordinary ones."""
    from hicembler_b200.bam2clusters import distance_func
    rng = np.random.default_rng(4)
    x = np.concatenate([1.0 + rng.random(4000) * 2.0 ** -10, 1.0 + rng.random(4000) * 2.0 ** -30, 1.0 + rng.random(4000) * 50])
    x = x.reshape(120, 100)
    want, got = O.distance_func(x, 0.37, b), distance_func(x, 0.37, b)
    assert got.dtype == np.float64 and _ulp_diff(got, want) <= 4


def test_distance_func_float32_matches_reference_dtype(golden_dir):
    """float32 contacts are evaluated in float32 by the reference (contigs2scaffold path)."""
    from hicembler_b200.bam2clusters import distance_func
    z = np.load(os.path.join(golden_dir, "reference_pinned.npz"))
    x32 = z["dist_x"].astype(np.float32) + np.float32(1)
    want = z["distance_func_f32"]                     # the reference's own distance_func on float32 input
    assert want.dtype == np.float32 and np.array_equal(O.distance_func(x32, 0.5, 1.1), want)
    got = distance_func(x32, 0.5, 1.1)
    assert got.dtype == np.float32
    assert np.max(np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32))) <= 4        # CUDA powf: 2 ulp documented
    # and through the working-matrix kernel (float32 block means -> ward linkage input)
    from hicembler_b200 import ops
    n = x32.shape[0]
    W, _ = ops.transform_to_working(torch.from_numpy(x32).cuda(), ops.KIND_POWER_RAW, 0.5, 1.1)
    w = W.cpu().numpy()[np.triu_indices(n, 1)]
    w32 = w.astype(np.float32)
    assert np.array_equal(w32.astype(np.float64), w)    # float32 values widened
    assert np.max(np.abs(w32.view(np.int32).astype(np.int64) - want[np.triu_indices(n, 1)].view(np.int32))) <= 4


def test_condense_and_working_matrix():
    from hicembler_b200 import ops
    d = np.random.default_rng(8).random((77, 77))
    X = torch.from_numpy(d).cuda()
    assert np.array_equal(ops.condense(X).cpu().numpy(), O.condense(d))
    W, flags = ops.transform_to_working(X, ops.KIND_IDENTITY)
    w = W.cpu().numpy()
    up = np.triu(d, 1)
    assert np.array_equal(np.triu(w, 1), up) and np.array_equal(np.tril(w, -1), up.T)
    assert np.all(np.isinf(w.diagonal())) and flags[0].item() == 0
    W2, _ = ops.expand_condensed(ops.condense(X), 77, square=True)
    assert np.array_equal(np.triu(W2.cpu().numpy(), 1), up * up)


# ------------------------------------------------------------------------------------------
# K7 linkage
# ------------------------------------------------------------------------------------------
def _same_tree(Z, want, rtol=1e-12):
    assert Z.shape == want.shape
    assert np.array_equal(Z[:, [0, 1, 3]], want[:, [0, 1, 3]])
    assert np.allclose(Z[:, 2], want[:, 2], rtol=rtol, atol=0)


@pytest.mark.parametrize("method", ["ward", "average", "complete", "weighted"])
@pytest.mark.parametrize("n", [2, 3, 5, 64, 333, 1500])
def test_linkage_vs_oracle_random(n, method):
    from hicembler_b200.fastcluster import linkage
    y = np.random.default_rng(n).random(n * (n - 1) // 2)
    Z = linkage(y, method, engine="exact")
    _same_tree(Z, O.linkage_fastcluster(y, method))
    assert np.array_equal(Z[:, 2], O.linkage_fastcluster(y, method)[:, 2])   # heights bit-identical
    _same_tree(linkage(y, method), O.linkage_fastcluster(y, method))          # default engine ("auto")


@pytest.mark.parametrize("method", ["ward", "average"])
def test_linkage_tie_heavy_and_jittered(golden_dir, method):
    from hicembler_b200.fastcluster import linkage
    g = np.load(os.path.join(golden_dir, "linkage.npz"))
    for n in (40, 200, 777):
        for tag in ("ties", "jitter"):
            want = g["n%d_%s_%s_fc" % (n, tag, method)]
            Z = linkage(g["n%d_%s_y" % (n, tag)], method, engine="exact")
            _same_tree(Z, want)
            assert np.array_equal(Z[:, 2], want[:, 2])
            _same_tree(linkage(g["n%d_%s_y" % (n, tag)], method), want)   # auto: certified or falls back
        # tie-free input: scipy (the real library) agrees as well
        _same_tree(linkage(g["n%d_jitter_y" % n], method), g["n%d_jitter_%s_scipy" % (n, method)])


def test_linkage_scipy_known_answers():
    from scipy.cluster.hierarchy.tests import hierarchy_test_data as T
    from scipy.spatial.distance import pdist
    from hicembler_b200.fastcluster import linkage
    for m in ("complete", "average", "weighted"):
        assert np.allclose(linkage(T.ytdist, m), getattr(T, "linkage_ytdist_" + m), atol=1e-10)
    assert np.allclose(linkage(pdist(T.X), "ward"), T.linkage_X_ward, atol=1e-6)


def test_linkage_errors():
    from hicembler_b200.fastcluster import linkage
    with pytest.raises(ValueError):
        linkage(np.ones(4), "ward")
    with pytest.raises(FloatingPointError):
        linkage(np.array([1.0, np.nan, 2.0]), "average")
    assert linkage(np.zeros(0), "ward").shape == (0, 4)


def test_linkage_hic_distances_tie_heavy_medium():
    """count-derived distances: few distinct values, ~40 % tied heights"""
    from hicembler_b200.bam2clusters import array2Z
    d = O.symmetrize(hic_like_counts(2500, 31, scale=6.0).astype(np.float64))
    dist = O.transform_distance(d, 0.5, 1.1)
    for m in ("ward", "average"):
        want, stats = O.linkage_fastcluster(O.condense(dist), m, return_stats=True)
        Z, st = array2Z(dist, m, return_stats=True, engine="exact")
        _same_tree(Z, want)
        assert st["scans"] == stats[0] and st["scan_active"] == stats[1]   # same chain, step for step
        Z, st = array2Z(dist, m, return_stats=True)   # auto: ties -> not certified -> exact engine
        _same_tree(Z, want)
        assert st["engine"] == "exact" and not st["fallback_from_rounds"]["certified"]


@pytest.mark.parametrize("method", ["ward", "average", "single", "complete", "weighted"])
def test_linkage_nn_cache_on_off_identical(method, monkeypatch):
    """the per-row nearest-neighbour cache must not change a single merge, ties included; with the
    cache on, part of the chain's NN searches are answered without scanning a row"""
    from hicembler_b200.bam2clusters import array2Z
    d = O.symmetrize(hic_like_counts(1800, 7, scale=5.0).astype(np.float64))
    dist = O.transform_distance(d, 0.5, 1.1)
    want, stats = O.linkage_fastcluster(O.condense(dist), method, return_stats=True)
    out = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("HC_LINKAGE_NNCACHE", flag)
        Z, st = array2Z(dist, method, return_stats=True, engine="exact")
        _same_tree(Z, want)
        assert np.array_equal(Z[:, 2], want[:, 2])
        assert st["scans"] == stats[0] and st["scan_active"] == stats[1]
        out[flag] = st
    assert out["0"]["cached_lookups"] == 0 and out["0"]["row_scans"] == stats[0]
    assert out["1"]["cached_lookups"] > 0
    assert out["1"]["row_scans"] + out["1"]["cached_lookups"] == stats[0]


# ---- K7b round-parallel engine ---------------------------------------------------------------
@pytest.mark.parametrize("method", ["ward", "average", "single", "complete", "weighted"])
@pytest.mark.parametrize("n", [2, 3, 4, 17, 257, 1000, 3001])
def test_linkage_rounds_certified_equals_nnchain(n, method):
    """tie-free input: the round-parallel engine must certify and give fastcluster's merge order and sizes,
    heights to rounding (different interleaving of the same Lance-Williams operations)"""
    from hicembler_b200 import ops
    y = np.random.default_rng(1000 + n).random(n * (n - 1) // 2) + 0.05
    W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), n, square=(method == "ward"))
    Z, st = ops.linkage_rounds(W, method, nan_flags=flags)
    assert Z is not None and st["certified"], st
    assert st["merges"] == n - 1
    _same_tree(Z, O.linkage_fastcluster(y, method), rtol=1e-12)


@pytest.mark.parametrize("method", ["ward", "average"])
def test_linkage_rounds_hic_balanced_matrix(method):
    """Hi-C shaped input (Sinkhorn-balanced counts -> distance_func): long phases with one or two
    reciprocal pairs per round, several compactions; must equal the serial chain"""
    from hicembler_b200 import ops
    n = 2600
    d = O.symmetrize(hic_like_counts(n, 11, scale=6.0).astype(np.float64))
    bal = O.normalize(np.triu(d).copy())
    dist = O.distance_func(bal + 1, 0.5, 1.1)
    y = O.condense(dist)
    want = O.linkage_fastcluster(y, method)
    W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), n, square=(method == "ward"))
    Z, st = ops.linkage_rounds(W, method, nan_flags=flags)
    assert Z is not None, st
    assert st["compactions"] >= 2 and st["rounds"] < n
    _same_tree(Z, want, rtol=1e-11)


@pytest.mark.parametrize("env", [{}, {"HC_ROUNDS_NOCOMPACT": "1"}, {"HC_ROUNDS_BATCH": "1"}, {"HC_ROUNDS_BULK_SEGS": "32"},
                                 {"HC_ROUNDS_KTIME": "1"}, {"HC_ROUNDS_TRACE": "1"}, {"HC_ROUNDS_REBUILD": "16"},
                                 {"HC_ROUNDS_REBUILD": "16", "HC_ROUNDS_REBUILD_V": "1"}, {"HC_ROUNDS_REBUILD": "16", "HC_ROUNDS_REBUILD_V": "2"},
                                 {"HC_ROUNDS_THIN": "0"}, {"HC_ROUNDS_COMPACT": "1/2"},
                                 {"HC_ROUNDS_REBUILD": "0"}, {"HC_ROUNDS_PERSIST": "0"}, {"HC_ROUNDS_CHAIN_MAX": "3"},
                                 {"HC_ROUNDS_PERSIST": "0", "HC_ROUNDS_KTIME": "1"}])
def test_linkage_rounds_compaction_and_batching_do_not_change_results(env, monkeypatch):
    from hicembler_b200 import ops
    n = 1200
    y = np.random.default_rng(5).random(n * (n - 1) // 2) + 0.1
    want = O.linkage_fastcluster(y, "average")
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), n)
    Z, st = ops.linkage_rounds(W, "average", nan_flags=flags)
    assert Z is not None
    _same_tree(Z, want, rtol=1e-12)
    assert (sum(st["kernel_us"].values()) > 0) == ("HC_ROUNDS_KTIME" in env)
    assert st["persistent"] == (env.get("HC_ROUNDS_PERSIST") != "0" and "HC_ROUNDS_TRACE" not in env)
    ref = getattr(test_linkage_rounds_compaction_and_batching_do_not_change_results, "_ref", None)
    if ref is None:
        test_linkage_rounds_compaction_and_batching_do_not_change_results._ref = Z
    else:
        assert np.array_equal(Z, ref)    # bit-identical across compaction / batching settings


@pytest.mark.parametrize("method", ["ward", "average", "single", "complete", "weighted"])
@pytest.mark.parametrize("n", [40, 257, 1000, 3001])
def test_linkage_rounds_bulk_rebuild(n, method, monkeypatch):
    """bulk rounds (a fifth or more of the columns merge): merge + compact in one pass (kb_rebuild) must give
    the very same records as kr_light + kr_heavy, and fastcluster's tree"""
    from hicembler_b200 import ops
    y = np.random.default_rng(2000 + n).random(n * (n - 1) // 2) + 0.05
    out = {}
    # version 3 = list-driven TMA kernel (default), 2 = TMA kernel with per-row role tests, 1 = gather kernel
    for flag, version in (("0", "2"), ("16", "2"), ("16", "1"), ("16", "3")):
        monkeypatch.setenv("HC_ROUNDS_REBUILD", flag)
        monkeypatch.setenv("HC_ROUNDS_REBUILD_V", version)
        W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), n, square=(method == "ward"))
        Z, st = ops.linkage_rounds(W, method, nan_flags=flags)
        assert Z is not None and st["certified"], st
        out[flag + version] = (Z, st)
    assert np.array_equal(out["02"][0], out["162"][0]) and np.array_equal(out["02"][0], out["161"][0])
    assert np.array_equal(out["02"][0], out["163"][0])
    out = {"0": out["02"], "16": out["162"]}
    # (which entries a candidate list keeps depends on how a row was cut into segments, so a pair can be found a round
    # earlier or later on another path: the records are identical, the round count may differ by one or two)
    assert abs(out["16"][1]["rounds"] - out["0"][1]["rounds"]) <= 2 and out["16"][1]["compactions"] >= 1
    _same_tree(out["16"][0], O.linkage_fastcluster(y, method), rtol=1e-12)


def test_linkage_rounds_bulk_rebuild_hic(monkeypatch):
    from hicembler_b200 import ops
    n = 2600
    d = O.symmetrize(hic_like_counts(n, 11, scale=6.0).astype(np.float64))
    bal = O.normalize(np.triu(d).copy())
    y = O.condense(O.distance_func(bal + 1, 0.5, 1.1))
    want = O.linkage_fastcluster(y, "ward")
    monkeypatch.setenv("HC_ROUNDS_REBUILD", "64")
    W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), n, square=True)
    Z, st = ops.linkage_rounds(W, "ward", nan_flags=flags)
    assert Z is not None, st
    _same_tree(Z, want, rtol=1e-11)


def test_linkage_rounds_refuses_ties_and_auto_falls_back():
    """count-derived distances are tie-heavy: the engine must say "not certified", never guess"""
    from hicembler_b200 import ops
    from hicembler_b200.fastcluster import linkage
    d = O.symmetrize(hic_like_counts(900, 3, scale=5.0).astype(np.float64))
    y = O.condense(O.transform_distance(d, 0.5, 1.1))
    for method in ("ward", "average", "complete"):
        W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), 900, square=(method == "ward"))
        Z, st = ops.linkage_rounds(W, method, nan_flags=flags)
        assert Z is None and not st["certified"]
        with pytest.raises(ArithmeticError):
            linkage(y, method, engine="rounds")
        want = O.linkage_fastcluster(y, method)
        Za = linkage(y, method)      # auto
        _same_tree(Za, want)
        assert np.array_equal(Za[:, 2], want[:, 2])
    # exact duplicates of the smallest distance between disjoint pairs: heights tie -> not certified
    y = np.random.default_rng(0).random(10 * 9 // 2) + 1.0
    y[0] = y[-1] = 0.5            # d(0,1) == d(8,9)
    W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), 10)
    Z, st = ops.linkage_rounds(W, "average", nan_flags=flags)
    assert Z is None and st["near_tied_heights"] >= 1
    _same_tree(linkage(y, "average"), O.linkage_fastcluster(y, "average"))


def test_linkage_rounds_chain_worst_case():
    """points on a line with growing gaps: one reciprocal pair per round (n-1 rounds) must still be right"""
    from hicembler_b200 import ops
    n = 300
    x = np.cumsum(1.0 + 0.01 * np.arange(n) + 1e-3 * np.random.default_rng(2).random(n))
    from scipy.spatial.distance import pdist
    y = pdist(x[:, None])
    for method in ("single", "average", "ward"):
        W, flags = ops.expand_condensed(torch.from_numpy(y).cuda(), n, square=(method == "ward"))
        Z, st = ops.linkage_rounds(W, method, nan_flags=flags)
        assert Z is not None, st
        _same_tree(Z, O.linkage_fastcluster(y, method), rtol=1e-11)


def test_array2tree_newick_matches_reference_recursion():
    import scipy.cluster.hierarchy as sch
    from hicembler_b200.bam2clusters import array2tree, newick_from_linkage

    def getNewick(node, newick, parentdist, leaf_names):   # bam2clusters.py:360-371
        if node.is_leaf():
            return "%s:%.2f%s" % (leaf_names[node.id], parentdist - node.dist, newick)
        if len(newick) > 0:
            newick = "):%.2f%s" % (parentdist - node.dist, newick)
        else:
            newick = ");"
        newick = getNewick(node.get_left(), newick, node.dist, leaf_names)
        newick = getNewick(node.get_right(), ",%s" % (newick), node.dist, leaf_names)
        return "(%s" % (newick)
    d = O.transform_distance(O.symmetrize(hic_like_counts(90, 3).astype(np.float64)), 0.5, 1.1)
    names = ["c%d" % i for i in range(90)]
    t = array2tree(d, names)
    Z = O.array2Z(d)
    tree = sch.to_tree(Z, False)
    assert t.write() == getNewick(tree, "", tree.dist, names) == newick_from_linkage(Z, names)


def test_cluster_contigs_with_injected_selections():
    from hicembler_b200.bam2clusters import cluster_contigs
    rng = np.random.default_rng(5)
    n, k = 240, 6
    lab = np.repeat(np.arange(k), n // k)
    base = np.where(lab[:, None] == lab[None, :], 30.0, 2.0)
    cnt = rng.poisson(base).astype(np.float64)
    d = O.transform_distance(O.symmetrize(np.triu(cnt)), 0.5, 1.1)
    subset = int(round(n * 0.66))
    sels = [rng.permutation(n)[:subset] for _ in range(30)]
    names = np.array(["c%d" % i for i in range(n)])
    want, want_n, want_Z, want_msl = O.cluster_contigs(d, names, selections=sels)
    got, got_n, got_Z, got_msl = cluster_contigs("", d, names, None, selections=sels, return_details=True)
    assert got_n == want_n == k and got == want
    assert np.array_equal(got_Z, want_Z) and np.array_equal(got_msl, want_msl)


# ------------------------------------------------------------------------------------------
# K8 + per-cluster sequence
# ------------------------------------------------------------------------------------------
def test_average_reduce_and_normalize_rows():
    from hicembler_b200.bam2scaffolds import _average_reduce_2d, normalize_rows
    g, rec = small_case(n_contigs=9, n_windows=70, window_bp=5000, n_pairs=30000)
    wsz, windows, chr2window, _, _ = _tables(g, (5,))
    d = O.symmetrize(O.bam2arrays(windows, wsz, chr2window, [sam_lines(g, rec)], 10)[0])
    keys = np.array([c for c, s, e in windows[0]])
    assert np.array_equal(_average_reduce_2d(d.copy(), keys), O._average_reduce_2d(d.copy(), keys))
    a = d.copy()
    want = O.normalize_rows(d.copy())
    assert np.allclose(normalize_rows(a), want, rtol=2e-7, atol=0)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_normalize_diagional_and_window_size(dt, capsys):
    """bam2clusters.py:51-66 and bam2scaffolds.py:28-41 against the numpy restatement."""
    from hicembler_b200.bam2clusters import normalize_diagional
    from hicembler_b200.bam2scaffolds import normalize_window_size
    n = 150
    d = np.triu(hic_like_counts(n, 7)).astype(dt)
    d[np.arange(0, n, 13), np.arange(0, n, 13)] = 0            # zero-diagonal windows are dropped
    bin_chr = np.array(["c%d" % (i // 10) for i in range(n)])
    ends = np.cumsum(np.where(np.arange(n) % 10 == 9, 700, 1000))
    bin_position = np.stack([ends - np.where(np.arange(n) % 10 == 9, 700, 1000), ends], axis=1)
    d0 = d.copy()
    got, gc, gp = normalize_diagional(d, bin_chr, bin_position)
    want, wc, wp = O.normalize_diagional(d0.copy(), bin_chr, bin_position)
    assert np.array_equal(d, d0)                                # the reference leaves the caller's array alone
    assert got.dtype == want.dtype and got.shape == want.shape == (n - len(range(0, n, 13)),) * 2
    assert np.array_equal(gc, wc) and np.array_equal(gp, wp)
    assert np.allclose(got, want, rtol=4 * np.finfo(dt).eps, atol=0)   # mean/diag, two products: <= 4 ulp
    for ws in (0, 1000):
        a, b = d0.copy(), d0.copy()
        got, _, _ = normalize_window_size(a, bin_chr, bin_position, ws)
        want, _, _ = O.normalize_window_size(b, bin_chr, bin_position, ws)
        assert got is a and np.array_equal(a, b)                # in place, bit exact (one f64 product per cell)
    assert "1000 135" in capsys.readouterr().out                # the reference prints the modal window size (:39)


def test_contigs2tree_sequence():
    """contigs2scaffold up to the tree (bam2scaffolds.py:234-266) for one cluster, ws sweep 5/10/2 kb"""
    from hicembler_b200.bam2scaffolds import contigs2tree
    from hicembler_b200.FastaIndex import FastaIndex
    g, rec = small_case(n_contigs=14, n_windows=150, window_bp=5000, n_chrom=2, n_pairs=60000, seed=12)
    cluster = g.names[:8]
    params = (0.5, 1.1)

    class FI(FastaIndex):
        def __init__(self, sizes):
            self.id2stats = {c: [s] for c, s in sizes}
            self.genomeSize = sum(s for _, s in sizes)
    for ws in (5, 10, 2):
        wsz, windows, chr2window, _, _ = O.fasta2windows(g.contig_sizes, [ws], contigs=set(cluster), filterwindows=0)
        d = O.bam2arrays(windows, wsz, chr2window, [sam_lines(g, rec)], 10, regions=cluster)[0]
        O.symmetrize(d)
        keys = np.array([c for c, s, e in windows[0]])
        red = O._average_reduce_2d(d, keys)
        Z, uniq, dsym, bin_chr = contigs2tree([_bam(g, rec)], FI(g.contig_sizes), [ws], cluster, 10, params)
        assert np.array_equal(dsym.cpu().numpy(), d) and list(uniq) == list(np.unique(keys))
        # the reference evaluates distance_func on the float32 block means in float32 (value-based casting); CUDA
        # powf and glibc powf differ by float32 ulps, so heights agree to float32 precision, merge order exactly
        dist32 = O.distance_func(red + 1, 0.5, 1.1)
        assert dist32.dtype == np.float32
        want = O.linkage_fastcluster(O.condense(dist32).astype(np.float64), "ward")
        _same_tree(Z, want, rtol=1e-5)


def test_pairs_to_dendrogram_c1_end_to_end():
    """configs[0] end to end with Sinkhorn switched on: pairs -> counts -> balance -> distance -> Z"""
    from hicembler_b200 import ops, synth
    from hicembler_b200.pipeline import pairs_to_dendrogram
    cfg = synth.CONFIGS["C1"]
    g = synth.make_genome(cfg["n_contigs"], cfg["n_windows"], cfg["window_bp"], cfg["n_chrom"], seed=2)
    rec_t = synth.make_pairs(g, 300000, seed=2, device="cuda")
    rec = synth.records_numpy(rec_t)
    off = g.window_offsets()
    res = pairs_to_dendrogram(rec_t, torch.from_numpy(off.astype(np.int32)).cuda(), g.window_bp, cfg["n_windows"],
                              mapq=10, params=(0.5, 1.1), method="ward", balance=True, keep=True)
    up = O.bin_records(rec, g.window_bp, off, cfg["n_windows"], 10)
    sk = O.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
    bal = O.normalize(up.copy(), sk=sk, materialize=False)
    assert res["sk_iterations"] == sk._iterations
    r, c = res["r"].cpu().numpy(), res["c"].cpu().numpy()
    assert np.allclose(r, sk._r, rtol=1e-9) and np.allclose(c, sk._c, rtol=1e-9)
    # distances from the GPU's own r, c so that last-bit differences of pow() cannot flip ties
    sym1 = O.symmetrize(up.copy()) + 1
    P = sym1 / sym1.max()
    bal_gpu = (r[:, None] * P.astype(np.float64)) * c[None, :]
    assert np.max(np.abs(bal_gpu / bal - 1)) <= 1e-5
    W = res["working"].cpu().numpy()
    want_d = O.distance_func(bal_gpu + 1, 0.5, 1.1) ** 2
    iu = np.triu_indices(cfg["n_windows"], 1)
    assert np.allclose(W[iu], want_d[iu], rtol=1e-13)
    want = O.linkage_fastcluster(np.sqrt(W[iu]), "ward")
    _same_tree(res["Z"], want, rtol=1e-9)


def test_bam_file_to_dendrogram_through_the_host_entry(tmp_path):
    """BAM on disk -> native decoder (16-byte wire records, pinned) -> host_pairs_to_dendrogram: the same Z as
    the device-resident path from the 24-byte records, and the same counts as the oracle's SAM-text loop."""
    from hicembler_b200 import pairs, synth
    from hicembler_b200.pipeline import pairs_to_dendrogram, host_pairs_to_dendrogram
    g, rec = small_case(n_contigs=20, n_windows=240, window_bp=5000, n_pairs=40000, with_mates=True,
                        bad_flag_frac=0.02, vary_len=True)
    path = str(tmp_path / "hic.bam")
    pairs.write_bam(path, rec, g.names, [int(s) for s in g.sizes])
    names, sizes, wire, _, _ = pairs.read_bam_native(path, wire=16, pinned=True, side_arrays=False)
    assert names == g.names
    off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
    kw = dict(mapq=10, params=(0.5, 1.1), method="ward", balance=True)
    a = host_pairs_to_dendrogram(wire, off, g.window_bp, g.n_windows(), **kw)
    b = pairs_to_dendrogram(synth.records_tensor(rec, "cuda"), off, g.window_bp, g.n_windows(), **kw)
    assert np.array_equal(a["Z"], b["Z"])
    assert a["stats"]["binned"] == b["stats"]["binned"] == int(O.bin_records(rec, g.window_bp, g.window_offsets(),
                                                                             g.n_windows(), 10).sum())
    h24 = torch.from_numpy(rec.view(np.int32).reshape(-1, 6).copy())
    c = host_pairs_to_dendrogram(h24, off, g.window_bp, g.n_windows(), **kw)
    assert np.array_equal(c["Z"], b["Z"])


def test_host_entry_chunked_overlap_equals_one_piece():
    """host_pairs_to_dendrogram bins contig-aligned pieces under the H2D copy: same counts and Z as one piece;
    input that is not sorted by contig, absent contigs and trailing unmapped records are handled."""
    from hicembler_b200 import ops, pairs, synth
    from hicembler_b200.pipeline import pairs_to_dendrogram, host_pairs_to_dendrogram, _chunk_plan
    g, rec = small_case(n_contigs=40, n_windows=300, window_bp=5000, n_pairs=120000)
    rec = rec.copy()
    unm = rec[:50].copy()
    unm["refID"], unm["flag"] = -1, 4 | 1                       # unplaced unmapped reads close a sorted BAM
    rec = np.concatenate([rec, unm])
    off_np = g.window_offsets().astype(np.int32)
    off = torch.from_numpy(off_np).cuda()
    kw = dict(mapq=10, params=(0.5, 1.1), method="average", balance=True)
    want = pairs_to_dendrogram(synth.records_tensor(rec, "cuda"), off, g.window_bp, g.n_windows(), keep=True, **kw)
    for width in (4, 6):
        host = torch.from_numpy(pairs.pack_pairs16(rec).view(np.int32).reshape(-1, 4).copy() if width == 4
                                else rec.view(np.int32).reshape(-1, 6).copy())
        st = {"chunks": 6, "min_chunk_records": 1000}
        plan = _chunk_plan(host, off_np, g.n_windows(), 6)
        assert plan is not None and len(plan) >= 4 and plan[0][0] == 0 and plan[-1][1] == len(rec)
        got = host_pairs_to_dendrogram(host, off, g.window_bp, g.n_windows(), staging=st, keep=True, **kw)
        assert got["stats"]["chunks"] == len(plan) and got["stats"]["binned"] == want["stats"]["binned"]
        assert torch.equal(got["sym"], want["sym"]) and np.array_equal(got["Z"], want["Z"])
    # contigs swapped out of order: the guard sees it and the call bins in one piece
    order = np.argsort(-rec["refID"], kind="stable")
    host = torch.from_numpy(rec[order].view(np.int32).reshape(-1, 6).copy())
    got = host_pairs_to_dendrogram(host, off, g.window_bp, g.n_windows(),
                                   staging={"chunks": 6, "min_chunk_records": 1000}, keep=True, **kw)
    assert "chunks" not in got["stats"] and torch.equal(got["sym"], want["sym"])
    # a window table with absent contigs (offset -1) still tiles the rows
    off2 = off_np.copy()
    off2[5] = off2[17] = -1
    want2 = pairs_to_dendrogram(synth.records_tensor(rec, "cuda"), torch.from_numpy(off2).cuda(), g.window_bp,
                                g.n_windows(), keep=True, **kw)
    host = torch.from_numpy(rec.view(np.int32).reshape(-1, 6).copy())
    got2 = host_pairs_to_dendrogram(host, torch.from_numpy(off2).cuda(), g.window_bp, g.n_windows(),
                                    staging={"chunks": 6, "min_chunk_records": 1000}, keep=True, **kw)
    assert got2["stats"]["skipped_absent"] == want2["stats"]["skipped_absent"] > 0
    assert torch.equal(got2["sym"], want2["sym"])
