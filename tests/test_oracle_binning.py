"""Binning oracle: the literal per-line loop (bam2clusters.py:128-166 restated) against its
vectorised numpy form, the SAM/BAM encoders/decoders the tests feed it with, and the committed
golden fixture.  The reference holds no test or golden vector for this part ("parity unpinned")."""
import os

import numpy as np
import pytest

import hicembler_oracle as O
from hicembler_b200 import pairs, synth
from util import sam_lines, small_case


def _tables(g, ws_kb=(5,), **kw):
    return O.fasta2windows(g.contig_sizes, list(ws_kb), minSize=2000, filterwindows=0, **kw)


def test_get_window_py2_semantics():
    # Py2: int(start + round((end-start)/2)) with integer division
    assert O.get_window("c", 1, 101, 10, {"c": 0}) == (1 + 50) // 10
    assert O.get_window("c", 9990, 1, 10000, {"c": 7}) == 7        # len("*") == 1 -> half 0
    assert O.get_window("c", 9990, 100, 10000, {"c": 7}) == 8      # spills into the next window


def test_literal_loop_vs_vectorised():
    g, rec = small_case(with_mates=True, bad_flag_frac=0.02, vary_len=True)
    wsz, windows, chr2window, _, _ = _tables(g, (5, 10, 2))
    arrays = O.bam2arrays(windows, wsz, chr2window, [sam_lines(g, rec)], 10)
    member = np.array([chr2window[0].get(c, -1) for c in g.names], dtype=np.int64)
    for ii in range(len(wsz)):
        off = np.array([chr2window[ii].get(c, -1) for c in g.names], dtype=np.int64)
        vec = O.bin_records(rec, wsz[ii], off, len(windows[ii]), 10, member=member)
        assert np.array_equal(vec, arrays[ii])
        assert np.array_equal(np.tril(arrays[ii], -1), np.zeros_like(arrays[ii]))
    kept = ((rec["flag"] & 3980) == 0) & (rec["mapq"] >= 10)
    assert arrays[0].sum() == kept.sum()


def test_upto_and_regions_and_absent_contigs():
    g, rec = small_case(n_pairs=3000)
    sub = g.names[2:9]
    wsz, windows, chr2window, _, _ = _tables(g, (5,), contigs=sub)
    lines = sam_lines(g, rec)
    a = O.bam2arrays(windows, wsz, chr2window, [lines], 10, upto=700, regions=sub)[0]
    off = np.array([chr2window[0].get(c, -1) for c in g.names], dtype=np.int64)
    mask = np.array([c in sub for c in g.names], dtype=np.uint8)
    v = O.bin_records(rec, wsz[0], off, len(windows[0]), 10, upto=700, region_mask=mask)
    assert np.array_equal(a, v) and 0 < a.sum() < 700
    # two BAMs share the running line index (enumerate(proc.stdout, i))
    two = O.bam2arrays(windows, wsz, chr2window, [lines, lines], 10, upto=700, regions=sub)[0]
    assert np.array_equal(two, a)


def test_out_of_range_window_raises():
    g, rec = small_case(n_pairs=200)
    wsz, windows, chr2window, _, _ = _tables(g, (5,))
    rec = rec.copy()
    last = len(g.names) - 1
    rec["refID"][:] = last
    rec["next_refID"][:] = last
    rec["pos"][:] = g.sizes[last] + 10000
    rec["mapq"][:] = 60
    rec["flag"][:] = 99
    with pytest.raises(IndexError):
        O.bam2arrays(windows, wsz, chr2window, [sam_lines(g, rec)], 10)


def test_sam_and_bam_round_trip(tmp_path):
    g, rec = small_case(n_pairs=500, with_mates=True, vary_len=True)
    rec2, tlen, span = pairs.sam_to_records(sam_lines(g, rec), {n: i for i, n in enumerate(g.names)})
    assert np.array_equal(rec2, rec) and np.array_equal(span, np.maximum(rec["l_seq"], 1))
    path = str(tmp_path / "x.bam")
    pairs.write_bam(path, rec, g.names, [int(s) for s in g.sizes], tlen=tlen)
    names, sizes, rec3, tlen3, span3 = pairs.read_bam(path)
    assert names == g.names and sizes == [int(s) for s in g.sizes]
    assert np.array_equal(rec3, rec) and np.array_equal(tlen3, tlen) and np.array_equal(span3, span)


@pytest.mark.skipif(not os.path.exists("/root/reference/bin/snap/tests/datatest/datatest.bam"),
                    reason="reference checkout not present (GPU box)")
def test_bam_decoder_against_snap_fixture():
    # known answer from bin/snap/tests/datatest/datatest.sam
    names, sizes, rec, tlen, span = pairs.read_bam("/root/reference/bin/snap/tests/datatest/datatest.bam")
    r = rec[0]
    assert (r["refID"], r["pos"], r["mapq"], r["flag"], r["l_seq"], r["next_pos"]) == (0, 3907, 69, 99, 101, 4441)
    assert tlen[0] == 635


def test_contig_level_matrix():
    g, rec = small_case(n_contigs=10, n_windows=80, window_bp=5000, n_pairs=6000)
    minSize, windows, chr2window, contig2size = O.contigs2windows(g.contig_sizes, minSize=2000)
    arr, c2d = O.bam2array(windows, contig2size, chr2window, [sam_lines(g, rec)], mapq=10, minSize=minSize, threads=3)
    assert arr.dtype == np.float64 and arr.shape == (len(windows),) * 2
    assert np.array_equal(np.triu(arr, 1), np.zeros_like(arr))     # transposed -> lower triangular (:244)
    assert sum(len(v) for v in c2d.values()) > 0


def test_golden_binning(golden_dir):
    gld = np.load(os.path.join(golden_dir, "binning.npz"))
    rec = np.ascontiguousarray(gld["records"]).view(pairs.PAIR_DTYPE).reshape(-1)
    member = gld["offsets0"].astype(np.int64)
    for ii, ws in enumerate(gld["window_bp"]):
        off = gld["offsets%d" % ii].astype(np.int64)
        a = gld["array%d" % ii]
        assert np.array_equal(O.bin_records(rec, int(ws), off, a.shape[0], 10, member=member), a)
    a = gld["array0_upto1500"]
    assert np.array_equal(O.bin_records(rec, int(gld["window_bp"][0]), member, a.shape[0], 10, upto=1500), a)


def test_normalize_window_size_restatement():
    """bam2scaffolds.py:28-41 on a 3-window case worked by hand."""
    d = np.array([[2., 3., 0.], [0., 4., 5.], [0., 0., 6.]])
    pos = np.array([[0, 10], [10, 20], [20, 25]])
    out, _, _ = O.normalize_window_size(d, np.array(["a"] * 3), pos)       # modal size 10
    want = np.array([[2., 3., 0.], [3., 4., 10.], [0., 10., 24.]])
    assert out is d and np.array_equal(out, want)
    d2 = np.array([[1., 1.], [0., 1.]])
    out2, _, _ = O.normalize_window_size(d2, None, np.array([[0, 5], [5, 15]]), windowSize=10)
    assert np.array_equal(out2, np.array([[4., 2.], [2., 1.]]))


def test_pair16_wire_format_keeps_windows_and_filters():
    """hc_pair16_t (include/hicembler_b200.h): the midpoint form bins exactly like the 24-byte records."""
    import torch
    from hicembler_b200.pairs import pack_pairs16, unpack_pairs16, PAIR16_DTYPE, records_from_columns
    from hicembler_b200 import synth
    from util import small_case
    g, rec = small_case(n_pairs=3000, with_mates=True, bad_flag_frac=0.05, vary_len=True)
    rec = rec.copy()
    rec["l_seq"][::7] = 0            # SEQ "*": len == 1 in the reference, 1//2 == 0
    rec["refID"][5] = -1             # RNAME "*"
    rec["next_refID"][9] = -1
    packed = pack_pairs16(rec)
    assert packed.dtype == PAIR16_DTYPE and packed.itemsize == 16
    back = unpack_pairs16(packed)
    for f in ("refID", "next_refID", "flag", "mapq"):
        assert np.array_equal(back[f], rec[f])
    half = np.where(rec["l_seq"] > 0, rec["l_seq"], 1) // 2
    assert np.array_equal(back["pos"], rec["pos"] + half) and np.array_equal(back["next_pos"], rec["next_pos"] + half)
    assert not back["l_seq"].any()
    # same counts through the oracle's record binner
    for ws in (5000, 1700):
        off = g.window_offsets(ws)
        ok = rec["refID"] >= 0
        a = O.bin_records(rec[ok & (rec["next_refID"] >= 0)], ws, off, g.n_windows(ws), 10)
        b = O.bin_records(back[ok & (rec["next_refID"] >= 0)], ws, off, g.n_windows(ws), 10)
        assert np.array_equal(a[0] if isinstance(a, tuple) else a, b[0] if isinstance(b, tuple) else b)
    # the torch packer (used for device-resident synthetic data) writes the same bytes
    t = pack_pairs16(synth.records_tensor(rec)).numpy().view(PAIR16_DTYPE).reshape(-1)
    assert np.array_equal(t, packed)
    big = records_from_columns([1 << 20], [0], [0], [0], [10], [99], [60])
    with pytest.raises(ValueError):
        pack_pairs16(big)
