"""Native BAM/BGZF decoder (csrc/bam_decode.cu, host only) against the pure-Python restatement
(pairs.read_bam: layouts of bin/snap/SNAPLib/Bam.h:43-105,322-350) and the SNAP fixture's known answer."""
import os

import numpy as np
import pytest

from hicembler_b200 import pairs
from util import small_case

SNAP_BAM = "/root/reference/bin/snap/tests/datatest/datatest.bam"


def _write(tmp_path, n_pairs, **kw):
    g, rec = small_case(n_contigs=30, n_windows=300, n_pairs=n_pairs, with_mates=True, bad_flag_frac=0.02,
                        vary_len=True, **kw)
    rec = rec.copy()
    rec["l_seq"][::9] = 0
    tlen = (np.arange(len(rec)) % 977 - 400).astype(np.int32)
    path = str(tmp_path / "t.bam")
    pairs.write_bam(path, rec, g.names, [int(s) for s in g.sizes], tlen=tlen)
    return g, rec, tlen, path


@pytest.mark.parametrize("threads", [1, 4])
def test_native_equals_python_restatement(tmp_path, threads):
    g, rec, tlen, path = _write(tmp_path, 20000)          # ~70 BGZF blocks, records straddle block boundaries
    names, sizes, r_py, t_py, s_py = pairs.read_bam(path)
    n2, s2, r_c, t_c, s_c = pairs.read_bam_native(path, threads=threads, chunk=3000)   # several hc_bam_next calls
    assert names == n2 == g.names and sizes == s2
    assert r_c.tobytes() == r_py.tobytes() == rec.tobytes()
    assert np.array_equal(t_c, t_py) and np.array_equal(t_c, tlen) and np.array_equal(s_c, s_py)
    # 16-byte wire form straight from the decoder == pack_pairs16 of the 24-byte records
    _, _, w16, _, _ = pairs.read_bam_native(path, wire=16, threads=threads, side_arrays=False)
    assert w16.tobytes() == pairs.pack_pairs16(rec).tobytes()
    t = pairs.read_bam_native(path, wire=16, pinned=True, side_arrays=False)[2]
    assert tuple(t.shape) == (len(rec), 4) and t.numpy().tobytes() == w16.tobytes()


def test_load_alignments_uses_the_native_decoder(tmp_path):
    g, rec, tlen, path = _write(tmp_path, 500)
    r, tl, sp, names = pairs.load_alignments(path)
    assert r.tobytes() == rec.tobytes() and names == g.names and np.array_equal(tl, tlen)


def test_empty_and_corrupt_files(tmp_path):
    g, rec, tlen, path = _write(tmp_path, 300)
    empty = str(tmp_path / "e.bam")
    pairs.write_bam(empty, rec[:0], g.names, [int(s) for s in g.sizes])
    assert len(pairs.read_bam_native(empty)[2]) == 0
    raw = bytearray(open(path, "rb").read())
    raw[len(raw) // 2] ^= 0x55                                   # flip a byte inside a deflate stream
    bad = str(tmp_path / "bad.bam")
    open(bad, "wb").write(bytes(raw))
    with pytest.raises(ValueError):
        pairs.read_bam_native(bad)
    trunc = str(tmp_path / "trunc.bam")
    open(trunc, "wb").write(bytes(open(path, "rb").read()[:-40]))
    with pytest.raises(ValueError):
        pairs.read_bam_native(trunc)
    with pytest.raises(OSError):
        pairs.read_bam_native(str(tmp_path / "missing.bam"))


@pytest.mark.skipif(not os.path.exists(SNAP_BAM), reason="reference checkout not present (GPU box)")
def test_snap_fixture_known_answer():
    # bin/snap/tests/datatest/datatest.sam, first record
    names, sizes, rec, tlen, span = pairs.read_bam_native(SNAP_BAM)
    r = rec[0]
    assert (r["refID"], r["pos"], r["mapq"], r["flag"], r["l_seq"], r["next_pos"]) == (0, 3907, 69, 99, 101, 4441)
    assert tlen[0] == 635
    _, _, r_py, t_py, s_py = pairs.read_bam(SNAP_BAM)
    assert rec.tobytes() == r_py.tobytes() and np.array_equal(span, s_py) and np.array_equal(tlen, t_py)
