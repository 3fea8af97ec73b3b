"""The C-ABI library loads and exports every symbol include/hicembler_b200.h declares (no compute)."""
import ctypes
import os
import re

from hicembler_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "hicembler_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        from hicembler_b200.build import build
        build()
    h = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(h, n), "missing export " + n
    # and the Python binding covers exactly the declared surface
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_buffer():
    L = _lib.lib()
    assert L.hc_version() == 200
    buf = ctypes.create_string_buffer(64)
    assert L.hc_last_error(buf, 64) == 0


def test_record_layout():
    from hicembler_b200.pairs import PAIR_DTYPE
    assert PAIR_DTYPE.itemsize == 24
    assert [PAIR_DTYPE.fields[k][1] for k in ("refID", "pos", "next_refID", "next_pos", "l_seq", "flag", "mapq")] == \
        [0, 4, 8, 12, 16, 20, 22]
