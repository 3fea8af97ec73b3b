"""ctypes binding of libhicembler_b200.so (the C ABI declared in include/hicembler_b200.h).

There is no CPU fallback: if the CUDA library is missing or a call fails, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhicembler_b200.so")

HC_OK, HC_EINVAL, HC_ERANGE, HC_EKEY, HC_ENAN, HC_ENEG, HC_ECUDA, HC_EWORKSPACE, HC_ENOTRI = \
    0, -1, -2, -3, -4, -5, -6, -7, -8
HC_EAMBIGUOUS = 1

_p, _i32, _i64, _u32, _dbl, _flt, _sz = (C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_double,
                                         C.c_float, C.c_size_t)

# name -> (restype, argtypes); every symbol include/hicembler_b200.h declares
SIGNATURES = {
    "hc_version": (C.c_int, []),
    "hc_last_error": (C.c_int, [C.c_char_p, _sz]),
    "hc_set_device": (C.c_int, [C.c_int]),
    "hc_device_info": (C.c_int, [_p, _p, _p, _p]),
    "hc_bin_pairs_workspace_bytes": (_sz, [_i64, _i64]),
    "hc_bin_pairs": (C.c_int, [_p, _i64, _p, _p, _p, _i32, _u32, _i32, _i32, _u32, _i32, _i32, _p,
                               _p, _sz, C.c_int, _p, _p]),
    "hc_bin_pairs_stats": (C.c_int, [_p, _p, _p]),
    "hc_bin_pairs_multi": (C.c_int, [_p, _i64, _i32, _p, _p, _i32, _p, _p, _i32, _u32, _p, C.c_int, _p, _sz, _p, _p]),
    "hc_upto_cutoff": (C.c_int, [_p, _i64, _p, _i32, _i32, _u32, _i64, _p, _sz, _p, _p, _p]),
    "hc_contig_region_filter": (C.c_int, [_p, _i64, _p, _p, _p, _p, _i32, _i32, _i32, _i32, _u32,
                                          _p, _p, _p]),
    "hc_fold_counts": (C.c_int, [_p, _i32, C.c_int, _flt, _p, _p, _p]),
    "hc_symmetrize": (C.c_int, [_p, _i32, C.c_int, _dbl, _p, _p]),
    "hc_contig_normalize": (C.c_int, [_p, _i32, _p, _p, _p]),
    "hc_sk_num_row_chunks": (_i32, [_i32]),
    "hc_sk_colpass": (C.c_int, [_p, C.c_int, _i32, _i32, _dbl, _p, _p, _p, _p]),
    "hc_sk_colfinish": (C.c_int, [_p, _i32, _i32, _p, _p]),
    "hc_sk_sym_num_blocks": (_i32, [_i32]),
    "hc_sk_sympass": (C.c_int, [_p, C.c_int, _i32, _dbl, _p, _p, _p, _p]),
    "hc_sk_symfinish": (C.c_int, [_p, _i32, _p, _p, _p, _p]),
    "hc_selftest_scaled_division": (C.c_int, [_p, _i32, _p, _p]),
    "hc_sk_update_c": (C.c_int, [_p, _p, _i32, _dbl, _p, _p, _p]),
    "hc_sk_outside": (C.c_int, [_p, _i32, _dbl, _p, _p]),
    "hc_sk_rowpass": (C.c_int, [_p, C.c_int, _i32, _i32, _dbl, _p, _p, _p, _p, _p]),
    "hc_sk_apply": (C.c_int, [_p, C.c_int, _i32, _i32, _dbl, _p, _p, _p, C.c_int, _p]),
    "hc_transform_to_working": (C.c_int, [_p, C.c_int, _i32, C.c_int, _dbl, _dbl, _p, _p, _dbl,
                                          C.c_int, _p, _p, _p]),
    "hc_transform_gather": (C.c_int, [_p, _i32, _i32, C.c_int, _i32, C.c_int, _dbl, _dbl, _p, _p, _dbl, C.c_int, _p, _p, _p]),
    "hc_sharded_preload": (C.c_int, []),
    "hc_shared_alloc": (C.c_int, [_sz, _p, _p]),
    "hc_shared_free": (C.c_int, [_p]),
    "hc_shared_open": (C.c_int, [_p, _p]),
    "hc_shared_close": (C.c_int, [_p]),
    "hc_fold_block": (C.c_int, [_p, _i32, _i32, _i32, _i32, _flt, _p, _p, _p]),
    "hc_sk_exchange_bytes": (_sz, [_i32]),
    "hc_sk_fit_block_workspace_bytes": (_sz, [_i32, _i32]),
    "hc_sk_fit_block": (C.c_int, [_p, C.c_int, _i32, _i32, _dbl, _dbl, _i32, _i32, _p, _i32, _i32, _p, _p, _sz, _p, _p, _p, _p]),
    "hc_transform": (C.c_int, [_p, C.c_int, _i64, C.c_int, _dbl, _dbl, _p, C.c_int, _p]),
    "hc_condense": (C.c_int, [_p, C.c_int, _i32, _p, _p]),
    "hc_expand_condensed": (C.c_int, [_p, _i32, C.c_int, _p, _p, _p]),
    "hc_linkage_workspace_bytes": (_sz, [_i32]),
    "hc_linkage": (C.c_int, [_p, _i32, C.c_int, _p, _sz, _p, _p, _p]),
    "hc_linkage_launch": (C.c_int, [_p, _i32, C.c_int, _p, _sz, _p]),
    "hc_linkage_finish": (C.c_int, [_i32, C.c_int, _p, _sz, _p, _p, _p]),
    "hc_linkage_rounds_workspace_bytes": (_sz, [_i32]),
    "hc_linkage_rounds": (C.c_int, [_p, _i32, C.c_int, _p, _sz, _p, _p, _p]),
    "hc_merges_to_linkage": (C.c_int, [_i32, C.c_int, _p, _p, _p, _p, C.c_double, _p]),
    "hc_segment_mean_2d": (C.c_int, [_p, _i32, _p, _p, _i32, _p, _p, _p]),
    "hc_normalize_rows": (C.c_int, [_p, _i32, _p, _sz, _p]),
    "hc_block_sums4": (C.c_int, [_p, _i32, _p, _p, _p, _p, _i32, _p, _p]),
    "hc_segment_mean_2d_static": (C.c_int, [_p, _i32, _p, _p, _i32, _p, _p]),
    "hc_matrix_max": (C.c_int, [_p, C.c_int, _i64, _p, _i32, _p, _p]),
    "hc_gather_elements": (C.c_int, [_p, _i32, _p, _p, _i64, _p, _p]),
    "hc_unpack_pairs16": (C.c_int, [_p, C.c_int64, _p, _p]),
    "hc_bam_open": (C.c_int, [C.c_char_p, _i32, _p, _p, _p]),
    "hc_bam_next": (C.c_int, [_p, _i32, _p, C.c_int64, _p, _p, _p]),
    "hc_bam_close": (C.c_int, [_p]),
    "hc_gather_square": (C.c_int, [_p, C.c_int, _i32, _p, _i32, _p, _p]),
    "hc_diag_normalize": (C.c_int, [_p, C.c_int, _i32, C.c_double, _p, _p]),
    "hc_window_size_normalize": (C.c_int, [_p, C.c_int, _i32, _p, C.c_double, _p]),
}

_lib = None


def lib():
    """Load the shared library (once).  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "hicembler_b200: %s is missing -- build it with `python -m hicembler_b200.build` "
                "(nvcc, sm_100a); there is no CPU fallback" % LIB_PATH)
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        _lib = h
    return _lib


def last_error() -> str:
    buf = C.create_string_buffer(512)
    lib().hc_last_error(buf, 512)
    return buf.value.decode(errors="replace")


def check(rc: int):
    """Map a status code to the exception type the reference would raise."""
    if rc == HC_OK:
        return
    msg = last_error()
    if rc == HC_ERANGE:
        raise IndexError(msg or "window index out of range")
    if rc == HC_EKEY:
        raise KeyError(msg or "contig missing from chr2window")
    if rc == HC_ENAN:
        raise FloatingPointError(msg or "NaN dissimilarity value.")
    if rc == HC_ENEG:
        raise AssertionError(msg or "P must be non-negative")
    if rc == HC_ENOTRI:
        raise ValueError(msg or "The length of the condensed distance matrix must be (k choose 2)")
    if rc == HC_EINVAL:
        raise ValueError(msg or "invalid argument")
    raise RuntimeError("hicembler_b200 error %d: %s" % (rc, msg))
