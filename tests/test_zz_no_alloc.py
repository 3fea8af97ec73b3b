"""hc_segment_mean_2d_static: the allocation-free variant of hc_segment_mean_2d (scratch in a file-scope
``static __device__`` array, include/hicembler_b200.h).  No counterpart in the reference: the check is against the
existing entry point, bit for bit (same kernels, same summation order), on two calls in a row that share the array.

Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

MAX_ELEMS = 1 << 21          # HC_SEGMEAN_STATIC_MAX_ELEMS


def _case(n, n_seg, seed):
    rng = np.random.default_rng(seed)
    cuts = np.sort(rng.choice(np.arange(1, n), n_seg - 1, replace=False))
    begin = np.concatenate([[0], cuts]).astype(np.int32)
    end = np.concatenate([cuts, [n]]).astype(np.int32)
    i, j = np.indices((n, n))
    A = rng.poisson(40.0 / (1.0 + np.abs(i - j)) + 0.3).astype(np.float32)
    return A, begin, end


def _both(L, A, begin, end):
    """(existing entry point with a caller-provided tmp, static entry point) on the same device inputs"""
    from hicembler_b200 import _lib
    n, k = A.shape[0], len(begin)
    dA, db, de = torch.from_numpy(A).cuda(), torch.from_numpy(begin).cuda(), torch.from_numpy(end).cuda()
    tmp = torch.empty((k, n), dtype=torch.float32, device="cuda")
    want = torch.full((k, k), -1.0, dtype=torch.float32, device="cuda")
    got = torch.full((k, k), -2.0, dtype=torch.float32, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    _lib.check(L.hc_segment_mean_2d(p(dA), n, p(db), p(de), k, p(tmp), p(want), stream))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    _lib.check(L.hc_segment_mean_2d_static(p(dA), n, p(db), p(de), k, p(got), stream))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    return want.cpu().numpy(), got.cpu().numpy(), free0, free1


def test_static_scratch_matches_existing_entry_point_twice_in_a_row():
    import hicembler_oracle as O
    from hicembler_b200 import _lib
    L = _lib.lib()
    torch.cuda.set_device(0)
    # first call fills a large part of the shared array, the second (smaller, other segment layout) runs on top of
    # what the first left behind
    frees = []
    for n, k, seed in ((1500, 96, 1), (700, 41, 2)):
        A, begin, end = _case(n, k, seed)
        want, got, free0, free1 = _both(L, A, begin, end)
        assert np.array_equal(got, want)
        keys = np.repeat(np.arange(k), end - begin)
        assert np.array_equal(got, O._average_reduce_2d(A, keys).astype(np.float32))
        frees.append((free0, free1))
    # no device memory taken by the static call (the first call may load the module lazily; the second may not)
    assert frees[1][0] == frees[1][1]


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from hicembler_b200 import _lib
    L = _lib.lib()
    n, k = 4096, MAX_ELEMS // 4096 + 1
    dA = torch.zeros(1, dtype=torch.float32, device="cuda")          # never dereferenced: the size check comes first
    db = torch.zeros(k, dtype=torch.int32, device="cuda")
    out = torch.zeros(1, dtype=torch.float32, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = L.hc_segment_mean_2d_static(p(dA), n, p(db), p(db), k, p(out), None)
    assert rc == -7                                                   # HC_EWORKSPACE
    buf = C.create_string_buffer(512)
    L.hc_last_error(buf, 512)
    assert b"exceeds the static scratch" in buf.value
    # the largest size that fits is accepted
    n, k = 2048, MAX_ELEMS // 2048
    A, begin, end = _case(n, k, 3)
    want, got, _, _ = _both(L, A, begin, end)
    assert np.array_equal(got, want)
