"""Host-only tail of both linkage engines (csrc/dendrogram.cu, hc_merges_to_linkage) against a numpy
restatement of fastcluster's generate_SciPy_dendrogram: stable sort by height, union-find labels
(the reference reaches it through fastcluster.linkage, bam2clusters.py:385)."""
import ctypes as C

import numpy as np
import pytest

from hicembler_b200 import _lib


def _reference(n, ward, a, b, d):
    d = np.sqrt(d) if ward else d.copy()
    order = np.argsort(d, kind="stable")
    parent = list(range(2 * n - 1))

    def find(x):
        while parent[x] != x:
            x = parent[x]
        return x
    size = [1] * (2 * n - 1)
    Z = np.empty((n - 1, 4))
    nxt = n
    for row, k in enumerate(order):
        x, y = find(a[k]), find(b[k])
        parent[x] = parent[y] = nxt
        size[nxt] = size[x] + size[y]
        Z[row] = (min(x, y), max(x, y), d[k], size[nxt])
        nxt += 1
    near = int(np.sum(~(np.diff(d[order]) > 1e-9 * np.abs(d[order][1:]))))
    return Z, near


def _random_merges(n, rng, levels):
    """A valid agglomeration: non-decreasing heights drawn from `levels` distinct values (ties)."""
    clusters = [[i] for i in range(n)]
    a, b = [], []
    for _ in range(n - 1):
        i, j = rng.choice(len(clusters), 2, replace=False)
        a.append(int(rng.choice(clusters[i])))
        b.append(int(rng.choice(clusters[j])))
        clusters[i] = clusters[i] + clusters[j]
        clusters.pop(j)
    d = np.sort(rng.integers(1, levels + 1, n - 1)).astype(np.float64)
    return np.array(a, np.int32), np.array(b, np.int32), d


def _call(n, ward, a, b, d):
    L = _lib.lib()
    Z = np.empty((max(n - 1, 0), 4))
    dd = d.copy()
    near = C.c_int64(-1)
    rc = L.hc_merges_to_linkage(n, int(ward), a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p),
                                dd.ctypes.data_as(C.c_void_p), Z.ctypes.data_as(C.c_void_p), 1e-9, C.byref(near))
    return rc, Z, near.value


@pytest.mark.parametrize("n", [2, 3, 50, 1000, 20000])
@pytest.mark.parametrize("ward", [False, True])
def test_matches_numpy_restatement(n, ward):
    rng = np.random.default_rng(n + ward)
    for levels in (3, 10 * n):
        a, b, d = _random_merges(n, rng, levels)
        if levels > n:
            d = d + rng.random(n - 1) * 0.5  # mostly distinct, still sorted enough to stay a valid order
            d.sort()
        rc, Z, near = _call(n, ward, a, b, d)
        want, want_near = _reference(n, ward, a, b, d)
        assert rc == 0
        assert np.array_equal(Z, want)
        assert near == want_near
        assert Z[-1, 3] == n


def test_unsorted_sequence_of_independent_subtrees():
    # two subtrees whose merges interleave out of height order (what a round-parallel run emits)
    a = np.array([0, 2, 0, 4, 0], np.int32)
    b = np.array([1, 3, 2, 5, 4], np.int32)
    d = np.array([5.0, 1.0, 6.0, 2.0, 7.0])
    rc, Z, _ = _call(6, False, a, b, d)
    want, _ = _reference(6, False, a, b, d)
    assert rc == 0 and np.array_equal(Z, want)
    assert Z[:, 2].tolist() == [1.0, 2.0, 5.0, 6.0, 7.0]


def test_negative_zero_and_bad_index():
    a = np.array([0, 0], np.int32)
    b = np.array([1, 2], np.int32)
    rc, Z, _ = _call(3, False, a, b, np.array([0.0, -0.0]))
    assert rc == 0 and Z[:, :2].tolist() == [[0.0, 1.0], [2.0, 3.0]]   # -0.0 == 0.0: sequence order kept
    rc, _, _ = _call(3, False, a, np.array([1, 7], np.int32), np.array([1.0, 2.0]))
    assert rc == _lib.HC_EINVAL


@pytest.mark.parametrize("case", ["near_equal_reversed", "two_levels_shuffled", "tiny_differences"])
def test_sort_fallback_paths(case):
    """heights that agree in their top 33 bits: the insertion pass after the three top radix digits meets many inversions
    and the full six-digit sort takes over; the result is still numpy's stable order"""
    n = 6000
    rng = np.random.default_rng(7)
    a, b, _ = _random_merges(n, rng, 5)
    m = n - 1
    if case == "near_equal_reversed":
        d = 1.0 + np.arange(m)[::-1] * 1e-13          # all inside one bucket of the top digits, reversed
    elif case == "two_levels_shuffled":
        d = np.where(rng.random(m) < 0.5, 2.0, 2.0 + 2e-12) + rng.integers(0, 3, m) * 1e-14
    else:
        d = 5.0 + rng.random(m) * 1e-10
    d = np.ascontiguousarray(d, dtype=np.float64)
    # (heights out of merge order make no dendrogram, but the tail only sorts and labels: compare with the restatement)
    rc, Z, near = _call(n, False, a, b, d)
    want, _ = _reference(n, False, a, b, d)
    assert rc == 0
    assert np.array_equal(Z[:, 2], want[:, 2])
    assert np.array_equal(Z, want)
