"""The Sinkhorn-Knopp oracle against the reference's own known answers: the vendored library's
docstring example (sinkhorn_knopp.py:61-69), its unit tests (tests/test_sinkhorn_knopp.py:9-99,
restated 1:1) and golden outputs produced by the vendored library itself (oracle/make_golden.py)."""
import os
import warnings

import numpy as np
import pytest

import hicembler_oracle as O
from hicembler_oracle import SinkhornKnopp


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "sinkhorn.npz"))


def test_docstring_known_answer():
    sk = SinkhornKnopp()
    P_ds = sk.fit([[.011, .15], [1.71, .1]])
    assert np.allclose(P_ds, [[0.06102561, 0.93897439], [0.93809928, 0.06190072]], atol=1e-8)
    assert np.allclose(P_ds.sum(0), [0.99912489, 1.00087511], atol=1e-8)
    assert np.allclose(P_ds.sum(1), [1., 1.])


def test_init():
    sk = SinkhornKnopp()
    assert sk._max_iter == 1000 and sk._epsilon == 1e-3
    for kw in (dict(max_iter='1'), dict(epsilon='1'), dict(max_iter=-1), dict(epsilon=1), dict(epsilon=0)):
        with pytest.raises(AssertionError):
            SinkhornKnopp(**kw)


def test_fit_inputs():
    sk = SinkhornKnopp()
    for bad in (1, [[1, 2], [1, 2], [1, 2]], [[1, 2], [1, -1]]):
        with pytest.raises(AssertionError):
            sk.fit(bad)


def test_stopping_condition():
    sk = SinkhornKnopp(max_iter=5)
    assert sk._stopping_condition is None
    sk.fit(np.eye(2))
    assert sk._stopping_condition == 'epsilon'
    sk.fit(np.array([[.011, .15], [1.71, .1]]))
    assert sk._stopping_condition == 'max_iter'
    assert sk._iterations == 5


def test_support_warning():
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        SinkhornKnopp().fit(np.array([[0, 0], [1, 0]]))
        assert w[0].category is UserWarning


def test_sinkhorn_knopp_sums():
    for P, eps in (([[1, 2], [3, 4]], 1e-3), ([[1.4, .2, 4], [3, 4, .7], [.4, 6, 1]], 1e-8)):
        sk = SinkhornKnopp(epsilon=eps)
        Pt = sk.fit(np.asarray(P))
        assert sk._stopping_condition == 'epsilon'
        for s in (Pt.sum(0), Pt.sum(1)):
            assert np.all(s < 1 + eps) and np.all(s > 1 - eps)


def test_diagonal_matrices():
    sk = SinkhornKnopp()
    P = np.eye(2)
    assert np.all(sk.fit(P) == P)
    assert sk._D1.ndim == 2 and sk._D2.ndim == 2
    assert np.all(sk._D1 == P) and np.all(sk._D2 == P)


@pytest.mark.parametrize("name", ["doc", "t2x2", "t3x3", "eye2", "rand64", "rand97_f32"])
def test_against_vendored_library(gold, name):
    sk = SinkhornKnopp(epsilon=float(gold[name + "_eps"]))
    out = sk.fit(gold[name + "_P"])
    assert np.array_equal(out, gold[name + "_out"])          # same arithmetic, bit for bit
    assert sk._iterations == int(gold[name + "_iter"])
    assert sk._stopping_condition == str(gold[name + "_stop"])
    # the broadcasting form used for large N is the same product
    sk2 = SinkhornKnopp(epsilon=float(gold[name + "_eps"]))
    assert np.allclose(sk2.fit(gold[name + "_P"], materialize=False), out, rtol=1e-14, atol=0)


@pytest.mark.parametrize("key", ["norm150_float32", "norm257_float32", "norm120_float64"])
def test_normalize_against_vendored_library(gold, key):
    d = gold[key + "_in"].copy()
    res = O.normalize(d)
    assert np.array_equal(d, gold[key + "_mutated"])          # in-place symmetrise + 1 (bam2clusters.py:42-43)
    assert res.dtype == np.float64 and np.array_equal(res, gold[key + "_out"])
