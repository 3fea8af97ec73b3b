"""The linkage oracle (oracle/nnchain.c: restated fastcluster NN-chain) against scipy -- the
reference's own call at array2scaffolds.py:163 -- and scipy's published known-answer vectors."""
import os

import numpy as np
import pytest
import scipy.cluster.hierarchy as sch

import hicembler_oracle as O


def test_scipy_known_answers():
    from scipy.cluster.hierarchy.tests import hierarchy_test_data as T
    from scipy.spatial.distance import pdist
    for m in ("single", "complete", "average", "weighted"):
        Z = O.linkage_fastcluster(T.ytdist, m)
        assert np.allclose(Z, getattr(T, "linkage_ytdist_" + m), atol=1e-10)
    Z = O.linkage_fastcluster(pdist(T.X), "ward")
    assert np.allclose(Z, T.linkage_X_ward, atol=1e-6)


@pytest.mark.parametrize("n", [2, 3, 17, 120])
@pytest.mark.parametrize("method", ["complete", "average", "weighted", "ward"])
def test_c_vs_python_restatement_and_scipy(n, method):
    rng = np.random.default_rng(n)
    y = rng.random(n * (n - 1) // 2)
    Z = O.linkage_fastcluster(y, method)
    assert np.array_equal(Z, O.linkage_fastcluster_py(y, method))
    Zs = sch.linkage(y, method)
    assert np.array_equal(Z[:, [0, 1, 3]], Zs[:, [0, 1, 3]])
    assert np.allclose(Z[:, 2], Zs[:, 2], rtol=1e-12)


def test_python_restatement_with_ties():
    rng = np.random.default_rng(1)
    y = rng.integers(1, 5, 60 * 59 // 2).astype(float)
    for m in ("average", "ward", "complete"):
        assert np.array_equal(O.linkage_fastcluster(y, m), O.linkage_fastcluster_py(y, m))


def test_errors():
    with pytest.raises(ValueError):
        O.linkage_fastcluster(np.ones(4), "ward")
    with pytest.raises(FloatingPointError):
        O.linkage_fastcluster(np.array([1.0, np.nan, 2.0]), "average")


def test_golden_linkage(golden_dir):
    g = np.load(os.path.join(golden_dir, "linkage.npz"))
    for n in (40, 200, 777):
        for m in ("average", "ward", "complete", "weighted"):
            y = g["n%d_jitter_y" % n]
            Z = O.linkage_fastcluster(y, m)
            Zs = g["n%d_jitter_%s_scipy" % (n, m)]
            assert np.array_equal(Z[:, [0, 1, 3]], Zs[:, [0, 1, 3]]), (n, m)
            assert np.allclose(Z[:, 2], Zs[:, 2], rtol=1e-12)
            # tie-heavy: heights are library independent, the oracle is deterministic
            yt = g["n%d_ties_y" % n]
            Zt = O.linkage_fastcluster(yt, m)
            assert np.array_equal(Zt, g["n%d_ties_%s_fc" % (n, m)])
            assert np.allclose(np.sort(Zt[:, 2]), np.sort(g["n%d_ties_%s_scipy" % (n, m)][:, 2]), rtol=1e-9)
