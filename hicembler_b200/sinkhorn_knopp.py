"""Drop-in for ``sinkhorn_knopp.SinkhornKnopp`` (reference bin/sinkhorn_knopp/sinkhorn_knopp/
sinkhorn_knopp.py:7-163): same constructor checks, attributes, warnings and return value; the
matvec / rescale passes run as sm_100a kernels (csrc/sinkhorn.cu)."""
from __future__ import annotations

import numpy as np
import torch

from . import ops


class SinkhornKnopp:
    def __init__(self, max_iter=1000, epsilon=1e-3):
        # sinkhorn_knopp.py:74-86 (IntType/FloatType == int/float)
        assert type(max_iter) is int or type(max_iter) is float, \
            "max_iter is not of type int or float: %r" % max_iter
        assert max_iter > 0, "max_iter must be greater than 0: %r" % max_iter
        self._max_iter = int(max_iter)
        assert type(epsilon) is float or type(epsilon) is int, \
            "epsilon is not of type float or int: %r" % epsilon
        assert epsilon > 0 and epsilon < 1, \
            "epsilon must be between 0 and 1 exclusive: %r" % epsilon
        self._epsilon = epsilon
        self._stopping_condition = None
        self._iterations = 0
        self._D1 = np.ones(1)
        self._D2 = np.ones(1)

    def fit_device(self, X, vmax=1.0, out_dtype=torch.float64, materialize=True, allreduce=None, symmetric=False):
        """Device-resident fit: X is a CUDA float32/float64 (row block of a) square matrix, P = X/vmax.
        Returns the balanced matrix as a CUDA tensor (or None when ``materialize`` is False); the
        scaling vectors stay available as ``self._r`` / ``self._c`` (CUDA float64).  ``symmetric=True``: the
        caller knows X == X.T bitwise; the passes then read the upper triangle only."""
        res = ops.sinkhorn_fit(X, vmax, self._max_iter, self._epsilon, self._iterations, allreduce, symmetric)
        self._iterations = res.iterations
        # sticky, as in the reference (:154-157): only the first non-None value survives
        if res.stopping_condition == "max_iter":
            self._stopping_condition = "max_iter"
        if not self._stopping_condition:
            self._stopping_condition = "epsilon"
        self._r, self._c = res.r, res.c
        if not materialize:
            return None
        return ops.sinkhorn_apply(X, res.r, res.c, vmax, out_dtype)

    def fit(self, P):
        P = np.asarray(P)
        # sinkhorn_knopp.py:110-112
        assert P.ndim == 2
        assert P.shape[0] == P.shape[1]
        if P.dtype not in (np.float32, np.float64):
            P = P.astype(np.float64)
        X = torch.from_numpy(np.ascontiguousarray(P)).cuda()
        out = self.fit_device(X)  # raises AssertionError on negative entries (:110)
        # the reference keeps dense diagonal matrices; built lazily from the vectors here
        self._D1 = np.diag(self._r.cpu().numpy())
        self._D2 = np.diag(self._c.cpu().numpy())
        return out.cpu().numpy()
