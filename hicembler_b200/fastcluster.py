"""Drop-in for ``fastcluster.linkage`` as the reference calls it (bam2clusters.py:385,403,411,430):
``linkage(y, method)`` with a condensed distance vector.  The NN-chain runs on the GPU
(csrc/linkage.cu) and reproduces fastcluster's tie-breaking."""
from __future__ import annotations

import numpy as np
import torch

from . import ops


def linkage(X, method="single", metric="euclidean", preserve_input=True, engine="auto"):
    """Hierarchical clustering of a condensed dissimilarity vector; returns the scipy-style
    linkage array (n-1, 4) float64.  Observation matrices (2-D X) are outside the hot path."""
    if isinstance(X, torch.Tensor):
        y = X.to(torch.float64)
    else:
        X = np.asarray(X)
        if X.ndim != 1:
            raise NotImplementedError("only condensed (1-D) input is on the HiCembler hot path")
        y = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float64))
    NN = y.numel()
    N = int(np.ceil(np.sqrt(NN * 2)))
    if (N * (N - 1) // 2) != NN:
        raise ValueError("The length of the condensed distance matrix must be (k \\choose 2) for "
                         "k data points!")
    if method not in ops.METHODS:
        raise ValueError("unsupported method %r" % (method,))
    if method == "single":
        # fastcluster uses an MST algorithm for single linkage; the NN-chain gives the same
        # dendrogram when heights are distinct, tie order is not reproduced
        pass
    if N < 2:
        return np.empty((0, 4))
    y = y.cuda()
    # engine="auto": round-parallel engine when every decision is certified tie-free, otherwise the serial
    # NN-chain whose scan order reproduces fastcluster's tie-breaking (csrc/linkage_rounds.cu, linkage.cu)
    return ops.linkage_auto(lambda: ops.expand_condensed(y, N, square=(method == "ward")), method, engine=engine)


def linkage_square(D, method="ward", kind=ops.KIND_IDENTITY, p0=0.0, p1=1.0, r=None, c=None, vmax=1.0,
                   return_stats=False, engine="auto"):
    """Device fast path: linkage of ``transform(D)[np.triu_indices(n,1)]`` for a square CUDA matrix
    D without materialising the condensed vector (K6 fused into the working-matrix build)."""
    return ops.linkage_auto(lambda: ops.transform_to_working(D, kind, p0, p1, r, c, vmax, square=(method == "ward")),
                            method, engine=engine, return_stats=return_stats)
