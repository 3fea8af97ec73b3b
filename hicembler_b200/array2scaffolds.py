"""Drop-ins for the hot-path functions of the reference's ``array2scaffolds.py``: ``transform``
(:31), ``normalize_rows`` (:33-41), the symmetrisation of ``load_matrix`` (:116-119) and the linkage
call of ``array2tree`` (:160-179)."""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .bam2clusters import HCTree, array2Z
from .bam2scaffolds import _average_reduce_2d, normalize_rows  # noqa: F401


def transform(x):
    """``np.log(np.max(x+1)) - np.log(x+1)`` -- array2scaffolds.py:31 (computed in x's dtype)."""
    dev_in = isinstance(x, torch.Tensor)
    X = x if dev_in else torch.from_numpy(np.ascontiguousarray(x)).cuda()
    if X.dtype not in (torch.float32, torch.float64):
        X = X.double()
    vmax2 = float(X.max().item()) + 1.0 if X.dtype == torch.float64 else float(np.float32(X.max().item()) + np.float32(1))
    out = ops.transform(X, ops.KIND_LOGMAX, vmax2)
    return out if dev_in else out.cpu().numpy()


def symmetrize(d):
    """``d += d.T; d -= np.diag(d.diagonal()/2)`` -- array2scaffolds.py:117-118 (in place)."""
    if isinstance(d, torch.Tensor):
        return ops.symmetrize_(d)[0]
    X = torch.from_numpy(np.ascontiguousarray(d)).cuda()
    ops.symmetrize_(X)
    d[...] = X.cpu().numpy()
    return d


def array2tree(d, names, outbase="", method="ward"):
    """Linkage of d[np.triu_indices(n,1)] -- array2scaffolds.py:160-179.  The reference calls
    scipy's linkage here; the GPU chain follows fastcluster's tie-breaking, which gives the same
    tree as scipy whenever merge heights are distinct."""
    Z = array2Z(d, method)
    t = HCTree(Z, names)
    if outbase:
        with open(outbase + ".nw", "w") as out:
            out.write(t.write())
    return t
