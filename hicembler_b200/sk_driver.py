"""Control flow of SinkhornKnopp.fit (reference sinkhorn_knopp.py:114-163), independent of where
the passes run.  ``passes`` supplies the per-rank matrix passes over its row block (CUDA kernels in
ops._CudaSinkhornPasses; a numpy stand-in in the gloo tests); ``allreduce`` sums a tensor over the
ranks of a row-sharded fit (the NCCL all-reduce of the column-sum vector).  Every rank takes the same
decisions because every decision is taken on all-reduced data."""
from __future__ import annotations

import warnings

_SUPPORT = "Matrix P must have total support. See documentation"


def sinkhorn_driver(passes, max_iter, iterations0=0, allreduce=None):
    """Returns (iterations, stopping_condition, warned); the scaling vectors stay in ``passes``."""
    def colpass(use_r, check=False):
        y = passes.colpass(use_r, check)
        if allreduce is not None:
            allreduce(y)

    # r = ones; c = 1 / P^T r                                                        (:120-129)
    colpass(False, check=True)
    # first evaluation of the while condition is on P_eps = copy(P)                  (:137-141)
    outside = passes.raw_outside()
    if allreduce is not None:
        allreduce(outside)
    passes.update_c(test=False)
    # r = 1 / (P c)                                                                  (:130-134)
    passes.rowpass()
    fl = passes.read_flags()
    if fl["negative"]:
        raise AssertionError("P must be non-negative (sinkhorn_knopp.py:110)")
    warned = False
    if fl["zero_col"]:
        warnings.warn(_SUPPORT, UserWarning)
        warned = True
    if fl["zero_row"]:
        warnings.warn(_SUPPORT, UserWarning)
        warned = True
    cond = passes.flag_value(outside)
    iterations = int(iterations0)
    stop = None
    if cond:
        colpass(True)                      # y = P^T r
    first = True
    while cond:
        # c = 1/(P^T r); r = 1/(P c)                                                  (:143-144)
        if first:
            passes.update_c(test=False)
            first = False
        passes.rowpass()
        iterations += 1
        if iterations >= max_iter:                                                   # (:152-154)
            stop = "max_iter"
            break
        # column sums of D1.P.D2 are c * (P^T r): the next column pass IS the test    (:138-141);
        # row sums are r * (P c) = 1 by construction
        colpass(True)
        cond = passes.update_c(test=True)  # swaps in the new c only when the loop continues
    return iterations, (stop or "epsilon"), warned
