"""Control flow of SinkhornKnopp.fit (reference sinkhorn_knopp.py:114-163), independent of where
the passes run.  ``passes`` supplies the per-rank matrix passes over its row block (CUDA kernels in
ops._CudaSinkhornPasses; a numpy stand-in in the gloo tests); ``allreduce`` sums a tensor over the
ranks of a row-sharded fit (the NCCL all-reduce of the column-sum vector).  Every rank takes the same
decisions because every decision -- the flags for the AssertionError / UserWarning included -- is taken on
all-reduced data.  (On an NVLink node the product path is hc_sk_fit_block, which keeps this loop on the device;
this driver serves the symmetric single-GPU fit and process groups without peer memory.)"""
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
    if allreduce is not None:
        # the flags are rank-local (a negative entry or an empty row sits in ONE row block): all ranks must raise /
        # warn together, or the others would wait in the next all-reduce for ever
        ft = passes.y.new_tensor([float(fl["negative"]), float(fl["zero_col"]), float(fl["zero_row"])])
        allreduce(ft)
        neg, zc, zr = (bool(v) for v in ft.tolist())
        fl = dict(negative=neg, zero_col=zc, zero_row=zr)
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
