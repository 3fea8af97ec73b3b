"""Drop-ins for the hot-path helpers of the reference's ``bam2scaffolds.py``: the window table the
binning kernel consumes (fasta2windows, :73-124), the block mean (_average_reduce_2d, :53-71) and
the per-cluster call sequence of contigs2scaffold (:234-266) up to the contig-order linkage."""
from __future__ import annotations

import sys

import numpy as np
import torch

from . import ops
from .FastaIndex import FastaIndex
from .bam2clusters import (array2Z, array2tree, bam2arrays, bam2counts, distance_func, logger,  # noqa: F401
                           normalize, HCTree)


def fasta2windows(fasta, windowSize, verbose, skipShorter=1, minSize=2000, contigs=[], filterwindows=1):
    """Generate windows over chromosomes -- bam2scaffolds.py:73-124 (windowSize in kb)."""
    faidx = fasta if isinstance(fasta, FastaIndex) else FastaIndex(fasta)
    if verbose:
        logger("Parsing FastA file...")
    windowSize = list(windowSize)
    if filterwindows:
        maxchrlen = max(faidx.id2stats[c][0] for c in faidx)
        windowSize = [x for x in windowSize if 1000 * x < maxchrlen and 1000 * x < 0.01 * faidx.genomeSize
                      and 1000 * x > 0.000001 * faidx.genomeSize]
        if verbose:
            logger(" selected %s windows [kb]: %s" % (len(windowSize), str(windowSize)))
    windowSize = [w * 1000 for w in windowSize]
    windows, chr2window = [[] for w in windowSize], [{} for w in windowSize]
    genomeSize = 0
    base2chr = {}
    skipped = []
    for i, c in enumerate(faidx, 1):
        if contigs and c not in contigs:
            continue
        size = faidx.id2stats[c][0]
        if skipShorter and size < min(windowSize) or size < minSize:
            skipped.append(size)
            continue
        for ii in range(len(windowSize)):
            if skipShorter and size < windowSize[ii]:
                continue
            chr2window[ii][c] = len(windows[ii])
            start = 0
            for start in range(0, size, windowSize[ii]):
                windows[ii].append((c, start, start + windowSize[ii]))
            if not skipShorter:
                windows[ii][-1] = (c, start, size)
        base2chr[genomeSize + size // 2] = c
        genomeSize += size
    if verbose:
        logger(' %s bases in %s contigs divided in %s-%s windows. ' % (faidx.genomeSize, len(faidx), len(windows[0]), len(windows[-1])))
        if skipped:
            logger('  %s bases in %s contigs skipped.' % (sum(skipped), len(skipped)))
    return windowSize, windows, chr2window, base2chr, faidx.genomeSize


def _segments(keys):
    """Window runs per key in np.unique(keys) order; every key must cover one contiguous run (true
    for window tables, where a contig's windows are consecutive)."""
    keys = np.asarray(keys)
    uniq = np.unique(keys)
    begin, end = np.empty(len(uniq), np.int32), np.empty(len(uniq), np.int32)
    for k, u in enumerate(uniq):
        idx = np.flatnonzero(keys == u)
        if idx[-1] - idx[0] + 1 != len(idx):
            raise NotImplementedError("windows of key %r are not contiguous" % (u,))
        begin[k], end[k] = idx[0], idx[-1] + 1
    return uniq, begin, end


def _average_reduce_2d(A, keys):
    """Block mean windows x windows -> contigs x contigs -- bam2scaffolds.py:53-71; float32 in/out,
    rows/columns ordered as np.unique(keys)."""
    _, begin, end = _segments(keys)
    dev_in = isinstance(A, torch.Tensor)
    X = A if dev_in else torch.from_numpy(np.ascontiguousarray(A, dtype=np.float32)).cuda()
    out = ops.segment_mean_2d(X.contiguous(), torch.from_numpy(begin).to(X.device), torch.from_numpy(end).to(X.device))
    return out if dev_in else out.cpu().numpy()


def normalize_window_size(d, bin_chr, bin_position, windowSize=0):
    """Return symmetric and normalised matrix by window size -- bam2scaffolds.py:28-41 (mutates ``d``)."""
    from collections import Counter
    logger("Window-size normalisation...")
    X = torch.from_numpy(np.ascontiguousarray(d)).cuda()
    ops.symmetrize_(X)                                            # :32-33
    sizes = np.diff(bin_position, axis=1)                         # :36
    if not windowSize:
        c = Counter(sizes.reshape(len(sizes)).tolist())
        windowSize, occurencies = c.most_common(1)[0]
        print(windowSize, occurencies)                            # :39
    ops.window_size_normalize_(X, torch.from_numpy(sizes.reshape(-1).astype(np.int64)).cuda(), windowSize)  # :40
    d[...] = X.cpu().numpy()
    return d, bin_chr, bin_position


def normalize_rows(a):
    """Normalise rows so the sums among rows are identical -- bam2scaffolds.py:43-51 (in place)."""
    if isinstance(a, torch.Tensor):
        return ops.normalize_rows_(a)
    X = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()
    ops.normalize_rows_(X)
    a[...] = X.cpu().numpy()
    return a


def contigs2tree(bam, fasta, windowSize, contigs, mapq, params, names=None, method="ward"):
    """The matrix part of contigs2scaffold (bam2scaffolds.py:234-266), device resident: window table
    -> binning -> symmetrise -> block mean -> distance transform -> linkage.  Returns
    (Z, contig names in np.unique order, symmetrised window matrix (CUDA float32), bin_chr)."""
    windowSize, windows, chr2window, base2chr, genomeSize = fasta2windows(
        fasta, windowSize, verbose=0, skipShorter=1, contigs=set(contigs), filterwindows=0)
    counts, _ = bam2counts(windows, windowSize, chr2window, bam, mapq, regions=list(contigs), names=names)
    bin_chr = np.array([c for c, s, e in windows[0]])
    if not len(bin_chr):
        return None, [], None, bin_chr
    d, _ = ops.fold_counts(counts[0], "sym", inplace=True)
    uniq = np.unique(bin_chr)
    if len(uniq) == 1:
        return None, uniq, d, bin_chr
    reduced = _average_reduce_2d(d, bin_chr)
    Z = array2Z(reduced, method, transform=("power", params[0], params[1]))
    return Z, uniq, d, bin_chr
