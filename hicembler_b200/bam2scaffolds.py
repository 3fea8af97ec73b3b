"""Drop-ins for the hot-path helpers of the reference's ``bam2scaffolds.py``: the window table the
binning kernel consumes (fasta2windows, :73-124), the block mean (_average_reduce_2d, :53-71), the
per-cluster call sequence of contigs2scaffold (:234-278) -- contig-order linkage, scaffold joining along
the tree with the corner-block sums on the GPU, gap estimation -- and the bam2scaffolds() driver."""
from __future__ import annotations

import os
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
    bin_chr = np.array([c for c, s, e in windows[0]])
    if not len(bin_chr):
        return None, [], None, bin_chr
    counts, _ = bam2counts(windows, windowSize, chr2window, bam, mapq, regions=list(contigs), names=names)
    d, _ = ops.fold_counts(counts[0], "sym", inplace=True)
    uniq = np.unique(bin_chr)
    if len(uniq) == 1:
        return None, uniq, d, bin_chr
    reduced = _average_reduce_2d(d, bin_chr)
    Z = array2Z(reduced, method, transform=("power", params[0], params[1]))
    return Z, uniq, d, bin_chr


# ---------------------------------------------------------------------------------------------
# scaffolding: order and orient the contigs of one cluster along the contig tree
# ---------------------------------------------------------------------------------------------
def get_reversed(scaffold):
    """Return reversed scaffold, updating orientation -- bam2scaffolds.py:126-128."""
    return [(name, not orientation) for name, orientation in reversed(scaffold)]


def get_indices(scaffold, contig2indices):
    """Return list with window indices representing given scaffold -- bam2scaffolds.py:130-140."""
    indices = []
    for name, reverse in scaffold:
        run = contig2indices.get(name)
        if run is None:
            continue
        indices.extend(reversed(run) if reverse else run)
    return indices


def get_contig2indices(bin_chr):
    """Return contig -> window indices -- bam2scaffolds.py:176-181."""
    contig2indices = {}
    for i, c in enumerate(bin_chr):
        contig2indices.setdefault(c, []).append(i)
    return contig2indices


def _device_matrix(d):
    if isinstance(d, torch.Tensor):
        return d if d.dtype == torch.float32 and d.is_contiguous() else d.float().contiguous()
    return torch.from_numpy(np.ascontiguousarray(d, dtype=np.float32)).cuda()


def _plan_join(scaffold1, scaffold2, contig2indices, minWindows):
    """The host half of join_scaffolds (:144-153): the longer scaffold goes first; scaffolds with fewer
    than minWindows windows are dropped.  Returns (scaffold1, scaffold2, job) -- job is None when no
    orientation has to be decided (then scaffold1 is the result)."""
    indices1 = get_indices(scaffold1, contig2indices)
    indices2 = get_indices(scaffold2, contig2indices)
    if len(indices1) < len(indices2):
        scaffold1, indices1, scaffold2, indices2 = scaffold2, indices2, scaffold1, indices1
    if len(indices2) < minWindows:
        return (scaffold1 if len(indices1) >= minWindows else []), None, None
    return scaffold1, scaffold2, (indices1, indices2)


def _oriented(scaffold1, scaffold2, sums):
    """:157-173 -- 0: start-start, 1: start-end, 2: end-start, 3: end-end (first maximum wins, like np.argmax)"""
    orientation = int(np.argmax(sums))
    if orientation == 0:
        return get_reversed(scaffold2) + scaffold1
    if orientation == 1:
        return scaffold2 + scaffold1
    if orientation == 2:
        return scaffold1 + scaffold2
    return scaffold1 + get_reversed(scaffold2)


def join_scaffolds(scaffold1, scaffold2, d, contig2indices, minWindows=3):
    """Join two adjacent scaffolds -- bam2scaffolds.py:142-174.  The four corner sums of the window matrix are
    taken on the GPU (``hc_block_sums4``), nothing N x N is gathered on the host."""
    scaffold1, scaffold2, job = _plan_join(scaffold1, scaffold2, contig2indices, minWindows)
    if job is None:
        return scaffold1
    return _oriented(scaffold1, scaffold2, ops.block_sums4(_device_matrix(d), [job])[0])


def tree2scaffold(t, d, bin_chr, bin_position, minWindows):
    """Scaffold contigs based on contact matrix and tree -- bam2scaffolds.py:183-200.

    ``t`` = the HCTree of ``array2tree`` (or a ``(Z, names)`` pair).  The joins of one tree level are
    independent, so each level is ONE ``hc_block_sums4`` launch and one read-back instead of a Python
    fancy-index gather per node.  Children order is the reference's: the tree parsed from getNewick's
    string lists the right child first (bam2clusters.py:368-370)."""
    Z, names = (t.Z, t.names) if isinstance(t, HCTree) else t
    contig2indices = get_contig2indices(bin_chr)
    n = len(names)
    scaffolds = [[(names[i], 0)] for i in range(n)] + [None] * (n - 1)
    if n == 1:
        return scaffolds[0]
    A = _device_matrix(d)
    left, right = Z[:, 0].astype(np.int64), Z[:, 1].astype(np.int64)
    level = np.zeros(2 * n - 1, dtype=np.int64)
    for k in range(n - 1):
        level[n + k] = 1 + max(level[left[k]], level[right[k]])
    order = np.argsort(level[n:], kind="stable")
    bounds = np.flatnonzero(np.diff(level[n:][order])) + 1
    for nodes in np.split(order, bounds):
        pending, jobs = [], []
        for k in nodes:
            s1, s2, job = _plan_join(scaffolds[right[k]], scaffolds[left[k]], contig2indices, minWindows)
            if job is None:
                scaffolds[n + k] = s1
            else:
                pending.append((k, s1, s2))
                jobs.append(job)
        for (k, s1, s2), sums in zip(pending, ops.block_sums4(A, jobs)):
            scaffolds[n + k] = _oriented(s1, s2, sums)
    return scaffolds[-1]


def scaffold2gaps(d, bin_chr, scaffold, params, fasta, mingap=100, maxgap=1000000, gapfrac=.33):
    """Add gaps to scaffolds -- bam2scaffolds.py:202-232.  ``d`` = contig x contig mean contacts (numpy),
    ``bin_chr`` its row labels; the gap is distance_func(contacts) minus half of either contig (Python-2
    integer halves), clipped to [mingap, min(gapfrac * (size1 + size2), maxgap)]."""
    faidx = fasta if isinstance(fasta, FastaIndex) else FastaIndex(fasta)
    contig2size = {c: stats[0] for c, stats in faidx.id2stats.items()}
    first = {}
    for i, c in enumerate(bin_chr):
        first.setdefault(c, i)
    d = d.cpu().numpy() if isinstance(d, torch.Tensor) else d
    scaffold_with_gaps = []
    contigbp = gapbp = 0
    for (c, o), (c2, o2) in zip(scaffold[:-1], scaffold[1:]):
        contacts = d[first[c], first[c2]]
        with np.errstate(divide="ignore"):
            distance = distance_func(contacts, *params) * 1000
        both = contig2size[c] + contig2size[c2]
        gapSize = distance - contig2size[c] // 2 - contig2size[c2] // 2
        if gapSize < mingap:
            gapSize = mingap
        elif gapSize > gapfrac * both:
            gapSize = gapfrac * both
        if gapSize > maxgap:
            gapSize = maxgap
        gapSize = int(round(gapSize))
        contigbp += contig2size[c]
        gapbp += gapSize
        scaffold_with_gaps.append((c, o, gapSize))
    c, o = scaffold[-1]  # IndexError on an empty scaffold, like :229
    scaffold_with_gaps.append((c, o, 0))
    return scaffold_with_gaps, contigbp, gapbp


def contigs2scaffold(args):
    """Combine contigs into scaffold -- bam2scaffolds.py:234-278; ``args`` = (i, bam, fasta, [windowSize kb],
    contigs, mapq, minWindows, params).  Everything N x N stays on the GPU."""
    i, bam, fasta, windowSize, contigs, mapq, minWindows, params = args
    Z, uniq, d, bin_chr = contigs2tree(bam, fasta, windowSize, contigs, mapq, params)
    if not len(bin_chr):
        return []
    if len(uniq) == 1:
        return [(bin_chr[0], 0, 0)]
    scaffold = tree2scaffold((Z, list(uniq)), d, bin_chr, None, minWindows)
    scaffold, contigbp, gapbp = scaffold2gaps(_average_reduce_2d(d, bin_chr), uniq, scaffold, params, fasta)
    info = " cluster_%s with %s windows for %s out of %s contigs; %s bp in gaps [%2.2f%s]"
    logger(info % (i, d.shape[0], len(uniq), len(contigs), gapbp, 100. * gapbp / contigbp, '%'))
    return scaffold


def get_scaffolds(out, bam, fasta, clusters, mapq, minWindows, threads, verbose, windowSize=10):
    """Return scaffolds -- bam2scaffolds.py:280-303.  The reference fits the distance parameters once
    outside its try block (:287), so a failing fit propagates; the Pool (:294-298) becomes a loop in this
    process (CUDA contexts do not survive fork) over alignments decoded once."""
    from .bam2clusters import Alignments, estimate_distance_parameters
    faidx = fasta if isinstance(fasta, FastaIndex) else FastaIndex(fasta)
    contig2size = {c: stats[0] for c, stats in faidx.id2stats.items()}
    bam = [Alignments.load(b) for b in bam]
    params = estimate_distance_parameters(out, bam, mapq, contig2size, windowSize * 1000)
    return [contigs2scaffold([i, bam, faidx, [windowSize], contigs, mapq, minWindows, params])
            for i, contigs in enumerate(clusters, 1)]


def report_scaffolds(outbase, scaffolds, faidx, w=60):
    """Save scaffolds (FASTA + one line of oriented contig names per scaffold) -- bam2scaffolds.py:305-331."""
    totsize = 0
    fastafn = outbase + ".scaffolds.fa"
    with open(fastafn, "w") as out, open(outbase + ".scaffolds.tab", "w") as outscaffolds:
        for i, scaffold in enumerate(scaffolds, 1):
            if not scaffold:
                continue
            seqs, elements = [], []
            for c, o, gapSize in scaffold:
                seqs.append(faidx.get_sequence(c, reverse=o))
                seqs.append('N' * gapSize)
                elements.append("%s%s" % (c, "-" if o else "+"))
            seq = "".join(seqs)
            seq = "\n".join(seq[s:s + w] for s in range(0, len(seq), w))
            out.write(">scaffold%s %s bp in %s contigs\n%s\n" % (i, len(seq), len(elements), seq))
            outscaffolds.write("%s\n" % "\t".join(elements))
            totsize += len(seq)
    logger(" %s in %s scaffolds reported to %s" % (totsize, len(scaffolds), fastafn))
    return fastafn


def bam2scaffolds(bam, fasta, outdir, minSize, windowSizes, mapq, threads, dpi, upto, minWindows, nchr,
                  verbose, metagenomics=0, contigs=[], selections=None):
    """Report scaffolds based on BAM file -- bam2scaffolds.py:333-366: clusters from the contig-level
    matrix (bam2clusters), then one ordered, oriented and gapped scaffold per cluster and window size."""
    from .bam2clusters import Alignments, bam2clusters
    fasta = fasta if isinstance(fasta, str) else fasta.name
    faidx = FastaIndex(fasta)
    maxwindow = max(stats[0] for stats in faidx.id2stats.values()) / 10e3
    windowSizes = [x for x in windowSizes if x <= maxwindow]
    if not windowSizes:
        info = "Select window sizes (-w) <= %s kb (at least 10x smaller than the longest contig)!\n"
        sys.stderr.write(info % int(maxwindow))
        sys.exit(1)
    logger(" Selected %s window sizes <= %s kb: %s" % (len(windowSizes), int(maxwindow), str(windowSizes)))
    bam = [Alignments.load(b) for b in bam]
    clusters = bam2clusters(bam, fasta, outdir, minSize, mapq, threads, dpi, upto, nchr, verbose, contigs=contigs,
                            selections=selections)
    if metagenomics:
        return clusters
    for windowSize in windowSizes:
        outbase = os.path.join(outdir, "%s.%sk" % ("chr_%s" % nchr if nchr else "auto", windowSize))
        logger("=== %sk windows ===" % windowSize)
        logger("Constructing scaffolds...")
        scaffolds = get_scaffolds(outbase, bam, fasta, clusters, mapq, minWindows, threads, verbose, windowSize)
        if scaffolds:
            logger("Reporting %s scaffolds..." % len(scaffolds))
            report_scaffolds(outbase, scaffolds, faidx)
    logger("Done!")
