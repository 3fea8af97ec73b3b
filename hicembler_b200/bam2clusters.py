"""Drop-ins for the hot-path functions of the reference's ``bam2clusters.py`` -- same names,
argument order, return types and error behaviour; the work runs in sm_100a kernels through the
C ABI (``_lib`` / ``ops``).  numpy at the surface like the reference; every function also accepts /
offers device-resident torch tensors so N x N data never has to cross PCIe.

Differences forced by the environment (documented in DESIGN.md):
  * ``bam`` entries are BAM paths (decoded by ``pairs.read_bam``; no samtools needed), SAM text
    paths, PAIR_DTYPE arrays, or dicts {records, names[, tlen, ref_span]} holding numpy arrays or
    CUDA tensors -- instead of paths handed to a ``samtools view`` subprocess (:108-117).
  * ``threads`` is ignored (it already is in the reference's bam2arrays, :128); CUDA contexts are not
    fork-safe, so nothing here uses multiprocessing.
"""
from __future__ import annotations

import gzip
import os
import pickle
import resource
import sys
from datetime import datetime

import numpy as np
import torch

from . import ops
from .FastaIndex import FastaIndex
from .fastcluster import linkage, linkage_square
from .pairs import PAIR_DTYPE, SKIP_FLAG, load_alignments
from .sinkhorn_knopp import SinkhornKnopp


def logger(message, log=sys.stdout):
    """bam2clusters.py:68-71"""
    memory = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024
    log.write("[%s] %s    [memory: %6i Mb]\n" % (datetime.ctime(datetime.now()), message, memory))


# ---------------------------------------------------------------------------------------------
# alignments
# ---------------------------------------------------------------------------------------------
class Alignments:
    """Decoded alignment records of one BAM, resident on the GPU."""

    def __init__(self, records, names, tlen=None, ref_span=None):
        self.records, self.names, self.tlen, self.ref_span = records, list(names), tlen, ref_span
        self.name2id = {n: i for i, n in enumerate(self.names)}

    @classmethod
    def load(cls, src, names=None, device="cuda"):
        if isinstance(src, Alignments):
            return src
        rec, tlen, span, nm = load_alignments(src) if not isinstance(src, torch.Tensor) else (src, None, None, None)
        if isinstance(src, dict):
            nm = src.get("names", nm)
        nm = nm or names
        if nm is None:
            raise ValueError("contig names of the alignments are unknown: pass a BAM/SAM path or a "
                             "dict with 'names'")
        as_dev = lambda a: None if a is None else (a if isinstance(a, torch.Tensor)
                                                   else torch.from_numpy(np.ascontiguousarray(a))).to(device)
        return cls(ops.to_device_records(rec, device), nm, as_dev(tlen), as_dev(span))

    def offsets(self, chr2window):
        """chr2window (name -> first window) as an int32 table indexed by refID, -1 = absent."""
        off = np.full(len(self.names), -1, dtype=np.int32)
        for c, w in chr2window.items():
            i = self.name2id.get(c)
            if i is not None:
                off[i] = w
        return torch.from_numpy(off).to(self.records.device)

    def region_mask(self, regions):
        if not len(regions):
            return None
        m = np.zeros(len(self.names), dtype=np.uint8)
        for r in regions:
            if ":" in r:
                raise NotImplementedError("bam2arrays supports whole-contig regions only")
            i = self.name2id.get(r)
            if i is not None:
                m[i] = 1
        return torch.from_numpy(m).to(self.records.device)


# ---------------------------------------------------------------------------------------------
# window arithmetic / tables (host side, tiny)
# ---------------------------------------------------------------------------------------------
def get_window(chrom, start, length, windowSize, chr2window):
    """bam2clusters.py:119-126 (Python-2 integer division)."""
    end = start + length
    pos = int(start + round((end - start) // 2))
    return pos // windowSize + chr2window[chrom]


def contigs2windows(fasta, minSize=2000, verbose=0, genomeFrac=0.95, contigs=[]):
    """bam2clusters.py:73-106: one window per contig >= max(minSize, N95)."""
    contigs = set(contigs)
    windows, skipped, chr2window = [], [], {}
    faidx = fasta if isinstance(fasta, FastaIndex) else FastaIndex(fasta)
    contig2size = {c: faidx.id2stats[c][0] for c in faidx}
    n90 = faidx.get_N_and_L(genomeFrac)
    if n90 > minSize:
        minSize = n90
        if verbose:
            logger(" updated --minSize to N%s: %s" % (int(100 * genomeFrac), minSize))
    for c in faidx:
        size = contig2size[c]
        if size < minSize:
            skipped.append(size)
            continue
        elif contigs and c not in contigs:
            continue
        chr2window[c] = len(windows)
        windows.append((c, 0, size))
    if verbose:
        logger(' %s bases in %s contigs divided in %s windows. ' % (faidx.genomeSize, len(faidx), len(windows)))
        if skipped:
            logger('  %s bases in %s contigs skipped.' % (sum(skipped), len(skipped)))
    return minSize, windows, chr2window, contig2size


# ---------------------------------------------------------------------------------------------
# K1: binning
# ---------------------------------------------------------------------------------------------
_ONE_PASS_BYTES = 96 << 20      # all matrices of a multi-window-size call fit the 126 MB L2 together


def _last_line_mapq(records, mapq, rmask, n_pairs):
    """MAPQ of the last record among the first n_pairs that passes -q/-F/regions (None if there is none)."""
    n = int(n_pairs)
    step = 1 << 16
    hi = n
    while hi > 0:
        lo = max(0, hi - step)
        blk = records[lo:hi].cpu().numpy()
        flag, q = blk[:, 5] & 0xffff, (blk[:, 5] >> 16) & 0xff
        ok = ((flag & SKIP_FLAG) == 0) & (q >= int(mapq))
        if rmask is not None:
            m = rmask.cpu().numpy().astype(bool)
            ref = blk[:, 0]
            ok &= (ref >= 0) & (ref < len(m)) & m[np.clip(ref, 0, len(m) - 1)]
        idx = np.flatnonzero(ok)
        if len(idx):
            return int(q[idx[-1]])
        hi = lo
    return None


def bam2counts(windows, windowSize, chr2window, bam, mapq, upto=0, regions=[], names=None, mapq_shadowing=False):
    """Device fast path of bam2arrays: list of u32 count matrices (CUDA int32 tensors, FULL
    orientation: row = window of read 1) plus per-window-size stats.  ``ops.fold_counts`` turns
    them into the reference's arrays or straight into the symmetrised matrix."""
    nws = len(windowSize)
    sizes = [len(w) for w in windows]
    counts, stats = [None] * nws, [None] * nws
    i = 1  # the reference's running line index (:132,139): shared by all BAMs
    for _bam in bam:
        al = Alignments.load(_bam, names)
        rmask = al.region_mask(regions)
        n_pairs = al.records.shape[0]
        if upto:
            # enumerate(proc.stdout, i) restarts at the last index used; a line with index > upto breaks
            budget = max(0, upto - i + 1)
            n_eff, lines = ops.upto_cutoff(al.records, mapq, budget, rmask) if budget else (0, 0)
            more = n_eff < n_pairs and ops.upto_cutoff(al.records[n_eff:], mapq, 1, rmask)[1] > 0
            if more:
                i = max(i, upto + 1) if budget else i
            elif lines:
                i = i + lines - 1
            n_pairs = n_eff
        member = al.offsets(chr2window[0])
        offs = [member] + [al.offsets(chr2window[ii]) for ii in range(1, nws)]

        def add_stats(ii, st):
            stats[ii] = st if stats[ii] is None else {k: (stats[ii][k] + v if isinstance(v, int) and not isinstance(v, bool) else v)
                                                     for k, v in st.items()}
        if 1 < nws <= 4 and sum(4 * n * n for n in sizes) <= _ONE_PASS_BYTES:
            # the reference's inner loop (:154-161): every kept line goes into all window sizes -- one read of the
            # records, matrices resident in L2 (the per-cluster 5 / 10 / 2 kb matrices of the scaffolding stage)
            acc = counts[0] is not None
            outs, st = ops.bin_pairs_multi(al.records, offs, windowSize, sizes, mapq, region_mask=rmask, n_pairs=n_pairs,
                                           outs=counts, accumulate=acc)
            for ii in range(nws):
                counts[ii] = outs[ii]
                add_stats(ii, st)
        else:
            for ii in range(nws):
                mode = 2 if counts[ii] is not None else 0
                counts[ii], st = ops.bin_pairs(al.records, offs[ii], windowSize[ii], sizes[ii], mapq, member=member,
                                               region_mask=rmask, n_pairs=n_pairs, out=counts[ii], mode=mode,
                                               return_stats=True)
                add_stats(ii, st)
        if mapq_shadowing and al.records.shape[0]:
            # bam2clusters.py:145 unpacks the line's MAPQ into ``mapq`` itself, so the reference filters the NEXT
            # BAM with -q <MAPQ of the last line it read from this one>
            last = _last_line_mapq(al.records, mapq, rmask, n_pairs)
            if last is not None:
                mapq = last
    for ii in range(nws):
        if counts[ii] is None:
            counts[ii] = torch.zeros((sizes[ii], sizes[ii]), dtype=torch.int32, device="cuda")
    return counts, stats


def bam2arrays(windows, windowSize, chr2window, bam, mapq, upto=0, regions=[], verbose=1, threads=1,
               device_out=False, names=None, mapq_shadowing=False):
    """Return contact matrices based on BAM -- bam2clusters.py:128-166.

    One float32 (N_ii, N_ii) upper-triangular (incl. diagonal) array per window size.  Raises
    IndexError / KeyError where the reference's loop would (:155,161).  ``mapq_shadowing=True`` reproduces the
    reference's accidental re-use of the last line's MAPQ as ``-q`` for every BAM after the first (:145 overwrites the
    ``mapq`` argument); the default keeps the caller's threshold for all BAMs."""
    counts, stats = bam2counts(windows, windowSize, chr2window, bam, mapq, upto, regions, names, mapq_shadowing)
    arrays = []
    for cnt in counts:
        a, _ = ops.fold_counts(cnt, "upper", inplace=True)
        arrays.append(a if device_out else a.cpu().numpy())
    if verbose:
        logger(" %s alignments parsed" % (stats[0]["lines"] if stats and stats[0] else 0))
    return arrays


def bam2array(windows, contig2size, chr2window, bam, mapq=10, upto=0, regions=[], minSize=2000, threads=4,
              verbose=1, nth=5, names=None, device_out=False):
    """Return contig-level contact matrix and insert sizes -- bam2clusters.py:208-246.

    Only alignments overlapping the sampled regions (s+1, s+minSize), s = 0, nth*minSize, ... of the
    clustered contigs are used (:222); the result is the float64, transposed, size-normalised matrix
    of :244 and ``c2dists`` = {contig: [abs(TLEN) or -1]} (:183-189).  ``upto`` only stops the
    reference between (unordered) worker results (:239-240); it is not applied here."""
    K = len(windows)
    if not regions:
        regions = list(chr2window.keys())
    counts = None
    c2dists = {}
    lines = 0
    region_set = set(regions)
    for _bam in bam:
        al = Alignments.load(_bam, names)
        dev = al.records.device
        member = al.offsets({c: w for c, w in chr2window.items()})
        # samtools is queried for the regions of the worker's contigs only
        rmember = al.offsets({c: chr2window[c] for c in chr2window if c in region_set})
        csize = np.zeros(len(al.names), dtype=np.int32)
        for c, s in contig2size.items():
            if c in al.name2id:
                csize[al.name2id[c]] = s
        keep, dist = ops.contig_region_filter(al.records, al.tlen, al.ref_span, torch.from_numpy(csize).to(dev),
                                              rmember, minSize, nth, mapq)
        if int(keep.max().item() if keep.numel() else 0) > 1:
            raise NotImplementedError("alignment spanning several sampled regions")
        sel = keep.bool()
        kept = al.records[sel].contiguous()
        lines += kept.shape[0]
        counts = ops.bin_pairs(kept, member, None, K, 0, member=member, skip_flag=0, out=counts,
                               mode=2 if counts is not None else 0)
        ref = al.records[sel][:, 0].cpu().numpy()
        dv = dist[sel].cpu().numpy()
        order = np.argsort(ref, kind="stable")
        ref, dv = ref[order], dv[order]
        cut = np.flatnonzero(np.diff(ref)) + 1
        for r, chunk in zip(ref[np.r_[0, cut]] if len(ref) else [], np.split(dv, cut)):
            c2dists.setdefault(al.names[r], [])
            c2dists[al.names[r]] += chunk.tolist()
    for c in regions:
        c2dists.setdefault(c, [])
    if verbose:
        logger(" %s alignments parsed" % lines)
    sizes = np.array([len(range(0, contig2size[c] + 1 - minSize, nth * minSize)) for c, s, e in windows],
                     dtype=np.int32)
    if counts is None:
        counts = torch.zeros((K, K), dtype=torch.int32, device="cuda")
    out = torch.empty((K, K), dtype=torch.float64, device=counts.device)
    from . import _lib
    _lib.check(_lib.lib().hc_contig_normalize(ops._ptr(counts), K, ops._ptr(torch.from_numpy(sizes).to(counts.device)),
                                              ops._ptr(out), ops._stream()))
    return (out if device_out else out.cpu().numpy()), c2dists


# ---------------------------------------------------------------------------------------------
# K2-K5: normalisation
# ---------------------------------------------------------------------------------------------
def normalize(d, max_iter=1000, epsilon=0.00001):
    """Return fully balanced matrix -- bam2clusters.py:36-49.

    Like the reference it symmetrises ``d`` and adds the pseudo-count IN PLACE (:42-43), and returns
    a new float64 array D1.(d/vmax).D2."""
    sk = SinkhornKnopp(max_iter=max_iter, epsilon=epsilon)
    if isinstance(d, torch.Tensor):
        _, vmax = ops.symmetrize_(d, add=1.0)
        return sk.fit_device(d, float(vmax.item()), symmetric=True)    # d + d.T is symmetric bitwise
    X = torch.from_numpy(d).cuda() if d.flags.c_contiguous else torch.from_numpy(np.ascontiguousarray(d)).cuda()
    _, vmax = ops.symmetrize_(X, add=1.0)
    d[...] = X.cpu().numpy()
    return sk.fit_device(X, float(vmax.item()), symmetric=True).cpu().numpy()


def normalize_diagional(d, bin_chr, bin_position):
    """Return symmetric and diagonal normalised matrix -- bam2clusters.py:51-66 (the caller's ``d`` is left
    untouched: the reference's fancy indexing copies it first)."""
    logger("Diagonal normalisation...")
    d = np.asarray(d)
    if d.dtype not in (np.float32, np.float64):
        d = d.astype(np.float64)
    indices = d.diagonal() != 0                                   # :55
    keep = torch.from_numpy(np.flatnonzero(indices).astype(np.int32)).cuda()
    X = ops.gather_square(torch.from_numpy(np.ascontiguousarray(d)).cuda(), keep)   # :56-57
    bin_chr = bin_chr[indices]
    bin_position = bin_position[indices, :]
    if X.shape[0]:
        ops.symmetrize_(X)                                        # :61-62 (d += d.T; d -= diag/2: same values)
        mean = np.mean(X.diagonal().cpu().numpy())                # :64, numpy's own summation
        X = ops.diag_normalize(X, mean)                           # :65
    return X.cpu().numpy(), bin_chr, bin_position


def distance_func(x, a, b, maxv=float('inf')):
    """Distance from normalised contact frequency -- bam2clusters.py:248-251."""
    if isinstance(x, torch.Tensor):
        return ops.transform(x, ops.KIND_POWER_RAW, a, b)
    if isinstance(x, np.ndarray) and x.ndim > 0:
        xt = torch.from_numpy(np.ascontiguousarray(x if x.dtype in (np.float32, np.float64)
                                                   else x.astype(np.float64))).cuda()
        return ops.transform(xt, ops.KIND_POWER_RAW, a, b).cpu().numpy()
    return (1. / (a * x)) ** (1. / b)


def contact_func(x, a, b):
    """bam2clusters.py:253-255"""
    return 1. / (a * x ** b)


# ---------------------------------------------------------------------------------------------
# distance-decay fit (host side: a few thousand numbers; the matrices behind it come from K1)
# ---------------------------------------------------------------------------------------------
def get_distances_contacts(bam, mapq, contig2size, windowSize, icontigs=5, upto=1e7, names=None):
    """Return contacts grouped by genomic distance for the longest contigs -- bam2clusters.py:257-288.

    One intra-contig matrix per contig is binned on the GPU (``bam2arrays`` with ``regions=[c]``,
    :273); the per-distance groups are the matrix diagonals.  The reference's double loop (:281-286)
    visits offsets 0..n-2 of the leading (n-1) x (n-1) block, so distance 0 holds the diagonal
    twice (:279 and j == i) and distances are ``k*windowSize // 1000`` (Python-2 integer division)."""
    logger(" Estimating distance parameters...")
    d2c = {0: []}
    longest_contigs = sorted(contig2size, key=lambda c: contig2size[c], reverse=True)
    maxdist = 0.75 * contig2size[longest_contigs[0]] / 1000.
    bam = [Alignments.load(b, names) for b in bam]
    for c in longest_contigs[:icontigs]:
        n = contig2size[c] // windowSize + 1
        d = bam2arrays([range(n)], [windowSize], [{c: 0}], bam, mapq, upto=int(upto), regions=[c], verbose=0)[0]
        d += d.T - np.diag(d.diagonal())
        d2c[0].append(d.diagonal()[:n - 1])
        m = n - 1  # len(positions)
        blk = d[:m, :m]
        for k in range(m):
            # rows i = 0..m-2 only (:281): offset k has m-k cells, the last of them (i = m-1) only for k = 0
            diag = blk.diagonal(k)
            diag = diag[:m - 1] if k == 0 else diag
            if len(diag):
                d2c.setdefault(k * windowSize // 1000, []).append(diag)
    dists = np.array([x for x in sorted(d2c) if x < maxdist])
    contacts = [np.concatenate(d2c[x]) if len(d2c[x]) else np.empty(0, dtype=np.float32) for x in dists]
    return dists, contacts


def _insert_size_profile(contig2size, windowSize, c2dists, limit, contigs):
    """x / yn / dists / contacts of bam2clusters.py:304-330 from the insert sizes bam2array collects
    (abs(TLEN), -1 for mates on another contig).  Bins are visited in ascending order: CPython 2's
    ``Counter.iteritems()`` order for small-int keys once the table outgrows the largest key."""
    edges = np.arange(-windowSize, (limit + 11) * windowSize, windowSize)
    nb = len(edges) + 1
    per_bin = [[] for _ in range(nb)]
    x, yn = [], []
    if contigs:
        contig2size = {c: s for c, s in contig2size.items() if c in contigs}
    for c in sorted(contig2size, key=lambda c: contig2size[c], reverse=True)[:50]:
        v = np.asarray(c2dists[c], dtype=np.int64)
        if not v.max():  # ValueError on an empty list, like max([]) at :308
            continue
        hist = np.bincount(np.searchsorted(edges, v, side="left"), minlength=nb)
        total = int(v.sum())
        for b in np.flatnonzero(hist):
            nc = 1e6 * int(hist[b]) / total
            per_bin[b - 1].append(nc)
            x.append(edges[b - 1] / 1000.)
            yn.append(nc)
    keep = [(e / 1000., cs) for e, cs in zip(edges[1:-1], per_bin[1:-1]) if cs]
    return x, yn, [e for e, _ in keep][:limit], [cs for _, cs in keep][:limit]


def estimate_distance_parameters(outbase, bam=None, mapq=10, contig2size=None, windowSize=2000,
                                 skipfirst=5, icontigs=5, c2dists={}, limit=30, upto=1e7, contigs=[]):
    """Return estimated fit parameters (a, b) of contact_func -- bam2clusters.py:290-358.

    ``bam``, ``contig2size`` & ``windowSize`` or ``c2dists`` have to be provided.  The boxplot / fit
    figures (:326-335,350-356) are left out (no graphics stack on this box); ``outbase +
    ".distance.params"`` is written like :346-348."""
    from scipy.optimize import curve_fit
    x, yn = [], []
    if bam and not c2dists:
        dists, contacts = get_distances_contacts(bam, mapq, contig2size, windowSize, icontigs, upto)
        dists, contacts = dists[:limit], contacts[:limit]
    else:
        logger(" Estimating distance parameters...")
        x, yn, dists, contacts = _insert_size_profile(contig2size, windowSize, c2dists, limit, contigs)
    if not x:
        x, yn = dists, np.array([np.median(c) for c in contacts])
    params, pcov = curve_fit(contact_func, x[skipfirst:], yn[skipfirst:], maxfev=1000)
    if outbase:
        with open(outbase + ".distance.params", "wb") as out:
            pickle.dump(params, out)
    return params


# ---------------------------------------------------------------------------------------------
# K6-K7: tree
# ---------------------------------------------------------------------------------------------
class HCTree:
    """What array2tree returns when ete3 is not installed: the linkage array, leaf names and the
    Newick string the reference builds with sch.to_tree + getNewick (bam2clusters.py:360-371,386)."""

    def __init__(self, Z, names):
        self.Z, self.names = Z, list(names)
        self._newick = None

    def write(self):
        if self._newick is None:
            self._newick = newick_from_linkage(self.Z, self.names)
        return self._newick

    def get_leaf_names(self):
        import re
        return [m for m in re.findall(r"[(,]([^(),:;]+):", self.write())]


def newick_from_linkage(Z, leaf_names):
    """Iterative equivalent of ``getNewick(sch.to_tree(Z), "", tree.dist, names)``
    (bam2clusters.py:360-371): the reference recursion overflows Python's stack on deep trees."""
    n = Z.shape[0] + 1
    if n == 1:
        return "%s:%.2f;" % (leaf_names[0], 0.0)
    left = Z[:, 0].astype(np.int64)
    right = Z[:, 1].astype(np.int64)
    dist = np.concatenate([np.zeros(n), Z[:, 2]])
    root = 2 * n - 2
    out = []
    # emit tokens left-to-right: "(" right-subtree "," left-subtree ")" [":%.2f"]
    stack = [("node", root, None)]
    while stack:
        kind, node, pdist = stack.pop()
        if kind == "text":
            out.append(node)
        elif node < n:
            out.append("%s:%.2f" % (leaf_names[node], pdist - dist[node]))
        else:
            k = node - n
            close = ");" if pdist is None else "):%.2f" % (pdist - dist[node])
            stack.append(("text", close, None))
            stack.append(("node", int(left[k]), dist[node]))
            stack.append(("text", ",", None))
            stack.append(("node", int(right[k]), dist[node]))
            out.append("(")
    return "".join(out)


def array2Z(d, method="ward", transform=None, return_stats=False, engine="auto"):
    """Linkage array of ``d[np.triu_indices(n,1)]`` (bam2clusters.py:385) for a numpy or CUDA matrix.
    ``transform`` = None, ("power", a, b) for distance_func(x+1, a, b) or ("logmax", vmax)."""
    X = d if isinstance(d, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(d)).cuda()
    if X.dtype not in (torch.float32, torch.float64):
        X = X.double()
    kind, p0, p1 = ops.KIND_IDENTITY, 0.0, 1.0
    if transform is not None:
        if transform[0] == "power":
            kind, p0, p1 = ops.KIND_POWER, transform[1], transform[2]
        elif transform[0] == "logmax":
            kind, p0 = ops.KIND_LOGMAX, transform[1]
        else:
            raise ValueError(transform)
    if X.shape[0] < 2:
        return (np.empty((0, 4)), {}) if return_stats else np.empty((0, 4))
    return linkage_square(X.contiguous(), method, kind, p0, p1, return_stats=return_stats, engine=engine)


def array2tree(d, names, outbase="", method="ward"):
    """Return tree representation for array -- bam2clusters.py:382-399 (PDF rendering left out:
    no graphics stack)."""
    Z = array2Z(d, method)
    t = HCTree(Z, names)
    try:
        import ete3
        t = ete3.Tree(t.write())
    except ImportError:
        pass
    if outbase:
        with open(outbase + ".nw", "w") as out:
            out.write(t.write())
    return t


def _dist_rank_world():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except ImportError:
        pass
    return 0, 1


def cluster_contigs(outbase, d, bin_chr, bin_position, threads=4, frac=0.66, iterations=30, maxnumchr=500,
                    seed=0, dpi=100, minchr=3, nchr=0, selections=None, return_details=False):
    """Estimate number of chromosomes and assign contigs to chromosomes -- bam2clusters.py:406-457.

    Like the reference, ``seed`` and ``threads`` are ignored and the sub-samples come from the global
    ``np.random`` (:410,429) unless ``selections`` (list of index arrays) injects them.  The 1 + 30
    ward linkages run on the GPU, the bootstrap ones overlapped on separate streams."""
    import scipy.cluster.hierarchy as sch
    prng = np.random
    D = d if isinstance(d, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(d, dtype=np.float64)).cuda()
    Z = linkage_square(D, "ward")
    mean_step_len = None
    if not nchr:
        n = D.shape[0]
        subset = int(round(n * frac))
        if subset < maxnumchr:
            maxnumchr = subset - 1
        # every draw is made on every rank (the global numpy RNG advances as in the reference, :429); the linkages
        # themselves are independent and are dealt round-robin over the ranks of a torch.distributed job
        sels = [selections[i] if selections is not None else prng.permutation(n)[:subset] for i in range(iterations)]
        rank, world = _dist_rank_world()
        from .parallel import assign_round_robin
        mine = assign_round_robin(iterations, rank, world)
        Ws = []
        for i in mine:
            sel = torch.from_numpy(np.ascontiguousarray(sels[i], dtype=np.int32)).to(D.device)
            sub = ops.gather_square(D, sel)                                  # d[selection, :][:, selection] (:430)
            Ws.append(ops.transform_to_working(sub, ops.KIND_IDENTITY, square=True)[0])
        part = {i: np.diff(iZ[:, 2]) for i, iZ in zip(mine, ops.linkage_batch(Ws, "ward"))}
        if world > 1:
            import torch.distributed as dist
            parts = [None] * world
            dist.all_gather_object(parts, part)
            part = {k: v for p in parts for k, v in p.items()}
        dZ = [part[i] for i in range(iterations)]
        mean_step_len = np.mean(np.array(dZ), 0)
        nchr = int(np.argmax(mean_step_len[-minchr + 1::-1][:maxnumchr]) + minchr)
    clusters = [[] for _ in range(nchr)]
    assignments = sch.fcluster(Z, t=nchr, criterion='maxclust')
    for i, c in zip(assignments - 1, bin_chr):
        clusters[i].append(c)
    if return_details:
        return clusters, nchr, Z, mean_step_len
    return clusters


def bam2clusters(bam, fasta, outdir, minSize=2000, mapq=10, threads=4, dpi=100, upto=0, nchr=0,
                 verbose=1, minchr=3, method="ward", contigs=[], selections=None, normalize_matrix=False):
    """Return clusters computed from the contig-level contact matrix -- bam2clusters.py:459-537.

    Same files as the reference under ``outdir``: ``auto|chr_N`` + ``.windows.tab.gz``, ``.npz``,
    ``.distance.params`` (cache, reused on the next call like :490-506) and ``.clusters.tab``.  The
    matrix stays on the GPU from ``bam2array`` through symmetrisation (:511), the distance transform
    (:516) and the 1 + 30 ward linkages of ``cluster_contigs``; ``selections`` is handed to
    ``cluster_contigs`` (the reference draws from the unseeded global RNG).  Like the reference,
    ``method`` is accepted and unused (cluster_contigs is ward only, :411).  ``normalize_matrix=True`` switches on
    the Sinkhorn-Knopp balancing that the reference carries commented out at :512 (default off = HEAD)."""
    logger("=== Clustering ===")
    outbase = os.path.join(outdir, "chr_%s" % nchr if nchr else "auto")
    # link with auto (:466-469)
    if nchr and os.path.isfile(os.path.join(outdir, "auto.npz")) and not os.path.isfile(outbase + ".npz"):
        logger(" reusing %s/auto.* ..." % outbase)
        for ext in (".npz", ".windows.tab.gz", ".distance.params"):
            if os.path.lexists(outbase + ext):
                os.unlink(outbase + ext)
            os.symlink("auto" + ext, outbase + ext)
    # load clusters (:472-476)
    fname = outbase + ".clusters.tab"
    if os.path.isfile(fname):
        clusters = [l[:-1].split('\t') for l in open(fname)]
        logger("  %s clusters loaded from %s." % (len(clusters), fname))
        return clusters
    minSize, windows, chr2window, contig2size = contigs2windows(fasta, minSize, verbose, contigs=contigs)
    bin_chr = np.array([c for c, s, e in windows])
    bin_position = np.array([(s, e) for c, s, e in windows])
    if not os.path.isfile(outbase + ".npz"):
        logger("Parsing BAM...")
        D, c2dists = bam2array(windows, contig2size, chr2window, bam, mapq, upto, threads=threads,
                               minSize=minSize, device_out=True)
        logger("Saving array...")
        with gzip.open(outbase + ".windows.tab.gz", "wt") as out:
            out.write("\n".join("\t".join(map(str, w)) for w in windows) + "\n")
        with open(outbase + ".npz", "wb") as out:
            np.savez_compressed(out, D.cpu().numpy())
        params = estimate_distance_parameters(outbase, contig2size=contig2size, c2dists=c2dists,
                                              windowSize=minSize, contigs=contigs)
    else:
        npy = np.load(outbase + ".npz")
        D = torch.from_numpy(np.ascontiguousarray(npy[npy.files[0]], dtype=np.float64)).cuda()
        with open(outbase + ".distance.params", "rb") as fh:
            params = pickle.load(fh)
    # make symmetric (:511) and get clusters on the transformed matrix (:516-522), all on the device
    if normalize_matrix:
        # the call the reference keeps commented out at :512 (``#d = normalize(d)``): Sinkhorn-Knopp balancing of
        # the symmetrised matrix before the distance transform.  normalize() symmetrises and adds the pseudo-count
        # itself (:42-43), so it replaces :511; default off = the reference at HEAD
        D = normalize(D)
    else:
        ops.symmetrize_(D)
    dist = ops.transform(D, ops.KIND_POWER, float(params[0]), float(params[1]))
    clusters = cluster_contigs(outbase, dist, bin_chr, bin_position, dpi=dpi, minchr=minchr, nchr=nchr,
                               selections=selections)
    # skip empty clusters (:525-535)
    clusters = [c for c in clusters if c]
    totsize = ncontigs = 0
    logger("Reporting %s clusters to %s ..." % (len(clusters), fname))
    with open(fname, "w") as out:
        for cluster in clusters:
            totsize += sum(contig2size[c] for c in cluster)
            ncontigs += len(cluster)
            out.write("\t".join(cluster) + "\n")
    logger("  %3s bp in %s clusters generated from %s contigs." % (totsize, len(clusters), ncontigs))
    return clusters
