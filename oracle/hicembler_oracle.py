"""CPU oracle for the HiCembler hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package
(``hicembler_b200``) never imports anything under ``oracle/``.

Every function is a literal Python-3 / numpy restatement of the reference lines it cites
(paths relative to the reference checkout).  Python-2 traps that were resolved while
restating: integer ``/`` is ``//``; ``round()`` of an already-floored integer is the
identity; dict iteration order is replaced by an explicit ordered contig list.

Parity pinning:
  * Sinkhorn-Knopp: pinned.  ``SinkhornKnopp`` below is checked against the vendored
    ``bin/sinkhorn_knopp`` (imported with a one-line py3 shim, see
    ``oracle/make_golden.py``), against its docstring known answer
    (``sinkhorn_knopp.py:61-69``) and its six unit tests (restated in
    ``tests/test_oracle_sinkhorn.py``).
  * linkage: pinned against ``scipy.cluster.hierarchy.linkage`` (the reference's own call at
    ``array2scaffolds.py:163``) and scipy's known-answer vectors; the fastcluster NN-chain
    restatement (``oracle/nnchain.c``) is pinned only through those (fastcluster itself is a
    PyPI dependency absent from the reference tree and from this image) -> under exact ties
    its merge ORDER is "parity unpinned".
  * binning / transform / block mean: the reference holds no test or golden vector for them
    -> "parity unpinned"; the restatement follows the cited lines literally.
"""
from __future__ import annotations

import ctypes
import os
import warnings
from collections import Counter

import numpy as np

# --------------------------------------------------------------------------------------
# SAM filtering and window arithmetic
# --------------------------------------------------------------------------------------
SKIP_FLAG = 3980  # bam2clusters.py:108  (4+8+128+256+512+1024+2048)


def samtools_view(lines, mapq=0, skipFlag=SKIP_FLAG):
    """Restates ``samtools view -q mapq -F skipFlag`` (bam2clusters.py:108-117).

    Yields the SAM text lines samtools would print: header lines dropped, records kept when
    ``MAPQ >= mapq`` and ``(FLAG & skipFlag) == 0``.
    """
    for line in lines:
        if line.startswith("@"):
            continue
        f = line.split("\t", 5)
        if int(f[1]) & skipFlag:
            continue
        if int(f[4]) < mapq:
            continue
        yield line


def get_window(chrom, start, length, windowSize, chr2window):
    """bam2clusters.py:119-126 (copies: fastq2array.py:146-153, bam2plot.py:96-103).

    Py2: ``pos = int(start+round((end-start)/2))`` with integer ``/``.
    """
    end = start + length
    pos = int(start + round((end - start) // 2))
    window = pos // windowSize + chr2window[chrom]
    return window


def bam2arrays(windows, windowSize, chr2window, sam, mapq, upto=0, regions=(), verbose=0, shadow_mapq=True):
    """bam2clusters.py:128-166.  ``sam`` is a list of iterables of SAM text lines (one per
    BAM); each stands for the ``samtools view`` pipe of that BAM *before* -q/-F filtering.

    ``regions`` restates ``samtools view bam <contig>...``: only records whose RNAME is one of
    the region contigs are printed (whole-contig regions, the only form the live pipeline
    passes to this function, bam2scaffolds.py:244).
    """
    arrays = [np.zeros((len(w), len(w)), dtype="float32") for w in windows]
    i = 1
    regions = set(regions)
    for _sam in sam:
        stream = samtools_view(_sam, int(mapq))
        if regions:
            stream = (l for l in stream if l.split("\t", 3)[2] in regions)
        for i, line in enumerate(stream, i):
            if upto and i > upto:
                break
            # :145 unpacks the line's MAPQ field into ``mapq`` itself: the argument is overwritten, and the NEXT
            # BAM is filtered with ``-q <MAPQ of the last line read>`` (:138).  shadow_mapq=False = the intent.
            ref1, start1, _mapq, cigar, ref2, start2, insertSize, seq = line.split("\t")[2:10]
            if shadow_mapq:
                mapq = _mapq
            start1, start2, seqlen = int(start1), int(start2), len(seq)
            if ref2 == "=":
                ref2 = ref1
            if ref1 not in chr2window[0] or ref2 not in chr2window[0]:
                continue
            for ii in range(len(windowSize)):
                w1 = get_window(ref1, start1, seqlen, windowSize[ii], chr2window[ii])
                w2 = get_window(ref2, start2, seqlen, windowSize[ii], chr2window[ii])
                if w2 < w1:
                    w1, w2 = w2, w1
                arrays[ii][w1][w2] += 1
    return arrays


def bin_records(rec, window_bp, offsets, n_windows, mapq, skip_flag=SKIP_FLAG, upto=0,
                member=None, region_mask=None):
    """Vectorised numpy form of the same arithmetic over decoded pair records
     -- used where the literal per-line loop is too slow.

    rec: structured array with refID,pos,next_refID,next_pos,l_seq,flag,mapq (0-based pos).
    offsets: int64[n_contigs], -1 = contig absent from chr2window.
    Returns float32 upper-triangular (n_windows, n_windows), as bam2arrays does.
    """
    keep = ((rec["flag"].astype(np.int64) & skip_flag) == 0) & (rec["mapq"] >= mapq)
    if region_mask is not None:
        keep &= region_mask[rec["refID"]].astype(bool)
    if upto:
        keep &= np.cumsum(keep) <= upto
    member = offsets if member is None else member
    r = rec[keep]
    ok = (member[r["refID"]] >= 0) & (member[r["next_refID"]] >= 0)
    r = r[ok]
    L = np.maximum(r["l_seq"].astype(np.int64), 1)
    p1 = r["pos"].astype(np.int64) + 1 + L // 2
    p2 = r["next_pos"].astype(np.int64) + 1 + L // 2
    o1, o2 = offsets[r["refID"]], offsets[r["next_refID"]]
    if np.any(o1 < 0) or np.any(o2 < 0):
        raise KeyError("contig present in chr2window[0] but absent for this window size")
    w1 = p1 // window_bp + o1
    w2 = p2 // window_bp + o2
    lo, hi = np.minimum(w1, w2), np.maximum(w1, w2)
    if len(hi) and hi.max() >= n_windows:
        raise IndexError("window index out of range")
    flat = np.bincount(lo * n_windows + hi, minlength=n_windows * n_windows)
    # the reference accumulates float32 += 1, which is exact below 2**24 per cell
    assert flat.max(initial=0) < (1 << 24)
    return flat.reshape(n_windows, n_windows).astype(np.float32)


# --------------------------------------------------------------------------------------
# contig-level matrix (bam2clusters.py:168-246)
# --------------------------------------------------------------------------------------
def contig_regions(contig2size, chr2window, minSize, nth=5):
    """bam2clusters.py:222 -- sampled regions (s+1, s+minSize) every nth*minSize bp."""
    return {c: np.array([(s + 1, s + minSize)
                         for s in range(0, contig2size[c] + 1 - minSize, nth * minSize)])
            for c in chr2window}


def samtools_region_lines(lines, regions_of):
    """Restates ``samtools view bam c:s-e ...`` on a coordinate-sorted BAM: a record is
    printed once per region its alignment [POS, POS+reflen-1] overlaps (regions here are
    disjoint).  reflen is taken from the CIGAR (M/D/N/=/X consume the reference)."""
    import re
    cig = re.compile(r"(\d+)([MIDNSHP=X])")
    for line in lines:
        f = line.split("\t", 7)
        reg = regions_of.get(f[2])
        if reg is None or not len(reg):
            continue
        pos = int(f[3])
        reflen = sum(int(n) for n, op in cig.findall(f[5]) if op in "MDN=X") or 1
        end = pos + reflen - 1
        hits = int(np.sum((reg[:, 0] <= end) & (reg[:, 1] >= pos)))
        for _ in range(hits):
            yield line


def _bam2array(sam, contigs, mapq, contig2size, chr2window, c2regions):
    """bam2clusters.py:168-206 (one worker)."""
    data = Counter()
    c2dists = {c: [] for c in contigs}
    i = 0
    stream = samtools_region_lines(samtools_view(sam, mapq), {c: c2regions[c] for c in contigs})
    for i, line in enumerate(stream, 1):
        ref1, start1, _mapq, cigar, ref2, start2, isize, seq = line.split("\t")[2:10]
        start1, start2, seqlen, isize = int(start1), int(start2), len(seq), abs(int(isize))
        if ref2 == "=":
            ref2 = ref1
        else:
            isize = -1
        c2dists[ref1].append(isize)
        start2 += seqlen // 2
        if ref2 in chr2window and np.any([c2regions[ref2][:, 0] <= start2,
                                          c2regions[ref2][:, 1] >= start2]):
            w1, w2 = chr2window[ref1], chr2window[ref2]
            if w2 < w1:
                w1, w2 = w2, w1
            data[(w1, w2)] += 1
    return data, i, c2dists


def bam2array(windows, contig2size, chr2window, sam, mapq=10, upto=0, regions=(),
              minSize=2000, threads=4, nth=5):
    """bam2clusters.py:208-246 with the Pool replaced by a serial loop over the same
    strided contig subsets (the result is a sum of Counters, order-independent)."""
    array = np.zeros((len(windows), len(windows)), dtype="float32")
    regions = list(regions) if len(regions) else list(chr2window.keys())
    threads = min(threads, len(regions))
    i = 0
    c2dists = {}
    c2regions = contig_regions(contig2size, chr2window, minSize, nth)
    sizes = [len(c2regions[c]) for c, s, e in windows]
    for _sam in sam:
        _sam = list(_sam)
        for ii in range(threads):
            wdata, algs, _c2dists = _bam2array(_sam, regions[ii::threads], mapq, contig2size,
                                               chr2window, c2regions)
            i += algs
            for (w1, w2), c in wdata.items():
                array[w1][w2] += c
            for c in _c2dists:
                c2dists.setdefault(c, [])
                c2dists[c] += _c2dists[c]
    array = (array / sizes).T / sizes
    return array, c2dists


# --------------------------------------------------------------------------------------
# window tables
# --------------------------------------------------------------------------------------
def get_N_and_L(sizes, genomeFrac=0.5):
    """bin/FastaIndex/FastaIndex.py:304-325 (N only)."""
    genomeSize = sum(sizes)
    totsize = 0
    size = 0
    for size in sorted(sizes, reverse=True):
        totsize += size
        if totsize >= genomeFrac * genomeSize:
            break
    return size


def contigs2windows(contig_sizes, minSize=2000, genomeFrac=0.95, contigs=()):
    """bam2clusters.py:73-106.  ``contig_sizes`` is an ORDERED list of (name, size) standing
    for FastaIndex iteration order (explicit)."""
    contigs = set(contigs)
    windows, chr2window = [], {}
    contig2size = dict(contig_sizes)
    n90 = get_N_and_L([s for _, s in contig_sizes], genomeFrac)
    if n90 > minSize:
        minSize = n90
    for c, size in contig_sizes:
        if size < minSize:
            continue
        elif contigs and c not in contigs:
            continue
        chr2window[c] = len(windows)
        windows.append((c, 0, size))
    return minSize, windows, chr2window, contig2size


def fasta2windows(contig_sizes, windowSize, skipShorter=1, minSize=2000, contigs=(),
                  filterwindows=1):
    """bam2scaffolds.py:73-124; windowSize in kb.  Returns (windowSize_bp, windows,
    chr2window, base2chr, genomeSize)."""
    sizes = [s for _, s in contig_sizes]
    genome_total = sum(sizes)
    windowSize = list(windowSize)
    if filterwindows:
        maxchrlen = max(sizes)
        windowSize = [x for x in windowSize if 1000 * x < maxchrlen
                      and 1000 * x < 0.01 * genome_total and 1000 * x > 0.000001 * genome_total]
    windowSize = [w * 1000 for w in windowSize]
    windows, chr2window = [[] for w in windowSize], [{} for w in windowSize]
    genomeSize = 0
    base2chr = {}
    contigs = set(contigs)
    for c, size in contig_sizes:
        if contigs and c not in contigs:
            continue
        if skipShorter and size < min(windowSize) or size < minSize:
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
    return windowSize, windows, chr2window, base2chr, genome_total


# --------------------------------------------------------------------------------------
# symmetrise / normalise / transform
# --------------------------------------------------------------------------------------
def symmetrize(d):
    """bam2clusters.py:42,274,511; bam2scaffolds.py:262 (in place)."""
    d += d.T - np.diag(d.diagonal())
    return d


def symmetrize_halfdiag(d):
    """array2scaffolds.py:117-118, bam2scaffolds.py:32-33 (in place)."""
    d += d.T
    d -= np.diag(d.diagonal() / 2)
    return d


class SinkhornKnopp:
    """Literal restatement of bin/sinkhorn_knopp/sinkhorn_knopp/sinkhorn_knopp.py:73-163
    (``IntType, FloatType`` -> ``int, float``)."""

    def __init__(self, max_iter=1000, epsilon=1e-3):
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

    def fit(self, P, materialize=True):
        """``materialize=False`` skips the dense diag GEMMs (O(N^3)) and builds the same
        P_eps by broadcasting -- identical products, used only for large-N oracle runs."""
        P = np.asarray(P)
        assert np.all(P >= 0)
        assert P.ndim == 2
        assert P.shape[0] == P.shape[1]
        N = P.shape[0]
        max_thresh = 1 + self._epsilon
        min_thresh = 1 - self._epsilon
        r = np.ones((N, 1))
        pdotr = P.T.dot(r)
        msg = "Matrix P must have total support. See documentation"
        if not np.all(pdotr != 0):
            warnings.warn(msg, UserWarning)
        with np.errstate(divide="ignore", invalid="ignore"):
            c = 1 / pdotr
            pdotc = P.dot(c)
            if not np.all(pdotc != 0):
                warnings.warn(msg, UserWarning)
            r = 1 / pdotc
            del pdotr, pdotc

            def scaled(r, c):
                if materialize:
                    self._D1 = np.diag(np.squeeze(r))
                    self._D2 = np.diag(np.squeeze(c))
                    return self._D1.dot(P).dot(self._D2)
                self._D1, self._D2 = np.squeeze(r), np.squeeze(c)
                return (r * P) * c.T

            P_eps = np.copy(P)
            while np.any(np.sum(P_eps, axis=1) < min_thresh) \
                    or np.any(np.sum(P_eps, axis=1) > max_thresh) \
                    or np.any(np.sum(P_eps, axis=0) < min_thresh) \
                    or np.any(np.sum(P_eps, axis=0) > max_thresh):
                c = 1 / P.T.dot(r)
                r = 1 / P.dot(c)
                P_eps = scaled(r, c)
                self._iterations += 1
                if self._iterations >= self._max_iter:
                    self._stopping_condition = "max_iter"
                    break
            if not self._stopping_condition:
                self._stopping_condition = "epsilon"
            P_eps = scaled(r, c)
        self._r, self._c = np.squeeze(r), np.squeeze(c)
        return P_eps


def normalize(d, max_iter=1000, epsilon=0.00001, materialize=True, sk=None):
    """bam2clusters.py:36-49 (mutates ``d`` in place, returns a new float64 array)."""
    sk = sk or SinkhornKnopp(max_iter=max_iter, epsilon=epsilon)
    d += d.T - np.diag(d.diagonal())
    d += 1
    vmax = d.max()
    d = sk.fit(d / vmax, materialize=materialize)
    return d


def distance_func(x, a, b, maxv=float("inf")):
    """bam2clusters.py:248-251.  dtype: under the reference's NumPy (value-based casting) the scalars a and 1./b --
    np.float64 from curve_fit -- do not promote a float32 array, so a float32 ``x`` is evaluated entirely in
    float32.  NumPy >= 2 (NEP 50) promotes for np.float64 scalars but not for Python floats; the cast below makes
    the restatement independent of both the NumPy version and the scalar type."""
    if isinstance(x, np.ndarray) and x.dtype == np.float32:
        return (np.float32(1.) / (np.float32(a) * x)) ** np.float32(1. / b)
    return (1. / (a * x)) ** (1. / b)


def contact_func(x, a, b):
    """bam2clusters.py:253-255."""
    return 1. / (a * x ** b)


def transform_distance(d, a, b):
    """bam2clusters.py:516, bam2scaffolds.py:265: ``distance_func(x+1, *params)``."""
    return distance_func(d + 1, a, b)


def transform_logmax(x):
    """array2scaffolds.py:31."""
    return np.log(np.max(x + 1)) - np.log(x + 1)


def normalize_rows(a):
    """array2scaffolds.py:33-41 / bam2scaffolds.py:43-51 (in place)."""
    rows, cols = a.shape
    maxv = a.sum(axis=0).max()
    for i in range(rows):
        if a[i].max():
            a[i] *= 1. * maxv / a[i].sum()
    return a


def normalize_diagional(d, bin_chr, bin_position):
    """bam2clusters.py:51-66."""
    indices = d.diagonal() != 0
    d = d[indices, :]
    d = d[:, indices]
    bin_chr = bin_chr[indices]
    bin_position = bin_position[indices, :]
    d += d.T
    d -= np.diag(d.diagonal() / 2)
    n2 = np.mean(d.diagonal()) / d.diagonal()
    d *= (d * n2).T * n2
    return d, bin_chr, bin_position


def normalize_window_size(d, bin_chr, bin_position, windowSize=0):
    """bam2scaffolds.py:28-41."""
    from collections import Counter
    d += d.T
    d -= np.diag(d.diagonal() / 2)
    sizes = np.diff(bin_position, axis=1)
    if not windowSize:
        c = Counter(sizes.reshape(len(sizes)).tolist())
        windowSize, occurencies = c.most_common(1)[0]
    d *= 1. * windowSize ** 2 / (sizes * sizes.T)
    return d, bin_chr, bin_position


def _func_reduce(A, keys, func, allkeys=None):
    """bam2scaffolds.py:53-65."""
    unique_keys = np.unique(keys)
    if allkeys is None:
        allkeys = unique_keys
    newshape = (len(allkeys),) + A.shape[1:]
    newA = np.zeros(newshape, dtype=A.dtype)
    for i, k in enumerate(allkeys):
        indices = (keys == k)
        newA[i] = func(A[indices], axis=0)
    return newA


def _average_reduce(A, keys):
    return _func_reduce(A, keys, func=np.mean)


def _average_reduce_2d(A, keys):
    """bam2scaffolds.py:70-71."""
    return _average_reduce(_average_reduce(A, keys).T, keys).T


def condense(d):
    """``d[np.triu_indices(n, 1)]`` -- bam2clusters.py:385,403,411,430; array2scaffolds.py:163."""
    return d[np.triu_indices(d.shape[0], 1)]


# --------------------------------------------------------------------------------------
# linkage
# --------------------------------------------------------------------------------------
METHODS = {"single": 0, "complete": 1, "average": 2, "weighted": 3, "ward": 4}
_HERE = os.path.dirname(os.path.abspath(__file__))
_nn = None


def _nnchain_lib():
    global _nn
    if _nn is None:
        path = os.path.join(_HERE, "libnnchain.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-s", "-C", _HERE, "libnnchain.so"])
        _nn = ctypes.CDLL(path)
        _nn.fc_linkage.restype = ctypes.c_int
        _nn.fc_linkage.argtypes = [ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                   ctypes.c_void_p, ctypes.c_void_p]
    return _nn


def linkage_fastcluster(y, method="ward", return_stats=False):
    """``fastcluster.linkage(y, method)`` (call sites bam2clusters.py:385,403,411,430) through
    the C restatement of its NN-chain (oracle/nnchain.c; fastcluster is a PyPI dependency,
    1.1.x, not vendored in the reference)."""
    y = np.array(y, dtype=np.float64, order="C", copy=True)
    NN = len(y)
    N = int(np.ceil(np.sqrt(NN * 2)))
    if (N * (N - 1) // 2) != NN:
        raise ValueError("The length of the condensed distance matrix must be (k \\choose 2) "
                         "for k data points!")
    Z = np.empty((max(N - 1, 0), 4))
    stats = np.zeros(4, dtype=np.int64)
    if N > 1:
        rc = _nnchain_lib().fc_linkage(N, y.ctypes.data, METHODS[method], Z.ctypes.data,
                                       stats.ctypes.data)
        if rc == -1:
            raise FloatingPointError("NaN dissimilarity value.")
        if rc:
            raise RuntimeError("nnchain oracle failed: %d" % rc)
    return (Z, stats) if return_stats else Z


def linkage_fastcluster_py(y, method="ward"):
    """Pure-Python restatement of the same algorithm for small N;
    cross-checks the C restatement."""
    y = np.array(y, dtype=np.float64)
    NN = len(y)
    N = int(np.ceil(np.sqrt(NN * 2)))
    assert N * (N - 1) // 2 == NN
    if np.isnan(y).any():
        raise FloatingPointError("NaN dissimilarity value.")
    D = np.zeros((N, N))
    D[np.triu_indices(N, 1)] = y
    if method == "ward":
        D = D * D

    def g(i, j):
        return D[i, j] if i < j else D[j, i]

    def s(i, j, v):
        if i < j:
            D[i, j] = v
        else:
            D[j, i] = v

    members = [1] * N
    active = list(range(N))
    chain = [0] * N
    tip = 0
    rec = []
    for _ in range(N - 1):
        if tip <= 3:
            a = active[0]
            chain[0] = a
            tip = 1
            b = active[1]
            m = g(a, b)
            for i in active[2:]:
                if g(a, i) < m:
                    m = g(a, i)
                    b = i
        else:
            tip -= 3
            a = chain[tip - 1]
            b = chain[tip]
            m = g(a, b)
        while True:
            chain[tip] = b
            for i in active:
                if i != b and g(i, b) < m:
                    m = g(i, b)
                    a = i
            b = a
            a = chain[tip]
            tip += 1
            if b == chain[tip - 2]:
                break
        rec.append((a, b, m))
        i1, i2 = (a, b) if a < b else (b, a)
        n1, n2 = float(members[i1]), float(members[i2])
        members[i2] += members[i1]
        active.remove(i1)
        for k in active:
            if k == i2:
                continue
            if method == "average":
                sa, t = n1 / (n1 + n2), n2 / (n1 + n2)
                s(k, i2, sa * g(k, i1) + t * g(k, i2))
            elif method == "ward":
                v = float(members[k])
                s(k, i2, ((v + n1) * g(k, i1) - v * m + (v + n2) * g(k, i2)) / (n1 + n2 + v))
            elif method == "complete":
                s(k, i2, max(g(k, i1), g(k, i2)))
            elif method == "single":
                s(k, i2, min(g(k, i1), g(k, i2)))
            elif method == "weighted":
                s(k, i2, (g(k, i1) + g(k, i2)) * .5)
    h = np.array([r[2] for r in rec])
    if method == "ward":
        h = np.sqrt(h)
    order = np.argsort(h, kind="stable")
    parent = list(range(2 * N - 1))
    size = [1] * (2 * N - 1)

    def find(x):
        root = x
        while parent[root] != root:
            root = parent[root]
        while parent[x] != root:
            parent[x], x = root, parent[x]
        return root

    Z = np.empty((N - 1, 4))
    nxt = N
    for row, k in enumerate(order):
        a, b = find(rec[k][0]), find(rec[k][1])
        parent[a] = parent[b] = nxt
        size[nxt] = size[a] + size[b]
        Z[row] = (min(a, b), max(a, b), h[k], size[nxt])
        nxt += 1
    return Z


def linkage_scipy(y, method="ward"):
    """``sch.linkage(y, method)`` -- array2scaffolds.py:163 (the real library, importable)."""
    import scipy.cluster.hierarchy as sch
    return sch.linkage(y, method=method)


def array2Z(d, method="ward", lib="fastcluster"):
    """Linkage part of array2tree: bam2clusters.py:385 (fastcluster) /
    array2scaffolds.py:163 (scipy)."""
    y = condense(d)
    return linkage_fastcluster(y, method) if lib == "fastcluster" else linkage_scipy(y, method)


def cluster_contigs(d, bin_chr, frac=0.66, iterations=30, maxnumchr=500, minchr=3, nchr=0,
                    selections=None, prng=None):
    """bam2clusters.py:406-457 without the plots.  The reference draws from the global
    unseeded ``np.random`` (:410,429); ``selections`` (list of index arrays) or ``prng``
    make the draw injectable.  Returns (clusters, nchr, Z, mean_step_len)."""
    import scipy.cluster.hierarchy as sch
    prng = prng or np.random
    Z = linkage_fastcluster(condense(d), "ward")
    mean_step_len = None
    if not nchr:
        dZ = []
        n = d.shape[0]
        subset = int(round(n * frac))
        if subset < maxnumchr:
            maxnumchr = subset - 1
        for i in range(iterations):
            selection = selections[i] if selections is not None else prng.permutation(n)[:subset]
            iZ = linkage_fastcluster(condense(d[selection, :][:, selection]), "ward")
            dZ.append(np.diff(iZ[:, 2]))
        mean_step_len = np.mean(np.array(dZ), 0)
        nchr = int(np.argmax(mean_step_len[-minchr + 1::-1][:maxnumchr]) + minchr)
    clusters = [[] for n in range(nchr)]
    assignments = sch.fcluster(Z, t=nchr, criterion="maxclust")
    for i, c in zip(assignments - 1, bin_chr):
        clusters[i].append(c)
    return clusters, nchr, Z, mean_step_len


# ---------------------------------------------------------------------------------------------
# distance-decay fit and the bam2clusters() driver -- plots left out
# ---------------------------------------------------------------------------------------------
def get_distances_contacts(sam, mapq, contig2size, windowSize, icontigs=5, upto=1e7):
    """bam2clusters.py:257-288.  ``sam`` = list of SAM line lists (one per BAM); the
    ``regions=[c]`` query of :273 is restated as "read-1 reference is c".  Py2 traps: ``s /
    windowSize`` (:268) and ``(positions[j]-positions[i]) / 1000`` (:283) are integer divisions;
    ``sorted(contig2size, ...)`` (:262) is a stable sort of dict order -> callers pass an
    ordered dict."""
    d2c = {0: []}
    longest_contigs = sorted(contig2size, key=lambda x: contig2size[x], reverse=True)
    maxdist = 0.75 * contig2size[longest_contigs[0]] / 1000.
    for c in longest_contigs[:icontigs]:
        s = contig2size[c]
        n = s // windowSize + 1
        positions = list(range(windowSize, n * windowSize, windowSize))
        arrays = [np.zeros((n, n), dtype='float32')]
        chr2window = [{c: 0}]
        sub = [[l for l in lines if l.split('\t')[2] == c] for lines in sam]
        arrays = bam2arrays(arrays, [windowSize], chr2window, sub, mapq, upto=upto, verbose=0)
        d = arrays[0]
        d += d.T - np.diag(d.diagonal())
        d2c[0] += [d[i][i] for i in range(d.shape[0] - 1)]
        for i in range(len(positions) - 1):
            for j in range(i, len(positions)):
                dist = (positions[j] - positions[i]) // 1000
                if dist not in d2c:
                    d2c[dist] = []
                d2c[dist].append(d[i][j])
    dists = np.array([d for d in sorted(d2c) if d < maxdist])
    contacts = [d2c[d] for d in dists]
    return dists, contacts


def distance_fit_inputs(contig2size=None, windowSize=2000, c2dists=None, limit=30, contigs=()):
    """The x / yn / dists / contacts that bam2clusters.py:304-330 hands to curve_fit when insert
    sizes are given.  ``Counter(...).iteritems()`` (:315) runs in ascending key order here: that is
    CPython 2's order for small-int keys once the dict table is larger than the largest key."""
    from collections import Counter
    x, yn = [], []
    dists = list(range(-windowSize, (limit + 11) * windowSize, windowSize))
    contacts = [[] for i in range(len(dists) + 1)]
    if contigs:
        contig2size = {c: s for c, s in contig2size.items() if c in contigs}
    longest_contigs = sorted(contig2size, key=lambda x: contig2size[x], reverse=True)[:50]
    for c in longest_contigs:
        maxc = max(c2dists[c])
        if not maxc:
            continue
        for i, nc in sorted(Counter(np.digitize(c2dists[c], dists, right=True).tolist()).items()):
            nc = 1e6 * nc / sum(c2dists[c])
            contacts[i - 1].append(nc)
            x.append(dists[i - 1] / 1000.)
            yn.append(nc)
    _dists, _contacts = dists[1:-1], contacts[1:-1]
    dists, contacts = [], []
    for d, c in zip(_dists, _contacts):
        if c:
            dists.append(d / 1000.)
            contacts.append(c)
    return x, yn, dists[:limit], contacts[:limit]


def estimate_distance_parameters(sam=None, mapq=10, contig2size=None, windowSize=2000, skipfirst=5,
                                 icontigs=5, c2dists=None, limit=30, upto=1e7, contigs=()):
    """bam2clusters.py:290-358 without the figures and the pickle: returns ``params``."""
    from scipy.optimize import curve_fit
    x, yn = [], []
    if sam and not c2dists:
        dists, contacts = get_distances_contacts(sam, mapq, contig2size, windowSize, icontigs, upto)
        dists, contacts = dists[:limit], contacts[:limit]
    else:
        x, yn, dists, contacts = distance_fit_inputs(contig2size, windowSize, c2dists, limit, contigs)
    if not x:
        x, yn = dists, np.array([np.median(c) for c in contacts])
    params, pcov = curve_fit(contact_func, x[skipfirst:], yn[skipfirst:], maxfev=1000)
    return params


# ---------------------------------------------------------------------------------------------
# scaffolding stage: bam2scaffolds.py:126-278,305-331 (= array2scaffolds.py:258-334)
# ---------------------------------------------------------------------------------------------
def get_reversed(scaffold):
    """bam2scaffolds.py:126-128."""
    return [(name, not orientation) for name, orientation in reversed(scaffold)]


def get_indices(scaffold, contig2indices):
    """bam2scaffolds.py:130-140."""
    indices = []
    for name, reverse in scaffold:
        if name not in contig2indices:
            continue
        if reverse:
            indices += reversed(contig2indices[name])
        else:
            indices += contig2indices[name]
    return indices


def join_scaffolds(scaffold1, scaffold2, d, contig2indices, minWindows=3):
    """bam2scaffolds.py:142-174 (Py2 ``/`` on ints is ``//``)."""
    indices1 = get_indices(scaffold1, contig2indices)
    indices2 = get_indices(scaffold2, contig2indices)
    if len(indices1) < len(indices2):
        scaffold1, indices1, scaffold2, indices2 = scaffold2, indices2, scaffold1, indices1
    if len(indices2) < minWindows:
        if len(indices1) < minWindows:
            return []
        return scaffold1
    _d = d[:, indices1 + indices2][indices1 + indices2, :]
    n1, n2 = len(indices1), len(indices2)
    i = n2 // 2
    d1, d2 = _d[:n1 // 2, n1:], _d[n1 - n1 // 2:n1, n1:]
    orientation = np.argmax(list(map(np.sum, (d1[:, :i], d1[:, -i:], d2[:, :i], d2[:, -i:]))))
    if orientation == 0:
        scaffold = get_reversed(scaffold2) + scaffold1
    elif orientation == 1:
        scaffold = scaffold2 + scaffold1
    elif orientation == 2:
        scaffold = scaffold1 + scaffold2
    else:
        scaffold = scaffold1 + get_reversed(scaffold2)
    return scaffold


def get_contig2indices(bin_chr):
    """bam2scaffolds.py:176-181."""
    contig2indices = {c: [] for c in np.unique(bin_chr)}
    for i, c in enumerate(bin_chr):
        contig2indices[c].append(i)
    return contig2indices


def join_scaffolds_diag(scaffold1, scaffold2, d, contig2indices, minWindows=3):
    """array2scaffolds.py:274-309: the OTHER join_scaffolds -- orientation from four anti/diagonal sums of the
    rectangle between the two scaffolds instead of corner-block sums, and no empty result for two short scaffolds
    (Py2 ``n2/2`` is ``//``).  ``d1[:i]`` / ``d3[-i:]`` with i == 0: ``[:0]`` is empty, ``[-0:]`` is everything."""
    indices1 = get_indices(scaffold1, contig2indices)
    indices2 = get_indices(scaffold2, contig2indices)
    if len(indices1) < len(indices2):
        scaffold1, indices1, scaffold2, indices2 = scaffold2, indices2, scaffold1, indices1
    if len(indices2) < minWindows:
        return scaffold1
    _d = d[:, indices1 + indices2][indices1 + indices2, :]
    n1, n2 = len(indices1), len(indices2)
    i = n2 // 2
    dflip = np.fliplr(_d[:n1, n2:])
    d1, d2, d3, d4 = np.diag(_d, k=n1), np.diag(dflip), np.diag(dflip, k=n2 - n1), np.diag(_d, k=n2)
    orientation = np.argmax(list(map(sum, (d1[:i], d2[:i], d3[-i:], d4[-i:]))))
    if orientation == 0:
        scaffold = get_reversed(scaffold2) + scaffold1
    elif orientation == 1:
        scaffold = scaffold2 + scaffold1
    elif orientation == 2:
        scaffold = scaffold1 + scaffold2
    else:
        scaffold = scaffold1 + get_reversed(scaffold2)
    return scaffold


def tree2scaffold(Z, names, d, bin_chr, minWindows, join=None, children="newick"):
    """bam2scaffolds.py:183-200 (= array2scaffolds.py:319-338) on the linkage array instead of the ete3 tree: rows
    of Z are a postorder.  ``n.get_children()`` is [right, left] = [Z[k, 1], Z[k, 0]] for the tree parsed from
    getNewick's string (bam2clusters.py:368-370: "(" right "," left ")"; ``children="newick"``) and
    [Z[k, 0], Z[k, 1]] for the tree distance_matrix2tree builds (array2scaffolds.py:141-142; ``children="dm2t"``)."""
    join = join or join_scaffolds
    contig2indices = get_contig2indices(bin_chr)
    n = len(names)
    scaffolds = [[(names[i], 0)] for i in range(n)]
    for k in range(n - 1):
        n1, n2 = (int(Z[k, 1]), int(Z[k, 0])) if children == "newick" else (int(Z[k, 0]), int(Z[k, 1]))
        scaffolds.append(join(scaffolds[n1], scaffolds[n2], d, contig2indices, minWindows))
    return scaffolds[-1]


# ---------------------------------------------------------------------------------------------
# array2scaffolds.py: distance_matrix2tree (:123-145) and get_clusters (:199-256) on plain arrays.
# ete3 semantics used by the reference (third-party, unpinned; documented in oracle/ete3_min.py): a node made by
# ``ete3.Tree()`` has dist 1.0 and name ""; get_farthest_leaf returns the FIRST leaf in pre-order with the
# strictly largest path length; get_ancestors lists the parent first; len(tree) / iteration = leaves in pre-order;
# newick format 0 prints "%0.6g" distances and the support value (1) after every internal ")".
# ---------------------------------------------------------------------------------------------
class _DM2Tree(object):
    def __init__(self, Z, names):
        n = Z.shape[0] + 1
        self.n = n
        nn = 2 * n - 1
        self.children = [[] for _ in range(nn)]
        self.up = [-1] * nn
        self.dist = [1.0] * nn                    # ete3 default branch length (the root keeps it)
        self.name = [names[i] if i < n else "" for i in range(nn)]
        for i in range(n - 1):
            a, b = int(Z[i, 0]), int(Z[i, 1])
            h = float(Z[i, 2])
            k = n + i
            self.dist[a] = h - self.farthest_leaf(a)[1]      # :137-138 dist - child.get_farthest_leaf()[1]
            self.dist[b] = h - self.farthest_leaf(b)[1]
            self.children[k] = [a, b]
            self.up[a] = self.up[b] = k
        self.root = nn - 1

    def leaves(self, node):
        out, stack = [], [node]
        while stack:
            k = stack.pop()
            if self.children[k]:
                stack.extend(reversed(self.children[k]))
            else:
                out.append(k)
        return out

    def farthest_leaf(self, node):
        """ete3 TreeNode.get_farthest_leaf: a single running distance, += on entering an internal node, -= on
        leaving it; first leaf with the strictly largest value wins (rounding history included, see ete3_min.py)."""
        if not self.children[node]:
            return node, 0.0
        best, best_d = None, None
        d = 0.0
        stack = [(c, False) for c in reversed(self.children[node])]
        while stack:
            k, leaving = stack.pop()
            if leaving:
                d -= self.dist[k]
            elif self.children[k]:
                d += self.dist[k]
                stack.append((k, True))
                stack.extend((c, False) for c in reversed(self.children[k]))
            else:
                tot = d + self.dist[k]
                if best_d is None or tot > best_d:
                    best, best_d = k, tot
        return best, best_d

    def ancestors(self, node):
        out, k = [], self.up[node]
        while k >= 0:
            out.append(k)
            k = self.up[k]
        return out

    def newick(self):
        out, stack = [], [("n", self.root)]
        while stack:
            kind, k = stack.pop()
            if kind == "t":
                out.append(k)
            elif self.children[k]:
                out.append("(")
                stack.append(("t", ")" if k == self.root else ")1:%0.6g" % self.dist[k]))
                ch = self.children[k]
                for q, c in enumerate(reversed(ch)):
                    stack.append(("n", c))
                    if q != len(ch) - 1:
                        stack.append(("t", ","))
            else:
                out.append("%s:%0.6g" % (self.name[k], self.dist[k]))
        return "".join(out) + ";"


def distance_matrix2tree_newick(Z, names):
    """``distance_matrix2tree(Z, names).write()`` (array2scaffolds.py:123-145,171-172): child branch length =
    merge height - height of the child's farthest leaf."""
    return _DM2Tree(Z, names).newick()


def get_clusters(Z, names, contig2size, bin_chr):
    """array2scaffolds.py:199-256 on the tree of distance_matrix2tree(Z, names) (without the file output).
    Restated quirks: ``clusters`` is sized by the leaf count of the LAST subtree (:232, ``len(subtree)`` instead of
    ``len(subtrees)``) and indexed with the 1-based subtree number (:241), so cluster k lands in slot k, slot 0 stays
    empty and an IndexError is raised when a subtree number reaches that leaf count; an IndexError is also raised
    when the longest branch above the farthest leaf is the root's own default branch (:209-216).  Contigs are
    visited in np.unique order (Python-3 dict order of :220; hash order under Python 2)."""
    get_name = lambda c: c.split()[0]
    T = _DM2Tree(Z, names)
    subtrees = []
    while len(T.leaves(T.root)) > 2:
        n, dist = T.farthest_leaf(T.root)
        anc = T.ancestors(n)
        dists = [T.dist[a] for a in anc]
        ai = dists.index(max(dists))
        a = anc[ai]
        if T.name[n]:
            subtrees.append(a)
        p = anc[ai + 1]
        T.children[p].remove(a)
        T.up[a] = -1
    contig2cluster = {get_name(c): Counter() for c in np.unique(bin_chr)}
    subtree = None
    for i, subtree in enumerate(subtrees, 1):
        lv = T.leaves(subtree)
        c = Counter(get_name(T.name[k]) for k in lv if T.name[k])
        c.most_common(1)[0][1]
        for k, v in c.items():
            if not k:
                continue
            contig2cluster[get_name(k)][i] += v
    clusters = [[] for i in range(len(T.leaves(subtree)))]
    for c, counter in contig2cluster.items():
        if not counter:
            continue
        clusteri, count = counter.most_common(1)[0]
        clusters[clusteri].append(c)
    return [x for x in clusters if x]


def scaffold2gaps(d, bin_chr, scaffold, params, contig2size, mingap=100, maxgap=1000000, gapfrac=.33):
    """bam2scaffolds.py:202-232 (``contig2size[c]/2`` is integer division in Py2)."""
    contig2indices = get_contig2indices(bin_chr)
    scaffold_with_gaps = []
    contigbp = gapbp = 0
    for i, (c, o) in enumerate(scaffold[:-1]):
        c2, o2 = scaffold[i + 1]
        csize = contig2size[c]
        contacts = d[contig2indices[c][0], contig2indices[c2][0]]
        with np.errstate(divide="ignore"):
            distance = distance_func(contacts, *params) * 1000
        gapSize = distance - contig2size[c] // 2 - contig2size[c2] // 2
        if gapSize < mingap:
            gapSize = mingap
        elif gapSize > gapfrac * (contig2size[c] + contig2size[c2]):
            gapSize = gapfrac * (contig2size[c] + contig2size[c2])
        if gapSize > maxgap:
            gapSize = maxgap
        gapSize = int(round(gapSize))
        contigbp += csize
        gapbp += gapSize
        scaffold_with_gaps.append((c, o, gapSize))
    c, o = scaffold[-1]
    scaffold_with_gaps.append((c, o, 0))
    return scaffold_with_gaps, contigbp, gapbp


def contigs2scaffold(sam, contig_sizes, windowSize, contigs, mapq, minWindows, params):
    """bam2scaffolds.py:234-278 (``windowSize`` = [kb])."""
    windowSize, windows, chr2window, base2chr, genomeSize = fasta2windows(
        contig_sizes, windowSize, skipShorter=1, contigs=set(contigs), filterwindows=0)
    d = bam2arrays(windows, windowSize, chr2window, sam, mapq, regions=contigs)[0]
    bin_chr = np.array([c for c, s, e in windows[0]])
    if not d.shape[0]:
        return []
    if len(np.unique(bin_chr)) == 1:
        return [(bin_chr[0], 0, 0)]
    d += d.T - np.diag(d.diagonal())
    red = _average_reduce_2d(d, bin_chr)
    uniq = np.unique(bin_chr)
    Z = linkage_fastcluster(condense(distance_func(red + 1, *params)), "ward")
    scaffold = tree2scaffold(Z, uniq, d, bin_chr, minWindows)
    scaffold, contigbp, gapbp = scaffold2gaps(red, uniq, scaffold, params, dict(contig_sizes))
    return scaffold


def get_reverse_complement(seq):
    """bin/FastaIndex/FastaIndex.py get_reverse_complement: complement table over ACGT/acgt (+IUPAC
    left as is), reversed."""
    comp = str.maketrans("ACGTacgt", "TGCAtgca")
    return seq.translate(comp)[::-1]


def report_scaffolds(scaffolds, seqs, w=60):
    """bam2scaffolds.py:305-331: returns (FASTA text, scaffolds.tab text)."""
    fa, tab = [], []
    for i, scaffold in enumerate(scaffolds, 1):
        if not scaffold:
            continue
        parts, elements = [], []
        for c, o, gapSize in scaffold:
            parts.append(get_reverse_complement(seqs[c]) if o else seqs[c])
            parts.append('N' * gapSize)
            elements.append("%s-" % c if o else "%s+" % c)
        seq = "".join(parts)
        seq = "\n".join(seq[s:s + w] for s in range(0, len(seq), w))
        fa.append(">scaffold%s %s bp in %s contigs\n%s\n" % (i, len(seq), len(elements), seq))
        tab.append("%s\n" % "\t".join(elements))
    return "".join(fa), "".join(tab)


# ---------------------------------------------------------------------------------------------
# array2scaffolds.py: matrix cache loader (:43-121)
# ---------------------------------------------------------------------------------------------
def get_contig2size(bin_chr, bin_position):
    """array2scaffolds.py:43-49 (get_name = first word of the contig name, :181-182)."""
    contig2size = {c.split()[0]: 0 for c in np.unique(bin_chr)}
    for c, (s, e) in zip(bin_chr, bin_position):
        contig2size[c.split()[0]] += e - s
    return contig2size


def load_matrix(d, windows, chrs=(), remove_shorter=True, scaffolds=()):
    """array2scaffolds.py:51-121 after the file reads: ``d`` = the cached upper-triangular array,
    ``windows`` = rows of the .windows.tab.gz.  (:108-109 extend bin_chr by the RUNNING index count per
    scaffold -- a reference bug that makes bin_chr longer than the matrix for >1 scaffold; restated
    literally.)"""
    from collections import Counter
    if len(scaffolds):
        remove_shorter = True
    d = np.array(d)
    bin_chr = np.array([w[0] for w in windows])
    bin_position = np.array([[int(w[1]), int(w[2])] for w in windows])
    contig2size = get_contig2size(bin_chr, bin_position)
    if remove_shorter:
        c = Counter(np.diff(bin_position, axis=1)[:, 0].tolist())
        windowSize, occurencies = c.most_common(1)[0]
        valid_rowcols = ~(np.diff(bin_position, axis=1)[:, 0] != windowSize)
        d = d[valid_rowcols, :]
        d = d[:, valid_rowcols]
        bin_chr = bin_chr[valid_rowcols]
        bin_position = bin_position[valid_rowcols, :]
    if len(chrs):
        indices = np.any(bin_chr[None].T == list(chrs), 1)
        d = d[indices, :]
        d = d[:, indices]
        bin_chr = bin_chr[indices]
        bin_position = bin_position[indices, :]
    if len(scaffolds):
        contig2indices = get_contig2indices(bin_chr)
        indices, bin_chr, bin_position = [], [], []
        for i, scaffold in enumerate(scaffolds, 1):
            name = "scaffold%s" % i
            indices += get_indices(scaffold, contig2indices)
            bin_chr += [name] * len(indices)
            bin_position += [(s, s + windowSize) for s in range(0, windowSize * len(indices), windowSize)]
        d = d[:, indices][indices, :]
        bin_chr = np.array(bin_chr)
        bin_position = np.array(bin_position)
        contig2size = get_contig2size(bin_chr, bin_position)
    d += d.T
    d -= np.diag(d.diagonal() / 2)
    d = normalize_rows(d)
    return d, bin_chr, bin_position, contig2size
