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


def get_name(contig):
    """array2scaffolds.py:181-182"""
    return contig.split()[0]


def get_contig2size(bin_chr, bin_position):
    """Return contig2size (sum of window lengths per contig) -- array2scaffolds.py:43-49."""
    contig2size = {}
    for c, (s, e) in zip(bin_chr, bin_position):
        contig2size[get_name(c)] = contig2size.get(get_name(c), 0) + int(e - s)
    return contig2size


def load_matrix(fname, chrs=[], remove_shorter=True, scaffolds=[], verbose=0, device_out=False):
    """Load Hi-C interaction matrix from the ``.npz`` + ``.windows.tab.gz`` cache -- array2scaffolds.py:51-121.

    Returns (d, bin_chr, bin_position, contig2size).  The row/column selections (most common window size
    :90-99, chromosomes :102-107, scaffold order :110-117) are composed on the host into ONE index vector and
    applied by a single device gather; symmetrisation (:117-118) and ``normalize_rows`` (:119) run on the
    device too.  The reference's running ``len(indices)`` at :114-115 over-extends bin_chr / bin_position when
    more than one scaffold is given; that is reproduced."""
    import gzip
    import sys
    from collections import Counter
    from .bam2scaffolds import get_contig2indices, get_indices
    if scaffolds:
        remove_shorter = True
    npy = np.load(fname)
    d = npy[npy.files[0]]
    bin_chr, bin_position = [], []
    with gzip.open(fname[:-4] + '.windows.tab.gz', "rt") as fh:
        for l in fh:
            chrom, start, end = l[:-1].split('\t')
            bin_chr.append(chrom)
            bin_position.append((int(start), int(end)))
    bin_position = np.array(bin_position)
    bin_chr = np.array(bin_chr)
    contig2size = get_contig2size(bin_chr, bin_position)
    keep = np.arange(len(bin_chr))
    if remove_shorter:
        sizes = bin_position[:, 1] - bin_position[:, 0]
        windowSize, occurencies = Counter(sizes.tolist()).most_common(1)[0]
        if verbose:
            sys.stderr.write(" most common window: %s bp [%5.2f%s]\n" % (windowSize, occurencies * 100. / len(bin_chr), '%'))
        keep = keep[sizes == windowSize]
    if len(chrs):
        keep = keep[np.isin(bin_chr[keep], list(chrs))]
    bin_chr, bin_position = bin_chr[keep], bin_position[keep]
    if scaffolds:
        contig2indices = get_contig2indices(bin_chr)
        indices, names, positions = [], [], []
        for i, scaffold in enumerate(scaffolds, 1):
            indices += get_indices(scaffold, contig2indices)
            names += ["scaffold%s" % i] * len(indices)
            positions += [(s, s + windowSize) for s in range(0, windowSize * len(indices), windowSize)]
        keep = keep[np.asarray(indices, dtype=np.int64)]
        bin_chr, bin_position = np.array(names), np.array(positions)
        contig2size = get_contig2size(bin_chr, bin_position)
    X = torch.from_numpy(np.ascontiguousarray(d, dtype=np.float32)).cuda()
    X = ops.gather_square(X, torch.from_numpy(keep.astype(np.int32)).cuda())
    if X.shape[0]:
        ops.symmetrize_(X)
        ops.normalize_rows_(X)
    return (X if device_out else X.cpu().numpy()), bin_chr, bin_position, contig2size
