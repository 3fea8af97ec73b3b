"""TEST INFRASTRUCTURE ONLY -- executes the REFERENCE's own Python sources (read from /root/reference where they lie,
never copied into the repo) so that the CPU restatement ``oracle/hicembler_oracle.py`` can be pinned against what the
reference itself computes.  Used by ``oracle/make_golden.py`` (fixtures under ``tests/golden/``) and by the live
checks of ``tests/test_oracle_pinned.py`` (skipped where /root/reference does not exist, i.e. on the GPU box).

The reference is Python 2.  ``array2scaffolds.py`` / ``FastaIndex.py`` parse under Python 3; ``bam2clusters.py`` and
``bam2scaffolds.py`` do not (print statements).  Each module is compiled from its source text after the shims below,
all applied IN MEMORY and all listed here -- nothing else is touched:

generic (every module)
  G1  ``print X`` statement            -> ``print(X)``                    (syntax only; output is discarded)
  G2  ``.iteritems()/.itervalues()``    -> ``.items()/.values()``          (same pairs; order = insertion order, the
                                                                             Py2 order is hash order: see DESIGN 8)
  G3  names injected into the module namespace: ``xrange = range``; ``map`` / ``filter`` / ``zip`` return lists
      (Py2 semantics); ``file`` = io.TextIOWrapper (``type(handle) is file``, FastaIndex.py:50)
  G4  ``gzip.open(fn)`` opens in text mode (Py2 strings are bytes)
  G5  ``except E, e:``                  -> ``except E as e:``              (syntax only)
line-anchored (Py2 integer ``/`` on ints -> ``//``; each entry is asserted to match the line it names)
  bam2clusters.py:123   (end-start)/2          get_window midpoint
  bam2clusters.py:125   pos / windowSize       get_window window index
  bam2clusters.py:192   seqlen/2               _bam2array mate midpoint
  bam2clusters.py:267   s / windowSize + 1     get_distances_contacts window count
  bam2clusters.py:280   (..) / 1000            get_distances_contacts kb distance
  bam2clusters.py:216   chr2window.keys()      -> list(...) (sliced at :226)
  bam2clusters.py:347   open(.., "w") for pickle -> "wb"
  bam2scaffolds.py:116  size/2                 fasta2windows base2chr (unused downstream)
  bam2scaffolds.py:159-160  n2/2, n1/2         join_scaffolds corner sizes
  bam2scaffolds.py:216  contig2size[c]/2       scaffold2gaps
  array2scaffolds.py:291 n2/2                  join_scaffolds (diagonal variant)
stand-ins for what the reference imports but this image lacks
  S1  ``samtools view`` (bam2clusters.py:108-117): ``_get_samtools_proc`` is replaced by an iterator over SAM TEXT
      lines that applies ``-q mapq -F 3980`` and the region list the same way (htslib semantics restated in
      ``samtools_view`` below); the reference loop then runs unchanged on those lines
  S2  ``fastcluster.linkage``: fastcluster is a PyPI dependency that is neither vendored nor installable here ->
      bound to the C restatement ``oracle/nnchain.c`` (fastcluster tie order stays "parity unpinned")
  S3  ``ete3``: ``oracle/ete3_min.py`` (documented subset);  S4 ``matplotlib`` / ``pysam``: inert stubs (figures are
      out of scope: no graphics stack);  S5 ``multiprocessing.Pool``: serial in-process stand-in (fork + CUDA context)
  S6  ``sinkhorn_knopp``: the vendored library itself with its one-line ``types`` shim (as in round 1)
"""
import builtins
import gzip as _gzip
import io
import os
import re
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("HC_REFERENCE_DIR", "/root/reference")

LINE_SHIMS = {
    "bam2clusters.py": [
        (123, "(end-start)/2", "(end-start)//2"),
        (125, "pos / windowSize", "pos // windowSize"),
        (192, "seqlen/2", "seqlen//2"),
        (216, "regions = chr2window.keys()", "regions = list(chr2window.keys())"),
        (267, "s / windowSize + 1", "s // windowSize + 1"),
        (280, "(positions[j]-positions[i]) / 1000", "(positions[j]-positions[i]) // 1000"),
        (347, 'open(outfn, "w")', 'open(outfn, "wb")'),
    ],
    "bam2scaffolds.py": [
        (116, "size/2", "size//2"),
        (159, "n2/2", "n2//2"),
        (160, "_d[:n1/2, n1:], _d[n1-n1/2:n1, n1:]", "_d[:n1//2, n1:], _d[n1-n1//2:n1, n1:]"),
        (216, "contig2size[c]/2 - contig2size[c2]/2", "contig2size[c]//2 - contig2size[c2]//2"),
    ],
    "array2scaffolds.py": [
        (291, "i = n2/2", "i = n2//2"),
    ],
}


def available():
    return os.path.isfile(os.path.join(REF, "bam2clusters.py"))


def samtools_view(lines, mapq=0, skipFlag=3980, regions=()):
    """``samtools view -q mapq -F skipFlag bam [regions]`` over SAM text (S1): keeps records with MAPQ >= mapq and
    (FLAG & skipFlag) == 0; with regions ("chr" or "chr:beg-end", 1-based inclusive) a record is printed once per
    region its alignment [POS, POS+reflen) overlaps, regions in the order given, records in file order within each
    (htslib iterates region by region)."""
    def reflen(cigar):
        if cigar == "*":
            return 1
        n = sum(int(k) for k, op in re.findall(r"(\d+)([MIDNSHP=X])", cigar) if op in "MDN=X")
        return max(n, 1)
    recs = []
    for l in lines:
        if not l or l[0] == "@":
            continue
        f = l.split("\t")
        if int(f[4]) < int(mapq) or (int(f[1]) & int(skipFlag)):
            continue
        recs.append((f[2], int(f[3]), int(f[3]) + reflen(f[5]) - 1, l))
    if not regions:
        for r in recs:
            yield r[3]
        return
    for reg in regions:
        chrom, _, span = reg.partition(":")
        if span:
            b, _, e = span.partition("-")
            b, e = int(b.replace(",", "")), int(e.replace(",", "")) if e else 1 << 60
        else:
            b, e = 1, 1 << 60
        for c, s, t, l in recs:
            if c == chrom and s <= e and t >= b:
                yield l


class _Proc(object):
    def __init__(self, it):
        self.stdout = it

    def terminate(self):
        pass


class _SerialPool(object):
    def __init__(self, *a, **k):
        pass

    def imap(self, f, it, *a, **k):
        return [f(x) for x in it]

    imap_unordered = imap
    map = imap

    def close(self):
        pass

    def join(self):
        pass


class _Inert(types.ModuleType):
    """module stub whose every attribute is a callable returning another inert object"""
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _InertObj()


class _InertObj(object):
    def __call__(self, *a, **k):
        return _InertObj()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _InertObj()

    def __iter__(self):
        return iter(())


def _vendored_sinkhorn():
    path = os.path.join(REF, "bin/sinkhorn_knopp/sinkhorn_knopp/sinkhorn_knopp.py")
    src = open(path).read().replace("from types import IntType, FloatType", "IntType, FloatType = int, float")
    mod = types.ModuleType("sinkhorn_knopp.sinkhorn_knopp")
    exec(compile(src, path, "exec"), mod.__dict__)
    pkg = types.ModuleType("sinkhorn_knopp")
    pkg.sinkhorn_knopp = mod
    return pkg, mod


def _fastcluster_stub():
    if HERE not in sys.path:
        sys.path.insert(0, HERE)
    import hicembler_oracle as O
    m = types.ModuleType("fastcluster")
    m.linkage = lambda y, method="single", metric="euclidean", preserve_input=True: O.linkage_fastcluster(
        np.asarray(y, dtype=np.float64), method)
    return m


def _shim_source(fname, src):
    lines = src.split("\n")
    for lineno, old, new in LINE_SHIMS.get(fname, []):
        assert old in lines[lineno - 1], "%s:%d does not contain %r (reference changed?)" % (fname, lineno, old)
        lines[lineno - 1] = lines[lineno - 1].replace(old, new)
    out = []
    for l in lines:
        if not l.lstrip().startswith("#"):
            l = re.sub(r"\bprint (?!\()(.+)$", r"print(\1)", l)                       # G1
        l = l.replace(".iteritems()", ".items()").replace(".itervalues()", ".values()")  # G2
        l = re.sub(r"\bexcept (\w+), (\w+):", r"except \1 as \2:", l)                    # G5
        out.append(l)
    return "\n".join(out)


_CACHE = {}


def load(fname, sam_sources=None):
    """Module object for /root/reference/<fname> after the shims above.  ``sam_sources``: dict path -> list of SAM
    text lines; ``_get_samtools_proc(path, mapq, regions)`` then iterates ``samtools_view`` over them (S1)."""
    if not available():
        raise RuntimeError("reference sources not found under %s" % REF)
    key = fname
    if key in _CACHE:
        mod = _CACHE[key]
    else:
        path = os.path.join(REF, fname)
        src = _shim_source(fname, open(path).read())
        mod = types.ModuleType("ref_" + fname[:-3])
        mod.__file__ = path
        ns = mod.__dict__
        ns.update(xrange=range, file=io.TextIOWrapper,                                     # G3
                  map=lambda f, *a: list(builtins.map(f, *a)), filter=lambda f, a: list(builtins.filter(f, a)),
                  zip=lambda *a: list(builtins.zip(*a)))
        pkg, skmod = _vendored_sinkhorn()
        stubs = {"matplotlib": _Inert("matplotlib"), "matplotlib.pyplot": _Inert("matplotlib.pyplot"),
                 "pysam": _Inert("pysam"), "commands": _Inert("commands"), "fastcluster": _fastcluster_stub(),
                 "sinkhorn_knopp": pkg, "sinkhorn_knopp.sinkhorn_knopp": skmod}
        import ete3_min
        stubs["ete3"] = ete3_min
        gz = types.ModuleType("gzip")
        gz.open = lambda fn, mode="r", *a, **k: _gzip.open(fn, mode if "b" in mode or "t" in mode else mode + "t", *a, **k)
        # modules this image lacks stay installed: the reference imports some of them at CALL time
        # (bam2clusters.py:39 sinkhorn_knopp, :384 ete3)
        for k in ("sinkhorn_knopp", "sinkhorn_knopp.sinkhorn_knopp", "ete3", "fastcluster"):
            sys.modules.setdefault(k, stubs[k])
        saved = {k: sys.modules.get(k) for k in list(stubs) + ["FastaIndex", "fastq2array", "bam2clusters"]}
        saved_path, saved_argv0 = list(sys.path), sys.argv[0]
        try:
            sys.modules.update(stubs)
            if fname != "FastaIndex.py":
                sys.modules["FastaIndex"] = load("FastaIndex.py")
            if fname == "array2scaffolds.py":
                fq = types.ModuleType("fastq2array")
                fq.logger = lambda *a, **k: None
                sys.modules["fastq2array"] = fq
            if fname == "bam2scaffolds.py":
                sys.modules["bam2clusters"] = load("bam2clusters.py")
            sys.argv[0] = path        # bam2clusters.py:30 derives its bin/ paths from argv[0]
            exec(compile(src, path, "exec"), ns)
        finally:
            sys.argv[0] = saved_argv0
            sys.path[:] = saved_path
            for k, v in saved.items():
                if v is None:
                    sys.modules.pop(k, None)
                else:
                    sys.modules[k] = v
        ns["gzip"] = gz                                                                     # G4
        if "Pool" in ns:
            ns["Pool"] = _SerialPool                                                        # S5
        if "logger" in ns:
            ns["logger"] = lambda *a, **k: None
        ns["print"] = lambda *a, **k: None
        _CACHE[key] = mod
    if sam_sources is not None and hasattr(mod, "_get_samtools_proc"):
        def _proc(bam, mapq=0, regions=[], skipFlag=3980):
            return _Proc(samtools_view(sam_sources[bam], mapq, skipFlag, regions))
        mod._get_samtools_proc = _proc
        for dep in ("bam2clusters",):
            pass
    return mod
