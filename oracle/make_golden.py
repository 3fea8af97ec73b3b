"""Generates tests/golden/*.npz -- run in the BUILD container only (it imports the reference).

  python oracle/make_golden.py

Sources of truth:
  * sinkhorn_*.npz : the reference's own vendored library, /root/reference/bin/sinkhorn_knopp,
    imported unmodified except for the Python-3 shim of its line 2
    (``from types import IntType, FloatType`` -> ``IntType, FloatType = int, float``), applied in
    memory; plus ``normalize()`` of bam2clusters.py:36-49 evaluated with that library.
  * linkage_*.npz  : scipy.cluster.hierarchy.linkage, the reference's call at array2scaffolds.py:163.
  * binning.npz, reference_pinned.npz : the reference's OWN functions, executed from /root/reference by
    oracle/ref_exec.py (source compiled in memory after the documented Py2->Py3 shims; ``samtools view`` replaced
    by an iterator over SAM text) -- the restatement is asserted equal while the fixtures are written.
The committed fixtures travel to the GPU box; /root/reference does not.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
GOLD = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"


def vendored_sinkhorn():
    path = os.path.join(REF, "bin/sinkhorn_knopp/sinkhorn_knopp/sinkhorn_knopp.py")
    src = open(path).read().replace("from types import IntType, FloatType", "IntType, FloatType = int, float")
    mod = types.ModuleType("ref_sinkhorn_knopp")
    exec(compile(src, path, "exec"), mod.__dict__)
    return mod


def hic_like_counts(n, seed, scale=40.0, dtype=np.float32):
    """upper-triangular count matrix with power-law decay off the diagonal + noise"""
    rng = np.random.default_rng(seed)
    i, j = np.indices((n, n))
    lam = scale / (1.0 + np.abs(i - j)) ** 1.08 + 0.05
    m = rng.poisson(lam).astype(dtype)
    return np.triu(m)


def main():
    os.makedirs(GOLD, exist_ok=True)
    import hicembler_oracle as O
    sk_mod = vendored_sinkhorn()

    # --- Sinkhorn: vendored library on fixed inputs -------------------------------------------
    cases = {
        "doc": (np.array([[.011, .15], [1.71, .1]]), dict()),                  # sinkhorn_knopp.py:61-69
        "t2x2": (np.array([[1., 2.], [3., 4.]]), dict()),                      # tests/test_sinkhorn_knopp.py:60
        "t3x3": (np.array([[1.4, .2, 4], [3, 4, .7], [.4, 6, 1]]), dict(epsilon=1e-8)),  # :76-78
        "eye2": (np.eye(2), dict()),                                           # :31, :93
    }
    rng = np.random.default_rng(7)
    cases["rand64"] = (rng.random((64, 64)), dict(epsilon=1e-6))
    cases["rand97_f32"] = (rng.random((97, 97)).astype(np.float32), dict(epsilon=1e-5))
    out = {}
    for name, (P, kw) in cases.items():
        sk = sk_mod.SinkhornKnopp(**kw)
        res = sk.fit(P)
        out[name + "_P"] = P
        out[name + "_out"] = res
        out[name + "_iter"] = np.array(sk._iterations)
        out[name + "_stop"] = np.array(sk._stopping_condition)
        out[name + "_eps"] = np.array(kw.get("epsilon", 1e-3))
        # our restatement must agree with the vendored library bit for bit on the CPU
        sk2 = O.SinkhornKnopp(**kw)
        res2 = sk2.fit(P)
        assert np.array_equal(res, res2) and sk2._iterations == sk._iterations, name
    # sticky counters across fits (tests/test_sinkhorn_knopp.py:28-39)
    sk = sk_mod.SinkhornKnopp(max_iter=5)
    sk.fit(np.eye(2))
    a = (sk._stopping_condition, sk._iterations)
    sk.fit(np.array([[.011, .15], [1.71, .1]]))
    out["sticky"] = np.array([a[0], str(a[1]), sk._stopping_condition, str(sk._iterations)])
    # normalize() of bam2clusters.py:36-49 with the vendored library
    for n, seed, dt in [(150, 3, np.float32), (257, 4, np.float32), (120, 5, np.float64)]:
        d = hic_like_counts(n, seed, dtype=dt)
        d0 = d.copy()
        sk = sk_mod.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
        d += d.T - np.diag(d.diagonal())
        d += 1
        vmax = d.max()
        res = sk.fit(d / vmax)
        key = "norm%d_%s" % (n, np.dtype(dt).name)
        out[key + "_in"] = d0
        out[key + "_mutated"] = d
        out[key + "_out"] = res
        out[key + "_iter"] = np.array(sk._iterations)
        d1 = d0.copy()
        sk2 = O.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
        res2 = O.normalize(d1, sk=sk2)
        assert np.array_equal(res, res2) and np.array_equal(d1, d) and sk2._iterations == sk._iterations
    np.savez_compressed(os.path.join(GOLD, "sinkhorn.npz"), **out)

    # --- linkage: scipy on fixed inputs (tie-free and tie-heavy) ---------------------------------
    import scipy.cluster.hierarchy as sch
    out = {}
    for n, seed in [(40, 11), (200, 12), (777, 13)]:
        d = hic_like_counts(n, seed).astype(np.float64)
        d += d.T - np.diag(d.diagonal())
        dist = O.transform_distance(d, 0.5, 1.1)
        jit = dist + np.random.default_rng(seed).random(dist.shape) * 1e-3
        jit = np.triu(jit, 1) + np.triu(jit, 1).T
        for tag, D in (("ties", dist), ("jitter", jit)):
            y = O.condense(D)
            out["n%d_%s_y" % (n, tag)] = y
            for m in ("average", "ward", "complete", "weighted"):
                out["n%d_%s_%s_scipy" % (n, tag, m)] = sch.linkage(y, m)
                out["n%d_%s_%s_fc" % (n, tag, m)] = O.linkage_fastcluster(y, m)
    np.savez_compressed(os.path.join(GOLD, "linkage.npz"), **out)

    # --- binning: the REFERENCE'S OWN bam2arrays (bam2clusters.py:128-166, executed by oracle/ref_exec.py with the
    # documented Py2 shims) over SAM text; the restatement must reproduce it bit for bit ---------------------------
    import ref_exec as R
    from hicembler_b200 import synth
    from hicembler_b200.pairs import records_to_sam
    g = synth.make_genome(12, 60, 5000, 3, seed=5)
    rec = synth.records_numpy(synth.make_pairs(g, 4000, seed=5, with_mates=True, bad_flag_frac=0.02,
                                               vary_len=True))
    sam = records_to_sam(rec, g.names)
    ws_kb = [5, 10, 2]
    wsz, windows, chr2window, _, _ = O.fasta2windows(g.contig_sizes, ws_kb, minSize=2000, filterwindows=0)
    ref = R.load("bam2clusters.py", sam_sources={"a.bam": sam})
    arrays = ref.bam2arrays(windows, wsz, chr2window, ["a.bam"], 10, verbose=0)
    for a, b in zip(arrays, O.bam2arrays(windows, wsz, chr2window, [sam], 10)):
        assert np.array_equal(a, b)
    out = dict(records=rec.view(np.int32).reshape(-1, 6), sizes=g.sizes, window_bp=np.array(wsz))
    for ii, a in enumerate(arrays):
        out["array%d" % ii] = a
        out["offsets%d" % ii] = np.array([chr2window[ii].get(c, -1) for c in g.names], dtype=np.int32)
    arrays_upto = ref.bam2arrays(windows[:1], wsz[:1], chr2window[:1], ["a.bam"], 10, upto=1500, verbose=0)
    out["array0_upto1500"] = arrays_upto[0]
    np.savez_compressed(os.path.join(GOLD, "binning.npz"), **out)

    # --- reference-executed outputs of the elementwise / block functions (tests/test_oracle_pinned.py re-checks the
    # restatement against them where /root/reference does not exist; the GPU parity tests use the same inputs) -----
    b2c, b2s, a2s = ref, R.load("bam2scaffolds.py"), R.load("array2scaffolds.py")
    n = 96
    hic = hic_like_counts(n, 31)
    hic[[7, 50]] = 0
    hic[:, [7, 50]] = 0
    sym = hic + hic.T - np.diag(hic.diagonal())
    keys = np.array(["ctg%02d" % (k // 8) for k in range(n)])
    bp = np.array([(0, 5000) if k % 8 else (0, 3100) for k in range(n)])
    x = np.abs(np.random.default_rng(3).normal(3, 2, (40, 40))) + 0.01
    out = dict(hic_f32=hic, sym_f32=sym, keys=keys, bin_position=bp, dist_x=x,
               transform_logmax=a2s.transform(hic.copy()),
               transform_logmax_f64=a2s.transform(hic.astype(np.float64)),
               normalize_rows=a2s.normalize_rows(sym.copy()),
               average_reduce_2d=a2s._average_reduce_2d(sym, keys),
               distance_func=b2c.distance_func(x, 0.5, 1.1),
               distance_func_f32=b2c.distance_func(x.astype(np.float32) + np.float32(1), 0.5, 1.1),
               normalize_diagional=b2c.normalize_diagional(hic.copy(), keys, bp)[0],
               normalize_window_size=b2s.normalize_window_size(hic.copy(), keys, bp)[0])
    # contigs2scaffold-shaped case: join_scaffolds (both variants), tree2scaffold, get_clusters, newick
    rng = np.random.default_rng(3)
    nwin = rng.integers(1, 7, 9)
    bin_chr = np.array([("ctg%02d" % c) for c, k in enumerate(nwin) for _ in range(k)])
    m = len(bin_chr)
    i, j = np.indices((m, m))
    d = np.triu(rng.poisson(30.0 / (1.0 + np.abs(i - j)) + 0.2).astype(np.float32))
    d += d.T - np.diag(d.diagonal())
    names = list(np.unique(bin_chr))
    red = a2s._average_reduce_2d(d, bin_chr)
    import scipy.cluster.hierarchy as sch
    Z = sch.linkage((a2s.transform(red).astype(np.float64) + rng.random(red.shape) * 1e-9)[np.triu_indices(len(names), 1)], "ward")
    out.update(sc_d=d, sc_bin_chr=bin_chr, sc_Z=Z, sc_newick=np.array(a2s.distance_matrix2tree(Z, names).write()))
    for mw in (1, 2, 3):
        for tag, mod in (("b2s", b2s), ("a2s", a2s)):
            sc = mod.tree2scaffold(a2s.distance_matrix2tree(Z, names), d, bin_chr, None, mw)
            out["sc_%s_mw%d" % (tag, mw)] = np.array(["%s%s" % (c, "-" if o else "+") for c, o in sc])
    np.savez_compressed(os.path.join(GOLD, "reference_pinned.npz"), **out)
    print("golden vectors written to", GOLD)


if __name__ == "__main__":
    main()
