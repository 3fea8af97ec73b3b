"""Shared builders for the parity tests."""
import numpy as np

from hicembler_b200 import synth
from hicembler_b200.pairs import records_to_sam


def hic_like_counts(n, seed, scale=40.0, dtype=np.float32):
    rng = np.random.default_rng(seed)
    i, j = np.indices((n, n))
    lam = scale / (1.0 + np.abs(i - j)) ** 1.08 + 0.05
    return np.triu(rng.poisson(lam).astype(dtype))


def small_case(n_contigs=12, n_windows=60, window_bp=5000, n_chrom=3, n_pairs=4000, seed=5, **kw):
    g = synth.make_genome(n_contigs, n_windows, window_bp, n_chrom, seed=seed)
    rec = synth.records_numpy(synth.make_pairs(g, n_pairs, seed=seed, **kw))
    return g, rec


def sam_lines(g, rec):
    return records_to_sam(rec, g.names)
