"""Synthetic genomes and power-law Hi-C read pairs.

Used by tests and bench.py to build inputs of the shapes BASELINE.json names; written with
torch ops so the same code produces small cases on the CPU and the 500 M-pair case directly in
HBM.  Data generation is never inside a timed region.
"""
from __future__ import annotations

import numpy as np
import torch

from .pairs import PAIR_DTYPE


class Genome:
    """Ordered contigs (file order == refID order == window order) grouped into chromosomes."""

    def __init__(self, sizes, chrom_of, window_bp):
        self.sizes = np.asarray(sizes, dtype=np.int64)
        self.chrom_of = np.asarray(chrom_of, dtype=np.int64)
        self.window_bp = int(window_bp)
        self.names = ["ctg%06d" % i for i in range(len(self.sizes))]
        self.starts = np.concatenate([[0], np.cumsum(self.sizes)[:-1]])
        self.total = int(self.sizes.sum())
        nchrom = int(self.chrom_of.max()) + 1
        self.chrom_start = np.array([self.starts[self.chrom_of == c][0] for c in range(nchrom)])
        self.chrom_end = np.array([(self.starts + self.sizes)[self.chrom_of == c][-1]
                                   for c in range(nchrom)])

    @property
    def contig_sizes(self):
        return list(zip(self.names, (int(s) for s in self.sizes)))

    def n_windows(self, window_bp=None):
        w = window_bp or self.window_bp
        return int(np.sum(-(-self.sizes // w)))

    def window_offsets(self, window_bp=None):
        """chr2window as an int64 array indexed by refID (all contigs present)."""
        w = window_bp or self.window_bp
        n = -(-self.sizes // w)
        return np.concatenate([[0], np.cumsum(n)[:-1]]).astype(np.int64)


def make_genome(n_contigs, n_windows, window_bp, n_chrom, seed=1, sigma=0.8, min_tail=1):
    """Contig lengths ~ lognormal(sigma) with every contig >= 2 windows and
    sum(ceil(size/window_bp)) == n_windows exactly; contiguous runs of contigs form chromosomes.
    The trailing (partial) window of every contig is min_tail .. window_bp bases long.  A trailing
    window shorter than half a read can never receive a read midpoint (get_window,
    bam2clusters.py:119-126): with the default min_tail=1 about n_contigs * 50 / window_bp windows are
    structurally empty, i.e. identical rows and exact distance ties; min_tail >= read length makes
    every window mappable."""
    rng = np.random.Generator(np.random.PCG64(seed))
    assert n_windows >= 2 * n_contigs
    w = rng.lognormal(0.0, sigma, n_contigs)
    extra = n_windows - 2 * n_contigs
    n = 2 + np.floor(extra * w / w.sum()).astype(np.int64)
    short = n_windows - int(n.sum())
    n[rng.choice(n_contigs, short, replace=False)] += 1  # floor leftovers: short < n_contigs
    sizes = n * window_bp - rng.integers(1, window_bp, n_contigs)
    if min_tail > 1:  # same draws as min_tail=1 (fixtures keep their values), tails clipped from below
        sizes = np.maximum(sizes, (n - 1) * window_bp + min(int(min_tail), window_bp))
    bounds = np.linspace(0, sizes.sum(), n_chrom + 1)[1:-1]
    chrom_of = np.searchsorted(bounds, np.cumsum(sizes) - sizes // 2, side="right")
    # make sure chromosome ids are contiguous 0..C-1 even if a chromosome came out empty
    _, chrom_of = np.unique(chrom_of, return_inverse=True)
    return Genome(sizes, chrom_of, window_bp)


def make_pairs(genome, n_pairs, seed=1, device="cpu", read_len=100, cis_frac=0.85, alpha=1.08,
               smin=1000, lowq_frac=0.10, with_mates=False, bad_flag_frac=0.0, sort=True,
               vary_len=False):
    """Return an int32 tensor [n, 6] of 24-byte pair records (see pairs.PAIR_DTYPE) on ``device``.

    85 % cis pairs with separation ~ s^-alpha on [smin, chromosome length], the rest uniform
    trans; MAPQ 60 for 90 %, uniform 0-9 otherwise; FLAG 99/83 (read 1).  ``with_mates`` adds the
    FLAG 147/163 mate records and ``bad_flag_frac`` marks that share duplicate/supplementary so
    the ``-F 3980`` filter has something to drop.  Sorted by (refID, pos) like a BAM."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    dev = torch.device(device)
    P = int(n_pairs)
    t = lambda a, dt=torch.int64: torch.as_tensor(np.asarray(a), dtype=dt, device=dev)
    cstart, csize = t(genome.starts), t(genome.sizes)
    chs, che = t(genome.chrom_start), t(genome.chrom_end)
    u = lambda: torch.rand(P, generator=g, device=dev, dtype=torch.float64)
    g1 = (u() * genome.total).long().clamp_(0, genome.total - 1)
    ch = torch.searchsorted(chs, g1, right=True) - 1
    lo, hi = chs[ch], che[ch]
    L = (hi - lo).double()
    e = 1.0 - alpha
    s = ((smin ** e) + u() * (L ** e - smin ** e)) ** (1.0 / e)
    s = s.long()
    up = g1 + s
    dn = g1 - s
    g2 = torch.where(u() < 0.5, up, dn)
    g2 = torch.where(g2 >= hi, dn, g2)
    g2 = torch.where(g2 < lo, up, g2)
    g2 = torch.minimum(torch.maximum(g2, lo), hi - 1)
    trans = u() >= cis_frac
    g2 = torch.where(trans, (u() * genome.total).long().clamp_(0, genome.total - 1), g2)
    del s, up, dn, L, lo, hi, ch, trans

    def locate(gpos):
        c = torch.searchsorted(cstart, gpos, right=True) - 1
        p = gpos - cstart[c]
        p = torch.minimum(p, csize[c] - read_len).clamp_(min=0)
        return c, p

    c1, p1 = locate(g1)
    c2, p2 = locate(g2)
    del g1, g2
    mapq = torch.where(u() < lowq_frac, (u() * 10).long(), torch.full((P,), 60, device=dev))
    flag = torch.where(u() < 0.5, torch.full((P,), 99, device=dev), torch.full((P,), 83, device=dev))
    lseq = torch.full((P,), read_len, device=dev, dtype=torch.int64)
    if vary_len:
        lseq = (read_len - (u() * (read_len // 2)).long())
    if with_mates:
        mflag = torch.where(flag == 99, torch.full_like(flag, 147), torch.full_like(flag, 163))
        c1, c2 = torch.cat([c1, c2]), torch.cat([c2, c1])
        p1, p2 = torch.cat([p1, p2]), torch.cat([p2, p1])
        mapq, flag, lseq = torch.cat([mapq, mapq]), torch.cat([flag, mflag]), torch.cat([lseq, lseq])
        P = 2 * P
    if bad_flag_frac > 0:
        r = torch.rand(P, generator=g, device=dev)
        flag = torch.where(r < bad_flag_frac / 2, flag | 1024, flag)
        flag = torch.where((r >= bad_flag_frac / 2) & (r < bad_flag_frac), flag | 2048, flag)
    if sort:
        key = (c1 << 32) | p1
        order = torch.argsort(key)
        del key
        c1, p1, c2, p2, mapq, flag, lseq = (x[order] for x in (c1, p1, c2, p2, mapq, flag, lseq))
        del order
    rec = torch.empty((P, 6), dtype=torch.int32, device=dev)
    rec[:, 0], rec[:, 1], rec[:, 2], rec[:, 3], rec[:, 4] = c1, p1, c2, p2, lseq
    rec[:, 5] = (flag | (mapq << 16)).to(torch.int32)
    return rec


def records_numpy(rec_tensor):
    """View an int32 [n,6] record tensor (CPU) as the structured PAIR_DTYPE array."""
    return rec_tensor.cpu().contiguous().numpy().view(PAIR_DTYPE).reshape(-1)


def records_tensor(rec_np, device="cpu"):
    a = np.ascontiguousarray(rec_np).view(np.int32).reshape(-1, 6)
    return torch.from_numpy(a).to(device)


# the BASELINE.json configurations as concrete shapes
CONFIGS = {
    "C1": dict(n_contigs=200, n_windows=2000, window_bp=10_000, n_chrom=5, n_pairs=1_000_000,
               method="ward"),
    "C2": dict(n_contigs=2000, n_windows=10_000, window_bp=50_000, n_chrom=20,
               n_pairs=50_000_000, method="average"),
    "C3": dict(n_contigs=10_000, n_windows=50_000, window_bp=60_000, n_chrom=23,
               n_pairs=500_000_000, method="ward"),
}

CONFIG_DESCRIPTIONS = {
    "C1": "synthetic 200-contig 20 Mb genome, 1M pairs, 10 kb windows (N=2000), ward",
    "C2": "synthetic 2k-contig 500 Mb assembly, 50M pairs, 50 kb windows (N=10000), average linkage",
    "C3": "synthetic 10k-contig 3 Gb human-like assembly, 500M pairs, 60 kb windows (N=50000), ward",
}
