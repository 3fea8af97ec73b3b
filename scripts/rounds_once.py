"""One pairs->dendrogram pass on a synthetic config (for ncu captures of the linkage kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hicembler_b200 import synth
from hicembler_b200.pipeline import pairs_to_dendrogram
cfgname = sys.argv[1]
cfg = synth.CONFIGS[cfgname]
dev = torch.device("cuda:0")
N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1, min_tail=2000)
rec = synth.make_pairs(g, P, seed=1, device=dev)
offsets = torch.as_tensor(g.window_offsets(), dtype=torch.int32, device=dev)
out = pairs_to_dendrogram(rec, offsets, ws, N, mapq=10, params=(0.5, 1.1), method=cfg["method"], balance=True)
print(out["linkage_stats"])
