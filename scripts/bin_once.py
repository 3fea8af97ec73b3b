"""K1 alone on a synthetic config: CUDA-event time of hc_bin_pairs (for ncu captures and kernel experiments)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hicembler_b200 import ops, synth
cfgname = sys.argv[1] if len(sys.argv) > 1 else "C3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
cfg = synth.CONFIGS[cfgname]
dev = torch.device("cuda:0")
N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1, min_tail=2000)
rec = synth.make_pairs(g, P, seed=1, device=dev)
offsets = torch.as_tensor(g.window_offsets(), dtype=torch.int32, device=dev)
out = torch.empty((N, N), dtype=torch.int32, device=dev)
need = ops._lib.lib().hc_bin_pairs_workspace_bytes(P, N)
wsp = torch.empty(need, dtype=torch.uint8, device=dev)
ts = []
for r in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _, st = ops.bin_pairs(rec, offsets, ws, N, 10, out=out, workspace=wsp, return_stats=True)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("bin ms", ["%.3f" % t for t in ts], "checksum", int(out.sum(dtype=torch.int64).item()), st)
