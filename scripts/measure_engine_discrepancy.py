"""Round-parallel vs serial NN-chain on the bench workloads: is the merge order identical, how far apart
are the heights, how close were the closest decisions?  (HC_ROUNDS_DIAG=1 lets the round-parallel engine
finish even when it cannot certify.)  Usage: python scripts/measure_engine_discrepancy.py C3 [min_tail]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["HC_ROUNDS_DIAG"] = "1"
import numpy as np  # noqa: E402
import torch  # noqa: E402
from hicembler_b200 import ops, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C2"
min_tail = int(sys.argv[2]) if len(sys.argv) > 2 else 100
cfg = synth.CONFIGS[name]
N, P, ws, method = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"], cfg["method"]
g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1, min_tail=min_tail)
rec = synth.make_pairs(g, P, seed=1, device="cuda")
off = torch.from_numpy(g.window_offsets().astype(np.int32)).cuda()
counts = ops.bin_pairs(rec, off, ws, N, 10)
del rec
S1, vmax = ops.fold_counts(counts, "sym", add=1.0, inplace=True)
vm = float(vmax.item())
res = ops.sinkhorn_fit(S1, vm, 1000, 1e-5)


def make():
    return ops.transform_to_working(S1, ops.KIND_POWER, 0.5, 1.1, res.r, res.c, vm, square=(method == "ward"))


W, fl = make()
Zr, st = ops.linkage_rounds(W, method, nan_flags=fl)
cert = Zr is not None
if Zr is None:
    Zr = st.pop("Z_uncertified")
del W
W, fl = make()
Ze = ops.linkage_working(W, method, nan_flags=fl)
same = np.array_equal(Zr[:, [0, 1, 3]], Ze[:, [0, 1, 3]])
rel = np.abs(Zr[:, 2] / Ze[:, 2] - 1)
hs = np.sort(Ze[:, 2])
gap = np.min(np.diff(hs) / hs[1:])
out = dict(config=name, min_tail=min_tail, method=method, certified=cert, merge_order_identical=bool(same),
           rows_differing=int((Zr[:, [0, 1, 3]] != Ze[:, [0, 1, 3]]).any(1).sum()),
           max_rel_height_diff=float(rel.max()), p99_rel_height_diff=float(np.quantile(rel, 0.99)),
           min_rel_gap_sorted_heights=float(gap), stats={k: v for k, v in st.items() if k != "reason"},
           reason=st.get("reason"))
print(json.dumps(out))
