"""torchrun --nproc-per-node N scripts/sharded_check.py : one row-sharded pairs -> Z job over N real GPUs (cudaIpc peer
mappings, NVLink reads inside the kernels) checked against the single-GPU pipeline on rank 0."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from hicembler_b200 import synth
from hicembler_b200.parallel import ProcessGroup, RowShard
from hicembler_b200.pipeline import pairs_to_dendrogram, sharded_pairs_to_dendrogram

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
g = ProcessGroup()
n = int(os.environ.get("HC_CHECK_N", "4096"))
genome = synth.make_genome(max(8, n // 20), n, 5000, 5, seed=4, min_tail=2000)
rec = synth.make_pairs(genome, 400 * n, seed=4, device=dev)
off = torch.from_numpy(genome.window_offsets().astype(np.int32)).to(dev)
sh = RowShard(n, g)
out = sharded_pairs_to_dendrogram(sh, rec, off, genome.window_bp, balance=True, method="ward")
sym = sh.sym.clone()
single = pairs_to_dendrogram(rec, off, genome.window_bp, n, balance=True, keep=True, method="ward")
ok = torch.equal(sym, single["sym"][sh.row0:sh.row1]) and out["sk_iterations"] == single["sk_iterations"]
if g.rank == 0:
    Z, want = out["Z"], single["Z"]
    ok = ok and np.array_equal(Z[:, [0, 1, 3]], want[:, [0, 1, 3]]) and np.allclose(Z[:, 2], want[:, 2], rtol=1e-9, atol=0)
flag = torch.tensor([0.0 if ok else 1.0], device=dev)
dist.all_reduce(flag)
sh.close()
if g.rank == 0:
    print("sharded check ok" if flag.item() == 0 else "sharded check FAILED", "ranks", g.world, "n", n, "iterations", out["sk_iterations"],
          "engine", out["linkage_stats"].get("engine"))
dist.destroy_process_group()
sys.exit(0 if flag.item() == 0 else 1)
