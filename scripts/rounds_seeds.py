"""Certification rate and stage times of the round-parallel linkage over synthetic seeds (GPU box).
usage: python scripts/rounds_seeds.py C3 1 2 3 ...   -> one JSON line per seed"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hicembler_b200 import synth
from hicembler_b200.pipeline import pairs_to_dendrogram, StageTimer

cfgname = sys.argv[1]
seeds = [int(x) for x in sys.argv[2:]] or [1]
cfg = synth.CONFIGS[cfgname]
dev = torch.device("cuda:0")
N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
for seed in seeds:
    g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=seed, min_tail=2000)
    rec = synth.make_pairs(g, P, seed=seed, device=dev)
    offsets = torch.as_tensor(g.window_offsets(), dtype=torch.int32, device=dev)
    torch.cuda.synchronize(); torch.cuda.empty_cache()
    for rep in range(2):
        t = StageTimer(True)
        out = pairs_to_dendrogram(rec, offsets, ws, N, mapq=10, params=(0.5, 1.1), method=cfg["method"],
                                  balance=True, timer=t)
        ms = t.result_ms()
    st = out["linkage_stats"]
    fb = st.get("fallback_from_rounds")
    print(json.dumps(dict(config=cfgname, seed=seed, engine=st["engine"], stage_ms=ms,
                          rounds=(fb or st).get("rounds"), pending=(fb or st).get("pending_near_ties"),
                          near_heights=(fb or st).get("near_tied_heights"),
                          reason=(fb or {}).get("reason"))), flush=True)
    del rec, out
    torch.cuda.empty_cache()
