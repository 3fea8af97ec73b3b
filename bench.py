#!/usr/bin/env python
"""bench.py -- pairs -> dendrogram on synthetic power-law Hi-C data (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C3] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of P synthetic read pairs:
  bin_pairs -> fold(+1,max) -> Sinkhorn-Knopp -> distance transform -> NN-chain linkage -> Z.
`value`  : pairs/s with the 24-byte records already resident in HBM.
`e2e`    : the same through the public call with HOST (pinned) records: H2D copy of the records and
           the linkage array coming back to the host are inside the timed region.
N > 1    : the single-dendrogram linkage does not shard ("replicas only"): every rank runs
           an independent job of the same shape (weak scaling, no data-path collective).  The
           row-sharded Sinkhorn with its NCCL all-reduce is measured by --mode sinkhorn.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "read pairs/s, BAM pairs -> dendrogram (bin + Sinkhorn-Knopp + distance + linkage) at 50k windows"
UNIT = "pairs/s"
PARAMS = (0.5, 1.1)   # distance_func parameters (a, b); the reference fits them per data set


# DRAM traffic of all launches of a kernel in one step, from the committed ncu pass
# (profiles/r01_ncu_dram_kr_heavy_C3.csv: dram__bytes_read.sum + dram__bytes_write.sum per launch)
NCU_TRAFFIC = {}
_p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
if os.path.exists(_p):
    NCU_TRAFFIC = json.load(open(_p))


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


# --------------------------------------------------------------------------------------------
# CPU reference arm (the oracle port; the one place besides tests allowed to run oracle/)
# --------------------------------------------------------------------------------------------
def cpu_reference_sample(cfg, seed=1, n_sample_pairs=1_000_000, n_small=4000):
    """Times the reference's CPU path on a bounded sample of the workload and extrapolates with the
    stated laws: binning is linear in P (the reference loop is a single process: `threads` is unused,
    bam2clusters.py:128); Sinkhorn (matvec form, O(N^2) per iteration -- favourable to the CPU: the
    literal D1.P.D2 form is O(N^3)), transform+condense and linkage scale as N^2."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import hicembler_oracle as O
    import scipy.cluster.hierarchy as sch
    from hicembler_b200 import synth
    from hicembler_b200.pairs import records_to_sam
    N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
    # (1) binning: literal per-line loop over SAM text, same window table as the GPU arm
    g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=seed, min_tail=cfg.get("min_tail", 1))
    S = min(P, n_sample_pairs)
    rec = synth.records_numpy(synth.make_pairs(g, S, seed=seed))
    lines = records_to_sam(rec, g.names)
    wsz, windows, chr2window, _, _ = O.fasta2windows(g.contig_sizes, [ws // 1000], minSize=2000, filterwindows=0)
    t0 = time.perf_counter()
    arr = O.bam2arrays(windows, wsz, chr2window, [lines], 10)[0]
    t_bin = time.perf_counter() - t0
    del arr
    # (2) Sinkhorn + (3) transform/condense + (4) linkage on an n_small-window Hi-C-like matrix
    n = min(N, n_small)
    gs = synth.make_genome(max(2, n // 20), n, ws, max(1, min(cfg["n_chrom"], n // 40)), seed=seed)
    recs = synth.records_numpy(synth.make_pairs(gs, int(2_000 * n), seed=seed))
    d = O.bin_records(recs, ws, gs.window_offsets(), n, 10)
    t0 = time.perf_counter()
    sk = O.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
    bal = O.normalize(d, sk=sk, materialize=False)
    t_sk = time.perf_counter() - t0
    t0 = time.perf_counter()
    y = O.condense(O.distance_func(bal + 1, *PARAMS))
    t_tr = time.perf_counter() - t0
    t0 = time.perf_counter()
    sch.linkage(y, cfg["method"])   # the reference's own library call (array2scaffolds.py:163)
    t_link = time.perf_counter() - t0
    f = (N / n) ** 2
    total = t_bin * (P / S) + (t_sk + t_tr + t_link) * f
    return dict(value=P / total, seconds_extrapolated=total,
                parts=dict(bin_s=t_bin, bin_pairs=S, sinkhorn_s=t_sk, sk_iterations=sk._iterations,
                           transform_condense_s=t_tr, linkage_s=t_link, n_small=n),
                sample=("%d-pair prefix binned by the literal bam2arrays loop (single process, as the reference), "
                        "linear in P; Sinkhorn (matvec form), distance_func+condense and scipy %s linkage on a "
                        "%d-window matrix, scaled by (N/n)^2 = %.0f" % (S, cfg["method"], n, f)))


def run_reference(args, cfg, rank, world):
    if rank != 0:
        return
    res = None
    t_all = []
    for _ in range(max(1, min(args.steps, 2))):
        t0 = time.perf_counter()
        res = cpu_reference_sample(cfg)
        t_all.append(time.perf_counter() - t0)
    cores = os.cpu_count()
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * res["seconds_extrapolated"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 counts / f64",
            "data": "synthetic", "config": {"workload": cfg["name"], **{k: cfg[k] for k in ("n_windows", "n_pairs", "window_bp", "method")}},
            "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": res["sample"], "parts": res["parts"]},
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "sample_wall_s": t_all}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def run_ours(args, cfg, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from hicembler_b200 import ops, synth
    from hicembler_b200.pipeline import StageTimer, pairs_to_dendrogram

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
    g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1 + rank, min_tail=cfg.get("min_tail", 1))
    records = synth.make_pairs(g, P, seed=1 + rank, device=dev)
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    offsets = torch.from_numpy(g.window_offsets().astype(np.int32)).to(dev)
    workspace = torch.empty(int(ops._lib.lib().hc_bin_pairs_workspace_bytes(P, N)), dtype=torch.uint8, device=dev)
    launches = {"n": 0}

    def step(rec, timer=None):
        return pairs_to_dendrogram(rec, offsets, ws, N, mapq=10, params=PARAMS, method=cfg["method"],
                                   balance=not args.no_balance, timer=timer, workspace=workspace, engine=args.engine)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        res = step(records)
    # ---- timed: device-resident ----
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    stage_ms = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        timer = StageTimer(True)
        res = step(records, timer)
        for k, v in timer.result_ms().items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1)
    # ---- timed: end to end from pinned host records (16-byte wire form by default, --wire 24 for hc_pair_t) ----
    from hicembler_b200.pipeline import host_pairs_to_dendrogram
    from hicembler_b200.pairs import pack_pairs16
    if args.wire == 16:
        packed = pack_pairs16(records)                      # (outside the timed region: the decoder's output)
        host_wire = torch.empty(packed.shape, dtype=torch.int32, pin_memory=True)
        host_wire.copy_(packed)
        del packed
    else:
        host_wire = torch.empty(records.shape, dtype=records.dtype, pin_memory=True)
        host_wire.copy_(records)
    staging = {"rec24": torch.empty_like(records)}
    torch.cuda.synchronize()

    def step_e2e():
        return host_pairs_to_dendrogram(host_wire, offsets, ws, N, staging=staging, mapq=10, params=PARAMS,
                                        method=cfg["method"], balance=not args.no_balance, workspace=workspace,
                                        engine=args.engine)
    for _ in range(min(args.warmup, 1)):
        step_e2e()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for _ in range(args.steps):
        res = step_e2e()             # res["Z"] is a host numpy array: D2H is inside the linkage call
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop() if sampler else None
    # one more (untimed) step with CUDA events around every launch of the round-parallel linkage: the per-kernel
    # times behind `roofline` (events are recorded on the launching stream, no synchronisation is added)
    kernel_us = None
    if rank == 0:
        os.environ["HC_ROUNDS_KTIME"] = "1"
        try:
            kernel_us = step(records)["linkage_stats"].get("kernel_us")
        finally:
            os.environ.pop("HC_ROUNDS_KTIME", None)
    t = torch.tensor([ms_dev, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e = t.tolist()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    K = args.steps
    peak, peak_src = load_peaks()
    stage_ms = {k: v / K for k, v in stage_ms.items()}
    lst = res["linkage_stats"]
    sk_it = res.get("sk_iterations", 0)
    # algorithmic bytes per launch (DESIGN.md section "Kernels and rooflines")
    sk_passes = res.get("sk_passes", 2 * sk_it + 4) if not args.no_balance else 0      # matrix reads of 4 N^2 bytes each
    bytes_alg = {
        "bin_pairs": 24.0 * P + 4.0 * N * N,
        "fold": 8.0 * N * N,
        "sinkhorn": sk_passes * 4.0 * N * N,
        "transform": 4.0 * N * (N - 1) / 2 + 8.0 * N * N,
        # serial NN-chain: one f64 row per NN search, two rows read + one written per Lance-Williams update;
        # round-parallel engine: 8 B per live column of a rescan, 24 B per live column of a merged-row build
        # and per (row, pair) update (two reads, one write) -- counted by the kernels themselves
        "linkage": (8.0 * lst["rescan_cols"] + 24.0 * lst["build_cols"] + 24.0 * lst["light_updates"])
        if lst.get("engine") == "rounds" else (8.0 * lst["scan_active"] + 24.0 * lst["update_active"]),
    }
    roof = {}
    for k, b in bytes_alg.items():
        ms = stage_ms.get(k)
        if ms and b:
            a = b / (ms * 1e-3) / 1e9
            roof[k] = {"bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                       "traffic": None, "ms": ms, "algorithmic_bytes": b, "share_of_step": ms / (ms_dev / K)}
    dominant = max(roof, key=lambda k: roof[k]["ms"])
    main_roof = dict(roof[dominant], kernel={"bin_pairs": "k_bin_tiles", "fold": "k_tilepair64",
                                             "sinkhorn": "k_colpass+k_rowpass", "transform": "k_transform_tile64",
                                             "linkage": "kr_light+kr_heavy" if lst.get("engine") == "rounds" else "k_nnchain"}[dominant],
                     peak_source=peak_src)
    if dominant == "linkage" and lst.get("engine") == "rounds" and kernel_us and kernel_us.get("heavy"):
        # dominant KERNEL = kr_heavy (merged-row builds + exact rescans + the initial scan) with kb_rebuild, the same
        # work done as one merge+compact pass in the bulk rounds: all their launches of one
        # step; achieved = algorithmic bytes of those launches / their summed CUDA-event durations
        heavy_ms = (kernel_us["heavy"] + kernel_us["initial_scan"]) * 1e-3
        b = 8.0 * lst["rescan_cols"] + 24.0 * lst["build_cols"]
        a_ = b / (heavy_ms * 1e-3) / 1e9
        launches = lst["rounds"] + 1
        ncu = NCU_TRAFFIC.get(args.config, {}).get("kr_heavy")
        main_roof = {"bound": "hbm", "achieved": a_, "peak": peak, "unit": "GB/s", "frac": a_ / peak,
                     "traffic": (ncu["dram_bytes"] / ncu["launches"]) if ncu else None,
                     "kernel": "kr_heavy+kb_rebuild", "launches_per_step": launches, "ms": heavy_ms,
                     "avg_launch_us": heavy_ms * 1e3 / launches, "algorithmic_bytes": b,
                     "algorithmic_bytes_per_launch": b / launches, "share_of_step": heavy_ms / (ms_dev / K),
                     "peak_source": peak_src,
                     "kernel_us_per_step": kernel_us,
                     "note": "latency-bound in the %d rounds of the chain phase (1-4 merges per round); the bulk rounds "
                             "and the initial scan move most of the bytes, see profiles/" % lst["rounds"]}
    # launches per step: bin 8, fold 1, sinkhorn (colpass+finish)=2/pass ... counted from the call graph
    sk_launches = 0 if args.no_balance else res["sk_launches"]      # counted by ops._CudaSinkhornPasses
    link_launches = lst["launches"] if lst.get("engine") == "rounds" else 1     # counted by hc_linkage_rounds
    gpu_launches = K * 2 * (8 + 1 + sk_launches + 1 + link_launches) + (K if args.wire == 16 else 0)  # + unpack
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference_sample(cfg)
        cpu = {"value": c["value"], "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": c["sample"],
               "parts": c["parts"]}
    total_pairs = float(P) * world * K
    line = {
        "metric": METRIC, "value": total_pairs / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 counts, f32 matrix, f64 Sinkhorn vectors / distances", "data": "synthetic",
        "config": {"workload": cfg["name"], "n_contigs": cfg["n_contigs"], "n_windows": N, "n_pairs": P,
                   "window_bp": ws, "method": cfg["method"], "balance": not args.no_balance,
                   "l2": "inputs larger than L2 (records %.1f GB, matrix %.1f GB)" % (24e-9 * P, 4e-9 * N * N),
                   "min_tail_bp": cfg.get("min_tail", 1),
                   "linkage_engine": lst.get("engine"),
                   "parallelism": "replicas x%d" % world if world > 1 else "1 gpu"},
        "e2e": {"value": total_pairs / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(args.wire * P),
                "d2h_bytes_per_step": int((N - 1) * 32 + 64 + 64), "ms_per_step": ms_e2e / K,
                "host_record_bytes": args.wire,
                "api": "hicembler_b200.pipeline.host_pairs_to_dendrogram (pinned host records -> Z on the host)"},
        "gpu_launches": gpu_launches,
        "clocks": clocks,
        "roofline": main_roof,
        "stage_rooflines": roof,
        "stage_ms": stage_ms,
        "seconds_pairs_to_dendrogram": ms_dev / K * 1e-3,
        "binning_pairs_per_s": P / (stage_ms["bin_pairs"] * 1e-3) if "bin_pairs" in stage_ms else None,
        "sinkhorn": {"iterations": sk_it, "stop": res.get("sk_stop")},
        "linkage": lst,
        "bin_stats": res["stats"],
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_sinkhorn_mode(args, cfg, rank, world, local_rank):
    """--mode sinkhorn: row-sharded Sinkhorn-Knopp (BASELINE configs 3 and 5): every rank owns a row
    block of an N x N float32 matrix generated on the device; column sums are all-reduced over NCCL."""
    import torch
    import torch.distributed as dist
    from hicembler_b200 import ops
    from hicembler_b200.parallel import row_block, make_allreduce

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N = args.n or cfg["n_windows"]
    r0, r1 = row_block(N, rank, world)
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    # Poisson band + background on the row block (symmetric by construction: value depends on |i-j|
    # and a symmetric hash), +1 pseudo-count
    i = torch.arange(r0, r1, device=dev, dtype=torch.float32)[:, None]
    X = torch.empty((r1 - r0, N), dtype=torch.float32, device=dev)
    step_rows = max(1, (1 << 28) // N)
    for s in range(0, r1 - r0, step_rows):
        j = torch.arange(N, device=dev, dtype=torch.float32)[None, :]
        ii = i[s:s + step_rows]
        lam = 40.0 / (1.0 + (ii - j).abs()) ** 1.08 + 0.05
        h = torch.frac(torch.sin((ii * j + ii + j) * 12.9898) * 43758.5453).abs()
        X[s:s + step_rows] = torch.floor(lam * (0.5 + h)) + 1.0
    vmax = X.max()
    if world > 1:
        dist.all_reduce(vmax, op=dist.ReduceOp.MAX)
    vmax = float(vmax.item())
    allreduce = make_allreduce(dist.group.WORLD) if world > 1 else None
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        res = ops.sinkhorn_fit(X, vmax, 1000, 1e-5, allreduce=allreduce)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        res = ops.sinkhorn_fit(X, vmax, 1000, 1e-5, allreduce=allreduce)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = t.item() / args.steps
        passes = res.matrix_passes
        peak, src = load_peaks()
        gbs = passes * 4.0 * N * N / (ms * 1e-3) / 1e9
        print(json.dumps({"mode": "sinkhorn", "n_windows": N, "n_gpus": world, "iterations": res.iterations,
                          "stop": res.stopping_condition, "ms_per_fit": ms, "ms_per_iteration": ms / passes * 2,
                          "aggregate_gbs": gbs, "frac_of_peak_per_gpu": gbs / world / peak, "peak_source": src}), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=os.environ.get("HC_BENCH_CONFIG", "C3"))
    ap.add_argument("--mode", default="pipeline", choices=["pipeline", "sinkhorn"])
    ap.add_argument("--n", type=int, default=0, help="matrix size for --mode sinkhorn")
    ap.add_argument("--no-balance", action="store_true", help="reference HEAD behaviour: normalize() off")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--wire", type=int, default=16, choices=[16, 24],
                    help="bytes per host record in the e2e leg: 16 = hc_pair16_t (midpoints, unpacked on the device), "
                         "24 = hc_pair_t")
    ap.add_argument("--engine", default="auto", choices=["auto", "exact", "rounds"],
                    help="linkage engine: round-parallel with certification (auto falls back to the serial NN-chain on ties)")
    ap.add_argument("--min-tail", type=int, default=2000,
                    help="shortest trailing (partial) window of a synthetic contig in bp.  Default 2000 = the reference's own "
                         "shortest-sequence cutoff (minSize, bam2clusters.py:459).  1 leaves ~n_contigs*50/window_bp windows "
                         "that no read midpoint can reach: identical empty rows = exact distance ties, which the "
                         "round-parallel linkage refuses to certify, so the serial NN-chain engine runs instead")
    args = ap.parse_args()
    from hicembler_b200 import synth
    cfg = dict(synth.CONFIGS[args.config], name=args.config + ": " + synth.CONFIG_DESCRIPTIONS[args.config],
               min_tail=args.min_tail)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, cfg, rank, world)
    if args.mode == "sinkhorn":
        return run_sinkhorn_mode(args, cfg, rank, world, local_rank)
    run_ours(args, cfg, rank, world, local_rank)


if __name__ == "__main__":
    main()
