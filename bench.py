#!/usr/bin/env python
"""bench.py -- pairs -> dendrogram on synthetic power-law Hi-C data (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C3] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of P synthetic read pairs:
  bin_pairs -> fold(+1,max) -> Sinkhorn-Knopp -> distance transform -> NN-chain linkage -> Z.
`value`  : pairs/s with the 24-byte records already resident in HBM.
`e2e`    : the same through the public call with HOST (pinned) records: H2D copy of the records and
           the linkage array coming back to the host are inside the timed region.
N > 1    : ONE job over N GPUs (strong scaling): matrix build, fold and Sinkhorn-Knopp are sharded by row block
           (every rank bins the slice of the record stream that falls into its rows; host records cross each rank's
           own PCIe link), the exchanges are peer-memory reads over NVLink inside the kernels, and the dendrogram --
           which does not shard ("replicas only") -- is built on rank 0.  --mode replicas runs N
           independent jobs instead (weak scaling, round-1 behaviour); --mode sinkhorn the row-sharded fit alone.
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
    # the LITERAL normalize() of the reference (O(N^3): two dense GEMMs with diagonal matrices per iteration,
    # sinkhorn_knopp.py:146-148) at two small sizes, next to the matvec form used for the extrapolation
    literal = {}
    for nl in (2000, 4000):
        if nl > n:
            continue
        dl = d[:nl, :nl].copy()
        t0 = time.perf_counter()
        skl = O.SinkhornKnopp(max_iter=1000, epsilon=0.00001)
        O.normalize(dl, sk=skl, materialize=True)
        literal[str(nl)] = {"seconds": time.perf_counter() - t0, "iterations": skl._iterations}
    f = (N / n) ** 2
    total = t_bin * (P / S) + (t_sk + t_tr + t_link) * f
    return dict(value=P / total, seconds_extrapolated=total,
                parts=dict(bin_s=t_bin, bin_pairs=S, sinkhorn_s=t_sk, sk_iterations=sk._iterations,
                           transform_condense_s=t_tr, linkage_s=t_link, n_small=n,
                           sinkhorn_literal_D1PD2=literal,
                           sinkhorn_literal_note="O(N^3) per iteration: ~%.0f h per fit at N=%d by the N^3 law; not used in the extrapolation"
                                                 % ((literal.get("4000", literal.get("2000", {"seconds": 0}))["seconds"]
                                                     * (N / (4000.0 if "4000" in literal else 2000.0)) ** 3) / 3600.0, N)),
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
            "extrapolated": True,
            "extrapolation": "value and ms_per_step are EXTRAPOLATED from the bounded sample described in cpu_baseline.sample "
                             "(linear in pairs for the binning loop, (N/n)^2 for the matrix stages); measured_wall_s is what the "
                             "sample itself took",
            "measured_wall_s": t_all, "sample_wall_s": t_all}
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
    main_roof = dict(roof[dominant], kernel={"bin_pairs": "k_bin_band", "fold": "k_fold_counts",
                                             "sinkhorn": "k_sympass+k_symfinish", "transform": "k_transform_bal64",
                                             "linkage": "kr_light+kr_heavy" if lst.get("engine") == "rounds" else "k_nnchain"}[dominant],
                     peak_source=peak_src)
    if dominant == "linkage" and lst.get("engine") == "rounds" and kernel_us and kernel_us.get("rebuild"):
        # dominant KERNEL of the dominant stage = kb_rebuild3 (the bulk rounds: merge + compact in one pass over the
        # matrix): all its launches of one step; achieved = algorithmic bytes of those launches (counted by the kernel's
        # own retire step) / their summed CUDA-event durations.  The chain phase (kr_chain) is latency-bound by construction and reported next to it.
        rb_ms = kernel_us["rebuild"] * 1e-3
        # algorithmic bytes of the pass = the old rows read once (one row for a free new row, both rows of the pair for a
        # merged one) + the new matrix written once; 8 B per cell.  The ncu DRAM pass (`traffic`) measures the same number.
        b = 8.0 * (lst["bulk_rescan_cols"] + 2.0 * lst["bulk_build_cols"] + lst.get("bulk_new_cells", 0))
        a_ = b / (rb_ms * 1e-3) / 1e9
        launches = max(1, lst["bulk_rounds"])
        ncu = NCU_TRAFFIC.get(args.config, {}).get("kb_rebuild3")
        main_roof = {"bound": "hbm", "achieved": a_, "peak": peak, "unit": "GB/s", "frac": a_ / peak,
                     "traffic": (ncu["dram_bytes"] / ncu["launches"]) if ncu else None,
                     "kernel": "kb_rebuild3 (TMA-staged, list-driven merge + compact pass of the bulk rounds)", "launches_per_step": launches, "ms": rb_ms,
                     "avg_launch_us": rb_ms * 1e3 / launches, "algorithmic_bytes": b,
                     "algorithmic_bytes_per_launch": b / launches, "share_of_step": rb_ms / (ms_dev / K),
                     "peak_source": peak_src, "kernel_us_per_step": kernel_us,
                     "chain_phase": {"kernel": "kr_chain (persistent, cooperative)", "ms": kernel_us.get("chain", 0) * 1e-3,
                                     "rounds": lst["rounds"], "us_per_round": kernel_us.get("chain", 0) / max(1, lst["rounds"] - lst["bulk_rounds"]),
                                     "note": "1-4 merges per round: three grid barriers per round, latency-bound"},
                     "linkage_stage_bytes": 8.0 * lst["rescan_cols"] + 24.0 * lst["build_cols"] + 24.0 * lst["light_updates"]}
    # launches per step: bin 8, fold 1, sinkhorn (colpass+finish)=2/pass ... counted from the call graph
    sk_launches = 0 if args.no_balance else res["sk_launches"]      # counted by ops._CudaSinkhornPasses
    link_launches = lst["launches"] if lst.get("engine") == "rounds" else 1     # counted by hc_linkage_rounds
    gpu_launches = K * 2 * (8 + 1 + sk_launches + 1 + link_launches) + (K if args.wire == 16 else 0)  # + unpack
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference_sample(cfg)
        cpu = {"value": c["value"], "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": c["sample"],
               "parts": c["parts"]}
    extra, bam = None, None
    if world == 1 and not args.no_extra_cases:
        # what the headline does not show: workloads with exact distance ties, where the round-parallel engine cannot
        # certify fastcluster's merge order and the exact serial engine takes over (one timed step each)
        extra = {}
        del host_wire, staging
        torch.cuda.empty_cache()

        def one_case(rec, off, **kw):
            pairs_to_dendrogram(rec, off, ws, N, mapq=10, params=PARAMS, method=cfg["method"], workspace=workspace,
                                engine=args.engine, **kw)
            torch.cuda.synchronize()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            tm = StageTimer(True)
            r_ = pairs_to_dendrogram(rec, off, ws, N, mapq=10, params=PARAMS, method=cfg["method"], workspace=workspace,
                                     engine=args.engine, timer=tm, **kw)
            c1.record()
            torch.cuda.synchronize()
            ls = r_["linkage_stats"]
            return {"ms_per_step": c0.elapsed_time(c1), "linkage_engine": ls.get("engine"), "stage_ms": tm.result_ms(),
                    "fallback_reason": (ls.get("fallback_from_rounds") or {}).get("reason"),
                    "certified_merges_before_fallback": (ls.get("fallback_from_rounds") or {}).get("merges")}
        extra["no_balance"] = dict(one_case(records, offsets, balance=False),
                                   what="normalize() off = the reference at HEAD (bam2clusters.py:512): count-derived distances")
        g1 = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1, min_tail=1)
        rec1 = synth.make_pairs(g1, P, seed=1, device=dev)
        off1 = torch.from_numpy(g1.window_offsets().astype(np.int32)).to(dev)
        extra["min_tail_1"] = dict(one_case(rec1, off1, balance=not args.no_balance),
                                   what="contigs with trailing windows too short to receive a read midpoint: identical empty rows")
        del rec1
        torch.cuda.empty_cache()
        bam = bam_leg(cfg, offsets, ws, N, min(args.bam_records, P), dev, workspace, args.engine)
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
        "tie_heavy_cases": extra,
        "e2e_bam": bam,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_sharded(args, cfg, rank, world, local_rank):
    """--gpus N > 1: ONE pairs -> Z job over N GPUs.  Matrix build, fold and Sinkhorn-Knopp are row-sharded
    (pipeline.sharded_pairs_to_dendrogram), the dendrogram is built on rank 0.  Total work is fixed: strong scaling."""
    import torch
    import torch.distributed as dist
    from hicembler_b200 import ops, synth
    from hicembler_b200.pairs import pack_pairs16
    from hicembler_b200.parallel import ProcessGroup, RowShard
    from hicembler_b200.pipeline import StageTimer, shard_plan, sharded_pairs_to_dendrogram

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    N, P, ws = cfg["n_windows"], cfg["n_pairs"], cfg["window_bp"]
    # every rank generates the same synthetic stream (same seed) and keeps only its slice of it
    g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1, min_tail=cfg.get("min_tail", 1))
    records = synth.make_pairs(g, P, seed=1, device=dev)
    offsets = torch.from_numpy(g.window_offsets().astype(np.int32)).to(dev)
    pg = ProcessGroup()
    plan = shard_plan(records, g.window_offsets(), ws, N, world)       # the job of the BAM index (.bai)
    a, b = plan[rank]
    local = records[a:b].clone()
    del records
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    shard = RowShard(N, pg, dev)
    workspace = torch.empty(int(ops._lib.lib().hc_bin_pairs_workspace_bytes(b - a, max(shard.nrows, 1))), dtype=torch.uint8, device=dev)
    balance = not args.no_balance

    def step(rec, timer=None):
        return sharded_pairs_to_dendrogram(shard, rec, offsets, ws, mapq=10, params=PARAMS, method=cfg["method"], balance=balance,
                                           plan=plan, timer=timer, engine=args.engine, workspace=workspace, local_slice=True)

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        res = step(local)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    stage_ms = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        timer = StageTimer(True)
        res = step(local, timer)
        for k, v in timer.result_ms().items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1)
    # end to end: every rank copies ITS slice of the 16-byte wire records from pinned host memory over its own PCIe link
    packed = pack_pairs16(local) if args.wire == 16 else local
    host_wire = torch.empty(packed.shape, dtype=torch.int32, pin_memory=True)
    host_wire.copy_(packed)
    del packed
    staging = {}
    torch.cuda.synchronize()

    def step_e2e():
        return sharded_pairs_to_dendrogram(shard, host_wire, offsets, ws, mapq=10, params=PARAMS, method=cfg["method"],
                                           balance=balance, plan=plan, engine=args.engine, workspace=workspace,
                                           local_slice=True, staging=staging)
    for _ in range(min(args.warmup, 1)):
        step_e2e()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for _ in range(args.steps):
        res_e = step_e2e()
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop() if sampler else None
    K = args.steps
    t = torch.tensor([ms_dev, ms_e2e] + [stage_ms.get(k, 0.0) for k in ("h2d", "bin_pairs", "fold", "sinkhorn", "transform", "linkage")],
                     dtype=torch.float64, device=dev)
    tmax = t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    seen = torch.ones(1, device=dev)
    dist.all_reduce(seen)
    ms_dev, ms_e2e = tmax[0].item(), tmax[1].item()
    if rank == 0:
        peak, peak_src = load_peaks()
        lst = res["linkage_stats"]
        stage0 = {k: v / K for k, v in stage_ms.items()}
        stage_max = {k: tmax[2 + i].item() / K for i, k in enumerate(("h2d", "bin_pairs", "fold", "sinkhorn", "transform", "linkage"))}
        sk_it = res.get("sk_iterations", 0)
        front = sum(stage_max[k] for k in ("bin_pairs", "fold", "sinkhorn"))
        # per-GPU rooflines of the sharded stages: algorithmic bytes of one rank's share / its time
        rows = shard.nrows
        by = {"bin_pairs": 24.0 * (b - a) + 4.0 * rows * N, "fold": 12.0 * rows * N,
              "sinkhorn": res.get("sk_passes", 0) * 4.0 * rows * N,
              "transform": 4.0 * N * (N - 1) / 2 + 8.0 * N * N}
        roof = {}
        for k, bb in by.items():
            ms = stage0.get(k)
            if ms and bb:
                aa = bb / (ms * 1e-3) / 1e9
                roof[k] = {"bound": "hbm" if k != "transform" else "hbm+nvlink", "achieved": aa, "peak": peak, "unit": "GB/s",
                           "frac": aa / peak, "traffic": None, "ms": ms, "algorithmic_bytes": bb, "scope": "rank 0"}
        link_ms = stage0.get("linkage", 0.0)
        line = {
            "metric": METRIC, "value": float(P) * K / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32 counts, f32 matrix, f64 Sinkhorn vectors / distances", "data": "synthetic",
            "config": {"workload": cfg["name"], "n_contigs": cfg["n_contigs"], "n_windows": N, "n_pairs": P, "window_bp": ws,
                       "method": cfg["method"], "balance": balance,
                       "l2": "inputs larger than L2 (records %.1f GB, matrix %.1f GB over %d ranks)" % (24e-9 * P, 4e-9 * N * N, world),
                       "min_tail_bp": cfg.get("min_tail", 1), "linkage_engine": lst.get("engine"),
                       "parallelism": "row-sharded x%d (bin + fold + Sinkhorn sharded by row block, peer-memory exchanges over "
                                      "NVLink inside the kernels; linkage on rank 0)" % world,
                       "rows_per_rank": shard.per},
            "e2e": {"value": float(P) * K / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(args.wire * P),
                    "h2d_bytes_per_step_per_rank": int(args.wire * (b - a)), "d2h_bytes_per_step": int((N - 1) * 32 + 128),
                    "ms_per_step": ms_e2e / K, "host_record_bytes": args.wire,
                    "api": "hicembler_b200.pipeline.sharded_pairs_to_dendrogram (each rank: its slice of the pinned host "
                           "records over its own PCIe link -> Z on rank 0's host)"},
            "gpu_launches": int(K * 2 * (12 + 1 + res.get("sk_launches", 0) + 1 + lst.get("launches", 1))),
            "clocks": clocks,
            "roofline": dict(roof.get("sinkhorn" if balance else "bin_pairs", {}), kernel="kb_colpass+kb_rowpass (row-sharded Sinkhorn passes)"
                             if balance else "k_bin_tiles", peak_source=peak_src,
                             note="dominant SHARDED kernel; the step is dominated by the single-GPU linkage (see amdahl)"),
            "stage_rooflines": roof,
            "stage_ms": stage0, "stage_ms_max_over_ranks": stage_max,
            "comm_nranks_seen": int(seen.item()),
            "amdahl": {"linkage_ms_on_one_gpu": link_ms, "sharded_front_end_ms": front,
                       "note": "the dendrogram of ONE matrix is a serial merge sequence and runs on one GPU: the step can never be "
                               "faster than linkage + transform-gather, whatever N"},
            "seconds_pairs_to_dendrogram": ms_dev / K * 1e-3,
            "sinkhorn": {"iterations": sk_it, "stop": res.get("sk_stop")},
            "linkage": lst, "bin_stats": res["stats"], "cpu_baseline": None,
        }
        print(json.dumps(line), flush=True)
    shard.close()
    dist.destroy_process_group()


def bam_leg(cfg, offsets, ws, N, n_records, dev, workspace, engine):
    """e2e_bam: BGZF file -> native multi-threaded decoder (hc_bam_next) -> pinned 16-byte records -> host entry -> Z,
    on a stated slice of the workload (writing the full 500 M-record file is ~100 GB)."""
    import torch
    from hicembler_b200 import synth
    from hicembler_b200.pairs import read_bam_native, write_bam_fixed
    from hicembler_b200.pipeline import host_pairs_to_dendrogram
    g = synth.make_genome(cfg["n_contigs"], N, ws, cfg["n_chrom"], seed=1, min_tail=cfg.get("min_tail", 1))
    rec = synth.records_numpy(synth.make_pairs(g, n_records, seed=7))
    path = os.path.join(tempfile.gettempdir(), "hc_bench_%d.bam" % os.getpid())
    try:
        raw = write_bam_fixed(path, rec, g.names, g.sizes)
        size = os.path.getsize(path)
        read_bam_native(path, wire=16, side_arrays=False, pinned=True)          # warm the page cache
        t0 = time.perf_counter()
        _, _, host, _, _ = read_bam_native(path, wire=16, side_arrays=False, pinned=True)
        t_dec = time.perf_counter() - t0
        staging = {}
        kw = dict(staging=staging, mapq=10, params=PARAMS, method=cfg["method"], balance=True, workspace=workspace, engine=engine)
        host_pairs_to_dendrogram(host, offsets, ws, N, **kw)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, _, host, _, _ = read_bam_native(path, wire=16, side_arrays=False, pinned=True)
        out = host_pairs_to_dendrogram(host, offsets, ws, N, **kw)
        torch.cuda.synchronize()
        t_all = time.perf_counter() - t0
    finally:
        if os.path.exists(path):
            os.unlink(path)
    return {"records": int(n_records), "bam_bytes": int(size), "uncompressed_bytes": int(raw), "decode_s": t_dec,
            "decoder_gb_per_s_uncompressed": raw / t_dec / 1e9, "decoder_records_per_s": n_records / t_dec,
            "threads": os.cpu_count(), "seconds_bam_to_dendrogram": t_all, "value": n_records / t_all, "unit": UNIT,
            "linkage_engine": out["linkage_stats"].get("engine"),
            "note": "BGZF inflate + CRC + record decode on the host cores, then the same host entry as `e2e`; a %d-record slice "
                    "(%.1f %% of the workload's pairs) of the %d-window job" % (n_records, 100.0 * n_records / cfg["n_pairs"], N)}


def run_sinkhorn_mode(args, cfg, rank, world, local_rank):
    """--mode sinkhorn: row-sharded Sinkhorn-Knopp alone (BASELINE configs 3 and 5): every rank owns a row block of an
    N x N float32 matrix generated on the device; the column sums are all-reduced by the exchange fused into the update
    kernel (peer memory over NVLink, csrc/sharded.cu), the loop condition stays on the device."""
    import torch
    import torch.distributed as dist
    from hicembler_b200.parallel import LocalGroup, ProcessGroup, RowShard

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        pg = ProcessGroup()
    else:
        pg = LocalGroup.Member(LocalGroup(1, dev), 0)
    N = args.n or cfg["n_windows"]
    shard = RowShard(N, pg, dev, with_counts=False)
    r0, r1 = shard.row0, shard.row1
    # Poisson band + background on the row block (symmetric by construction: value depends on |i-j|
    # and a symmetric hash), +1 pseudo-count
    i = torch.arange(r0, r1, device=dev, dtype=torch.float32)[:, None]
    X = shard.sym
    step_rows = max(1, (1 << 28) // N)
    for s in range(0, r1 - r0, step_rows):
        j = torch.arange(N, device=dev, dtype=torch.float32)[None, :]
        ii = i[s:s + step_rows]
        lam = 40.0 / (1.0 + (ii - j).abs()) ** 1.08 + 0.05
        h = torch.frac(torch.sin((ii * j + ii + j) * 12.9898) * 43758.5453).abs()
        X[s:s + step_rows] = torch.floor(lam * (0.5 + h)) + 1.0
    vmax = X.max() if shard.active else torch.zeros((), device=dev)
    vmax = float(pg.allreduce_max_float(vmax.reshape(1).float()).item())
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        res = shard.sinkhorn_fit(vmax, 1000, 1e-5)
    pg.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        res = shard.sinkhorn_fit(vmax, 1000, 1e-5)
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
                          "stop": res.stopping_condition, "ms_per_fit": ms, "matrix_passes": passes,
                          "ms_per_pass": ms / passes, "launches": res.launches,
                          "aggregate_gbs": gbs, "frac_of_peak_per_gpu": gbs / world / peak, "peak_source": src,
                          "exchange": "one-shot all-reduce through peer memory fused into the c = 1/y update (no NCCL call, "
                                      "no host round trip per iteration)",
                          "streaming_floor_ms": passes * 4.0 * N * N / world / (peak * 1e9) * 1e3}), flush=True)
    shard.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=os.environ.get("HC_BENCH_CONFIG", "C3"))
    ap.add_argument("--mode", default="pipeline", choices=["pipeline", "sinkhorn", "replicas"])
    ap.add_argument("--no-extra-cases", action="store_true", help="skip the tie-heavy cases and the BAM leg (N = 1 only)")
    ap.add_argument("--bam-records", type=int, default=5_000_000, help="records of the BGZF file of the e2e_bam leg")
    ap.add_argument("--matrix-n", dest="n", type=int, default=0, help="matrix size for --mode sinkhorn")
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
    if world > 1 and args.mode != "replicas":
        return run_sharded(args, cfg, rank, world, local_rank)
    run_ours(args, cfg, rank, world, local_rank)


if __name__ == "__main__":
    main()
