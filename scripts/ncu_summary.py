#!/usr/bin/env python
"""Summarise an .ncu-rep (ncu -i ... --page raw --csv) into one line per kernel launch."""
import csv, io, subprocess, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "smsp__inst_issued.sum", "smsp__average_warp_latency_per_inst_issued.ratio"] + [
        "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % k for k in (
            "long_scoreboard", "short_scoreboard", "wait", "barrier", "not_selected", "math_pipe_throttle", "mio_throttle",
            "lg_throttle", "branch_resolving", "no_instruction", "dispatch_stall", "membar", "drain", "sleeping")]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
idx = {}
for i, h in enumerate(hdr):
    idx.setdefault(h, i)
    for w in WANT:                      # some metrics carry a "UNIT.Section." prefix
        if h.endswith("." + w): idx.setdefault(w, i)
def num(s):
    try: return float(s.replace(",", ""))
    except ValueError: return float("nan")
print("# %s" % sys.argv[1])
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0]
    d = {}
    for w in WANT:
        if w in idx: d[w] = (num(r[idx[w]]), units[idx[w]])
    t, tu = d["gpu__time_duration.sum"]
    t_s = t * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(tu, 1e-9)
    def b(k):
        v, u = d.get(k, (float("nan"), ""))
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
    rd, wr = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
    print("%-34s grid=%-7d blk=%-5d regs=%-3d t=%10.3f ms  dram rd=%8.3f GB wr=%8.3f GB  %7.1f GB/s  dram%%=%5.1f sm%%=%5.1f  warps%%=%5.1f" % (
        name[-34:], d["launch__grid_size"][0], d["launch__block_size"][0], d["launch__registers_per_thread"][0], t_s * 1e3,
        rd / 1e9, wr / 1e9, (rd + wr) / t_s / 1e9, d.get("dram__cycles_active.avg.pct_of_peak_sustained_elapsed", (float("nan"),))[0],
        d.get("sm__throughput.avg.pct_of_peak_sustained_elapsed", (float("nan"),))[0], d.get("sm__warps_active.avg.pct_of_peak_sustained_active", (float("nan"),))[0]))
    # warp cycles per issued instruction, split by stall reason (sums to the average warp latency per instruction)
    lat = d.get("smsp__average_warp_latency_per_inst_issued.ratio", (float("nan"),))[0]
    st = ["%s=%.2f" % (k.split("stalled_")[1].split("_per")[0], v[0]) for k, v in d.items() if "stalled" in k and v[0] >= 0.05]
    print("      issue%%=%.1f inst=%.3fG cycles/inst=%.2f  stall cycles per instruction: %s" % (
        d.get("sm__inst_issued.avg.pct_of_peak_sustained_active", (float("nan"),))[0],
        d.get("smsp__inst_issued.sum", (float("nan"),))[0] / 1e9, lat, " ".join(st)))
