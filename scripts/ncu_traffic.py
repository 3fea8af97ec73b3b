#!/usr/bin/env python
"""profiles/ncu_traffic.json from an ncu --csv launch list with dram__bytes_read.sum / dram__bytes_write.sum /
gpu__time_duration.sum (one row per launch and metric): DRAM bytes summed over the launches of the linkage
row-pass kernels (kr_heavy and kb_rebuild), which bench.py reports per launch next to the algorithmic bytes.
usage: ncu_traffic.py <config> <launch list csv> [out json]"""
import csv, json, os, sys
cfg, path = sys.argv[1], sys.argv[2]
out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "profiles", "ncu_traffic.json")
hdr, acc, ids = None, {"dram__bytes_read.sum": 0.0, "dram__bytes_write.sum": 0.0, "gpu__time_duration.sum": 0.0}, set()
per = {}
for r in csv.reader(open(path)):
    if "Kernel Name" in r:
        hdr = r
        continue
    if not hdr or len(r) != len(hdr):
        continue
    d = dict(zip(hdr, r))
    name = d["Kernel Name"]
    if "kr_heavy" not in name and "kb_rebuild" not in name and "kr_chain" not in name:
        continue
    v = float(d["Metric Value"].replace(",", ""))
    u = d["Metric Unit"]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6}.get(u, 1.0)
    acc[d["Metric Name"]] += v * scale
    ids.add(d["ID"])
    k = "kb_rebuild3" if "kb_rebuild3" in name else "kb_rebuild2" if "kb_rebuild2" in name else ("kb_rebuild" if "kb_rebuild" in name else ("kr_chain" if "kr_chain" in name else "kr_heavy"))
    per.setdefault(k, {"launches": set(), "dram_bytes": 0.0, "ns": 0.0})
    per[k]["launches"].add(d["ID"])
    if d["Metric Name"].startswith("dram"):
        per[k]["dram_bytes"] += v * scale
    else:
        per[k]["ns"] += v * scale
js = json.load(open(out)) if os.path.exists(out) else {}
# top-level entries per kernel (bench.py reads "kb_rebuild2": DRAM bytes per launch next to the algorithmic bytes)
for k, v in per.items():
    js.setdefault(cfg, {})[k] = {"launches": len(v["launches"]), "dram_bytes": v["dram_bytes"], "ncu_time_ms": v["ns"] * 1e-6,
                                 "source": os.path.relpath(path)}
js.setdefault(cfg, {})["kr_heavy"] = {
    "launches": len(ids), "dram_bytes": acc["dram__bytes_read.sum"] + acc["dram__bytes_write.sum"],
    "dram_read": acc["dram__bytes_read.sum"], "dram_write": acc["dram__bytes_write.sum"],
    "ncu_time_ms": acc["gpu__time_duration.sum"] * 1e-6, "source": os.path.relpath(path),
    "kernels": {k: {"launches": len(v["launches"]), "dram_bytes": v["dram_bytes"], "ncu_time_ms": v["ns"] * 1e-6}
                for k, v in per.items()}}
json.dump(js, open(out, "w"), indent=1)
print(json.dumps(js[cfg]["kr_heavy"], indent=1))
