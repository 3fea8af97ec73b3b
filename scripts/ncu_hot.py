#!/usr/bin/env python
"""Hot SASS instructions of one kernel in an .ncu-rep (ncu -i rep --page source --csv): warp-stall samples and executed
warp instructions per SASS line, with the dominant stall reason.  usage: ncu_hot.py <rep> [top] [kernel-substring]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
want = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and cur["hdr"] and len(row) == len(cur["hdr"]):
        cur["rows"].append(row)
for b in blocks:
    if want not in b["name"]:
        continue
    h = b["hdr"]
    ci = {k: i for i, k in enumerate(h)}
    S, I = ci["# Samples"], ci["Instructions Executed"]
    stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    tot_s = sum(int(r[S] or 0) for r in b["rows"])
    tot_i = sum(int(r[I] or 0) for r in b["rows"])
    print("== %s: %d SASS lines, %d samples, %.3f G warp instructions" % (b["name"][:90], len(b["rows"]), tot_s, tot_i / 1e9))
    order = sorted(range(len(b["rows"])), key=lambda k: -int(b["rows"][k][S] or 0))[:top]
    for k in sorted(order):
        r = b["rows"][k]
        st = sorted(((int(r[ci[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
        print("%5d %5.1f%% smp %6.2f%% inst  %-70s %s" % (k, 100.0 * int(r[S] or 0) / max(tot_s, 1), 100.0 * int(r[I] or 0) / max(tot_i, 1),
                                                 r[ci["Source"]].strip()[:70], " ".join("%s=%d" % (n, v) for v, n in st if v)))
    # instruction mix by opcode
    mix = {}
    for r in b["rows"]:
        op = r[ci["Source"]].strip().split()
        op = [t for t in op if not t.startswith("@")]
        if op:
            key = op[0].split(".")[0]
            mix[key] = mix.get(key, 0) + int(r[I] or 0)
    print("   mix: " + " ".join("%s=%.1f%%" % (k, 100.0 * v / max(tot_i, 1)) for k, v in sorted(mix.items(), key=lambda kv: -kv[1])[:18]))
