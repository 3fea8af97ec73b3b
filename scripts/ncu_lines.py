#!/usr/bin/env python
"""Executed warp instructions and stall samples per CUDA SOURCE LINE of one kernel in an .ncu-rep captured with
--import-source on (ncu -i rep --page source --csv --print-source cuda,sass).  usage: ncu_lines.py <rep> [top] [kernel-substring]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
want = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fn, hdr, acc = "", None, {}
def flush():
    if not acc or want not in fn: return
    ti = sum(v[0] for v in acc.values()); ts = sum(v[1] for v in acc.values())
    print("== %s: %.3f G warp instructions, %d samples" % (fn[:100], ti / 1e9, ts))
    for (ln, src), (i, smp) in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%6s %6.2f%% inst %6.2f%% smp  %s" % (ln, 100.0 * i / max(ti, 1), 100.0 * smp / max(ts, 1), src.strip()[:120]))
for r in rows:
    if not r: continue
    if r[0] == "Function Name":
        flush(); fn, hdr, acc = r[1], None, {}
    elif r[0] == "Line No":
        hdr = {k: i for i, k in enumerate(r)}
    elif hdr and r[0] not in ("", "File Path", "File Name") and len(r) > hdr.get("Instructions Executed", 10**9):
        try:
            acc[(r[0], r[1])] = (int(r[hdr["Instructions Executed"]] or 0), int(r[hdr["# Samples"]] or 0))
        except ValueError:
            pass
flush()
