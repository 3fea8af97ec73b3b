"""SASS tell-table: per kernel of libhicembler_b200.so, how many instructions of the mnemonics that prove a hardware
path (B200_PROFILING.md): UBLKCP = cp.async.bulk (TMA 1-D bulk copy), UBLKPF = cp.async.bulk.prefetch.L2,
SYNCS = mbarrier, REDG = fire-and-forget global reduction, ATOMG/ATOMS = global / shared atomics, REDUX / CREDUX = redux.sync (warp reductions),
LDG.E.128 = 16-byte global loads, UCGABAR/ "BAR.*CLUSTER"-like = cluster barrier, DADD/DMUL/DFMA = FP64 pipe.
usage: python scripts/sass_tells.py > profiles/rNN_sass_tells.txt   (needs cuobjdump; no GPU)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "hicembler_b200", "libhicembler_b200.so")
TELLS = ["UBLKCP", "UBLKPF", "UTMALDG", "SYNCS", "REDG", "ATOMG", "ATOMS", "REDUX", "CREDUX", "LDG.E.128", "STG.E.128", "UCGABAR",
         "DFMA", "DADD", "DMUL", "MUFU", "BAR.SYNC"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.split("\n")
    counts, order, cur, k = collections.defaultdict(collections.Counter), [], None, 0
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = re.sub(r"\(.*", "", names[k].replace("(anonymous namespace)::", "")).replace("void ", "")
            k += 1
            while cur in counts:
                cur += "'"
            order.append(cur)
            counts[cur]["_n"] = 0
            continue
        if cur is None or "/*" not in line:
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_n"] += 1
        for t in TELLS:
            if op == t or op.startswith(t + ".") or (("." in t) and op.startswith(t)):
                counts[cur][t] += 1
    print("# SASS tell-table of hicembler_b200/libhicembler_b200.so (cuobjdump -sass, sm_100a); columns = instruction counts")
    print("kernel".ljust(64) + " instr " + " ".join(t.rjust(9) for t in TELLS))
    tot = collections.Counter()
    for name in order:
        c = counts[name]
        tot.update(c)
        print(name[:63].ljust(64) + str(c["_n"]).rjust(6) + " " + " ".join(str(c[t]).rjust(9) for t in TELLS))
    print("TOTAL".ljust(64) + str(tot["_n"]).rjust(6) + " " + " ".join(str(tot[t]).rjust(9) for t in TELLS))


if __name__ == "__main__":
    sys.exit(main())
