P=gpurun_out/${1:-bd}
mkdir -p gpurun_out
for d in 5 8 12 16 24 40; do
  HC_ROUNDS_BULK_DIV=$d HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_tmp.log 2>&1
  python - <<PY
import re,ast
t=open("${P}_tmp.log").read().strip().splitlines()[-1]
d=ast.literal_eval(t)
k=d["kernel_us"]
print("div $d: total %.1f ms  %s bulk_rounds %d rounds %d launches %d certified %s" % (sum(k.values())/1e3, {a:round(b/1e3,1) for a,b in k.items()}, d["bulk_rounds"], d["rounds"], d["launches"], d["certified"]))
PY
done
