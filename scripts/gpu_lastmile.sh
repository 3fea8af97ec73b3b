# last experiments of the round: select with two load levels (parity tests), bulk threshold with kb_rebuild3
P=gpurun_out/${1:-lm}
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
for d in 5 5 10 20 40; do
  HC_ROUNDS_BULK_DIV=$d HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_tmp.log 2>&1
  python - <<PY
import ast
d=ast.literal_eval(open("${P}_tmp.log").read().strip().splitlines()[-1])
k=d["kernel_us"]
print("div $d: total %.1f ms  %s bulk_rounds %d rounds %d certified %s" % (sum(k.values())/1e3, {a:round(b/1e3,1) for a,b in k.items()}, d["bulk_rounds"], d["rounds"], d["certified"]))
PY
done
