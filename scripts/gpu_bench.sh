# C3 bench line (+ optional args); usage: bash scripts/gpu_bench.sh <prefix> [bench args...]
P=gpurun_out/${1:-b}; shift
mkdir -p gpurun_out
timeout 900 python bench.py "$@" > ${P}.json 2> ${P}.err; echo "bench rc=$?"; tail -5 ${P}.err
python - <<PY
import json
d=json.loads(open("${P}.json").read().strip().splitlines()[-1])
print("ms/step", d["ms_per_step"], "e2e", d["e2e"].get("ms_per_step"))
print("stage_ms", {k: round(v,2) for k,v in d["stage_ms"].items()})
print("fracs", {k: round(v["frac"],3) for k,v in d.get("stage_rooflines",{}).items()})
print("roofline", {k: d["roofline"].get(k) for k in ("kernel","achieved","frac","ms","launches_per_step")}, d["roofline"].get("chain_phase"))
print("linkage", {k: d["linkage"].get(k) for k in ("engine","rounds","launches","kernel_us","persistent","compactions")})
print("extra", d.get("tie_heavy_cases")); print("bam", d.get("e2e_bam")); print("clocks", d.get("clocks"))
PY
