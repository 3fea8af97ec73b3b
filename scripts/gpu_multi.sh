# real multi-GPU run: sharded check + sharded bench line; usage: bash scripts/gpu_multi.sh <prefix> <ngpus> [bench args]
P=gpurun_out/${1:-m}; shift
N=${1:-2}; shift
mkdir -p gpurun_out
nvidia-smi topo -m > ${P}_topo.txt 2>&1
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/sharded_check.py > ${P}_check.log 2>&1; echo "check rc=$?"; tail -4 ${P}_check.log
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N "$@" > ${P}_bench.json 2> ${P}_bench.err; echo "bench rc=$?"; tail -5 ${P}_bench.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("${P}_bench.json").read().strip().splitlines() if l.startswith("{")][-1])
    print("ms/step", d["ms_per_step"], "e2e", d["e2e"].get("ms_per_step"), "scaling", d["scaling"], d["config"]["parallelism"][:40])
    print("stage_ms", {k: round(v,2) for k,v in d["stage_ms"].items()})
    print("stage_max", {k: round(v,2) for k,v in d.get("stage_ms_max_over_ranks",{}).items()})
    print("fracs", {k: round(v["frac"],3) for k,v in d.get("stage_rooflines",{}).items()}, "ranks seen", d.get("comm_nranks_seen"))
except Exception as e:
    print("no bench line:", e)
PY
if [ -n "$HC_SK_SIZES" ]; then
  for n in $HC_SK_SIZES; do
    timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --mode sinkhorn --matrix-n $n --steps 5 --warmup 2 > ${P}_sk_$n.json 2> ${P}_sk_$n.err; echo "sinkhorn n=$n rc=$?"; grep "^{" ${P}_sk_$n.json | tail -1; tail -2 ${P}_sk_$n.err
  done
fi
