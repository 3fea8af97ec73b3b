# row-sharded Sinkhorn alone on N GPUs; usage: bash scripts/gpu_sk.sh <prefix> <ngpus> <sizes...>
P=gpurun_out/${1:-sk}; shift
N=${1:-8}; shift
mkdir -p gpurun_out
for n in "$@"; do
  if [ "$N" = "1" ]; then
    timeout -s KILL 400 python bench.py --gpus 1 --mode sinkhorn --matrix-n $n --steps 5 --warmup 2 > ${P}_sk_${N}gpu_$n.json 2> ${P}_sk_${N}gpu_$n.err
  else
    timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --mode sinkhorn --matrix-n $n --steps 5 --warmup 2 > ${P}_sk_${N}gpu_$n.json 2> ${P}_sk_${N}gpu_$n.err
  fi
  echo "sinkhorn gpus=$N n=$n rc=$?"; grep "^{" ${P}_sk_${N}gpu_$n.json | tail -1; tail -2 ${P}_sk_${N}gpu_$n.err | grep -v OMP
done
