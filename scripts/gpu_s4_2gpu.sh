mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s23_C3_2gpu.json 2> gpurun_out/s23_C3_2gpu.err; echo "rc=$?"; tail -3 gpurun_out/s23_C3_2gpu.err
python -c "
import json
for l in open('gpurun_out/s23_C3_2gpu.json'):
    if l.startswith('{'):
        d=json.loads(l);print(d['n_gpus'],d['value'],d['ms_per_step'],d['stage_ms'],d['e2e'],d['linkage']['engine'],d.get('clocks'))"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --mode sinkhorn --steps 3 --warmup 3 > gpurun_out/s23_sk_2gpu.json 2> gpurun_out/s23_sk_2gpu.err; echo "rc=$?"; tail -3 gpurun_out/s23_sk_2gpu.err; cat gpurun_out/s23_sk_2gpu.json
timeout 300 python bench.py --mode sinkhorn --steps 3 --warmup 3 > gpurun_out/s23_sk_1gpu.json 2> gpurun_out/s23_sk_1gpu.err; cat gpurun_out/s23_sk_1gpu.json
