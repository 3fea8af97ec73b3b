mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
nvidia-smi -L | wc -l; free -g | head -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/s13_C3_4gpu.json 2> gpurun_out/s13_C3_4gpu.err; echo "rc=$?"; tail -3 gpurun_out/s13_C3_4gpu.err
python -c "
import json
for l in open('gpurun_out/s13_C3_4gpu.json'):
    if l.startswith('{'):
        d=json.loads(l);print(d['n_gpus'],d['value'],d['ms_per_step'],d['stage_ms'],d['e2e'],d['linkage']['engine'],d.get('clocks'))"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --mode sinkhorn --steps 3 --warmup 3 > gpurun_out/s13_sk_4gpu.json 2> gpurun_out/s13_sk_4gpu.err; echo "rc=$?"; cat gpurun_out/s13_sk_4gpu.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 4 --impl reference --steps 1 --warmup 0 > gpurun_out/s13_ref_4gpu.json 2> gpurun_out/s13_ref_4gpu.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/s13_ref_4gpu.json
