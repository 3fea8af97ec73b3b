set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s2_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s2_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s2_smoke.log 2>&1; tail -2 gpurun_out/s2_smoke.log
HC_LINKAGE_PROFILE=1 timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/s2_C3.json 2> gpurun_out/s2_C3.err; echo "rc=$?"; grep hc_linkage gpurun_out/s2_C3.err | tail -1
python -c "
import json;d=json.load(open('gpurun_out/s2_C3.json'));print(d['ms_per_step'],d['stage_ms']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['e2e'])"
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/s2_ref.json 2> gpurun_out/s2_ref.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s2_launches_C3.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s2_ncu_launches.log 2>&1; echo "ncu rc=$?"
nvidia-smi --query-gpu=name,memory.total --format=csv
nproc; free -g | head -2
