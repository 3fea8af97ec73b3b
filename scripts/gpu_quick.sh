set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/q_pytest.log
HC_LINKAGE_PROFILE=1 timeout 900 python bench.py --config C3 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/q_C3.json 2> gpurun_out/q_C3.err; echo "rc=$?"; grep hc_linkage gpurun_out/q_C3.err | tail -1
python -c "
import json;d=json.load(open('gpurun_out/q_C3.json'));print(d['ms_per_step'],d['stage_ms']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['linkage']); print(d['bin_stats'])"
