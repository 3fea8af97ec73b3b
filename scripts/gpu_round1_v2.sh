set -x
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/v2_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/v2_pytest.log
for c in C1 C2 C3; do
HC_LINKAGE_PROFILE=1 timeout 900 python bench.py --config $c --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/v2_$c.json 2> gpurun_out/v2_$c.err; echo "rc=$?"; grep hc_linkage gpurun_out/v2_$c.err | tail -1
python -c "
import json;d=json.load(open('gpurun_out/v2_$c.json'));print(d['ms_per_step'],d['stage_ms']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['linkage'])"
done
HC_LINKAGE_NNCACHE=0 HC_LINKAGE_PROFILE=1 timeout 900 python bench.py --config C3 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/v2_C3_nocache.json 2> gpurun_out/v2_C3_nocache.err; grep hc_linkage gpurun_out/v2_C3_nocache.err | tail -1
