mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s16_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/s16_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/s16_C3.json 2> gpurun_out/s16_C3.err; echo "bench rc=$?"; tail -3 gpurun_out/s16_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s16_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['sinkhorn'], d['gpu_launches'])"
