mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s6_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s6_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/s6_C3.json 2> gpurun_out/s6_C3.err; echo "bench rc=$?"; tail -3 gpurun_out/s6_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s6_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print(d['roofline']['frac'], d['roofline']['traffic'], d['gpu_launches'])"
timeout 900 python bench.py --no-cpu-baseline --wire 24 --steps 2 > gpurun_out/s6_C3_w24.json 2> gpurun_out/s6_C3_w24.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/s6_C3_w24.json'));print(d['ms_per_step'],d['e2e'])"
