set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s3_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s3_pytest.log
timeout 600 python bench.py --config C3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s3_C3.json 2> gpurun_out/s3_C3.err; echo "rc=$?"; tail -3 gpurun_out/s3_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s3_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['linkage'])"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/s3_launches_C3.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/s3_ncu.log 2>&1; echo "ncu rc=$?"
