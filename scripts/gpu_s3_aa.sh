mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s18_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/s18_pytest.log
HC_ROUNDS_TRACE=1 timeout 300 python scripts/rounds_trace.py C3 > gpurun_out/s18_trace.out 2> gpurun_out/s18_trace.err; head -12 gpurun_out/s18_trace.err; tail -3 gpurun_out/s18_trace.out
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/s18_C3.json 2> gpurun_out/s18_C3.err; echo "bench rc=$?"; tail -3 gpurun_out/s18_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s18_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']['ms_per_step']); print(d['roofline']['kernel_us_per_step'], d['linkage'])"
