mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s17_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/s17_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/s17_C3.json 2> gpurun_out/s17_C3.err; echo "bench rc=$?"; tail -3 gpurun_out/s17_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s17_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['sinkhorn'], d['gpu_launches'])"
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_sympass|k_symfinish" -c 20 --csv --log-file gpurun_out/s17_sym.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/s17_ncu.log 2>&1; echo "ncu rc=$?"
python scripts/ncu_summary.py gpurun_out/s17_sym.csv 2>/dev/null | head -20 || tail -30 gpurun_out/s17_sym.csv
