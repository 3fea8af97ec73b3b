mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s4_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s4_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s4_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s4_smoke.log
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/s4_ref.json 2> gpurun_out/s4_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/s4_ref.json
timeout 900 python bench.py > gpurun_out/s4_C3.json 2> gpurun_out/s4_C3.err; echo "bench rc=$?"; tail -3 gpurun_out/s4_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s4_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['roofline']); print(d['cpu_baseline'])"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:kr_heavy -c 800 --csv --log-file gpurun_out/s4_dram_kr_heavy_C3.csv python scripts/rounds_once.py C3 > gpurun_out/s4_ncu_d.log 2>&1; echo "ncu dram rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"kr_heavy|kr_light|k_transform_tile64|k_bin_tiles" -c 10 -o gpurun_out/s4_full_C3 -f python scripts/rounds_once.py C3 > gpurun_out/s4_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/s4_launches_C3.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/s4_ncu_l.log 2>&1; echo "ncu list rc=$?"
