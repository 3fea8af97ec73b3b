mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s12_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s12_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s12_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s12_smoke.log
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/s12_ref.json 2> gpurun_out/s12_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/s12_C3.json 2> gpurun_out/s12_C3.err; echo "bench rc=$?"; tail -3 gpurun_out/s12_C3.err
python -c "
import json;d=json.load(open('gpurun_out/s12_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['roofline']); print(d['cpu_baseline']['value'], d['clocks'])"
for c in C1 C2; do timeout 600 python bench.py --config $c --no-cpu-baseline > gpurun_out/s12_$c.json 2> gpurun_out/s12_$c.err; python -c "
import json;d=json.load(open('gpurun_out/s12_$c.json'));print('$c', d['ms_per_step'],d['stage_ms'],d['e2e']['ms_per_step'])"; done
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:kr_heavy -c 800 --csv --log-file gpurun_out/s12_dram_kr_heavy_C3.csv python scripts/rounds_once.py C3 > gpurun_out/s12_ncu_d.log 2>&1; echo "ncu dram rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"kr_heavy|kr_light|k_transform_tile64|k_bin_tiles|k_unpack" -c 10 -o gpurun_out/s12_full_C3 -f python scripts/rounds_once.py C3 > gpurun_out/s12_ncu_full.log 2>&1; echo "ncu full rc=$?"
HC_ROUNDS_NOGRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/s12_launches_C3.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/s12_ncu_l.log 2>&1; echo "ncu list rc=$?"
HC_ROUNDS_TRACE=1 timeout 300 python scripts/rounds_trace.py C3 > gpurun_out/s12_trace.out 2> gpurun_out/s12_trace.err
