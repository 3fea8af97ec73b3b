mkdir -p gpurun_out
P=gpurun_out/s22
timeout 1200 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 ${P}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > ${P}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 ${P}_smoke.log
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > ${P}_ref.json 2> ${P}_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > ${P}_C3.json 2> ${P}_C3.err; echo "bench rc=$?"; tail -3 ${P}_C3.err
python -c "
import json;d=json.load(open('${P}_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['roofline']); print(d['cpu_baseline']['value'], d['clocks'], d['gpu_launches'])"
for c in C1 C2; do timeout 600 python bench.py --config $c --no-cpu-baseline > ${P}_$c.json 2> ${P}_$c.err; python -c "
import json;d=json.load(open('${P}_$c.json'));print('$c', d['ms_per_step'],d['stage_ms'],d['e2e']['ms_per_step'])"; done
timeout 600 python bench.py --wire 24 --no-cpu-baseline > ${P}_wire24.json 2> ${P}_wire24.err; python -c "
import json;d=json.load(open('${P}_wire24.json'));print('wire24', d['ms_per_step'],d['e2e']['ms_per_step'])"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"kr_heavy|kb_rebuild" -c 800 --csv --log-file ${P}_dram_rowpass_C3.csv python scripts/rounds_once.py C3 > ${P}_ncu_d.log 2>&1; echo "ncu dram rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"kb_rebuild|kr_heavy|kr_light|k_transform_tile64|k_bin_tiles|k_unpack|k_sympass" -c 12 -o ${P}_full_C3 -f python scripts/rounds_once.py C3 > ${P}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py ${P}_full_C3.ncu-rep > ${P}_full_C3_summary.txt 2>&1; head -30 ${P}_full_C3_summary.txt
HC_ROUNDS_NOGRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file ${P}_launches_C3.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > ${P}_ncu_l.log 2>&1; echo "ncu list rc=$?"
HC_ROUNDS_TRACE=1 timeout 300 python scripts/rounds_trace.py C3 > ${P}_trace.out 2> ${P}_trace.err
rm -f ${P}_full_C3.ncu-rep
