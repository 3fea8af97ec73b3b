mkdir -p gpurun_out
P=gpurun_out/s5f
timeout 1200 python -m pytest tests -m gpu -q > ${P}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 ${P}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > ${P}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 ${P}_smoke.log
timeout 900 python bench.py > ${P}_C3.json 2> ${P}_C3.err; echo "bench rc=$?"; tail -3 ${P}_C3.err
python -c "
import json;d=json.load(open('${P}_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['roofline']); print(d['cpu_baseline']['value'], d['clocks'], d['gpu_launches'])"
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > ${P}_ref.json 2> ${P}_ref.err; echo "ref rc=$?"
for c in C1 C2; do timeout 600 python bench.py --config $c --no-cpu-baseline > ${P}_$c.json 2> ${P}_$c.err; python -c "
import json;d=json.load(open('${P}_$c.json'));print('$c', d['ms_per_step'],d['stage_ms'],d['e2e']['ms_per_step'])"; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_bin_tiles" -c 1 -o ${P}_bin_full -f python scripts/bin_once.py C3 1 > ${P}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py ${P}_bin_full.ncu-rep > ${P}_bin_full_summary.txt 2>&1; head -6 ${P}_bin_full_summary.txt
HC_ROUNDS_NOGRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file ${P}_launches_C3.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > ${P}_ncu_l.log 2>&1; echo "ncu list rc=$?"
rm -f ${P}_bin_full.ncu-rep
