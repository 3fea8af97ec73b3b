mkdir -p gpurun_out
P=gpurun_out/s5g
timeout 400 python bench.py > ${P}_C3.json 2> ${P}_C3.err; echo "bench rc=$?"; tail -3 ${P}_C3.err
python -c "
import json;d=json.load(open('${P}_C3.json'));print(d['ms_per_step'],d['stage_ms'],d['e2e']['ms_per_step']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['clocks'], d['gpu_launches'])"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:"k_bin_tiles" -c 1 -o ${P}_bin_full -f python scripts/bin_once.py C3 1 > ${P}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py ${P}_bin_full.ncu-rep > ${P}_bin_full_summary.txt 2>&1; head -3 ${P}_bin_full_summary.txt
rm -f ${P}_bin_full.ncu-rep
HC_ROUNDS_NOGRAPH=1 timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file ${P}_launches_C3.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > ${P}_ncu_l.log 2>&1; echo "ncu list rc=$?"
