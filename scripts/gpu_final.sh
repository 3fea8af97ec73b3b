# round-end evidence on one B200: full GPU test log, smoke, bench line, ncu launch list of the bench command, ncu DRAM pass of
# the linkage row-pass kernels, ncu --set full of the hot kernels; usage: bash scripts/gpu_final.sh <prefix>
P=gpurun_out/${1:-final}
mkdir -p gpurun_out
bash scripts/gpu_tests.sh ${1:-final} 400
python -c "import __graft_entry__ as g; g.smoke()" > ${P}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 ${P}_smoke.log
bash scripts/gpu_bench.sh ${1:-final}_bench
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file ${P}_launches.csv python bench.py --steps 1 --warmup 1 --no-extra-cases --no-cpu-baseline > ${P}_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"kb_rebuild|kr_heavy|kr_chain" -c 80 --csv --log-file ${P}_dram_rowpass.csv python scripts/rounds_once.py C3 > ${P}_ncu_dram.log 2>&1; echo "dram pass rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_bin_band|k_tilepair64|k_sympass<float, 1>|k_transform_bal64|kb_rebuild3" -c 7 -o ${P}_hot_full -f python scripts/rounds_once.py C3 > ${P}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py ${P}_hot_full.ncu-rep > ${P}_hot_summary.txt 2>&1; cat ${P}_hot_summary.txt
