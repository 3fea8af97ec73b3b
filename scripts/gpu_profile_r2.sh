# round-2 discovery pass: FP64 rates of the part + one ncu --set full capture each of k_bin_band, k_transform_tile64 and the
# first (largest) kb_rebuild of a C3 run; usage: bash scripts/gpu_profile_r2.sh <prefix>
P=gpurun_out/${1:-r2prof}
mkdir -p gpurun_out
./scripts/micro/fp64_rate > ${P}_fp64_rate.txt 2>&1; cat ${P}_fp64_rate.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_bin_band|k_transform_tile64|kb_rebuild" -c 3 \
  -o ${P}_three_full -f python scripts/rounds_once.py C3 > ${P}_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 ${P}_ncu.log
python scripts/ncu_summary.py ${P}_three_full.ncu-rep > ${P}_three_summary.txt 2>&1; cat ${P}_three_summary.txt
ls -la gpurun_out
