mkdir -p gpurun_out
P=gpurun_out/s5b
python scripts/bin_once.py C3 5 > ${P}_bin.log 2>&1; echo "rc=$?"; tail -2 ${P}_bin.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_bin_tiles" -c 1 -o ${P}_bin_full -f python scripts/bin_once.py C3 1 > ${P}_ncu.log 2>&1; echo "ncu rc=$?"
ls -la ${P}_bin_full.ncu-rep
