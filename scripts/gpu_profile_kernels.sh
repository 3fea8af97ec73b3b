# ncu --set full captures (with source) of the hot kernels at C3; usage: bash scripts/gpu_profile_kernels.sh <prefix>
P=gpurun_out/${1:-prof}
mkdir -p gpurun_out
python scripts/bin_once.py C3 3 > ${P}_bin.log 2>&1; echo "bin rc=$?"; tail -1 ${P}_bin.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_bin_tiles" -c 1 -o ${P}_bin_full -f python scripts/bin_once.py C3 1 > ${P}_ncu_bin.log 2>&1; echo "ncu bin rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"kb_rebuild" -c 1 -o ${P}_rebuild_full -f python scripts/rounds_once.py C3 > ${P}_ncu_rebuild.log 2>&1; echo "ncu rebuild rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"kr_(select|light|heavy)" -s 90 -c 6 -o ${P}_chain_full -f python scripts/rounds_once.py C3 > ${P}_ncu_chain.log 2>&1; echo "ncu chain rc=$?"
for f in bin rebuild chain; do python scripts/ncu_summary.py ${P}_${f}_full.ncu-rep > ${P}_${f}_summary.txt 2>&1; cat ${P}_${f}_summary.txt; done
