# ncu --set full (with source) of the chain-phase kernels in the one-kernel-per-phase path; usage: <prefix>
P=gpurun_out/${1:-chain}
mkdir -p gpurun_out
HC_ROUNDS_PERSIST=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"kr_(select|light|heavy)" -s 150 -c 6 -o ${P}_chain_full -f python scripts/rounds_once.py C3 > ${P}_ncu_chain.log 2>&1; echo "ncu chain rc=$?"
python scripts/ncu_summary.py ${P}_chain_full.ncu-rep > ${P}_chain_summary.txt 2>&1; cat ${P}_chain_summary.txt
