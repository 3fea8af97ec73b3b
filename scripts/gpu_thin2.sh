# clean timing (no phase diagnostics) of the chain phase with / without the thin-round path
P=gpurun_out/${1:-thin2}
mkdir -p gpurun_out
for t in 1 0 1 0; do
  HC_ROUNDS_THIN=$t HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_t${t}.log 2>&1; echo "thin=$t rc=$? $(grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_t${t}.log)"
done
