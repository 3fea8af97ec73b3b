P=gpurun_out/${1:-l8}
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
for l in 1 0 1; do
  HC_ROUNDS_LIGHT8=$l HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_tmp.log 2>&1
  echo "light8=$l: $(grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_tmp.log) $(grep -o "'rounds': [0-9]*" ${P}_tmp.log) $(grep -o "'rescans': [0-9]*" ${P}_tmp.log)"
done
HC_ROUNDS_PHASES=1 timeout 300 python scripts/rounds_once.py C3 2>&1 | grep "rounds phases" | cut -c1-300
