# thin-round path of kr_chain: rounds-engine parity tests, then C3 with and without it; usage: bash scripts/gpu_thin.sh <prefix>
P=gpurun_out/${1:-thin}
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -3 ${P}_pytest.log
for t in 1 0; do
  HC_ROUNDS_THIN=$t HC_ROUNDS_KTIME=1 HC_ROUNDS_PHASES=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_t${t}.log 2>&1; echo "thin=$t rc=$?"
  grep "rounds phases" ${P}_t${t}.log | cut -c1-330; grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_t${t}.log; grep -o "'rescans': [0-9]*" ${P}_t${t}.log
done
