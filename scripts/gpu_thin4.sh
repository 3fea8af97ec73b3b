# rounds-engine parity tests on the current build + compaction threshold sweep with the thin path
P=gpurun_out/${1:-thin4}
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
for t in 2/3 3/4 4/5; do
  HC_ROUNDS_COMPACT=$t HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_tmp.log 2>&1
  echo "$t: $(grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_tmp.log)"
done
