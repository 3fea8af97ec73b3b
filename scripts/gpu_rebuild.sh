# kb_rebuild2 check: rounds-engine parity tests, then one C3 run per rebuild version with per-kernel event times and the
# phase timers of the persistent kernel; usage: bash scripts/gpu_rebuild.sh <prefix>
P=gpurun_out/${1:-rb}
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage_rounds" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -4 ${P}_pytest.log
for v in 2 1; do
  HC_ROUNDS_REBUILD_V=$v HC_ROUNDS_KTIME=1 HC_ROUNDS_PHASES=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_v${v}.log 2>&1; echo "v$v rc=$?"; tail -3 ${P}_v${v}.log
done
HC_ROUNDS_TRACE=1 timeout 300 python scripts/rounds_once.py C3 2>&1 | grep -E "bulk round|initial scan|compaction|records_to_Z" > ${P}_trace.log; cat ${P}_trace.log
