# kb_rebuild2 check: rounds-engine parity tests, then one C3 run per rebuild variant with per-kernel event times;
# usage: bash scripts/gpu_rebuild.sh <prefix> [variants "v:cfg ..."]
P=gpurun_out/${1:-rb}
VARS=${2:-"2:0 2:1 1:0"}
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage_rounds" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
HC_ROUNDS_RB_CFG=1 timeout -s KILL 600 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "bulk_rebuild" > ${P}_pytest_cfg1.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest_cfg1.log
for vc in $VARS; do
  v=${vc%%:*}; c=${vc##*:}
  HC_ROUNDS_REBUILD_V=$v HC_ROUNDS_RB_CFG=$c HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_v${v}c${c}.log 2>&1; echo "v$v cfg$c rc=$?"; grep -o "'rebuild': [0-9]*" ${P}_v${v}c${c}.log
done
