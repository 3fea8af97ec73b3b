# mixed measurement pass; usage: bash scripts/gpu_mix.sh <prefix>
P=gpurun_out/${1:-mix}
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "bulk_rebuild" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_rounds.log 2>&1; echo "rounds rc=$?"; grep -o "'rebuild': [0-9]*" ${P}_rounds.log
for ef in 1 0; do HC_BIN_EVICT_FIRST=$ef timeout 300 python scripts/bin_once.py C3 5 > ${P}_bin_ef${ef}.log 2>&1; echo "bin ef=$ef rc=$?"; tail -1 ${P}_bin_ef${ef}.log | cut -c1-200; done
