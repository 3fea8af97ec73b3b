# binning: parity tests + event time at C3; usage: bash scripts/gpu_bin.sh <prefix> [env assignments...]
P=gpurun_out/${1:-bin}; shift
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "bin or bam2arrays or pair16 or counts or unsorted or accumulate or band or hot" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
env "$@" timeout 300 python scripts/bin_once.py C3 6 > ${P}_time.log 2>&1; echo "rc=$?"; tail -1 ${P}_time.log | cut -c1-120
HC_BIN_BAND_THREADS=512 timeout 300 python scripts/bin_once.py C3 6 > ${P}_time512.log 2>&1; echo "rc=$?"; tail -1 ${P}_time512.log | cut -c1-120
