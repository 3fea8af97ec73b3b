P=gpurun_out/${1:-db}
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "linkage" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -2 ${P}_pytest.log
one() { HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_tmp.log 2>&1; echo "$1: $(grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_tmp.log) $(grep -o "'rounds': [0-9]*" ${P}_tmp.log)"; }
one s512; one s512
cp hicembler_b200/build/lib_s256.so hicembler_b200/libhicembler_b200.so
one s256; one s256
