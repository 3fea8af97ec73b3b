# K6 check: odd sizes for the guards, C3 size for the time; usage: bash scripts/gpu_transform.sh <prefix>
P=gpurun_out/${1:-tr}
mkdir -p gpurun_out
for n in 1 63 64 65 130 1000 4097; do python scripts/transform_once.py $n 1; done > ${P}_transform.txt 2>&1
python scripts/transform_once.py 50000 3 >> ${P}_transform.txt 2>&1
cat ${P}_transform.txt
timeout -s KILL 300 python -m pytest -m gpu -q -x tests/test_gpu_parity.py -k "transform or distance or working or condense or pipeline or smoke" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -5 ${P}_pytest.log
