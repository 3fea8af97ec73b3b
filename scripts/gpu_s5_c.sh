mkdir -p gpurun_out
P=gpurun_out/s5c
python scripts/bin_once.py C3 5 > ${P}_bin.log 2>&1; echo "rc=$?"; tail -2 ${P}_bin.log
timeout 900 python -m pytest tests -m gpu -q -k "bin or bam2array or pair16 or host_entry or full_size or upto or region" > ${P}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 ${P}_pytest.log
