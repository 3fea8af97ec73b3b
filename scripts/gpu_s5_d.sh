mkdir -p gpurun_out
python scripts/bin_once.py C3 5 2>&1 | tail -1 | cut -c1-80
timeout 600 python -m pytest tests -m gpu -q -k "bin or bam2array or pair16 or host_entry or upto or region" 2>&1 | tail -2
