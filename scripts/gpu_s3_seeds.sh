set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
timeout 900 python scripts/rounds_seeds.py C3 1 2 3 4 5 6 7 8 > gpurun_out/s3_seeds_C3.jsonl 2> gpurun_out/s3_seeds_C3.err; echo "rc=$?"; tail -3 gpurun_out/s3_seeds_C3.err
cut -c1-400 gpurun_out/s3_seeds_C3.jsonl
