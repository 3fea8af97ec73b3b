mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s3j_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s3j_pytest.log
timeout 600 python scripts/rounds_seeds.py C3 1 2 > gpurun_out/s3j_seeds.jsonl 2> gpurun_out/s3j_seeds.err; echo "rc=$?"; tail -3 gpurun_out/s3j_seeds.err
cut -c1-330 gpurun_out/s3j_seeds.jsonl
timeout 600 python scripts/rounds_seeds.py C2 1 > gpurun_out/s3j_seeds_C2.jsonl 2> gpurun_out/s3j_seeds_C2.err; cut -c1-330 gpurun_out/s3j_seeds_C2.jsonl
timeout 600 python scripts/rounds_seeds.py C1 1 > gpurun_out/s3j_seeds_C1.jsonl 2> gpurun_out/s3j_seeds_C1.err; cut -c1-330 gpurun_out/s3j_seeds_C1.jsonl
