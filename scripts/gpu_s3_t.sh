mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 1200 python -m pytest tests/test_gpu_full_size.py tests/test_gpu_parity.py -m gpu -x -q -k "full_size or bam_file or rounds" > gpurun_out/s11_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/s11_pytest.log
