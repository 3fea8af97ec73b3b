mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "linkage or rounds or dendrogram" > gpurun_out/s21_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s21_pytest.log
HC_ROUNDS_TRACE=1 timeout 300 python scripts/rounds_trace.py C3 > gpurun_out/s21_trace.out 2> gpurun_out/s21_trace.err; head -5 gpurun_out/s21_trace.err; tail -1 gpurun_out/s21_trace.out | cut -c1-200
