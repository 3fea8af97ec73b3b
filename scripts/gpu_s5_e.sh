mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "load_matrix or scaffold or distance" > gpurun_out/s5e_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/s5e_pytest.log
