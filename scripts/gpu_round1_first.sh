set -x
nvidia-smi -L | head -2; nproc; free -g | head -2
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r1_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r1_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r1_smoke.log
timeout 600 python bench.py --config C1 --steps 3 --warmup 3 > gpurun_out/r1_bench_C1.json 2> gpurun_out/r1_bench_C1.err; echo "C1 rc=$?"; tail -c 3000 gpurun_out/r1_bench_C1.json
timeout 900 python bench.py --config C2 --steps 3 --warmup 3 > gpurun_out/r1_bench_C2.json 2> gpurun_out/r1_bench_C2.err; echo "C2 rc=$?"; tail -c 3000 gpurun_out/r1_bench_C2.json
