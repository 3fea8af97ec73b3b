set -x
mkdir -p gpurun_out
timeout 1500 python bench.py --config C3 --steps 3 --warmup 3 > gpurun_out/r1_bench_C3.json 2> gpurun_out/r1_bench_C3.err; echo "C3 rc=$?"; tail -c 2500 gpurun_out/r1_bench_C3.json; tail -5 gpurun_out/r1_bench_C3.err
# launch list (C2 shape; ncu serialises, so keep it short)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches_C2.csv python bench.py --config C2 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1_ncu_C2.log 2>&1; echo "ncu list rc=$?"
tail -3 gpurun_out/r1_ncu_C2.log
