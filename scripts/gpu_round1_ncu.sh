set -x
mkdir -p gpurun_out
K='regex:k_nnchain|k_bin_tiles|k_colpass|k_rowpass|k_fold_counts|k_transform_working'
timeout 1200 ncu --set full --clock-control none --import-source on -k "$K" -c 14 -f -o gpurun_out/r1_full_C2 python bench.py --config C2 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1_ncu_full_C2.log 2>&1; echo "ncu C2 rc=$?"
tail -2 gpurun_out/r1_ncu_full_C2.log | cut -c1-300
K3='regex:k_bin_tiles|k_colpass|k_rowpass|k_fold_counts|k_transform_working'
timeout 1500 ncu --set full --clock-control none --import-source on -k "$K3" -c 10 -f -o gpurun_out/r1_full_C3 python bench.py --config C3 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1_ncu_full_C3.log 2>&1; echo "ncu C3 rc=$?"
tail -2 gpurun_out/r1_ncu_full_C3.log | cut -c1-300
ls -la gpurun_out
