mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
for bs in 1 32; do
HC_ROUNDS_BULK_SEGS=$bs timeout 600 ncu --set full --clock-control none --import-source on -k regex:kr_heavy -s 1 -c 1 -o gpurun_out/s7_heavy_r2_bs$bs -f python scripts/rounds_once.py C3 > gpurun_out/s7_ncu_bs$bs.log 2>&1; echo "ncu rc=$?"
done
