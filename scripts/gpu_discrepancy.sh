mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
for a in "C1 100" "C2 100" "C2 1" "C3 100" "C3 1"; do
timeout 600 python scripts/measure_engine_discrepancy.py $a > gpurun_out/disc_$(echo $a | tr ' ' '_').json 2> gpurun_out/disc_$(echo $a | tr ' ' '_').err; echo "rc=$?"; cat gpurun_out/disc_$(echo $a | tr ' ' '_').json; grep diag gpurun_out/disc_$(echo $a | tr ' ' '_').err
done
