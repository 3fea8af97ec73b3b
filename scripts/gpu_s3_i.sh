mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s3i_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s3i_pytest.log
for mc in 0 28000 14000; do
if [ $mc != 0 ]; then export HC_BIN_MAX_CELLS=$mc; fi
timeout 600 python scripts/rounds_seeds.py C3 1 > gpurun_out/s3i_seeds_$mc.jsonl 2> gpurun_out/s3i_seeds_$mc.err; echo "rc=$?"; tail -3 gpurun_out/s3i_seeds_$mc.err
cut -c1-330 gpurun_out/s3i_seeds_$mc.jsonl
done
