mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
timeout 600 ncu --set full --clock-control none --import-source on -k regex:kr_ -c 14 -o gpurun_out/s3_rounds_full -f python scripts/rounds_once.py C3 > gpurun_out/s3_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"kr_|kc_" -c 4000 --csv --log-file gpurun_out/s3_launches_rounds.csv python scripts/rounds_once.py C3 > gpurun_out/s3_ncu_l.log 2>&1; echo "ncu list rc=$?"
ls -la gpurun_out | tail -5
