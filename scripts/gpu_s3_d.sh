mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q -k "linkage or dendrogram or cluster_contigs or tree or rounds or pipeline" > gpurun_out/s3l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s3l_pytest.log
timeout 600 python scripts/rounds_seeds.py C3 1 2 > gpurun_out/s3l_seeds_C3.jsonl 2> gpurun_out/s3l_seeds_C3.err; echo "rc=$?"; tail -3 gpurun_out/s3l_seeds_C3.err
cut -c1-330 gpurun_out/s3l_seeds_C3.jsonl
timeout 300 python scripts/rounds_trace.py C3 > gpurun_out/s3l_trace.out 2> gpurun_out/s3l_trace.err; echo rc=$?; grep trace gpurun_out/s3l_trace.err | awk 'NR<14 || NR%60==0'; tail -1 gpurun_out/s3l_trace.out
grep trace gpurun_out/s3l_trace.err | awk '{r=$4; us=$14; q=$8; if(r<=11){b+=us} else if (r<=545) {t+=us; tq+=q; n++} else {e+=us; en++}} END{print "bulk(2-11)",b,"chain(12-545)",t,"rescans",tq,"n",n,"end",e,en, "total",b+t+e}'
