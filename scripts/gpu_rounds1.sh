set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q -k "linkage or dendrogram or cluster_contigs or tree" > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r1_pytest.log
for c in C1 C2 C3; do
timeout 600 python bench.py --config $c --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r1_$c.json 2> gpurun_out/r1_$c.err; echo "rc=$?"; tail -3 gpurun_out/r1_$c.err
python -c "
import json;d=json.load(open('gpurun_out/r1_$c.json'));print(d['ms_per_step'],d['stage_ms']); print({k:round(v['frac'],3) for k,v in d['stage_rooflines'].items()}); print(d['linkage'])"
done
