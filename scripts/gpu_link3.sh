set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
timeout 600 python -m pytest tests -m gpu -x -q -k "linkage or dendrogram or tree or cluster_contigs" > gpurun_out/l3_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/l3_pytest.log
for c in C1 C3; do
HC_LINKAGE_PROFILE=1 timeout 600 python bench.py --config $c --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/l3_$c.json 2> gpurun_out/l3_$c.err; echo "rc=$?"; grep hc_linkage gpurun_out/l3_$c.err | tail -1
python -c "
import json;d=json.load(open('gpurun_out/l3_$c.json'));print(d['ms_per_step'],d['stage_ms'])"
done
