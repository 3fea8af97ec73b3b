set -x
mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
timeout 600 python -m pytest tests -m gpu -x -q -k "linkage or dendrogram or tree or cluster_contigs" > gpurun_out/l2_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/l2_pytest.log
for c in C1 C2 C3; do
HC_LINKAGE_PROFILE=1 timeout 600 python bench.py --config $c --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/l2_$c.json 2> gpurun_out/l2_$c.err; echo "rc=$?"; grep hc_linkage gpurun_out/l2_$c.err | tail -1
python -c "
import json;d=json.load(open('gpurun_out/l2_$c.json'));print(d['ms_per_step'],d['stage_ms'])"
done
for k in 1 2 4; do
HC_LINKAGE_CLUSTERS=$k HC_LINKAGE_PROFILE=1 timeout 600 python bench.py --config C3 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/l2_C3_k$k.json 2> gpurun_out/l2_C3_k$k.err; grep hc_linkage gpurun_out/l2_C3_k$k.err | tail -1
done
