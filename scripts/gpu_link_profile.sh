set -x
mkdir -p gpurun_out
for c in C1 C2 C3; do
HC_LINKAGE_PROFILE=1 timeout 900 python bench.py --config $c --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/lp_$c.json 2> gpurun_out/lp_$c.err; echo "rc=$?"; grep hc_linkage gpurun_out/lp_$c.err | tail -1
done
