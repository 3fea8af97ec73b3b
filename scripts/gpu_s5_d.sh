mkdir -p gpurun_out
for c in C1 C2; do echo "config $c"; python scripts/bin_once.py $c 6 2>&1 | tail -1 | cut -c1-90; done
