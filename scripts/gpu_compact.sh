# chain phase against the compaction threshold; usage: bash scripts/gpu_compact.sh <prefix>
P=gpurun_out/${1:-cmp}
mkdir -p gpurun_out
for t in 1/2 2/3 3/4 7/8 15/16; do
  HC_ROUNDS_COMPACT=$t HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_$(echo $t | tr / _).log 2>&1
  echo "$t: $(grep -o "'compactions': [0-9]*" ${P}_$(echo $t | tr / _).log) $(grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_$(echo $t | tr / _).log) $(grep -o "'rescans': [0-9]*" ${P}_$(echo $t | tr / _).log)"
done
