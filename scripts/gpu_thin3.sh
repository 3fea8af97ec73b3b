# chain phase: staging size (library variants) x rows-per-CTA threshold
P=gpurun_out/${1:-thin3}
mkdir -p gpurun_out
one() { HC_ROUNDS_THIN=$1 HC_ROUNDS_KTIME=1 timeout 300 python scripts/rounds_once.py C3 > ${P}_tmp.log 2>&1; echo "$2 thin=$1 rc=$? $(grep -o "'compaction': [0-9]*, 'initial_scan': [0-9]*, 'rebuild': [0-9]*, 'chain': [0-9]*" ${P}_tmp.log)"; }
for t in 3 2 6 1 0; do one $t seg512; done
cp hicembler_b200/build/lib_seg1024.so hicembler_b200/libhicembler_b200.so
for t in 3 6 0; do one $t seg1024; done
