mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
run() { echo "== $1"; env $1 timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, round(d['e2e']['ms_per_step'],1))"; }
run A=1
run HC_ROUNDS_NOGRAPH=1
run A=2
env A=1 timeout 600 python scripts/rounds_seeds.py C3 1 2>/dev/null | cut -c1-300
