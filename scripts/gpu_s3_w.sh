python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
run() { echo "== $1"; env $1 timeout 600 python scripts/rounds_seeds.py C3 1 2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['engine'], round(d['stage_ms']['linkage'],2))"; }
run HC_ROUNDS_BULK3=0
run HC_ROUNDS_BULK3=2
run HC_ROUNDS_BULK3=3
run HC_ROUNDS_BULK3=0
run HC_ROUNDS_BULK3=2
