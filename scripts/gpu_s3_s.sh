mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s10_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s10_pytest.log
run() { echo "== $1"; env $1 timeout 600 python scripts/rounds_seeds.py $2 1 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['engine'], {k:round(v,2) for k,v in d['stage_ms'].items()})"; }
run A=1 C3
run HC_ROUNDS_NOGRAPH=1 C3
run "HC_ROUNDS_BATCH=16" C3
run A=1 C2
run HC_ROUNDS_NOGRAPH=1 C2
run A=1 C1
run HC_ROUNDS_NOGRAPH=1 C1
