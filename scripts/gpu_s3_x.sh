mkdir -p gpurun_out
python -m hicembler_b200.build > /dev/null 2>&1 || echo BUILD FAILED
make -s -C oracle libnnchain.so
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s15_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/s15_pytest.log
run() { echo "== $1"; env $1 timeout 600 python scripts/rounds_seeds.py C3 1 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['engine'], {k:round(v,2) for k,v in d['stage_ms'].items()})"; }
run HC_BIN_CELL16=0
run HC_BIN_CELL16=1
run "HC_BIN_CELL16=1 HC_BIN_THREADS=256"
