# GPU parity tests, one pytest process per test file under its own `timeout` (a hung kernel costs one file's limit,
# not the call's); usage: bash scripts/gpu_tests.sh <prefix> [per-file seconds] [pytest args...]
P=gpurun_out/${1:-t}; shift
LIM=${1:-240}; shift
mkdir -p gpurun_out
: > ${P}_pytest.log
rc_all=0
for f in tests/test_*.py; do
  timeout -s KILL $LIM python -m pytest $f -m gpu -q -x "$@" >> ${P}_pytest.log 2>&1; rc=$?
  if [ $rc -ne 0 ] && [ $rc -ne 5 ]; then echo "FAILED rc=$rc: $f"; rc_all=1; fi
done
grep -E "passed|failed|error" ${P}_pytest.log | tail -20
echo "gpu tests rc=$rc_all"
