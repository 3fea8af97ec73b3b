# one pytest file / expression under a hard limit; usage: bash scripts/gpu_one.sh <prefix> <seconds> <pytest args...>
P=gpurun_out/${1:-t}; shift
LIM=${1:-120}; shift
mkdir -p gpurun_out
timeout -s KILL $LIM python -m pytest -m gpu -q -x "$@" > ${P}_pytest.log 2>&1; echo "rc=$?"; tail -25 ${P}_pytest.log
