# GPU parity tests (+ optional pytest -k filter); usage: bash scripts/gpu_tests.sh <prefix> [-k expr]
P=gpurun_out/${1:-t}; shift
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 120 "$@" > ${P}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 ${P}_pytest.log
