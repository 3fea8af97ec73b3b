# one ncu --set full capture; usage: bash scripts/gpu_ncu_one.sh <prefix> <kernel regex> <count> <script and args...>
P=gpurun_out/$1; KR=$2; CNT=$3; shift 3
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$KR" -c $CNT -o ${P}_full -f python "$@" > ${P}_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 ${P}_ncu.log
python scripts/ncu_summary.py ${P}_full.ncu-rep > ${P}_summary.txt 2>&1; cat ${P}_summary.txt
