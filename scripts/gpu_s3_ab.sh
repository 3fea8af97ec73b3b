mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"kb_rebuild" -c 1 -o gpurun_out/s19_rebuild -f python scripts/rounds_once.py C3 > gpurun_out/s19_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/s19_rebuild.ncu-rep --page details > gpurun_out/s19_details.txt 2>&1
grep -E "Duration|DRAM Throughput|Memory Throughput|L2 Cache Throughput|L1/TEX Hit|L2 Hit|Achieved Occupancy|Registers|Issue Slots Busy|Executed Ipc|Stall|stall|FP64|Pipe" gpurun_out/s19_details.txt | head -40
