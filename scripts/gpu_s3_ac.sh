mkdir -p gpurun_out
HC_ROUNDS_TRACE=1 timeout 300 python scripts/rounds_trace.py C3 > gpurun_out/s20_trace.out 2> gpurun_out/s20_trace.err; head -5 gpurun_out/s20_trace.err; tail -1 gpurun_out/s20_trace.out | cut -c1-200
