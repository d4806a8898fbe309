mkdir -p gpurun_out
timeout 120 python scripts/trace_part0.py 1184 > gpurun_out/trace_part0.log 2>&1; cat gpurun_out/trace_part0.log
