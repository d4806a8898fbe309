mkdir -p gpurun_out
timeout 200 python scripts/trace_attn.py > gpurun_out/trace_attn.log 2>&1; cat gpurun_out/trace_attn.log
