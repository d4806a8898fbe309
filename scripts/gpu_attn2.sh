mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gpt.py -m gpu -q -x 2>&1 | tail -n 2
timeout 200 python scripts/trace_attn.py 2>&1 | sed -n 4,9p
for args in "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1
done
