# CTA-pair GEMM iteration: parity first (watchdog + timeout bound a hang), then the rest of the suite and the microbench.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_linear.py -m gpu -q -x -k "not throughput" 2>&1 | grep -E "^E |FAILED|passed|failed|Error" | head -20
timeout 1200 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_linear.py 2>&1 | grep -E "^E |FAILED|passed|failed|Error" | head -20
timeout 600 python scripts/micro_linear.py 2>&1 | tee gpurun_out/micro_linear_pair.log | grep -E "sq|8k|head|fc_gelu|qkv"
for args in "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1 | cut -c1-260
  B2P_NO_PAIR=1 timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1 | cut -c1-260
done
