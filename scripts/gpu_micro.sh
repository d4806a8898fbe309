set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_linear.py -m gpu -q -x > gpurun_out/micro_tests.log 2>&1; tail -n 3 gpurun_out/micro_tests.log
timeout 300 python scripts/micro_linear.py > gpurun_out/micro_linear.log 2>&1; cat gpurun_out/micro_linear.log
for args in "--model small --math bf16 --head ids" "--model medium --math bf16 --head ids"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1
done
