set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_linear.py tests/test_gpu_gpt.py -m gpu -q -x > gpurun_out/m2_tests.log 2>&1; tail -n 2 gpurun_out/m2_tests.log
timeout 300 python scripts/micro_linear.py > gpurun_out/micro_linear2.log 2>&1; tail -n 12 gpurun_out/micro_linear2.log
for args in "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1
done
timeout 200 python scripts/micro_linear.py sq4k > /dev/null 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:linear_tc -s 3 -c 1 -o gpurun_out/r1j_linear_sq4k python scripts/micro_linear.py sq4k > gpurun_out/ncu_sq.log 2>&1
