set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_linear.py -m gpu -q -s -x > gpurun_out/tc_linear.log 2>&1; echo "exit $?" >> gpurun_out/tc_linear.log
timeout 900 python -m pytest tests/test_gpu_gpt.py tests/test_gpu_cifar.py -m gpu -q -s > gpurun_out/tc_modes.log 2>&1; echo "exit $?" >> gpurun_out/tc_modes.log
timeout 300 python bench.py --steps 50 --warmup 5 --math tf32x3 --no-cpu-baseline > gpurun_out/bench_tf32x3.json 2> gpurun_out/bench_tf32x3.err
timeout 300 python bench.py --steps 50 --warmup 5 --math tf32 --no-cpu-baseline > gpurun_out/bench_tf32.json 2> gpurun_out/bench_tf32.err
grep -E "passed|failed|exit|Error|error" gpurun_out/tc_linear.log gpurun_out/tc_modes.log | tail -n 20
