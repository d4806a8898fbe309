set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cifar.py -m gpu -q -s -x > gpurun_out/p0tc_cifar.log 2>&1; echo "exit $?" >> gpurun_out/p0tc_cifar.log
tail -n 30 gpurun_out/p0tc_cifar.log
timeout 300 python bench.py --steps 50 --warmup 5 --math tf32x3 --no-cpu-baseline > gpurun_out/p0tc_bench.json 2> gpurun_out/p0tc_bench.err; echo "bench exit $?"
cat gpurun_out/p0tc_bench.json
timeout 900 python -m pytest tests/test_gpu_linear.py -m gpu -q -s -k throughput > gpurun_out/p0tc_linear_tput.log 2>&1
grep TFLOP gpurun_out/p0tc_linear_tput.log
