set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cifar.py tests/test_gpu_linear.py -m gpu -q -x > gpurun_out/iter_tests.log 2>&1; echo "exit $?" >> gpurun_out/iter_tests.log; tail -n 4 gpurun_out/iter_tests.log
timeout 120 python scripts/trace_part0.py 1184 > gpurun_out/trace_part0.log 2>&1; head -n 14 gpurun_out/trace_part0.log
timeout 300 python scripts/time_stages.py tf32x3 > gpurun_out/time_stages.log 2>&1; cat gpurun_out/time_stages.log
timeout 300 python bench.py --steps 50 --warmup 5 --math tf32x3 > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench exit $?"; cat gpurun_out/bench_iter.json; tail -3 gpurun_out/bench_iter.err
