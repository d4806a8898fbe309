set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_linear.py -m gpu -q -s -x > gpurun_out/linear_x.log 2>&1; echo "exit $?" >> gpurun_out/linear_x.log
timeout 900 python -m pytest tests/test_gpu_linear.py -m gpu -q -s > gpurun_out/linear_all.log 2>&1; echo "exit $?" >> gpurun_out/linear_all.log
tail -n 40 gpurun_out/linear_x.log
