set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone exit $?" >> gpurun_out/noalloc_alone.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/noalloc_suite.log 2>&1; echo "suite exit $?" >> gpurun_out/noalloc_suite.log
tail -5 gpurun_out/noalloc_alone.log gpurun_out/noalloc_suite.log
