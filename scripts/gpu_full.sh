set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/full_tests.log 2>&1; echo "exit $?" >> gpurun_out/full_tests.log; tail -n 4 gpurun_out/full_tests.log
timeout 300 python scripts/micro_linear.py > gpurun_out/micro_linear.log 2>&1; cat gpurun_out/micro_linear.log
timeout 300 python scripts/time_stages.py tf32x3 > gpurun_out/time_stages.log 2>&1; cat gpurun_out/time_stages.log
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_pdl.json 2> gpurun_out/bench_pdl.err; echo "bench exit $?"; cut -c1-700 gpurun_out/bench_pdl.json; tail -3 gpurun_out/bench_pdl.err
B2P_NO_PDL=1 timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline 2>/dev/null | cut -c1-200
for args in "--model small --math bf16 --head ids" "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1
done
