set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gpt.py -m gpu -q -x > gpurun_out/attn_tests.log 2>&1; echo "exit $?" >> gpurun_out/attn_tests.log; tail -n 4 gpurun_out/attn_tests.log
for args in "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8" "--model small --math bf16 --head ids --microbatch 2"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1
done
timeout 300 python scripts/bench_gpt.py --model small --math bf16 --head ids --batch 8 --microbatch 8 --steps 1 > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 120 --csv --log-file gpurun_out/gpt_mb8_launches.csv python scripts/bench_gpt.py --model small --math bf16 --head ids --batch 8 --microbatch 8 --steps 1 > gpurun_out/ncu_gpt8.log 2>&1
