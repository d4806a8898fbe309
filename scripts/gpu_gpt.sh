set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gpt.py -m gpu -q -s -x > gpurun_out/gpt_tests.log 2>&1; echo "exit $?" >> gpurun_out/gpt_tests.log; grep -E "passed|failed|exit|gpt2-small|Error|assert" gpurun_out/gpt_tests.log | tail -n 12
for args in "--model small --math bf16 --head logits" "--model small --math bf16 --head ids" "--model medium --math bf16 --head ids" "--model small --math tf32x3 --head ids --steps 1" "--model small --math tf32 --head ids --steps 2"; do
  timeout 600 python scripts/bench_gpt.py $args >> gpurun_out/gpt_bench.log 2>&1
done
cat gpurun_out/gpt_bench.log
timeout 300 python scripts/bench_gpt.py --model small --math bf16 --head ids --batch 4 --steps 1 > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 200 --csv --log-file gpurun_out/gpt_launches.csv python scripts/bench_gpt.py --model small --math bf16 --head ids --batch 4 --steps 1 > gpurun_out/ncu_gpt.log 2>&1
timeout 300 python bench.py --steps 5 --warmup 3 --math tf32x3 --no-cpu-baseline --no-graph > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/cifar_tc_launches.csv python bench.py --steps 5 --warmup 3 --math tf32x3 --no-cpu-baseline --no-graph > gpurun_out/ncu_cifar.log 2>&1
