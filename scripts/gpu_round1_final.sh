# Round-1 evidence run (1 GPU): full GPU test-suite, smoke, bench both arms, ncu launch list + full captures.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/final_pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/final_pytest_gpu.log; tail -n 3 gpurun_out/final_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/final_smoke.log; cat gpurun_out/final_smoke.log
timeout 600 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; cat gpurun_out/final_bench.json; tail -n 3 gpurun_out/final_bench.err
timeout 600 python bench.py --impl reference --steps 500 --warmup 5 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "ref exit $?"; cut -c1-400 gpurun_out/final_bench_reference.json
for args in "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1 | tee -a gpurun_out/final_gpt_1gpu.log
done
timeout 300 python scripts/time_stages_graph.py tf32x3 1 128 1024 4096 > gpurun_out/final_stage_times_graph.log 2>&1; cat gpurun_out/final_stage_times_graph.log
timeout 300 python scripts/micro_linear.py 2>&1 | grep -E "sq|8k|head" > gpurun_out/final_linear_microbench.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 120 --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_final1.log 2>&1
ls -la gpurun_out | tail -n 12
