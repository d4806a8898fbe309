mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | grep -E "^E |FAILED|passed|failed|Error" | head -20
timeout 300 python scripts/time_stages.py tf32x3 2>&1 | grep -E "batch +(1|128|1024|4096):"
timeout 300 python bench.py --steps 500 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('graph', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['p50_request_latency_ms'], d['roofline']['us_per_launch'], d['roofline']['frac'], d['roofline']['traffic'])"
for args in "--model small --math bf16 --head ids --microbatch 8" "--model medium --math bf16 --head ids --microbatch 8"; do
  timeout 600 python scripts/bench_gpt.py $args 2>&1 | tail -n 1 | cut -c1-260
done
