# fused FC stage iteration: parity under a timeout, then stage timings with and without the fused kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cifar.py -m gpu -q -x 2>&1 | grep -E "^E |FAILED|passed|failed|Error" | head -20
echo "--- fused"; timeout 300 python scripts/time_stages.py tf32x3 2>&1 | grep -E "batch +(1|128|1024|4096):"
echo "--- 3 kernels"; B2P_FC_FUSED=0 timeout 300 python scripts/time_stages.py tf32x3 2>&1 | grep -E "batch +(1|128|1024|4096):"
for f in 1 0; do
B2P_FC_FUSED=$f timeout 300 python bench.py --steps 500 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('fused=$f graph', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['p50_request_latency_ms'], d['roofline']['us_per_launch'], d['roofline']['frac'])"
done
