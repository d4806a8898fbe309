mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cifar.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -n 2
timeout 300 python scripts/time_stages.py tf32x3 2>&1 | grep -E "batch +(128|1024)"
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('graph', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['p50_request_latency_ms'])"
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-graph 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('nograph', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'])"
