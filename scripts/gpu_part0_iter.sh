#!/bin/bash
# part-0 kernel iteration: localised parity first, then the CIFAR GPU tests, then a kernel-time line
mkdir -p gpurun_out
timeout 120 python scripts/debug_part0.py > gpurun_out/p0_debug.log 2>&1; echo "debug rc=$?" >> gpurun_out/p0_debug.log
if grep -q "bad 0/" gpurun_out/p0_debug.log && ! grep -q "bad [1-9]" gpurun_out/p0_debug.log; then
  timeout 600 python -m pytest tests/test_gpu_cifar.py tests/test_gpu_handoff.py tests/test_service.py -m gpu -x -q > gpurun_out/p0_tests.log 2>&1; echo "rc=$?" >> gpurun_out/p0_tests.log
  timeout 300 python bench.py --steps 300 --warmup 10 --no-cpu-baseline > gpurun_out/p0_bench.json 2> gpurun_out/p0_bench.err
fi
tail -40 gpurun_out/p0_debug.log; tail -5 gpurun_out/p0_tests.log 2>/dev/null; python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/p0_bench.json'))
    print('bench', d['value'], 'us part0', d['roofline']['us_per_launch'], 'frac', d['roofline']['frac'], 'part1', d['roofline_stage1']['us_per_launch'], d['e2e'].get('max_abs_diff_vs_oracle'))
except Exception as e:
    print('no bench', e)
PY
timeout 120 python scripts/trace_part0.py 1024 > gpurun_out/p0_trace.log 2>&1; grep -E "^n=[23] " gpurun_out/p0_trace.log
