#!/bin/bash
# Round-2 GPU evidence pass: full GPU test-suite, smoke, both bench arms, the tcgen05 micro-probe.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2_smoke.log 2>&1
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
timeout 600 python bench.py --steps 1000 --warmup 10 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
(cd scripts/micro && timeout 120 ./mma_probe 1 > ../../gpurun_out/r2_mma_probe.log 2>&1)
tail -5 gpurun_out/r2_pytest_gpu.log; tail -c 600 gpurun_out/r2_bench.json
