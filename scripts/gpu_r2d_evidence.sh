#!/bin/bash
# Round-2 (final) single-GPU evidence: full GPU test-suite, smoke, both bench arms, then -- each only after its plain run
# exited 0 -- the ncu launch list of the bench command and one --set full capture of the two dominant CIFAR kernels.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2d_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2d_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2d_smoke.log
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2d_bench_reference.json 2> gpurun_out/r2d_bench_reference.err
timeout 600 python bench.py --steps 1000 --warmup 10 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; brc=$?
tail -4 gpurun_out/r2d_pytest_gpu.log; tail -2 gpurun_out/r2d_smoke.log; tail -c 300 gpurun_out/r2d_bench.json
if [ $brc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2d_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-gpt > gpurun_out/r2d_ncu_bench.log 2>&1
fi
timeout 120 python scripts/ncu_targets.py > gpurun_out/r2d_targets_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"part0_tc2|fc1_f16x3_splitk|fc_tail_f32" -s 9 -c 6 \
  -o gpurun_out/r2d_cifar_kernels python scripts/ncu_targets.py > gpurun_out/r2d_ncu_full.log 2>&1
ncu -i gpurun_out/r2d_cifar_kernels.ncu-rep --page raw --csv > gpurun_out/r2d_cifar_kernels_1024_full_raw.csv 2>/dev/null
ls -la gpurun_out | grep r2d
