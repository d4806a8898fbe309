# launch list of the bench (device-resident region first, then the microbatched host-facing region) and full ncu
# captures of the two dominant kernels at the bench's launch shape (1024 rows per launch)
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 220 --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_final1.log 2>&1
timeout 300 python scripts/ncu_targets.py > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"part0_tc|fc1_x3_splitk" -s 6 -c 4 -o gpurun_out/final_cifar_kernels python scripts/ncu_targets.py > gpurun_out/ncu_final2.log 2>&1
tail -n 5 gpurun_out/ncu_final2.log; ls -la gpurun_out | tail -n 5
