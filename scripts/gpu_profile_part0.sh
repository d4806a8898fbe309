set -x
mkdir -p gpurun_out
timeout 300 python scripts/time_stages.py tf32x3 > gpurun_out/time_stages.log 2>&1; cat gpurun_out/time_stages.log
timeout 300 python bench.py --steps 50 --warmup 5 --math tf32x3 --no-cpu-baseline > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench exit $?"; cat gpurun_out/bench_graph.json; tail -5 gpurun_out/bench_graph.err
timeout 300 python scripts/time_stages.py tf32x3 3 > /dev/null 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:part0_tc -s 6 -c 2 -o gpurun_out/r1e_part0_tc python scripts/time_stages.py tf32x3 3 > gpurun_out/ncu_p0.log 2>&1
tail -3 gpurun_out/ncu_p0.log
