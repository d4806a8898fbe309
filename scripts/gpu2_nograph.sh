#!/bin/bash
mkdir -p gpurun_out
for flags in "--no-graph" ""; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 5 --no-gpt $flags > gpurun_out/g2_bench_x.json 2> gpurun_out/g2_bench_x.err; python -c "import json; d=json.load(open('gpurun_out/g2_bench_x.json')); print('$flags', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['max_abs_diff_vs_single_gpu_path'], d['config']['launch'])"
done
