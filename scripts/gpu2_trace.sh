#!/bin/bash
mkdir -p gpurun_out
B2P_TRACE_RING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 5 --no-gpt > gpurun_out/g2_bench.json 2> gpurun_out/g2_bench.err; python -c "import json; d=json.load(open('gpurun_out/g2_bench.json')); print(d['value'], d['stage_times'])"; grep "ring trace (last" gpurun_out/g2_bench.err
