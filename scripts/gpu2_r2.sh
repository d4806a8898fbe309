#!/bin/bash
# 2-GPU check of the ring path: handoff tests, the multi-GPU bench line (bit-exactness vs the single-GPU path inside)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_handoff.py -m gpu -q -x > gpurun_out/g2_handoff_tests.log 2>&1; echo "exit $?" >> gpurun_out/g2_handoff_tests.log; tail -n 3 gpurun_out/g2_handoff_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 5 --no-gpt > gpurun_out/g2_bench.json 2> gpurun_out/g2_bench.err; echo "bench2 exit $?"; cut -c1-1500 gpurun_out/g2_bench.json; tail -n 6 gpurun_out/g2_bench.err
