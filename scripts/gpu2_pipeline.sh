set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/g2_topo.log 2>&1
timeout 600 python -m pytest tests/test_gpu_handoff.py -m gpu -q -x > gpurun_out/g2_handoff_tests.log 2>&1; echo "exit $?" >> gpurun_out/g2_handoff_tests.log; tail -n 5 gpurun_out/g2_handoff_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --math tf32x3 > gpurun_out/g2_bench.json 2> gpurun_out/g2_bench.err; echo "bench2 exit $?"; cat gpurun_out/g2_bench.json; tail -n 15 gpurun_out/g2_bench.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/handoff_sweep.py > gpurun_out/g2_sweep.json 2> gpurun_out/g2_sweep.err; echo "sweep exit $?"; cat gpurun_out/g2_sweep.json | head -120; tail -n 15 gpurun_out/g2_sweep.err
