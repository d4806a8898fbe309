set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_handoff.py -m gpu -q -x > gpurun_out/g2_handoff_tests.log 2>&1; echo "exit $?" >> gpurun_out/g2_handoff_tests.log; tail -n 3 gpurun_out/g2_handoff_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/g2_bench.json 2> gpurun_out/g2_bench.err; echo "bench2 exit $?"; cut -c1-1200 gpurun_out/g2_bench.json; tail -n 6 gpurun_out/g2_bench.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 100 --warmup 5 --no-graph 2>/dev/null | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/handoff_sweep.py > gpurun_out/g2_sweep.json 2> gpurun_out/g2_sweep.err; echo "sweep exit $?"; python -c "
import json; d=json.load(open('gpurun_out/g2_sweep.json'))
for r in d['handoff_sweep']: print(r['bytes'], round(r['ring_copy_engine_GBps'],1), round(r['ring_sm_copy_GBps'],1), round(r['nccl_sendrecv_GBps'],1), r['ring_sm_copy_ok'])
"
