#!/bin/bash
# N-GPU evidence (N = 4 or 8): both bench arms (CIFAR replicas + the GPT config of that N under the `gpt` key), handoff sweep
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29521 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/g${N}_bench.json 2> gpurun_out/g${N}_bench.err; echo "bench$N exit $?"
python -c "
import json; d=json.load(open('gpurun_out/g${N}_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['max_abs_diff_vs_single_gpu_path'], d.get('stage_times')); print(json.dumps(d['gpt'])[:1200])"
tail -n 3 gpurun_out/g${N}_bench.err | cut -c1-300
timeout 900 $TR --master-port 29522 scripts/handoff_sweep.py > gpurun_out/g${N}_sweep.json 2> gpurun_out/g${N}_sweep.err; echo "sweep exit $?"
python -c "
import json; d=json.load(open('gpurun_out/g${N}_sweep.json'))
for r in d['handoff_sweep']: print(r['bytes'], round(r['ring_copy_engine_GBps'],1), round(r['ring_sm_copy_GBps'],1), round(r['nccl_sendrecv_GBps'],1), r['ring_sm_copy_ok'], r['reference_grpc_GBps'])
print(d['clocks_rank0'])"
tail -n 3 gpurun_out/g${N}_sweep.err | cut -c1-300
