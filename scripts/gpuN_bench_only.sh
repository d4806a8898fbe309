# bench.py (both arms) at N GPUs, nothing else
mkdir -p gpurun_out
N=${1:-4}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/g${N}_bench.json 2> gpurun_out/g${N}_bench.err; echo "bench$N exit $?"; python -c "
import json; d=json.load(open('gpurun_out/g${N}_bench.json')); print('N=$N', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e'].get('max_abs_diff_vs_single_gpu_path'), 'launches', d['gpu_launches'], d['clocks'])"; tail -n 3 gpurun_out/g${N}_bench.err | grep -v "^\*\|OMP"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --steps 100 --warmup 3 --impl reference > gpurun_out/g${N}_bench_ref.json 2> gpurun_out/g${N}_bench_ref.err; echo "ref$N exit $?"; cut -c1-300 gpurun_out/g${N}_bench_ref.json
