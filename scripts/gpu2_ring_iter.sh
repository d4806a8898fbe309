mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_handoff.py tests/test_gpu_cifar.py -m gpu -q -x 2>&1 | tail -4
timeout 300 python bench.py --steps 500 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N=1 graph', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['p50_request_latency_ms'], d['roofline']['us_per_launch'], d['roofline']['frac'], d['gpu_launches'])"
for extra in "" "--no-graph"; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 5 $extra 2> gpurun_out/g2_bench.err | tee gpurun_out/g2_bench$extra.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('N=2 $extra', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e'].get('max_abs_diff_vs_single_gpu_path'), 'launches', d['gpu_launches'], d['roofline']['us_per_launch'])"
tail -n 4 gpurun_out/g2_bench.err | grep -v "^\*\|OMP_NUM"
done
