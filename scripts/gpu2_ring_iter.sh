mkdir -p gpurun_out
for extra in "" "--no-graph"; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 5 $extra 2> gpurun_out/g2_bench.err | tee gpurun_out/g2_bench$extra.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$extra', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e'].get('max_abs_diff_vs_single_gpu_path'), 'launches', d['gpu_launches'], d['roofline']['us_per_launch'])"
tail -n 4 gpurun_out/g2_bench.err
done
