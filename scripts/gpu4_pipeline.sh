set -x
mkdir -p gpurun_out
N=${1:-4}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/g${N}_bench.json 2> gpurun_out/g${N}_bench.err; echo "bench$N exit $?"; cut -c1-900 gpurun_out/g${N}_bench.json; tail -n 5 gpurun_out/g${N}_bench.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --steps 20 --warmup 3 --impl reference > gpurun_out/g${N}_bench_ref.json 2> gpurun_out/g${N}_bench_ref.err; echo "ref$N exit $?"; cut -c1-400 gpurun_out/g${N}_bench_ref.json
for args in "--model small --math bf16 --microbatch 2" "--model small --math bf16 --microbatch 8" "--model small --math bf16 --microbatch 8 --balance" "--model medium --math bf16 --microbatch 8 --batch 128 --balance" "--model small --math tf32x3 --microbatch 8 --steps 2"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 scripts/bench_gpt_ranks.py $args 2>gpurun_out/g${N}_gpt.err | tail -n 1 | tee -a gpurun_out/g${N}_gpt.log
  tail -n 3 gpurun_out/g${N}_gpt.err | cut -c1-300
done
