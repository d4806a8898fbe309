set -x
mkdir -p gpurun_out
N=8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/g${N}_bench.json 2> gpurun_out/g${N}_bench.err; echo "bench$N exit $?"; cut -c1-700 gpurun_out/g${N}_bench.json; tail -n 4 gpurun_out/g${N}_bench.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 100 --warmup 3 --impl reference > gpurun_out/g${N}_bench_ref.json 2> gpurun_out/g${N}_bench_ref.err; echo "ref$N exit $?"; cut -c1-300 gpurun_out/g${N}_bench_ref.json
for args in "--model medium --math bf16 --microbatch 2 --batch 32" "--model medium --math bf16 --microbatch 8 --batch 128" "--model medium --math bf16 --microbatch 8 --batch 128 --balance" "--model medium --math bf16 --microbatch 16 --batch 256 --balance" "--model small --math bf16 --microbatch 8 --batch 128 --balance"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 scripts/bench_gpt_ranks.py $args 2>gpurun_out/g${N}_gpt.err | tail -n 1 | tee -a gpurun_out/g${N}_gpt.log
  tail -n 2 gpurun_out/g${N}_gpt.err | cut -c1-300
done
