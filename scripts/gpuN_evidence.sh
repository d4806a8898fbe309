# bench.py (both arms) + the GPT chain of BASELINE configs 3/4 at N GPUs
mkdir -p gpurun_out
N=${1:-4}
bash scripts/gpuN_bench_only.sh $N
if [ "$N" = "4" ]; then
  RUNS=("--model small --math bf16 --microbatch 8 --balance" "--model small --math bf16 --microbatch 8" "--model small --math tf32x3 --microbatch 8 --steps 2")
else
  RUNS=("--model medium --math bf16 --microbatch 2 --batch 32" "--model medium --math bf16 --microbatch 8 --batch 128 --balance" "--model medium --math bf16 --microbatch 16 --batch 256 --balance")
fi
for args in "${RUNS[@]}"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 scripts/bench_gpt_ranks.py $args 2>gpurun_out/g${N}_gpt.err | tail -n 1 | tee -a gpurun_out/g${N}_gpt.log | cut -c1-400
  tail -n 2 gpurun_out/g${N}_gpt.err | grep -v "^\*\|OMP" | cut -c1-300
done
