"""GPT stage chain across ranks (BASELINE config 3: GPT-2 small, 4 stages; config 4: GPT-2 medium,
8 stages, bf16, 16 microbatches).  Launch:
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_gpt_ranks.py --model medium
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import dnn_b200 as engine
from importlib import import_module

rp = import_module("distributed-neural-networks_b200.rank_pipeline")
ap = argparse.ArgumentParser()
ap.add_argument("--model", default="small", choices=["small", "medium"])
ap.add_argument("--math", default="bf16")
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--seq", type=int, default=1024)
ap.add_argument("--microbatch", type=int, default=2)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--head", default="ids", choices=["ids", "logits"])
ap.add_argument("--balance", action="store_true", help="give the final stage fewer blocks (LM head is heavy)")
a = ap.parse_args()

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
rp.quiet_nccl()
dist.init_process_group("nccl", device_id=dev)
dims = dict(engine.GPT2_SMALL if a.model == "small" else engine.GPT2_MEDIUM)
cfg = engine.GPTConfig(block_size=1024, vocab_size=50257, **dims)
model = engine.GPT(cfg, seed=1234)
ranges = engine.balanced_ranges(cfg, world, a.seq) if (a.balance and world > 1) else None
role = rp.plan(world, world)[rank]
pipe = rp.GptRankPipeline(role, dev, model, a.math, a.microbatch, a.seq, ranges=ranges)
n_mb = a.batch // a.microbatch
g = torch.Generator().manual_seed(0)
idx = torch.randint(0, cfg.vocab_size, (a.batch, a.seq), generator=g).to(dev) if rank == 0 else None
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
for mode in ("eager launches with programmatic dependent launch", "one CUDA graph per rank per batch"):
    step = pipe.step if mode.startswith("eager") else pipe.step_graph
    for _ in range(3):
        step(idx, n_mb, a.head)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(pipe.stream)
    for _ in range(a.steps):
        step(idx, n_mb, a.head)
    e1.record(pipe.stream)
    torch.cuda.synchronize(); dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0]) / a.steps
    if rank == 0:
        tokens = a.batch * a.seq
        C_, L_, V_ = cfg.n_embd, cfg.n_layer, cfg.vocab_size
        flop_tok = L_ * (24 * C_ * C_ + 2 * a.seq * C_) + 2 * C_ * V_
        tf = flop_tok * tokens / (ms * 1e-3) / 1e12
        print(json.dumps({"workload": f"gpt2-{a.model} {world} stages B={a.batch} T={a.seq} microbatch={a.microbatch} seq head={a.head}",
                          "launch": mode, "steps": a.steps, "ranges": ranges or "even", "math": a.math, "n_gpus": world, "ms_per_batch": ms,
                          "tokens_per_s": tokens / (ms * 1e-3), "model_tflops_total": tf,
                          "frac_of_sustained_bf16_peak_per_gpu": tf / world / peaks["bf16_tflops_sustained"]}), flush=True)
dist.barrier()
dist.destroy_process_group()
