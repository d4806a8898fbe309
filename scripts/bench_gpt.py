"""GPT throughput on one GPU (BASELINE configs 3/4 geometry): tokens/s for the whole model run as
the reference's part chain first -> final on a single device, plus per-kernel-class roofline
numbers from CUDA events.  Multi-GPU stage chains are measured by scripts/bench_gpt_ranks.py."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="small", choices=["small", "medium"])
ap.add_argument("--math", default="bf16")
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--seq", type=int, default=1024)
ap.add_argument("--microbatch", type=int, default=2, help="sequences per microbatch")
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--layers", type=int, default=0, help="override n_layer (0 = model default)")
ap.add_argument("--head", default="logits", choices=["logits", "ids"])
a = ap.parse_args()

dims = dict(engine.GPT2_SMALL if a.model == "small" else engine.GPT2_MEDIUM)
if a.layers:
    dims["n_layer"] = a.layers
cfg = engine.GPTConfig(block_size=1024, vocab_size=50257, **dims)
dev = torch.device("cuda:0")
model = engine.GPT(cfg, seed=1234)
mb_tokens = a.microbatch * a.seq
parts = [p.to(dev).eval() for p in engine.gpt_parts_for(model, 1, math=a.math, max_tokens=mb_tokens)]
first, final = parts
g = torch.Generator().manual_seed(0)
idx = torch.randint(0, cfg.vocab_size, (a.batch, a.seq), generator=g).to(dev)
hidden = torch.empty(a.microbatch, a.seq, cfg.n_embd, device=dev)
logits = torch.empty(a.microbatch, a.seq, cfg.vocab_size, device=dev) if a.head == "logits" else None
ids = torch.empty(a.microbatch, a.seq, dtype=torch.int64, device=dev)
s = torch.cuda.current_stream().cuda_stream
import ctypes as C
L = engine._lib


def step():
    for lo in range(0, a.batch, a.microbatch):
        first.forward_ptr(idx[lo:].data_ptr(), hidden.data_ptr(), a.microbatch, a.seq, s)
        if a.head == "logits":
            final.forward_ptr(hidden.data_ptr(), logits.data_ptr(), a.microbatch, a.seq, s)
        else:
            L.check(L.lib().b2p_stage_forward_argmax(final._handle, C.c_void_p(hidden.data_ptr()),
                                                     C.c_void_p(ids.data_ptr()), a.microbatch, a.seq, C.c_void_p(s)))


step(); step()
torch.cuda.synchronize()
l0 = engine.launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
tokens = a.batch * a.seq
C_, L_, V_, T_ = cfg.n_embd, cfg.n_layer, cfg.vocab_size, a.seq
flop_tok = L_ * (24 * C_ * C_ + 2 * T_ * C_) + 2 * C_ * V_          # SURVEY 8d (causal attention at half)
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
tf = flop_tok * tokens / (ms * 1e-3) / 1e12
print(json.dumps({"workload": f"gpt2-{a.model} L={L_} B={a.batch} T={a.seq} mb={a.microbatch}seq head={a.head}",
                  "math": a.math, "ms_per_batch": ms, "tokens_per_s": tokens / (ms * 1e-3),
                  "model_tflops": tf, "frac_of_sustained_bf16_peak": tf / peaks["bf16_tflops_sustained"],
                  "mflop_per_token": flop_tok / 1e6, "launches_per_batch": (engine.launch_count() - l0) / a.steps}))
