"""Greedy decoding throughput with and without the KV cache (GPT-2 small geometry, random-init, one GPU).
argv: math batch prompt new"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
math = sys.argv[1] if len(sys.argv) > 1 else "bf16"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
T0 = int(sys.argv[3]) if len(sys.argv) > 3 else 512
NEW = int(sys.argv[4]) if len(sys.argv) > 4 else 64
model = engine.GPT(engine.GPTConfig(block_size=1024, vocab_size=50257, **engine.GPT2_SMALL), seed=1)
parts = [p.to("cuda:0").eval() for p in engine.gpt_parts_for(model, 1, math=math, max_tokens=B * 1024)]
idx = torch.randint(0, 50257, (B, T0)).cuda()
for kv, graph in ((True, True), (True, False), (False, False)):
    engine.greedy_generate(parts, idx, 4, kv_cache=kv, graph=graph)    # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = engine.greedy_generate(parts, idx, NEW, kv_cache=kv, graph=graph)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"gpt2-small {math} B={B} prompt={T0} new={NEW} kv_cache={kv} cuda_graph={graph}: {dt * 1e3:8.1f} ms  "
          f"{B * NEW / dt:9.0f} new tokens/s  ({dt / NEW * 1e3:.3f} ms per step, prefill and graph capture included)")
    if kv and graph:
        ids_graph = out
    elif kv:
        ids_kv = out
        print(f"  graph decoding == eager decoding: {bool(torch.equal(ids_graph, ids_kv))}")
    else:
        same = (ids_kv == out).float().mean().item()
        print(f"  token agreement cached vs recomputed: {same * 100:.1f} %")
