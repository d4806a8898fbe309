"""Micro-benchmark of b2p_linear on the GPT microbatch shapes, to separate K-loop, epilogue and
launch-fixed costs."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
L = engine._lib
dev = "cuda:0"
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
s = torch.cuda.current_stream().cuda_stream
only = sys.argv[1] if len(sys.argv) > 1 else None


def run(tag, M, N, K, bn=0, res=False, out_bf16=False, act=0, bias=True, reps=20):
    if only and only != tag:
        return
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = torch.randn(N, K, device=dev).to(torch.bfloat16)
    b = torch.randn(N, device=dev) if bias else None
    r = torch.randn(M, N, device=dev) if res else None
    o = torch.empty(M, N, device=dev, dtype=torch.bfloat16 if out_bf16 else torch.float32)
    f = lambda: L.lib().b2p_linear(0 | (bn << 8), p(a), None, K, p(w), None, K, p(b), p(r), N if res else 0, p(o), N,
                                   int(out_bf16), act, M, N, K, C.c_void_p(s))
    for _ in range(3):
        assert f() == 0
    torch.cuda.synchronize()
    # replay the launches from a CUDA graph so the host-side cost of the call is not what is measured
    gs = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    fs = lambda: L.lib().b2p_linear(0 | (bn << 8), p(a), None, K, p(w), None, K, p(b), p(r), N if res else 0, p(o), N,
                                    int(out_bf16), act, M, N, K, C.c_void_p(gs.cuda_stream))
    with torch.cuda.graph(g, stream=gs):
        for _ in range(reps):
            fs()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / reps
    print(f"{tag:28s} M={M} N={N} K={K} bn={bn or 'auto'} res={int(res)} bf16out={int(out_bf16)} act={act}: {us:8.2f} us  {2*M*N*K/us/1e6:7.1f} TFLOP/s")


run("tiny", 128, 128, 64)
run("tiny_k768", 128, 128, 768)
run("one_wave_k768", 2048, 1152, 768, bn=128)          # 144 tiles
run("qkv", 2048, 2304, 768, out_bf16=True)
run("qkv_f32out", 2048, 2304, 768)
run("proj", 2048, 768, 768, res=True)
run("proj_nores", 2048, 768, 768)
run("proj_nores_nobias", 2048, 768, 768, bias=False)
run("proj_bf16", 2048, 768, 768, out_bf16=True)
run("fc_gelu", 2048, 3072, 768, out_bf16=True, act=2)
run("fc_gelu_bn128", 2048, 3072, 768, bn=128, out_bf16=True, act=2)
run("fc_noact", 2048, 3072, 768, out_bf16=True)
run("fc2", 2048, 768, 3072, res=True)
run("fc2_nores", 2048, 768, 3072)
run("head", 2048, 50257, 768, bias=False, reps=5)
run("big_m", 32768, 768, 768, res=True, reps=5)
run("fc8k_gelu", 8192, 3072, 768, out_bf16=True, act=2, reps=5)
run("fc8k_gelu_bn128", 8192, 3072, 768, bn=128, out_bf16=True, act=2, reps=5)
run("fc8k_noact", 8192, 3072, 768, out_bf16=True, reps=5)
run("fc8k_f32out", 8192, 3072, 768, reps=5)
run("qkv8k", 8192, 2304, 768, out_bf16=True, reps=5)
run("fc2_8k", 8192, 768, 3072, res=True, reps=5)
run("proj8k", 8192, 768, 768, res=True, reps=5)
run("sq4k", 4096, 4096, 4096, bias=False, bn=256, reps=5)
# the CTA-pair kernel (bn code 2) on the shapes that fill 74 pairs
run("sq4k_pair", 4096, 4096, 4096, bias=False, bn=2, reps=5)
run("sq8k_pair", 8192, 8192, 8192, bias=False, bn=2, out_bf16=True, reps=3)
run("sq8k_bn256", 8192, 8192, 8192, bias=False, bn=256, out_bf16=True, reps=3)
run("fc8k_gelu_pair", 8192, 3072, 768, bn=2, out_bf16=True, act=2, reps=5)
run("qkv8k_pair", 8192, 2304, 768, bn=2, out_bf16=True, reps=5)
run("fc2_8k_pair", 8192, 768, 3072, bn=2, res=True, reps=5)
run("head_pair", 2048, 50257, 768, bn=2, bias=False, reps=5)
run("head_bn256", 2048, 50257, 768, bn=256, bias=False, reps=5)
