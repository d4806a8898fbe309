"""CIFAR stage kernels replayed from a CUDA graph (20 forwards per replay), so the host-side launch cost is not
what is measured.  argv: math [batch ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine

math = sys.argv[1] if len(sys.argv) > 1 else "tf32x3"
batches = [int(a) for a in sys.argv[2:]] or [1, 128, 1024]
reps = 20
dev = torch.device("cuda:0")
torch.manual_seed(1234)
full = engine.NeuralNetwork()
sd = full.state_dict()
parts = []
for k in (0, 1):
    p = engine.MODEL_PARTS_CLASSES[k](full, math=math).to(dev)
    p.load_state_dict(sd, strict=False)
    parts.append(p.eval())
for batch in batches:
    x = torch.randn(batch, 3, 32, 32, device=dev)
    act = torch.randn(batch, 4096, device=dev).abs()
    out = torch.empty(batch, 10, device=dev)
    for name, part, a, b in (("part0", parts[0], x, act), ("part1", parts[1], act, out)):
        s = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            part.forward_ptr(a.data_ptr(), b.data_ptr(), batch, 0, s)
        torch.cuda.synchronize()
        gs = torch.cuda.Stream()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=gs):
            for _ in range(reps):
                part.forward_ptr(a.data_ptr(), b.data_ptr(), batch, 0, gs.cuda_stream)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        us = 1e3 * e0.elapsed_time(e1) / (5 * reps)
        print(f"{math} {name} batch {batch:5d}: {us:9.2f} us/forward (graph)  {us / batch:8.3f} us/sample")
