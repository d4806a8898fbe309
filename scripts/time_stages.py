"""Times the CIFAR stage kernels alone (CUDA events on the launch stream) for several batch
sizes, to separate per-launch start-up cost from steady-state per-sample cost."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine

math = sys.argv[1] if len(sys.argv) > 1 else "tf32x3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dev = torch.device("cuda:0")
torch.manual_seed(1234)
full = engine.NeuralNetwork()
sd = full.state_dict()
parts = []
for k in (0, 1):
    p = engine.MODEL_PARTS_CLASSES[k](full, math=math).to(dev)
    p.load_state_dict(sd, strict=False)
    parts.append(p.eval())
s = torch.cuda.current_stream().cuda_stream
for batch in (1, 128, 148, 296, 1024, 4096):
    x = torch.randn(batch, 3, 32, 32, device=dev)
    act = torch.empty(batch, 4096, device=dev)
    out = torch.empty(batch, 10, device=dev)
    for name, part, a, b in (("part0", parts[0], x, act), ("part1", parts[1], act, out)):
        for _ in range(3):
            part.forward_ptr(a.data_ptr(), b.data_ptr(), batch, 0, s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            part.forward_ptr(a.data_ptr(), b.data_ptr(), batch, 0, s)
        e1.record()
        torch.cuda.synchronize()
        us = 1e3 * e0.elapsed_time(e1) / reps
        print(f"{math} {name} batch {batch:5d}: {us:9.2f} us/launch  {us / batch:8.3f} us/sample")
