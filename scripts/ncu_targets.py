"""The two dominant CIFAR kernels at the bench's launch shape (one launch per 1024-sample step), for ncu:
3 warm-up launches of each stage, then 2 more of each (the ones to capture: -k regex:"part0_tc2|fc1_f16x3_splitk|fc_tail_f32" -s 9 -c 6)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
dev = torch.device("cuda:0")
torch.manual_seed(1234)
full = engine.NeuralNetwork(); sd = full.state_dict()
parts = []
for k in (0, 1):
    p = engine.MODEL_PARTS_CLASSES[k](full, math="tf32x3").to(dev); p.load_state_dict(sd, strict=False); parts.append(p.eval())
B = 1024
x = torch.randn(B, 3, 32, 32, device=dev); act = torch.empty(B, 4096, device=dev); out = torch.empty(B, 10, device=dev)
s = torch.cuda.current_stream().cuda_stream
for reps in (3, 2):
    for _ in range(reps):
        parts[0].forward_ptr(x.data_ptr(), act.data_ptr(), B, 0, s)
    for _ in range(reps):
        parts[1].forward_ptr(act.data_ptr(), out.data_ptr(), B, 0, s)
    torch.cuda.synchronize()
print("ok", float(out.sum()))
