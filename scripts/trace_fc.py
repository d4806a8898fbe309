"""Timeline of the fused CIFAR part-1 kernel (B2P_FC_TRACE=1): per-event clock64 deltas, median / max over the 128 CTAs."""
import os, sys, ctypes as C
os.environ["B2P_FC_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import dnn_b200 as engine
from oracle import cifar_oracle as co
L = engine._lib
sd = co.init_state_dict(1234)
p1 = engine.ModelPart1_2Node(engine.NeuralNetwork(), math="tf32x3").to("cuda:0")
p1.load_state_dict(sd, strict=False); p1.eval()
x = torch.randn(128, 4096, device="cuda:0").abs()
for _ in range(5):
    p1(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); p1(x); e1.record(); torch.cuda.synchronize()
print("single launch (events):", e0.elapsed_time(e1) * 1e3, "us")
buf = (C.c_longlong * 2048)()
L.check(L.lib().b2p_debug_fc_trace(buf, 2048))
t = np.array(list(buf), dtype=np.int64).reshape(128, 16)
names = ["start", "first_tile_landed", "first_item_converted", "first_acc_ready", "first_item_stored"]
rel = t - t[:, :1]
for e, n in enumerate(names):
    col = rel[:, e][t[:, e] != 0]
    if len(col):
        print(f"{e:2d} {n:18s} n={len(col):3d} median={int(np.median(col)):7d} min={col.min():7d} max={col.max():7d} cycles")
