import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
dev = torch.device("cuda:0")
torch.manual_seed(1234)
full = engine.NeuralNetwork()
p = engine.ModelPart0_2Node(full, math="tf32x3").to(dev).eval()
buf = (C.c_longlong * 256)()
L = engine._lib
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 8
x = torch.randn(batch, 3, 32, 32, device=dev)
p(x)                                   # warm
L.check(L.lib().b2p_stage_trace(p._handle, buf, 256))     # arm
p(x); torch.cuda.synchronize()
L.check(L.lib().b2p_stage_trace(p._handle, buf, 256))
t = list(buf)
t0 = t[192]
names = {0: ["loop", "in_ready", "conv1_done", "act_empty_ok", "published"],
         1: ["loop", "tmem_empty_ok", "act_full_ok", "issued"],
         2: ["loop", "tmem_full_ok", "tmem_read", "stored"]}
print("cycles relative to end of set-up (CTA 0), batch", batch)
for n in range(8):
    for role in range(3):
        ev = [t[(role * 8 + n) * 8 + e] for e in range(len(names[role]))]
        print(f"n={n} {['prod','mma ','epi '][role]}", "  ".join(f"{nm}={v - t0 if v else None}" for nm, v in zip(names[role], ev)))
