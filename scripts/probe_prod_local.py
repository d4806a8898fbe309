"""Producer-only timing of the span path: part0 whole step into GPU 1's ring, consumer = release kernels only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
from importlib import import_module
Ring = import_module("distributed-neural-networks_b200.handoff").Ring
MB, NMB, SLOTS = 128, 8, 8
d0 = torch.device("cuda:0"); d1 = d0
torch.manual_seed(1234)
full = engine.NeuralNetwork(); sd = full.state_dict()
p0 = engine.MODEL_PARTS_CLASSES[0](full, math="tf32x3").to(d0); p0.load_state_dict(sd, strict=False); p0.eval()
x = torch.randn(NMB * MB, 3, 32, 32, device=d0)
ring = Ring.create(0, SLOTS, MB * 4096 * 4); prod = ring.open_local(0)
with torch.cuda.device(d0):
    sp = torch.cuda.Stream(device=d0)
with torch.cuda.device(d1):
    sc = torch.cuda.Stream(device=d1)
def prod_step():
    with torch.cuda.device(d0):
        p0.forward_ring(x.data_ptr(), None, NMB * MB, 0, None, prod, 0, sp.cuda_stream, MB)
def cons_eager():
    with torch.cuda.device(d1):
        for i in range(NMB):
            ring.recv(i, sc.cuda_stream); ring.release(i, sc.cuda_stream)
prod_step(); cons_eager()
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
with torch.cuda.device(d1):
    gc = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gc, stream=sc):
        for i in range(NMB):
            ring.recv(i, sc.cuda_stream); ring.release(i, sc.cuda_stream)
def cons_step():                         # the fake consumer as one graph launch: never the bottleneck
    with torch.cuda.device(d1), torch.cuda.stream(sc):
        gc.replay()
for _ in range(3):
    prod_step(); cons_step()
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
with torch.cuda.device(d0):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(sp)
for _ in range(50):
    prod_step(); cons_step()
with torch.cuda.device(d0):
    e1.record(sp)
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
print(f"producer into a ring on ITS OWN GPU (eager launches, fake consumer graph): {1e3 * e0.elapsed_time(e1) / 50:7.2f} us/step")
