"""Consumer-only timing of the span path: part1 whole step from a ring whose producer only publishes (no data)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
from importlib import import_module
Ring = import_module("distributed-neural-networks_b200.handoff").Ring
MB, NMB, SLOTS = 128, 8, 8
d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
torch.manual_seed(1234)
full = engine.NeuralNetwork(); sd = full.state_dict()
p1 = engine.MODEL_PARTS_CLASSES[1](full, math="tf32x3").to(d1); p1.load_state_dict(sd, strict=False); p1.eval()
out1 = torch.empty(NMB * MB, 10, device=d1)
xloc = torch.randn(NMB * MB, 4096, device=d1).abs()
ring = Ring.create(1, SLOTS, MB * 4096 * 4); prod = ring.open_local(0)
with torch.cuda.device(d0):
    sp = torch.cuda.Stream(device=d0)
with torch.cuda.device(d1):
    sc = torch.cuda.Stream(device=d1)
with torch.cuda.device(d0):
    gp = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gp, stream=sp):
        for i in range(NMB):
            prod.publish(i, sp.cuda_stream)      # publish only: the counters just run ahead of the consumer
def prod_step():
    with torch.cuda.device(d0), torch.cuda.stream(sp):
        gp.replay()
def cons_step():
    with torch.cuda.device(d1):
        p1.forward_ring(None, out1.data_ptr(), NMB * MB, 0, ring, None, 0, sc.cuda_stream, MB)
def local_step():
    with torch.cuda.device(d1):
        p1.forward_ptr(xloc.data_ptr(), out1.data_ptr(), NMB * MB, 0, sc.cuda_stream)
def timeit(fn, n=50):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
    with torch.cuda.device(d1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(sc)
    for _ in range(n):
        fn()
    with torch.cuda.device(d1):
        e1.record(sc)
    torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
    return 1e3 * e0.elapsed_time(e1) / n
t_local = timeit(local_step)
# the fake producer publishes 50 steps ahead is impossible (8 slots): interleave, the producer side is two tiny kernels per message
t_ring = timeit(lambda: (prod_step(), cons_step()))
print(f"consumer local {t_local:7.2f} us/step, from ring {t_ring:7.2f} us/step")
