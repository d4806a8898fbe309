"""Producer-only timing of the span path: part0 whole step into GPU 1's ring, consumer = release kernels only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
from importlib import import_module
Ring = import_module("distributed-neural-networks_b200.handoff").Ring
MB, NMB, SLOTS = 128, 8, 16          # two steps of slots, as in the bench
d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
torch.manual_seed(1234)
full = engine.NeuralNetwork(); sd = full.state_dict()
p0 = engine.MODEL_PARTS_CLASSES[0](full, math="tf32x3").to(d0); p0.load_state_dict(sd, strict=False); p0.eval()
x = torch.randn(NMB * MB, 3, 32, 32, device=d0)
ring = Ring.create(1, SLOTS, MB * 4096 * 4); prod = ring.open_local(0)
with torch.cuda.device(d0):
    sp = torch.cuda.Stream(device=d0)
with torch.cuda.device(d1):
    sc = torch.cuda.Stream(device=d1)
step_no = [0, 0]
def prod_step():
    with torch.cuda.device(d0):
        p0.forward_ring(x.data_ptr(), None, NMB * MB, 0, None, prod, (step_no[0] % 2) * NMB, sp.cuda_stream, MB)
    step_no[0] += 1
def cons_eager(first):
    with torch.cuda.device(d1):
        for i in range(NMB):
            ring.recv(first + i, sc.cuda_stream); ring.release(first + i, sc.cuda_stream)
for h in range(2):
    prod_step(); cons_eager(h * NMB)
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
gcs = []
with torch.cuda.device(d1):
    for h in range(2):
        gc = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gc, stream=sc):
            for i in range(NMB):
                ring.recv(h * NMB + i, sc.cuda_stream); ring.release(h * NMB + i, sc.cuda_stream)
        gcs.append(gc)
def cons_step():                         # the fake consumer as one graph launch per step
    with torch.cuda.device(d1), torch.cuda.stream(sc):
        gcs[step_no[1] % 2].replay()
    step_no[1] += 1
for _ in range(4):
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
print(f"producer into the ring (eager launches, fake consumer graph): {1e3 * e0.elapsed_time(e1) / 50:7.2f} us/step")

# clock64 timeline of CTA 0's publisher and epilogue in ring mode (last traced step)
import ctypes as C
L = engine._lib
buf = (C.c_longlong * 1024)()
L.check(L.lib().b2p_stage_trace(p0._handle, buf, 1024))
for _ in range(4):
    prod_step(); cons_step()
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
L.check(L.lib().b2p_stage_trace(p0._handle, buf, 1024))
t = list(buf); t0 = t[192]
for n in range(7):
    e2 = [t[(2 * 8 + n) * 8 + e] - t0 for e in range(8)]
    mm = [t[(0 * 8 + n) * 8 + e] - t0 for e in range(8)]
    print(f"n={n} mma d2_empty(h0)={mm[2]} d2_empty(h1)={mm[5]} | e2 loop_start={e2[0]} drain(h0)_done={e2[1]} d2(h1)_loaded={e2[2]} staged={e2[3]} | pub acquired={e2[4]} staged_seen={e2[5]} read_done={e2[6]} prev_counted={e2[7]}")
print("CTA 0: entry", t[710] - t0, "released", t[711] - t0, "exit", t[712] - t0)
ent = [t[256 + 3 * i] for i in range(148)]; rel = [t[257 + 3 * i] for i in range(148)]; ext = [t[258 + 3 * i] for i in range(148)]
e0 = min(ent)
us = lambda v: round((v - e0) / 1e3, 1)
print(f"per-CTA (us after the first entry): entry max {us(max(ent))}, released min/max {us(min(rel))}/{us(max(rel))}, exit min/median/max {us(min(ext))}/{us(sorted(ext)[74])}/{us(max(ext))}")
nl = t[719]
rel_k = [t[720 + (k & 15)] for k in range(max(0, nl - 16), nl)]
ext_k = [t[736 + (k & 15)] for k in range(max(0, nl - 16), nl)]
print("CTA 0 over the traced launches: release-to-release", [round((b - a) / 1e3, 1) for a, b in zip(rel_k, rel_k[1:])],
      "us; release-to-exit", [round((e - r) / 1e3, 1) for r, e in zip(rel_k, ext_k)], "us; exit-to-next-release",
      [round((r - e) / 1e3, 1) for e, r in zip(ext_k, rel_k[1:])], "us")
dur = [(round((ext[i] - rel[i]) / 1e3, 1), i) for i in range(148)]
print("exit - release per CTA (us):", " ".join(f"{i}:{d}" for d, i in dur))
