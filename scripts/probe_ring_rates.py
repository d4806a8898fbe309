"""Where does the 2-GPU CIFAR pipeline lose time?  One process, two GPUs:
  a) part 0 alone, output in local memory vs in GPU 1's memory (peer stores), graph of 16 forwards
  b) producer alone: part0.forward_ring into a ring whose consumer only releases (one-thread kernels)
  c) consumer alone: part1.forward_ring from a ring whose producer only publishes (no data written)
  d) both, the real pipeline
All rates in us per 128-sample microbatch (CUDA events on the stream of the side being measured)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
from importlib import import_module
Ring = import_module("distributed-neural-networks_b200.handoff").Ring

MB, NMB, SLOTS = 128, 8, 8
SPAN = True      # one call per step (b2p_stage_forward_ring_span) instead of one per microbatch
d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
torch.manual_seed(1234)
full = engine.NeuralNetwork()
sd = full.state_dict()
p0 = engine.MODEL_PARTS_CLASSES[0](full, math="tf32x3").to(d0); p0.load_state_dict(sd, strict=False); p0.eval()
p1 = engine.MODEL_PARTS_CLASSES[1](full, math="tf32x3").to(d1); p1.load_state_dict(sd, strict=False); p1.eval()
x = torch.randn(NMB * MB, 3, 32, 32, device=d0)
out1 = torch.empty(NMB * MB, 10, device=d1)


def graph_of(dev, fn, reps):
    with torch.cuda.device(dev):
        gs = torch.cuda.Stream(device=dev)
        fn(gs.cuda_stream); torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=gs):
            fn(gs.cuda_stream)
    return g, gs


def time_graph(dev, g, gs, n_units, replays=20):
    with torch.cuda.device(dev), torch.cuda.stream(gs):
        g.replay(); torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(gs)
        for _ in range(replays):
            g.replay()
        e1.record(gs)
        torch.cuda.synchronize(dev)
    return 1e3 * e0.elapsed_time(e1) / (replays * n_units)


# a) plain part 0, local vs peer output
act_local = torch.empty(NMB * MB, 4096, device=d0)
act_peer = torch.empty(NMB * MB, 4096, device=d1)
with torch.cuda.device(d0):
    can = torch.cuda.can_device_access_peer(0, 1)
print("peer access 0->1:", can)
ring_tmp = Ring.create(1, 1, 1024); prod_tmp = ring_tmp.open_local(0)      # enables peer access 0 -> 1
for name, act in (("local", act_local), ("peer", act_peer)):
    def fn(s, act=act):
        for i in range(NMB):
            p0.forward_ptr(x[i * MB:].data_ptr(), act[i * MB:].data_ptr(), MB, 0, s)
    g, gs = graph_of(d0, fn, 1)
    print(f"a) part0 -> {name:5s} memory: {time_graph(d0, g, gs, NMB):6.2f} us/microbatch")

# b) producer alone (consumer = release kernels only)
ring = Ring.create(1, SLOTS, MB * 4096 * 4); prod = ring.open_local(0)
seq = [0]
phase = {"p": 0, "c": 0, "fp": 0, "fc": 0}
def prod_step(s, k="p"):
    base = phase[k]; phase[k] = (base + NMB) % SLOTS
    if SPAN:
        return p0.forward_ring(x.data_ptr(), None, NMB * MB, 0, None, prod, base, s, MB)
    for i in range(NMB):
        p0.forward_ring(x[i * MB:].data_ptr(), None, MB, 0, None, prod, base + i, s)
def fake_cons_step(s):
    base = phase["fc"]; phase["fc"] = (base + NMB) % SLOTS
    for i in range(NMB):
        ring.recv(base + i, s); ring.release(base + i, s)
gp, sp = graph_of(d0, prod_step, 1)
with torch.cuda.device(d1):
    sc = torch.cuda.Stream(device=d1)
    fake_cons_step(sc.cuda_stream); torch.cuda.synchronize(d1)     # consumes what the eager warm-up of graph_of produced
# graph capture above did not execute; now run producer graph with a fake consumer enqueued eagerly each replay
def run_pair(prod_graph, sp, cons_fn, sc, replays=20):
    with torch.cuda.device(d0), torch.cuda.stream(sp):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(sp)
        for _ in range(replays):
            prod_graph.replay()
        e1.record(sp)
    with torch.cuda.device(d1):
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(sc)
        for _ in range(replays):
            cons_fn()
        c1.record(sc)
    torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
    return 1e3 * e0.elapsed_time(e1) / (replays * NMB), 1e3 * c0.elapsed_time(c1) / (replays * NMB)
tp, tc = run_pair(gp, sp, lambda: fake_cons_step(sc.cuda_stream), sc)
print(f"b) producer alone (fake consumer): producer {tp:6.2f} us/microbatch (fake consumer stream {tc:6.2f})")

# d) real pipeline: consumer graph
def cons_step(s):
    base = phase["c"]; phase["c"] = (base + NMB) % SLOTS
    if SPAN:
        return p1.forward_ring(None, out1.data_ptr(), NMB * MB, 0, ring, None, base, s, MB)
    for i in range(NMB):
        p1.forward_ring(None, out1[i * MB:].data_ptr(), MB, 0, ring, None, base + i, s)
# warm-up pair eagerly (producer first)
with torch.cuda.device(d0):
    prod_step(sp.cuda_stream)
with torch.cuda.device(d1):
    cons_step(sc.cuda_stream)
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
with torch.cuda.device(d1):
    gc = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gc, stream=sc):
        cons_step(sc.cuda_stream)
def cons_replay():
    with torch.cuda.device(d1), torch.cuda.stream(sc):
        gc.replay()
tp, tc = run_pair(gp, sp, cons_replay, sc)
print(f"d) real pipeline: producer {tp:6.2f}, consumer {tc:6.2f} us/microbatch")

# c) consumer alone: fake producer publishes without writing
def fake_prod_step(s):
    base = phase["fp"]; phase["fp"] = (base + NMB) % SLOTS
    for i in range(NMB):
        prod.acquire(base + i, s); prod.publish(base + i, s)
with torch.cuda.device(d0):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(20):
        fake_prod_step(sp.cuda_stream)
with torch.cuda.device(d1), torch.cuda.stream(sc):
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record(sc)
    for _ in range(20):
        gc.replay()
    c1.record(sc)
torch.cuda.synchronize(d0); torch.cuda.synchronize(d1)
print(f"c) consumer alone (fake producer): consumer {1e3 * c0.elapsed_time(c1) / (20 * NMB):6.2f} us/microbatch")
