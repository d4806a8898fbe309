"""Producer-only timing of the span path with NO consumer: part0 whole steps into a ring deep enough to hold them all."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, ctypes as C
import dnn_b200 as engine
from importlib import import_module
Ring = import_module("distributed-neural-networks_b200.handoff").Ring
MB, NMB, STEPS = 128, 8, 12
SLOTS = NMB * STEPS
d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
torch.manual_seed(1234)
full = engine.NeuralNetwork(); sd = full.state_dict()
p0 = engine.MODEL_PARTS_CLASSES[0](full, math="tf32x3").to(d0); p0.load_state_dict(sd, strict=False); p0.eval()
x = torch.randn(NMB * MB, 3, 32, 32, device=d0)
o = torch.empty(NMB * MB, 4096, device=d0)
ring = Ring.create(1, SLOTS, MB * 4096 * 4); prod = ring.open_local(0)
with torch.cuda.device(d0):
    sp = torch.cuda.Stream(device=d0)
    for _ in range(20):                                   # warm the clocks with local-output launches
        p0.forward_ptr(x.data_ptr(), o.data_ptr(), NMB * MB, 0, sp.cuda_stream)
    torch.cuda.synchronize(d0)
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    L = engine._lib
    buf = (C.c_longlong * 1024)()
    L.check(L.lib().b2p_stage_trace(p0._handle, buf, 1024))
    e0.record(sp)
    for _ in range(STEPS):
        p0.forward_ptr(x.data_ptr(), o.data_ptr(), NMB * MB, 0, sp.cuda_stream)
    e1.record(sp)
    torch.cuda.synchronize(d0)
    L.check(L.lib().b2p_stage_trace(p0._handle, buf, 1024))
    t = list(buf)
    rel = [t[257 + 3 * i] for i in range(148)]; ext = [t[258 + 3 * i] for i in range(148)]
    dur = sorted(((ext[i] - rel[i]) / 1e3, i) for i in range(148))
    print("LOCAL exit - release (us): min %.1f median %.1f max %.1f" % (dur[0][0], dur[74][0], dur[-1][0]))
    print("LOCAL CTA 0 cycles: released", t[711] - t[192], "exit", t[712] - t[192], "; per-sample d2_empty(h0):", [t[(0 * 8 + n) * 8 + 2] - t[192] for n in range(7)])
    L.check(L.lib().b2p_stage_trace(p0._handle, buf, 1024))
    e1.record(sp)
    if os.environ.get("B2P_RING_WAIT") != "1":
        p0.set_pdl(7)
    for i in range(STEPS):
        p0.forward_ring(x.data_ptr(), None, NMB * MB, 0, None, prod, i * NMB, sp.cuda_stream, MB)
    e2.record(sp)
    torch.cuda.synchronize(d0)
    L.check(L.lib().b2p_stage_trace(p0._handle, buf, 1024))
print(f"local output {1e3 * e0.elapsed_time(e1) / STEPS:6.2f} us/step, into the peer ring (no consumer, free slots) {1e3 * e1.elapsed_time(e2) / STEPS:6.2f} us/step")
t = list(buf)
ent = [t[256 + 3 * i] for i in range(148)]; rel = [t[257 + 3 * i] for i in range(148)]; ext = [t[258 + 3 * i] for i in range(148)]
dur = sorted(((ext[i] - rel[i]) / 1e3, i) for i in range(148))
print("exit - release (us): min %.1f median %.1f max %.1f; slowest:" % (dur[0][0], dur[74][0], dur[-1][0]), " ".join(f"{i}:{d:.1f}" for d, i in dur[-8:]))
print("RING CTA 0 cycles: released", t[711] - t[192], "exit", t[712] - t[192], "; per-sample d2_empty(h0):", [t[(0 * 8 + n) * 8 + 2] - t[192] for n in range(7)])
print("RING CTA 0 publisher (staged_seen, read_done, counted) - staged:", [(t[(2*8+n)*8+5] - t[(2*8+n)*8+3], t[(2*8+n)*8+6] - t[(2*8+n)*8+3], t[(2*8+n)*8+7] - t[(2*8+n)*8+3]) for n in range(7)])
nl = t[719]
rel_k = [t[720 + (k & 15)] for k in range(max(0, nl - 16), nl)]
print("release-to-release (CTA 0):", [round((b - a) / 1e3, 1) for a, b in zip(rel_k, rel_k[1:])])
for m in range(8):
    print(f"message {m}: published by CTA {t[760 + 4 * m] // 16} at its sample {t[760 + 4 * m] % 16}: copy completion {t[761 + 4 * m]} cycles, fence + ticket {t[762 + 4 * m]}, fence + system-scope release {t[763 + 4 * m]}")
