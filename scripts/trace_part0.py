"""Developer aid: clock64 timeline of CTA 0 of part0_tc2 (b2p_stage_trace), cycles relative to the end of set-up."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
dev = torch.device("cuda:0")
torch.manual_seed(1234)
full = engine.NeuralNetwork()
p = engine.ModelPart0_2Node(full, math="tf32x3").to(dev).eval()
buf = (C.c_longlong * 1024)()
L = engine._lib
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
x = torch.randn(batch, 3, 32, 32, device=dev)
p(x)                                   # warm
L.check(L.lib().b2p_stage_trace(p._handle, buf, 1024))     # arm
p(x); torch.cuda.synchronize()
L.check(L.lib().b2p_stage_trace(p._handle, buf, 1024))
t = list(buf)
t0 = t[192]
names = {0: ["iter", "act_ready0", "d2_empty(h0)", "Q2h0_issued", "C1(n+1,h0)_issued", "d2_empty(h1)", "Q2h1_issued", "C1(n+1,h1)_issued"],
         1: ["e1_read(n+1,h0)_done", "e1_write(h0)_done", "e1_read(n+1,h1)_done", "e1_write(h1)_done", "mma:C1h1_patch_full", "mma:C1h1_d1_empty", "mma:C1h1_mmas_issued", "mma:C1h1_commit1", ],
         2: ["loop_start", "drain(h0)_done", "d2(h1)_loaded", "staged", "-", "-", "-", "-"],
         3: ["-", "image_staged", "patch_empty(h0)", "patches(h0)", "patch_empty(h1)", "patches(h1)", "-", "-"]}
print("cycles relative to end of set-up (CTA 0), batch", batch)
ent = [t[256 + 3 * i] for i in range(148)]; rel = [t[257 + 3 * i] for i in range(148)]; ext = [t[258 + 3 * i] for i in range(148)]
e0 = min(ent)
us = lambda v: round((v - e0) / 1e3, 1)
print(f"per-CTA (us after the first entry): entry max {us(max(ent))}, released min/max {us(min(rel))}/{us(max(rel))}, exit min/median/max {us(min(ext))}/{us(sorted(ext)[74])}/{us(max(ext))}")
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for bsz in (128, 1024):
    xb = x[:bsz].contiguous()
    p(xb); torch.cuda.synchronize()
    ev0.record()
    for _ in range(50):
        p.forward_ptr(xb.data_ptr(), torch.empty(bsz, 4096, device=dev).data_ptr(), bsz, 0, torch.cuda.current_stream().cuda_stream)
    ev1.record(); torch.cuda.synchronize()
    print(f"batch {bsz}: {ev0.elapsed_time(ev1) / 50 * 1e3:.1f} us per launch (50 back to back)")
for n in range(8):
    for role in range(4):
        ev = [t[(role * 8 + n) * 8 + e] for e in range(8)]
        if role == 3 and n == 0:
            ev[0] = 0
        print(f"n={n} {['mma','e1 ','e2 ','pat'][role]}", "  ".join(f"{nm}={v - t0}" for nm, v in zip(names[role], ev) if v and nm != "-"))
