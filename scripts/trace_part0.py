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
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1      # > 1: the stamps are those of the last of `reps` back-to-back launches
o = torch.empty(batch, 4096, device=dev)
for _ in range(reps):
    p.forward_ptr(x.data_ptr(), o.data_ptr(), batch, 0, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
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
print(f"CTA 0 (cycles relative to end of set-up): entry {t[710] - t0}, released {t[711] - t0}, exit {t[712] - t0}; ns: entry {us(ent[0])}, released {us(rel[0])}, exit {us(ext[0])} us")
nl = t[719]
if nl > 1:
    rel_k = [t[720 + (k & 15)] for k in range(max(0, nl - 16), nl)]
    ext_k = [t[736 + (k & 15)] for k in range(max(0, nl - 16), nl)]
    print("CTA 0 over the traced launches: release-to-release", [round((b - a) / 1e3, 1) for a, b in zip(rel_k, rel_k[1:])],
          "us; release-to-exit", [round((e - r) / 1e3, 1) for r, e in zip(rel_k, ext_k)], "us; exit-to-next-release",
          [round((r - e) / 1e3, 1) for e, r in zip(ext_k, rel_k[1:])], "us")
if nl > 1:
    print("raw ns (CTA 0, minus first entry): entry", ent[0] - e0, "released", rel[0] - e0, "exit", ext[0] - e0, "| ring: released",
          [v - e0 for v in rel_k[-3:]], "exit", [v - e0 for v in ext_k[-3:]])
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
