import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dnn_b200 as engine
L = engine._lib
cfg = engine.GPTConfig(block_size=1024, vocab_size=512, n_layer=1, n_head=12, n_embd=768)
model = engine.GPT(cfg, seed=1)
parts = [p.to("cuda:0").eval() for p in engine.gpt_parts_for(model, 2, math="bf16", max_tokens=8 * 1024)]
idx = torch.randint(0, 512, (8, 1024)).cuda()
for _ in range(2):
    parts[0](idx)
torch.cuda.synchronize()
buf = (C.c_longlong * 193)()
L.check(L.lib().b2p_debug_attn_trace(buf, 193))
t = list(buf); t0 = t[192]
ev_m = ["w0:wait_p", "w0:p_ok", "w0:pv_issued", "w0:s_next", "w1:wait_p", "w1:p_ok", "w1:pv_issued", "w1:s_next"]
ev_s = ["loop", "s_ok", "max_done", "p_published", "rescaled", "o_ok", "o_added"]
for j in range(8):
    print(f"j={j} mma ", "  ".join(f"{n}={t[(0*8+j)*8+e]-t0}" for e, n in enumerate(ev_m) if t[(0*8+j)*8+e]))
    for w in (0, 1):
        print(f"j={j} wg{w} ", "  ".join(f"{n}={t[((1+w)*8+j)*8+e]-t0}" for e, n in enumerate(ev_s) if t[((1+w)*8+j)*8+e]))
