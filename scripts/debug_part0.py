#!/usr/bin/env python
"""Developer aid: part 0 on the GPU against the CPU oracle, with error localisation.
   python scripts/debug_part0.py            (run under a time-out on the GPU box)
Cases: the seeded default-init weights, and probe weights that isolate one convolution (conv2 = centre-tap identity
shows conv1's output pooled twice; conv1 = centre-tap copy of the input channels shows conv2 alone)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dnn_b200 as engine
from oracle import cifar_oracle as co


def run(sd, x, math="tf32x3"):
    p = engine.ModelPart0_2Node(engine.NeuralNetwork(), math=math).to("cuda:0")
    p.load_state_dict(sd, strict=False)
    p.eval()
    with torch.no_grad():
        y = p(x.cuda())
    torch.cuda.synchronize()
    return y.cpu()


def report(name, got, want):
    d = (got - want).abs()
    bad = d > 2e-5 * max(1.0, float(want.abs().max()))
    print(f"{name}: max|d|={float(d.max()):.3e} (|want| up to {float(want.abs().max()):.3f}), bad {int(bad.sum())}/{bad.numel()}, "
          f"finite={bool(torch.isfinite(got).all())}")
    if bad.any():
        b = bad.view(-1, 64, 8, 8)
        print("   bad per sample   :", b.sum((1, 2, 3)).tolist()[:16])
        print("   bad per channel  :", b.sum((0, 2, 3)).tolist())
        print("   bad per pooled h :", b.sum((0, 1, 3)).tolist())
        print("   bad per pooled w :", b.sum((0, 1, 2)).tolist())
        idx = bad.view(-1).nonzero()[:6].view(-1).tolist()
        for i in idx:
            print(f"   flat {i}: got {float(got.view(-1)[i]):.6f} want {float(want.view(-1)[i]):.6f}")


sd = co.init_state_dict(1234)
for B in (1, 2, 150, 1024):
    x = co.make_input(B, seed=7 + B)
    report(f"default weights, B={B}", run(sd, x), co.part0(x, sd))

# probe 1: conv2 = identity on the centre tap -> output channel oc shows conv1 channel oc%32 (pooled again)
sd1 = {k: v.clone() for k, v in sd.items()}
sd1["conv2.weight"].zero_(); sd1["conv2.bias"].zero_()
for oc in range(64):
    sd1["conv2.weight"][oc, oc % 32, 1, 1] = 1.0
x = co.make_input(3, seed=11)
report("probe conv1 (conv2 = centre identity)", run(sd1, x), co.part0(x, sd1))

# probe 2: conv1 = centre-tap copy of input channel oc%3 (bias 0) -> conv2 alone on a known activation
sd2 = {k: v.clone() for k, v in sd.items()}
sd2["conv1.weight"].zero_(); sd2["conv1.bias"].zero_()
for oc in range(32):
    sd2["conv1.weight"][oc, oc % 3, 1, 1] = 1.0 + oc / 64.0
report("probe conv2 (conv1 = centre copy)", run(sd2, x), co.part0(x, sd2))

# probe 3: single taps of conv2
for tap in (0, 4, 8):
    sd3 = {k: v.clone() for k, v in sd2.items()}
    w = torch.zeros_like(sd["conv2.weight"]); w[:, :, tap // 3, tap % 3] = sd["conv2.weight"][:, :, tap // 3, tap % 3]
    sd3["conv2.weight"] = w
    report(f"probe conv2 tap {tap} only", run(sd3, x), co.part0(x, sd3))
print("old kernel for comparison: B2P_PART0_V1=1")
