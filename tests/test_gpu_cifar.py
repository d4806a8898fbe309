"""CIFAR parity on the GPU: the engine's CUDA parts, called through the C ABI, against
(i) golden vectors produced by the reference's own classes and (ii) the CPU oracle on
seeded inputs.  Tolerances (fp32 arithmetic, differs from oneDNN only by summation
order): activations |d| <= 2e-5, probabilities |d| <= 1e-5, arg-max class bit-exact
wherever the oracle's top-2 probability gap exceeds 1e-5 (near-ties are counted and
reported, not hidden)."""
import os

import numpy as np
import pytest
import torch

from oracle import cifar_oracle as co

pytestmark = pytest.mark.gpu
ACT_TOL, PROB_TOL, TIE = 2e-5, 1e-5, 1e-5


@pytest.fixture(scope="module")
def parts(cuda_engine):
    sd = co.init_state_dict(1234)
    full = cuda_engine.NeuralNetwork()
    out = []
    for k in (0, 1):                       # the reference's exact sequence, node.py:300-317
        p = cuda_engine.MODEL_PARTS_CLASSES[k](full).to("cuda:0")
        missing, unexpected = p.load_state_dict(sd, strict=False)
        assert not missing and len(unexpected) == 4
        out.append(p.eval())
    return sd, out[0], out[1]


def test_golden_from_reference_classes(parts, golden_dir):
    _, p0, p1 = parts
    g = np.load(os.path.join(golden_dir, "cifar_b8.npz"))
    with torch.no_grad():
        act = p0(torch.from_numpy(g["x"]).cuda())
        probs = p1(act)
        probs_from_golden_act = p1(torch.from_numpy(g["part0"]).cuda())
    assert act.shape == (8, 4096) and act.dtype == torch.float32 and act.is_contiguous()
    np.testing.assert_allclose(act.cpu().numpy(), g["part0"], rtol=0, atol=ACT_TOL)
    np.testing.assert_allclose(probs.cpu().numpy(), g["probs"], rtol=0, atol=PROB_TOL)
    np.testing.assert_allclose(probs_from_golden_act.cpu().numpy(), g["probs"], rtol=0, atol=PROB_TOL)
    assert np.array_equal(probs.argmax(1).cpu().numpy(), g["probs"].argmax(1))


@pytest.mark.parametrize("batch", [1, 3, 64, 128, 130, 1024])
def test_parity_vs_oracle(parts, batch):
    sd, p0, p1 = parts
    x = co.make_input(batch, seed=100 + batch)
    with torch.no_grad():
        act = p0(x.cuda())
        probs = p1(act).cpu()
    want_act, want = co.part0(x, sd), co.full(x, sd)
    assert (act.cpu() - want_act).abs().max().item() <= ACT_TOL
    assert (probs - want).abs().max().item() <= PROB_TOL
    decided = co.top2_margin(want) > TIE
    assert torch.equal(probs.argmax(1)[decided], want.argmax(1)[decided])
    np.testing.assert_allclose(probs.sum(1).numpy(), np.ones(batch), atol=1e-6)


def test_argmax_bit_exact_over_10k_samples(parts):
    sd, p0, p1 = parts
    n, flips, ties = 10240, 0, 0
    for chunk in range(n // 1024):
        x = co.make_input(1024, seed=5000 + chunk)
        with torch.no_grad():
            probs = p1(p0(x.cuda())).cpu()
        want = co.full(x, sd)
        decided = co.top2_margin(want) > TIE
        ties += int((~decided).sum())
        flips += int((probs.argmax(1)[decided] != want.argmax(1)[decided]).sum())
        assert (probs - want).abs().max().item() <= PROB_TOL
    print(f"\n10240 samples: {flips} arg-max flips among decided rows, {ties} near-ties (gap<={TIE})")
    assert flips == 0
    assert ties <= 10          # near-ties are rare for random-init weights; report if not


def test_microbatching_is_bit_exact_at_full_size(parts):
    """BASELINE config 2: batch 1024 as 8 microbatches of 128 == one batch of 1024."""
    _, p0, p1 = parts
    x = co.make_input(1024, seed=77).cuda()
    with torch.no_grad():
        whole = p1(p0(x))
        pieces = torch.cat([p1(p0(x[i:i + 128])) for i in range(0, 1024, 128)])
    assert torch.equal(whole, pieces)


def test_pipeline_equals_full_model_and_is_deterministic(cuda_engine, parts):
    sd, p0, p1 = parts
    full = cuda_engine.NeuralNetwork()
    full.load_state_dict(sd)
    full = full.to("cuda:0")
    x = co.make_input(256, seed=9).cuda()
    with torch.no_grad():
        a, b, c = p1(p0(x)), full(x), p1(p0(x))
    assert torch.equal(a, b) and torch.equal(a, c)


def test_edge_inputs(parts):
    _, p0, p1 = parts
    with torch.no_grad():
        assert p0(torch.zeros(0, 3, 32, 32)).shape == (0, 4096)
        assert p1(torch.zeros(0, 4096)).shape == (0, 10)
        # all-zero image: conv output = bias path only, still a valid distribution
        z = p1(p0(torch.zeros(2, 3, 32, 32)))
        assert torch.allclose(z.sum(1), torch.ones(2, device=z.device), atol=1e-6)
        assert torch.equal(z[0], z[1])
        # non-contiguous (channels-last strided) input is accepted like torch would
        x = co.make_input(4, seed=1)
        nc = x.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
        assert not nc.is_contiguous()
        assert torch.equal(p0(nc.cuda()), p0(x.cuda()))
        # host tensors are moved to the stage's device (node.py:48 does .to(DEVICE))
        assert p0(x).is_cuda
    with pytest.raises(RuntimeError):
        p0(torch.zeros(2, 3, 31, 32).cuda())
    with pytest.raises(RuntimeError):
        p1(torch.zeros(2, 4095).cuda())
    with pytest.raises(RuntimeError):
        p0(torch.zeros(2, 3, 32, 32, dtype=torch.float64).cuda())


def test_weights_reload_takes_effect(cuda_engine):
    sd = co.init_state_dict(1234)
    sd2 = co.init_state_dict(4321)
    p1 = cuda_engine.ModelPart1_2Node(cuda_engine.NeuralNetwork()).to("cuda:0")
    x = torch.rand(5, 4096)
    p1.load_state_dict(sd, strict=False)
    a = p1.eval()(x).cpu()
    p1.load_state_dict(sd2, strict=False)
    b = p1(x).cpu()
    assert (a - co.part1(x, sd)).abs().max() <= PROB_TOL
    assert (b - co.part1(x, sd2)).abs().max() <= PROB_TOL
    assert not torch.equal(a, b)


def test_native_library_is_what_ran(cuda_engine, parts):
    _, p0, _ = parts
    before = cuda_engine.launch_count()
    p0(torch.zeros(1, 3, 32, 32).cuda())
    assert cuda_engine.launch_count() == before + 1
    maps = open("/proc/self/maps").read()
    assert "libb200pipe.so" in maps


@pytest.mark.parametrize("mode,tol", [("tf32x3", PROB_TOL), ("tf32", 2e-3)])
def test_tensor_core_fc_stage(cuda_engine, mode, tol):
    """Part 1 with fc1 on the tensor cores (split-K tcgen05 GEMM + the same fused tail).
    tf32x3 is held to the fp32 probability tolerance; single-pass tf32 to 2e-3."""
    sd = co.init_state_dict(1234)
    p1 = cuda_engine.ModelPart1_2Node(cuda_engine.NeuralNetwork(), math=mode).to("cuda:0")
    p1.load_state_dict(sd, strict=False)
    p1.eval()
    for batch in (1, 128, 130, 1024):
        x = co.part0(co.make_input(batch, seed=300 + batch), sd)
        with torch.no_grad():
            probs = p1(x.cuda()).cpu()
        want = co.part1(x, sd)
        err = (probs - want).abs().max().item()
        decided = co.top2_margin(want) > 2 * tol
        assert err <= tol, f"{mode} batch {batch}: {err}"
        assert torch.equal(probs.argmax(1)[decided], want.argmax(1)[decided])
    # micro-batching stays bit-exact on the tensor-core path too (fixed K split)
    x = co.part0(co.make_input(256, seed=5), sd).cuda()
    with torch.no_grad():
        assert torch.equal(p1(x), torch.cat([p1(x[:128]), p1(x[128:])]))


def test_fc1_in_kernel_split_rows_and_back_to_back_launches(cuda_engine):
    """Part 1 with the activation split inside the fc1 kernel (fc_stage_tc.cuh): ragged batches, rows at any
    offset inside a 128-row tile, the same answer call after call, and launches queued back to back on one
    stream (programmatic dependent launch must not let launch i+1 overwrite partial sums launch i still reads)."""
    sd = co.init_state_dict(1234)
    p1 = cuda_engine.ModelPart1_2Node(cuda_engine.NeuralNetwork(), math="tf32x3").to("cuda:0")
    p1.load_state_dict(sd, strict=False)
    p1.eval()
    x = co.part0(co.make_input(1000, seed=4242), sd)
    want = co.part1(x, sd)
    xd = x.cuda()
    with torch.no_grad():
        whole = p1(xd)
        assert (whole.cpu() - want).abs().max().item() <= PROB_TOL
        for lo, n in [(0, 1), (0, 5), (3, 127), (64, 129), (500, 257), (999, 1), (0, 1000)]:
            got = p1(xd[lo:lo + n].contiguous())
            assert torch.equal(got, whole[lo:lo + n]), f"rows {lo}..{lo + n}"
        # 64 launches queued without a host sync in between, alternating inputs
        a, b = xd[:128].contiguous(), xd[128:256].contiguous()
        outs = [p1(a if i % 2 == 0 else b) for i in range(64)]
        torch.cuda.synchronize()
        for i, o in enumerate(outs):
            assert torch.equal(o, whole[:128] if i % 2 == 0 else whole[128:256]), f"launch {i}"


@pytest.fixture(scope="module")
def tc_parts(cuda_engine):
    """Both parts in the tensor-core parity mode: conv2 as a split-fp16 implicit GEMM
    (tcgen05), fc1 as a 3xTF32 split-K GEMM."""
    sd = co.init_state_dict(1234)
    full = cuda_engine.NeuralNetwork()
    out = []
    for k in (0, 1):
        p = cuda_engine.MODEL_PARTS_CLASSES[k](full, math="tf32x3").to("cuda:0")
        p.load_state_dict(sd, strict=False)
        out.append(p.eval())
    return sd, out[0], out[1]


@pytest.mark.parametrize("batch", [1, 2, 3, 128, 149, 300, 1024])
def test_tensor_core_part0_parity(tc_parts, batch):
    """Same tolerances as the fp32 CUDA-core path: the split-fp16 contraction is fp32-grade."""
    sd, p0, p1 = tc_parts
    x = co.make_input(batch, seed=700 + batch)
    with torch.no_grad():
        act = p0(x.cuda())
        probs = p1(act).cpu()
    want_act, want = co.part0(x, sd), co.full(x, sd)
    assert act.shape == (batch, 4096)
    err = (act.cpu() - want_act).abs().max().item()
    assert err <= ACT_TOL, f"activation error {err}"
    assert (probs - want).abs().max().item() <= PROB_TOL
    decided = co.top2_margin(want) > TIE
    assert torch.equal(probs.argmax(1)[decided], want.argmax(1)[decided])


def test_tensor_core_golden_and_10k(tc_parts, golden_dir):
    sd, p0, p1 = tc_parts
    g = np.load(os.path.join(golden_dir, "cifar_b8.npz"))
    with torch.no_grad():
        act = p0(torch.from_numpy(g["x"]).cuda())
        probs = p1(act)
    np.testing.assert_allclose(act.cpu().numpy(), g["part0"], rtol=0, atol=ACT_TOL)
    np.testing.assert_allclose(probs.cpu().numpy(), g["probs"], rtol=0, atol=PROB_TOL)
    flips = ties = 0
    worst = 0.0
    for chunk in range(10):
        x = co.make_input(1024, seed=9000 + chunk)
        with torch.no_grad():
            probs = p1(p0(x.cuda())).cpu()
        want = co.full(x, sd)
        decided = co.top2_margin(want) > TIE
        ties += int((~decided).sum())
        flips += int((probs.argmax(1)[decided] != want.argmax(1)[decided]).sum())
        worst = max(worst, (probs - want).abs().max().item())
    print(f"\ntensor-core mode, 10240 samples: {flips} flips, {ties} near-ties, max|dprob|={worst:.2e}")
    assert flips == 0 and worst <= PROB_TOL


def test_tensor_core_microbatching_bit_exact_and_matches_fp32_path(tc_parts, parts):
    _, p0, p1 = tc_parts
    _, f0, f1 = parts
    x = co.make_input(1024, seed=78).cuda()
    with torch.no_grad():
        whole = p1(p0(x))
        pieces = torch.cat([p1(p0(x[i:i + 128])) for i in range(0, 1024, 128)])
        ref = f1(f0(x))
        a_tc, a_f32 = p0(x), f0(x)
    assert torch.equal(whole, pieces)
    assert (a_tc - a_f32).abs().max().item() <= ACT_TOL      # device-side check at full size
    assert (whole - ref).abs().max().item() <= PROB_TOL


# ---- edge semantics (VERDICT r1 "edge semantics differ from torch and are untested") ---------------------------------
def _tc_parts(cuda_engine, sd, math="tf32x3"):
    full = cuda_engine.NeuralNetwork()
    out = []
    for k in (0, 1):
        p = cuda_engine.MODEL_PARTS_CLASSES[k](full, math=math).to("cuda:0")
        p.load_state_dict(sd, strict=False)
        out.append(p.eval())
    return out


@pytest.mark.parametrize("math", ["fp32", "tf32x3"])
def test_nan_inputs_propagate_like_torch(cuda_engine, math):
    """nn.ReLU and nn.MaxPool2d propagate NaN (fmaxf would drop it): a NaN pixel poisons exactly the activations the
    reference's does, every probability of that sample is NaN, and the other samples of the batch are untouched."""
    sd = co.init_state_dict(1234)
    p0, p1 = _tc_parts(cuda_engine, sd, math)
    x = co.make_input(3, seed=41)
    x[1, 2, 10, 17] = float("nan")
    want_act, want = co.part0(x, sd), co.full(x, sd)
    with torch.no_grad():
        act = p0(x.cuda())
        probs = p1(act).cpu()
    act = act.cpu()
    assert torch.equal(torch.isnan(act), torch.isnan(want_act))           # same poisoned set (conv2 spreads it 3x3)
    assert torch.isnan(want_act[1]).any() and not torch.isnan(want_act[[0, 2]]).any()
    ok = ~torch.isnan(want_act)
    assert (act[ok] - want_act[ok]).abs().max().item() <= ACT_TOL
    assert torch.isnan(probs[1]).all() and torch.isnan(want[1]).all()
    assert (probs[[0, 2]] - want[[0, 2]]).abs().max().item() <= PROB_TOL
    p0.check()                                                            # NaN is data, not a range error


def test_unnormalised_images_stay_inside_the_tensor_core_range(cuda_engine):
    """Raw 0..255 pixel values (a client that skipped node.py:142-147's Normalize): |conv1 output| stays far below the
    fp16 limit of the split scheme, results agree with the oracle to the same RELATIVE accuracy."""
    sd = co.init_state_dict(1234)
    p0, p1 = _tc_parts(cuda_engine, sd, "tf32x3")
    g = torch.Generator().manual_seed(3)
    x = torch.randint(0, 256, (64, 3, 32, 32), generator=g).float()
    want_act, want = co.part0(x, sd), co.full(x, sd)
    with torch.no_grad():
        act = p0(x.cuda())
        probs = p1(act).cpu()
    scale = want_act.abs().max().item()
    assert scale > 50                                                     # the activations really are large
    assert (act.cpu() - want_act).abs().max().item() <= ACT_TOL * scale
    assert (probs - want).abs().max().item() <= 1e-4                      # logits are O(100): softmax amplifies 1e-5 relative
    decided = co.top2_margin(want) > 1e-3
    assert torch.equal(probs.argmax(1)[decided], want.argmax(1)[decided])
    p0.check()                                                            # no range error


def test_fp16_range_of_the_tensor_core_modes_is_guarded(cuda_engine):
    """conv1 weights x1000 on 0..255 images: conv1 outputs reach ~1e6, beyond what fp16 hi/lo parts can carry
    (include/b200pipe.h).  The tensor-core mode flags it (b2p_stage_check -> B2P_ELIMIT) instead of returning garbage
    silently; the fp32 mode has the reference's range and still agrees with the oracle."""
    sd = {k: v.clone() for k, v in co.init_state_dict(1234).items()}
    sd["conv1.weight"] *= 1000.0
    g = torch.Generator().manual_seed(5)
    x = torch.randint(0, 256, (8, 3, 32, 32), generator=g).float()
    want_act = co.part0(x, sd)
    assert torch.isfinite(want_act).all() and want_act.abs().max() > 65504
    L = cuda_engine._lib
    p0_tc = _tc_parts(cuda_engine, sd, "tf32x3")[0]
    with torch.no_grad():
        p0_tc(x.cuda())
    with pytest.raises(L.EngineError) as ei:
        p0_tc.check()
    assert ei.value.code == L.ELIMIT and "fp16 range" in str(ei.value)
    p0_tc.check()                                                         # reported once, then clear
    with torch.no_grad():
        ok = p0_tc(co.make_input(8, seed=1).cuda() * 1e-3)                # small inputs: back inside the range
    p0_tc.check()
    assert torch.isfinite(ok).all()
    p0_f32 = _tc_parts(cuda_engine, sd, "fp32")[0]
    with torch.no_grad():
        act = p0_f32(x.cuda()).cpu()
    assert (act - want_act).abs().max().item() <= 2e-5 * want_act.abs().max().item()


def test_fp16_range_of_the_tensor_core_fc1_is_guarded(cuda_engine):
    """Part 1 in tf32x3 runs fc1 in split fp16 (fc_stage_tc.cuh): an activation at or beyond 65504 cannot be carried by
    the fp16 hi/lo parts.  The stage flags it (b2p_stage_check -> B2P_ELIMIT); large but representable activations
    (the 0..255-image case, O(1000)) keep the oracle's relative accuracy; fp32 mode has the reference's range."""
    sd = co.init_state_dict(1234)
    L = cuda_engine._lib
    p1 = _tc_parts(cuda_engine, sd, "tf32x3")[1]
    x = co.part0(co.make_input(130, seed=77), sd)
    big = x * 2000.0                                                     # activations up to a few thousand
    want = co.part1(big, sd)
    with torch.no_grad():
        probs = p1(big.cuda()).cpu()
    p1.check()                                                           # inside the range: no error
    decided = co.top2_margin(want) > 1e-3
    assert torch.equal(probs.argmax(1)[decided], want.argmax(1)[decided])
    assert (probs - want).abs().max().item() <= 1e-3                     # logits are O(1000): softmax amplifies 1e-7 relative
    x2 = x.clone()
    x2[5, 100] = 70000.0
    with torch.no_grad():
        p1(x2.cuda())
    with pytest.raises(L.EngineError) as ei:
        p1.check()
    assert ei.value.code == L.ELIMIT and "fp16 range" in str(ei.value)
    p1.check()                                                           # reported once, then clear
    p1_f32 = _tc_parts(cuda_engine, sd, "fp32")[1]
    with torch.no_grad():
        got = p1_f32(x2.cuda()).cpu()
    assert (got - co.part1(x2, sd)).abs().max().item() <= PROB_TOL
