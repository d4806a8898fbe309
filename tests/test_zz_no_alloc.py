"""The allocation-free fc-stage entry point, b2p_cifar_part1_forward_noalloc.

It exists for callers that cannot let this library allocate device memory -- a serving
thread that must stay clear of the device-wide synchronisation cudaMalloc/cudaFree
imply, or a host that handed all GPU memory to its own pool at start-up -- and keeps its
split-K partial sums in a fixed array inside the library instead.  Here it is called
through the C ABI exactly as such a caller would (raw device pointers for inputs,
weights and outputs) and compared with the regular entry point, b2p_stage_forward on a
ModelPart1_2Node stage, on two calls in a row: the second call uses a smaller batch and
other inputs, so anything stale left in the shared array by the first would show.
Tolerance: the suite's probability tolerance, 1e-5 (the two paths run the same kernels,
so the results are in fact expected to be identical).

The file name sorts last on purpose: a run that stops at the first failure has then
already executed every other test.  It must pass both alone in a fresh process and after
the rest of the suite.
"""
import ctypes as C

import pytest
import torch

from oracle import cifar_oracle as co

pytestmark = pytest.mark.gpu
PROB_TOL = 1e-5


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _noalloc(engine, x, w, out):
    stream = torch.cuda.current_stream(x.device).cuda_stream
    return engine._lib.lib().b2p_cifar_part1_forward_noalloc(
        _ptr(x), _ptr(w["fc1.weight"]), _ptr(w["fc1.bias"]), _ptr(w["fc2.weight"]), _ptr(w["fc2.bias"]),
        _ptr(out), x.shape[0], C.c_void_p(stream))


def test_noalloc_matches_regular_entry_point_twice(cuda_engine):
    engine = cuda_engine
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sd = co.init_state_dict(1234)
    regular = engine.ModelPart1_2Node(engine.NeuralNetwork()).to(dev)
    regular.load_state_dict(sd, strict=False)
    regular.eval()
    w = {k: sd[k].to(dev).contiguous() for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")}

    g = torch.Generator().manual_seed(42)
    for batch in (engine._lib.NOALLOC_MAX_BATCH, 37):          # two calls in a row, full then ragged
        x = torch.rand(batch, 4096, generator=g).to(dev)
        out = torch.full((batch, 10), float("nan"), device=dev)
        rc = _noalloc(engine, x, w, out)
        assert rc == 0, engine._lib.lib().b2p_last_error()
        with torch.no_grad():
            want = regular(x)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        assert (out - want).abs().max().item() <= PROB_TOL
        assert torch.equal(out.argmax(1), want.argmax(1))
        # and against the CPU oracle, like every other parity test
        assert (out.cpu() - co.part1(x.cpu(), sd)).abs().max().item() <= PROB_TOL


def test_noalloc_rejects_oversized_and_bad_arguments(cuda_engine):
    engine = cuda_engine
    L = engine._lib
    dev = torch.device("cuda", 0)
    sd = co.init_state_dict(1234)
    w = {k: sd[k].to(dev).contiguous() for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")}
    big = L.NOALLOC_MAX_BATCH + 1
    x = torch.zeros(big, 4096, device=dev)
    out = torch.zeros(big, 10, device=dev)
    assert _noalloc(engine, x, w, out) == L.ELIMIT
    msg = L.lib().b2p_last_error().decode()
    assert "exceeds" in msg and str(L.NOALLOC_MAX_BATCH) in msg
    assert L.lib().b2p_cifar_part1_forward_noalloc(None, None, None, None, None, None, 1, None) == L.EINVAL
    assert _noalloc(engine, x[:0], w, out[:0]) == L.EINVAL     # empty batch: nothing to launch
    # the kernels read x and the weights with 128-bit loads: a pointer that is not 16-byte aligned is refused
    flat = torch.zeros(4 * 4096 + 4, device=dev)
    skew = flat[1:1 + 4 * 4096].view(4, 4096)
    assert skew.data_ptr() % 16 != 0
    assert _noalloc(engine, skew, w, out[:4]) == L.EINVAL
    assert "16-byte aligned" in L.lib().b2p_last_error().decode()


def test_noalloc_leaves_free_device_memory_unchanged(cuda_engine):
    """What the name promises, measured: once the GPU has seen its first call (which is when the driver loads the
    library image and sets the fixed array aside), further calls neither take nor return device memory."""
    engine = cuda_engine
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sd = co.init_state_dict(1234)
    w = {k: sd[k].to(dev).contiguous() for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")}
    xs = [torch.rand(b, 4096, device=dev) for b in (engine._lib.NOALLOC_MAX_BATCH, 5, 130)]
    outs = [torch.empty(x.shape[0], 10, device=dev) for x in xs]
    assert _noalloc(engine, xs[0], w, outs[0]) == 0             # first call on this GPU: one-time set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)                     # cudaMemGetInfo: driver-level, sees through torch's pool
    for _ in range(3):
        for x, out in zip(xs, outs):
            assert _noalloc(engine, x, w, out) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, f"free device memory moved by {free0 - free1} bytes across allocation-free calls"
