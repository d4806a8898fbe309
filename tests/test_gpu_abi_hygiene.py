"""Behaviour the C ABI promises beyond arithmetic (include/b200pipe.h): no allocation inside forward once scratch is
reserved (so a forward can be captured in a CUDA graph), per-stage programmatic-launch masks, ring waits that give up
with an error record instead of trapping, and the KV workspace growing before the cache exists."""
import ctypes as C
from importlib import import_module

import pytest
import torch

from oracle import cifar_oracle as co

pytestmark = pytest.mark.gpu


def _part1(engine, math):
    sd = co.init_state_dict(1234)
    p = engine.ModelPart1_2Node(engine.NeuralNetwork(), math=math).to("cuda:0")
    p.load_state_dict(sd, strict=False)
    return sd, p.eval()


@pytest.mark.parametrize("math", ["fp32", "tf32x3", "tf32"])
def test_forward_allocates_nothing_within_the_reserved_batch(cuda_engine, math):
    """finalize sizes the scratch for 1024 rows: forwards up to that size leave cudaMemGetInfo's free bytes alone."""
    sd, p1 = _part1(cuda_engine, math)
    dev = torch.device("cuda", 0)
    xs = [torch.rand(b, 4096, device=dev) for b in (1024, 7, 300)]
    outs = [torch.empty(x.shape[0], 10, device=dev) for x in xs]
    s = torch.cuda.current_stream().cuda_stream
    p1.forward_ptr(xs[0].data_ptr(), outs[0].data_ptr(), 1024, 0, s)      # first launch: module load, attributes
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    for x, out in zip(xs, outs):
        p1.forward_ptr(x.data_ptr(), out.data_ptr(), x.shape[0], 0, s)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free0 == free1
    assert (outs[1].cpu() - co.part1(xs[1].cpu(), sd)).abs().max().item() <= (2e-3 if math == "tf32" else 1e-5)


def test_forward_under_capture_refuses_to_grow_and_reserve_fixes_it(cuda_engine):
    sd, p1 = _part1(cuda_engine, "tf32x3")
    L = cuda_engine._lib
    dev = torch.device("cuda", 0)
    big = 1536                                             # beyond finalize's 1024 rows
    x = torch.rand(big, 4096, device=dev)
    out = torch.empty(big, 10, device=dev)
    stream = torch.cuda.Stream(device=dev)
    g = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with pytest.raises(L.EngineError) as ei:
        with torch.cuda.graph(g, stream=stream):
            p1.forward_ptr(x.data_ptr(), out.data_ptr(), big, 0, stream.cuda_stream)
    assert ei.value.code == L.ESTATE and "b2p_stage_reserve" in str(ei.value)
    torch.cuda.synchronize()
    p1.reserve(big)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        p1.forward_ptr(x.data_ptr(), out.data_ptr(), big, 0, stream.cuda_stream)
    out.fill_(float("nan"))
    g.replay()
    torch.cuda.synchronize()
    assert (out.cpu() - co.part1(x.cpu(), sd)).abs().max().item() <= 1e-5


def test_pdl_mask_is_per_stage(cuda_engine):
    _, a = _part1(cuda_engine, "tf32x3")
    _, b = _part1(cuda_engine, "tf32x3")
    lib = cuda_engine._lib.lib()
    default = lib.b2p_set_pdl(3)
    assert a.set_pdl(0) == 3            # returns what applied before: the process default
    assert b.set_pdl(1) == 3
    assert a.set_pdl(2) == 0 and b.set_pdl(-1) == 1
    assert b.set_pdl(-1) == 3           # back on the default
    x = torch.rand(64, 4096, device="cuda:0")
    ya, yb = a(x), b(x)                 # same bits whatever the launch attribute is
    assert torch.equal(ya, yb)
    lib.b2p_set_pdl(default)


def test_ring_wait_gives_up_with_a_record_instead_of_trapping(cuda_engine):
    """A consumer waiting for a message that never comes: after the time-out the wait returns, b2p_ring_check reports
    B2P_ETIMEOUT once, and the context is still alive (a trap would have poisoned it for every later test)."""
    H = import_module("distributed-neural-networks_b200.handoff")
    L = cuda_engine._lib
    ring = H.Ring.create(0, slots=2, slot_bytes=4096)
    s = torch.cuda.current_stream().cuda_stream
    H.check(0)                                            # clean to start with
    H.set_timeout(0, 0.2)
    try:
        ring.recv(0, s)                                   # nobody publishes message 0
        torch.cuda.synchronize()                          # returns after ~0.2 s instead of hanging
        with pytest.raises(L.EngineError) as ei:
            H.check(0)
        assert ei.value.code == L.ETIMEOUT and "gave up" in str(ei.value)
        H.check(0)                                        # the record was cleared
    finally:
        H.set_timeout(0, 20.0)
    # the context survived: ordinary work still runs
    t = torch.arange(1024, device="cuda:0").float()
    assert float(t.sum()) == 1023 * 1024 / 2
    ring.close()


def test_tail_kernel_on_a_second_stream(cuda_engine):
    """b2p_stage_set_aux_stream: fc1 on the caller's stream, the tail on a second stream behind an event, partial sums in
    alternating buffers.  Back-to-back forwards into separate outputs must each equal the single-stream result bit for
    bit (same kernels, same association), and b2p_stage_join is what makes them visible to the caller's stream."""
    sd, p1 = _part1(cuda_engine, "tf32x3")
    dev = torch.device("cuda", 0)
    xs = [torch.rand(b, 4096, device=dev) for b in (1024, 1024, 300, 1024, 7)]
    with torch.no_grad():
        want = [p1(x).clone() for x in xs]                       # single-stream results
    main, aux = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    p1.set_aux_stream(aux.cuda_stream)
    outs = [torch.full((x.shape[0], 10), float("nan"), device=dev) for x in xs]
    for x, out in zip(xs, outs):                                 # enqueued without any host synchronisation in between
        p1.forward_ptr(x.data_ptr(), out.data_ptr(), x.shape[0], 0, main.cuda_stream)
    p1.join(main.cuda_stream)
    main.synchronize()
    for got, w, x in zip(outs, want, xs):
        assert torch.equal(got, w)
    assert (outs[2].cpu() - co.part1(xs[2].cpu(), sd)).abs().max().item() <= 1e-5
    p1.set_aux_stream(None)                                      # back to one stream
    with torch.no_grad():
        assert torch.equal(p1(xs[4]), want[4])
