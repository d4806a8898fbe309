"""Handoff rings on one GPU (data path, sequence flags, wrap-around, both copy modes).
Producer and consumer share the device here, so every wait is already satisfied when it
is enqueued (no kernel ever spins on another stream of the same GPU); the cross-GPU and
cross-process behaviour is exercised by scripts/handoff_sweep.py and bench.py --gpus 2."""
from importlib import import_module

import pytest
import torch

pytestmark = pytest.mark.gpu


class _Raw:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _view(ptr, n):
    return torch.as_tensor(_Raw(ptr, n), device="cuda:0")


@pytest.mark.parametrize("mode", ["sm_copy", "copy_engine"])
def test_ring_roundtrip_and_wraparound(cuda_engine, mode):
    Ring = import_module("distributed-neural-networks_b200.handoff").Ring
    n = (1 << 20) + 13                                   # odd size: exercises the byte tail
    ring = Ring.create(0, slots=2, slot_bytes=n)
    prod = ring.open_local(0)
    prod.set_mode(mode)
    s = torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(3)
    msgs = [torch.randint(0, 256, (n,), dtype=torch.uint8, generator=g).cuda() for _ in range(6)]
    seq = 0
    for round_ in range(3):                              # 3 rounds x 2 slots: slots are reused
        for k in range(2):
            prod.send(msgs[seq + k].data_ptr(), n, seq + k, s)
        torch.cuda.synchronize()
        for k in range(2):
            ptr = ring.recv(seq + k, s)
            torch.cuda.synchronize()
            assert torch.equal(_view(ptr, n), msgs[seq + k])
            ring.release(seq + k, s)
        torch.cuda.synchronize()
        seq += 2


def test_zero_copy_acquire_publish_with_a_stage(cuda_engine):
    """The stage kernel writes its output straight into the ring slot."""
    Ring = import_module("distributed-neural-networks_b200.handoff").Ring
    from oracle import cifar_oracle as co
    sd = co.init_state_dict(1234)
    p0 = cuda_engine.ModelPart0_2Node(cuda_engine.NeuralNetwork(), math="tf32x3").to("cuda:0")
    p0.load_state_dict(sd, strict=False)
    p0.eval()
    ring = Ring.create(0, slots=2, slot_bytes=16 * 4096 * 4)
    prod = ring.open_local(0)
    s = torch.cuda.current_stream().cuda_stream
    x = co.make_input(16, seed=4).cuda()
    slot = prod.acquire(0, s)
    p0.forward_ptr(x.data_ptr(), slot, 16, 0, s)
    prod.publish(0, s)
    torch.cuda.synchronize()
    ptr = ring.recv(0, s)
    torch.cuda.synchronize()
    got = _view(ptr, 16 * 4096 * 4).view(torch.float32).view(16, 4096)
    assert (got.cpu() - co.part0(x.cpu(), sd)).abs().max().item() <= 2e-5


@pytest.mark.parametrize("math", ["tf32x3", "fp32"])
def test_forward_ring_handshake_inside_the_stage_kernels(cuda_engine, math):
    """b2p_stage_forward_ring: part 0 stores into the slot and publishes it, part 1 waits, reads and releases --
    with the handshake inside the tensor-core kernels (tf32x3) or composed from the one-thread kernels (fp32).
    Twelve microbatches through a 2-slot ring (every counter wraps six times), ragged sizes, one stream in
    producer-then-consumer order; results must equal the plain two-call path bit for bit."""
    Ring = import_module("distributed-neural-networks_b200.handoff").Ring
    from oracle import cifar_oracle as co
    sd = co.init_state_dict(1234)
    parts = []
    for k in (0, 1):
        p = cuda_engine.MODEL_PARTS_CLASSES[k](cuda_engine.NeuralNetwork(), math=math).to("cuda:0")
        p.load_state_dict(sd, strict=False)
        parts.append(p.eval())
    p0, p1 = parts
    L = cuda_engine._lib
    ring = Ring.create(0, slots=2, slot_bytes=128 * 4096 * 4)
    prod = ring.open_local(0)
    s = torch.cuda.current_stream().cuda_stream
    sizes = [128, 128, 1, 77, 128, 128, 128, 5, 128, 128, 64, 128]
    xs = [co.make_input(n, seed=100 + i).cuda() for i, n in enumerate(sizes)]
    outs = [torch.empty(n, 10, device="cuda:0") for n in sizes]
    for seq, (x, o) in enumerate(zip(xs, outs)):
        p0.forward_ring(x.data_ptr(), None, x.shape[0], 0, None, prod, seq, s)
        p1.forward_ring(None, o.data_ptr(), x.shape[0], 0, ring, None, seq, s)
    torch.cuda.synchronize()
    with torch.no_grad():
        for x, o in zip(xs, outs):
            assert torch.equal(o, p1(p0(x)))
    # a whole step per call: 1000 rows as 8 messages of 128 (the last one ragged), one launch per stage where fused
    ring8 = Ring.create(0, slots=8, slot_bytes=128 * 4096 * 4)
    prod8 = ring8.open_local(0)
    xb = co.make_input(1000, seed=77).cuda()
    ob = torch.empty(1000, 10, device="cuda:0")
    for step in range(3):                                    # slots are reused from step to step
        p0.forward_ring(xb.data_ptr(), None, 1000, 0, None, prod8, 8 * step, s, 128)
        p1.forward_ring(None, ob.data_ptr(), 1000, 0, ring8, None, 8 * step, s, 128)
        torch.cuda.synchronize()
        with torch.no_grad():
            assert torch.equal(ob, p1(p0(xb))), f"step {step}"
        ob.zero_()
    # messages that are not 128 rows: the producer kernel still publishes them itself, the consumer falls back to one
    # call per message through the one-thread handshake kernels -- same ring, same counters, same results
    for step in range(2):
        p0.forward_ring(xb.data_ptr(), None, 300, 0, None, prod8, 24 + 3 * step, s, 100)
        p1.forward_ring(None, ob.data_ptr(), 300, 0, ring8, None, 24 + 3 * step, s, 100)
        torch.cuda.synchronize()
        with torch.no_grad():
            assert torch.equal(ob[:300], p1(p0(xb[:300]))), f"100-row messages, step {step}"
    with pytest.raises(L.EngineError, match="fewer slots"):
        p0.forward_ring(xb.data_ptr(), None, 1000, 0, None, prod8, 0, s, 64)
    if math == "tf32x3":
        # PDL mask bit 2: the stage-0 kernels of consecutive steps do not wait for each other (different slots, resident
        # inputs).  Three producer steps back to back into a 32-slot ring, then the three consumer steps: same bits.
        ring32 = Ring.create(0, slots=32, slot_bytes=128 * 4096 * 4)
        prod32 = ring32.open_local(0)
        xs3 = [co.make_input(1024, seed=90 + i).cuda() for i in range(3)]
        prev = p0.set_pdl(7)
        for rep in range(2):                                 # second round: the slots' counters are no longer zero
            for i, x3 in enumerate(xs3):
                p0.forward_ring(x3.data_ptr(), None, 1024, 0, None, prod32, 24 * rep + 8 * i, s, 128)
            for i, x3 in enumerate(xs3):
                ob3 = torch.empty(1024, 10, device="cuda:0")
                p1.forward_ring(None, ob3.data_ptr(), 1024, 0, ring32, None, 24 * rep + 8 * i, s, 128)
                torch.cuda.synchronize()
                with torch.no_grad():
                    assert torch.equal(ob3, p1(p0(x3))), f"no-wait ring mode, round {rep}, step {i}"
        p0.set_pdl(prev)
    # wrong side / oversize / missing operands are refused before anything is enqueued
    with pytest.raises(L.EngineError, match="consumer side"):
        p1.forward_ring(None, outs[0].data_ptr(), 1, 0, prod, None, 0, s)
    with pytest.raises(L.EngineError, match="producer side"):
        p0.forward_ring(xs[0].data_ptr(), None, 1, 0, None, ring, 0, s)
    with pytest.raises(L.EngineError, match="exceeds the slot size"):
        big = torch.zeros(129, 3, 32, 32, device="cuda:0")
        p0.forward_ring(big.data_ptr(), None, 129, 0, None, prod, 0, s)
    with pytest.raises(L.EngineError, match="no input"):
        p1.forward_ring(None, outs[0].data_ptr(), 1, 0, None, None, 0, s)


def test_ring_argument_errors(cuda_engine):
    L = cuda_engine._lib
    Ring = import_module("distributed-neural-networks_b200.handoff").Ring
    with pytest.raises(L.EngineError):
        Ring.create(0, slots=0, slot_bytes=1024)
    ring = Ring.create(0, slots=1, slot_bytes=1024)
    prod = ring.open_local(0)
    src = torch.zeros(4096, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(L.EngineError, match="exceeds the slot size"):
        prod.send(src.data_ptr(), 4096, 0, 0)
    with pytest.raises(L.EngineError, match="only the consumer side"):
        prod.export()
    assert len(ring.export()) == L.RING_HANDLE_BYTES
