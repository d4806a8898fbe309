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
