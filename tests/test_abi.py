"""The C-ABI library loads and exports every symbol include/b200pipe.h declares.
No compute calls here (no GPU on the CPU test box)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "b200pipe.h")).read()
    return sorted(set(re.findall(r"B2P_API[^;]*?\b(b2p_\w+)\s*\(", text)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    import dnn_b200
    return dnn_b200


def test_header_and_binding_agree(built):
    assert _declared() == built._lib.exported_symbols()


def test_every_declared_symbol_is_exported(built):
    h = ctypes.CDLL(built._lib.LIB_PATH)
    for name in _declared():
        assert hasattr(h, name), name
    assert built._lib.lib().b2p_abi_version() == 1


def test_errors_are_reported_not_raised_across_the_abi(built):
    L = built._lib
    lib = L.lib()
    h = ctypes.c_void_p()
    d = L.StageDesc()
    d.kind = 99
    assert lib.b2p_stage_create(ctypes.byref(h), ctypes.byref(d)) == L.EINVAL
    assert b"unknown stage kind" in lib.b2p_last_error()
    d.kind, d.math = L.GPT_MIDDLE, L.MATH_FP32
    d.n_embd, d.n_head = 100, 2
    assert lib.b2p_stage_create(ctypes.byref(h), ctypes.byref(d)) == L.EINVAL
    assert lib.b2p_stage_forward(None, None, None, 1, 1, None) == L.EINVAL
    with pytest.raises(L.EngineError):
        L.check(lib.b2p_stage_finalize(None))


def test_no_cpu_fallback(built):
    """Without a GPU the product path must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    part = built.ModelPart1_2Node(built.NeuralNetwork())
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        part.to("cuda")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        part(torch.zeros(1, 4096))
    L = built._lib
    h, d = ctypes.c_void_p(), L.StageDesc()
    d.kind = L.CIFAR_PART0
    assert L.lib().b2p_stage_create(ctypes.byref(h), ctypes.byref(d)) == L.ECUDA


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "distributed-neural-networks_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f
                assert "/root/reference" not in text, f
