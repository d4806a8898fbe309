"""ctypes binding of libb200pipe.so (C ABI in include/b200pipe.h).

The product path has no CPU fallback: if the CUDA library is missing this module
raises at first use, loudly, instead of computing anything on the host.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200pipe.so")

# mirror of the header's constants
OK, EINVAL, ECUDA, ESTATE, EMISSING, ELIMIT, ETIMEOUT = 0, -1, -2, -3, -4, -5, -6
CIFAR_PART0, CIFAR_PART1, GPT_FIRST, GPT_MIDDLE, GPT_FINAL = range(5)
MATH_FP32, MATH_TF32X3, MATH_TF32, MATH_BF16 = range(4)
MATH_NAMES = {"fp32": MATH_FP32, "tf32x3": MATH_TF32X3, "tf32": MATH_TF32, "bf16": MATH_BF16}
DT_F32, DT_BF16, DT_I64 = range(3)
RING_HANDLE_BYTES = 256
NOALLOC_MAX_BATCH = 256


class StageDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("math", C.c_int32), ("device", C.c_int32),
                ("start_layer", C.c_int32), ("end_layer", C.c_int32), ("n_head", C.c_int32),
                ("n_embd", C.c_int32), ("vocab_size", C.c_int32), ("block_size", C.c_int32),
                ("max_tokens", C.c_int32), ("reserved", C.c_int32 * 6)]


class EngineError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


_PROTOTYPES = {
    "b2p_abi_version": (C.c_int, []),
    "b2p_last_error": (C.c_char_p, []),
    "b2p_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "b2p_launch_count": (C.c_int64, []),
    "b2p_set_pdl": (C.c_int, [C.c_int]),
    "b2p_stage_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(StageDesc)]),
    "b2p_stage_destroy": (C.c_int, [C.c_void_p]),
    "b2p_stage_load_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int32,
                                        C.POINTER(C.c_int64), C.c_int32]),
    "b2p_stage_missing": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "b2p_stage_finalize": (C.c_int, [C.c_void_p]),
    "b2p_stage_reserve": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64]),
    "b2p_stage_set_pdl": (C.c_int, [C.c_void_p, C.c_int]),
    "b2p_stage_set_aux_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b2p_stage_join": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b2p_ring_set_timeout": (C.c_int, [C.c_int, C.c_double]),
    "b2p_ring_check": (C.c_int, [C.c_int]),
    "b2p_stage_out_elems": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.POINTER(C.c_int64)]),
    "b2p_stage_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "b2p_stage_trace": (C.c_int, [C.c_void_p, C.POINTER(C.c_longlong), C.c_int]),
    "b2p_stage_kv_reset": (C.c_int, [C.c_void_p, C.c_int64]),
    "b2p_stage_forward_kv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                       C.c_void_p]),
    "b2p_stage_forward_kv_at": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b2p_debug_attn_trace": (C.c_int, [C.POINTER(C.c_longlong), C.c_int]),
    "b2p_debug_fc_trace": (C.c_int, [C.POINTER(C.c_longlong), C.c_int]),
    "b2p_stage_check": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b2p_stage_forward_argmax": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "b2p_linear": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                             C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                             C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "b2p_split_tf32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "b2p_cifar_part1_forward_noalloc": (C.c_int, [C.c_void_p] * 6 + [C.c_int64, C.c_void_p]),
    "b2p_ring_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_size_t]),
    "b2p_ring_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b2p_ring_open": (C.c_int, [C.POINTER(C.c_void_p), C.c_void_p, C.c_int]),
    "b2p_ring_open_local": (C.c_int, [C.POINTER(C.c_void_p), C.c_void_p, C.c_int]),
    "b2p_ring_destroy": (C.c_int, [C.c_void_p]),
    "b2p_ring_set_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "b2p_ring_acquire": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p), C.c_void_p]),
    "b2p_ring_publish": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "b2p_ring_send": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p]),
    "b2p_ring_recv": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p), C.c_void_p]),
    "b2p_ring_release": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "b2p_stage_forward_ring": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                         C.c_void_p, C.c_uint64, C.c_void_p]),
    "b2p_stage_forward_ring_span": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                              C.c_void_p, C.c_uint64, C.c_int64, C.c_void_p]),
}

_lib = None


def exported_symbols() -> list[str]:
    return sorted(_PROTOTYPES)


def lib() -> C.CDLL:
    """The loaded library; raises (never falls back) when it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension is not built and there is no CPU "
                "fallback. Run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(needs nvcc) from the repository root.")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOTYPES.items():
            fn = getattr(h, name)          # AttributeError if the .so is stale
            fn.restype, fn.argtypes = res, args
        if h.b2p_abi_version() != 1:
            raise RuntimeError("libb200pipe.so ABI version mismatch; rebuild it")
        _lib = h
    return _lib


def check(rc: int) -> int:
    if rc < 0:
        msg = lib().b2p_last_error()
        raise EngineError(rc, msg.decode(errors="replace") if msg else f"b200pipe error {rc}")
    return rc


def launch_count() -> int:
    return int(lib().b2p_launch_count())
