"""Python face of the activation-handoff rings (include/b200pipe.h, b2p_ring_*).

Replaces the reference's per-request gRPC hop (node.py:73-89): the consumer stage owns a
ring of receive slots in its GPU's memory, the producer stage maps it (CUDA IPC between
processes) and writes activations straight into it over NVLink.
"""
from __future__ import annotations

import ctypes as C

from . import _lib


class Ring:
    def __init__(self, handle: C.c_void_p, slots: int, slot_bytes: int, role: str):
        self._h, self.slots, self.slot_bytes, self.role = handle, slots, slot_bytes, role

    # ---- construction ------------------------------------------------------------------
    @classmethod
    def create(cls, device: int, slots: int, slot_bytes: int) -> "Ring":
        """Consumer side: allocate the ring on `device`."""
        h = C.c_void_p()
        _lib.check(_lib.lib().b2p_ring_create(C.byref(h), device, slots, slot_bytes))
        return cls(h, slots, slot_bytes, "consumer")

    def export(self) -> bytes:
        buf = C.create_string_buffer(_lib.RING_HANDLE_BYTES)
        _lib.check(_lib.lib().b2p_ring_export(self._h, buf))
        return buf.raw

    @classmethod
    def open(cls, handle: bytes, producer_device: int, slots: int, slot_bytes: int) -> "Ring":
        """Producer side in another process."""
        h = C.c_void_p()
        buf = C.create_string_buffer(handle, _lib.RING_HANDLE_BYTES)
        _lib.check(_lib.lib().b2p_ring_open(C.byref(h), buf, producer_device))
        return cls(h, slots, slot_bytes, "producer")

    def open_local(self, producer_device: int) -> "Ring":
        """Producer side in the same process (peer access instead of IPC)."""
        h = C.c_void_p()
        _lib.check(_lib.lib().b2p_ring_open_local(C.byref(h), self._h, producer_device))
        return Ring(h, self.slots, self.slot_bytes, "producer")

    def set_mode(self, mode: str):
        _lib.check(_lib.lib().b2p_ring_set_mode(self._h, {"copy_engine": 0, "sm_copy": 1}[mode]))

    # ---- producer ----------------------------------------------------------------------------
    def acquire(self, seq: int, stream: int) -> int:
        p = C.c_void_p()
        _lib.check(_lib.lib().b2p_ring_acquire(self._h, seq, C.byref(p), C.c_void_p(stream)))
        return p.value

    def publish(self, seq: int, stream: int):
        _lib.check(_lib.lib().b2p_ring_publish(self._h, seq, C.c_void_p(stream)))

    def send(self, src_ptr: int, nbytes: int, seq: int, stream: int):
        _lib.check(_lib.lib().b2p_ring_send(self._h, C.c_void_p(src_ptr), nbytes, seq, C.c_void_p(stream)))

    # ---- consumer ------------------------------------------------------------------------------
    def recv(self, seq: int, stream: int) -> int:
        p = C.c_void_p()
        _lib.check(_lib.lib().b2p_ring_recv(self._h, seq, C.byref(p), C.c_void_p(stream)))
        return p.value

    def release(self, seq: int, stream: int):
        _lib.check(_lib.lib().b2p_ring_release(self._h, seq, C.c_void_p(stream)))

    def close(self):
        if self._h is not None:
            _lib.lib().b2p_ring_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def set_timeout(device: int, seconds: float):
    """How long a device-side ring wait on GPU ``device`` waits before it gives up (default 20 s; 0 = for ever, for a
    serving process whose front door may idle).  b2p_ring_set_timeout."""
    _lib.check(_lib.lib().b2p_ring_set_timeout(device, float(seconds)))


def check(device: int):
    """Raises ``EngineError`` (code ETIMEOUT) when a ring wait on GPU ``device`` has given up since the last check:
    the steps enqueued since then produced meaningless results.  Nothing trapped -- stages, rings and graphs stay
    usable.  b2p_ring_check."""
    _lib.check(_lib.lib().b2p_ring_check(device))
