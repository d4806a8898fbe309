"""Wire contract of the reference's ``NodeService`` (node_service.proto:1-41),
built at import time -- no protoc, no generated ``*_pb2.py``.

The reference ships generated code (node_service_pb2.py, node_service_pb2_grpc.py)
that cannot be regenerated offline and must not be copied, so the seven
messages and three unary RPCs are declared here as a ``FileDescriptorProto`` in
a *private* descriptor pool (a process that also imports the reference's own
``node_service_pb2`` therefore sees no duplicate-symbol clash).  The bytes on
the wire are identical: field numbers, types and the proto3 ``optional``
presence bit of ``TensorResponse.result_tensor`` (tested with ``HasField`` at
node.py:88,184) follow the .proto exactly, and tests/test_wire.py checks the
serialised form against golden bytes produced with the reference's gencode.

Public names mirror the reference's generated modules so client code reads the
same: ``Tensor``, ``TensorRequest``, ``TensorResponse``, ``Empty``,
``HealthCheckResponse``, ``MessageRequest``, ``MessageReply``,
``NodeServiceStub``, ``NodeServiceServicer``, ``add_NodeServiceServicer_to_server``.
"""
from __future__ import annotations

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto
_PKG = "node_service"
SERVICE_NAME = f"{_PKG}.NodeService"


def _field(msg, name, number, ftype, label=_F.LABEL_OPTIONAL, type_name=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    return f


def _build_file() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "node_service.proto"
    fd.package = _PKG
    fd.syntax = "proto3"

    m = fd.message_type.add(); m.name = "MessageRequest"          # proto:11-14
    _field(m, "sender_id", 1, _F.TYPE_STRING)
    _field(m, "message_text", 2, _F.TYPE_STRING)

    m = fd.message_type.add(); m.name = "MessageReply"            # proto:16-18
    _field(m, "confirmation_text", 1, _F.TYPE_STRING)

    m = fd.message_type.add(); m.name = "Empty"                   # proto:20

    m = fd.message_type.add(); m.name = "HealthCheckResponse"     # proto:22-24
    _field(m, "is_healthy", 1, _F.TYPE_BOOL)

    m = fd.message_type.add(); m.name = "Tensor"                  # proto:26-30
    _field(m, "tensor_data", 1, _F.TYPE_BYTES)
    _field(m, "shape", 2, _F.TYPE_INT32, _F.LABEL_REPEATED)
    _field(m, "dtype", 3, _F.TYPE_STRING)

    m = fd.message_type.add(); m.name = "TensorRequest"           # proto:32-35
    _field(m, "request_id", 1, _F.TYPE_STRING)
    _field(m, "tensor", 2, _F.TYPE_MESSAGE, type_name=f".{_PKG}.Tensor")

    m = fd.message_type.add(); m.name = "TensorResponse"          # proto:38-41
    _field(m, "status", 1, _F.TYPE_STRING)
    f = _field(m, "result_tensor", 2, _F.TYPE_MESSAGE, type_name=f".{_PKG}.Tensor")
    f.proto3_optional = True                                      # `optional` keyword
    f.oneof_index = 0
    m.oneof_decl.add().name = "_result_tensor"

    svc = fd.service.add(); svc.name = "NodeService"              # proto:4-8
    for name, req, rep in (("SendMessage", "MessageRequest", "MessageReply"),
                           ("HealthCheck", "Empty", "HealthCheckResponse"),
                           ("SendTensor", "TensorRequest", "TensorResponse")):
        meth = svc.method.add()
        meth.name, meth.input_type, meth.output_type = name, f".{_PKG}.{req}", f".{_PKG}.{rep}"
    return fd


_POOL = descriptor_pool.DescriptorPool()
_FILE = _POOL.Add(_build_file()) if hasattr(_POOL, "Add") else None
if _FILE is None:                       # very old/new protobuf: AddSerializedFile
    _POOL.AddSerializedFile(_build_file().SerializeToString())


def _cls(name):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(f"{_PKG}.{name}"))


MessageRequest = _cls("MessageRequest")
MessageReply = _cls("MessageReply")
Empty = _cls("Empty")
HealthCheckResponse = _cls("HealthCheckResponse")
Tensor = _cls("Tensor")
TensorRequest = _cls("TensorRequest")
TensorResponse = _cls("TensorResponse")

_METHODS = {
    "SendMessage": (MessageRequest, MessageReply),
    "HealthCheck": (Empty, HealthCheckResponse),
    "SendTensor": (TensorRequest, TensorResponse),
}


# --------------------------------------------------------------------------- #
# ndarray <-> Tensor message, exactly as the reference encodes it
# --------------------------------------------------------------------------- #
def ndarray_to_tensor(arr: np.ndarray):
    """node.py:64-68 / :77-81 / :174-178: raw C-order bytes, int32 dims, numpy
    dtype name."""
    arr = np.require(arr, requirements="C")      # (ascontiguousarray would promote 0-d to 1-d)
    for d in arr.shape:
        if d >= 2 ** 31:
            raise ValueError(f"dimension {d} does not fit the proto's int32 shape field")
    return Tensor(tensor_data=arr.tobytes(), shape=list(arr.shape), dtype=str(arr.dtype))


def tensor_to_ndarray(t) -> np.ndarray:
    """node.py:45-47 / :186-189.  Returns a read-only view over the message's
    bytes (the reference then ``.copy()``s it, node.py:48; the engine instead
    copies straight into pinned staging memory).

    One dtype string beyond what ``np.dtype()`` knows is accepted: ``"bfloat16"`` -- numpy has no
    such type, so the payload comes back as a ``uint16`` array of bit patterns, ``bfloat16_bits_to_float32`` widens it
    (``str(torch.bfloat16_tensor.dtype)`` is how such a client would label its bytes)."""
    name = "uint16" if t.dtype in ("bfloat16", "torch.bfloat16") else t.dtype
    return np.frombuffer(t.tensor_data, dtype=np.dtype(name)).reshape(tuple(t.shape))


def bfloat16_bits_to_float32(bits: np.ndarray) -> np.ndarray:
    """uint16 bit patterns of bfloat16 values -> float32 (exact: bf16 is the top half of an fp32)."""
    return (bits.astype(np.uint32) << 16).view(np.float32)


def float32_to_bfloat16_bits(x: np.ndarray) -> np.ndarray:
    """float32 -> bfloat16 bit patterns, round to nearest even (NaN stays NaN)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    rounded = (u + 0x7FFF + ((u >> 16) & 1)) >> 16
    return np.where(np.isnan(x), 0x7FC0, rounded).astype(np.uint16)


def bfloat16_tensor(x: np.ndarray):
    """A ``Tensor`` message carrying ``x`` as bfloat16 (payload = bit patterns, dtype string "bfloat16")."""
    bits = float32_to_bfloat16_bits(x)
    return Tensor(tensor_data=bits.tobytes(), shape=list(x.shape), dtype="bfloat16")


# --------------------------------------------------------------------------- #
# gRPC plumbing with the same surface as node_service_pb2_grpc.py
# --------------------------------------------------------------------------- #
class NodeServiceStub:
    """Client stub; same attribute names as node_service_pb2_grpc.py:28-51.
    Works with both ``grpc.insecure_channel`` and ``grpc.aio.insecure_channel``."""

    def __init__(self, channel):
        for name, (req, rep) in _METHODS.items():
            setattr(self, name, channel.unary_unary(
                f"/{SERVICE_NAME}/{name}",
                request_serializer=req.SerializeToString,
                response_deserializer=rep.FromString))


class NodeServiceServicer:
    """Base servicer; subclasses override the three RPCs
    (node_service_pb2_grpc.py:54-73)."""

    def SendMessage(self, request, context):
        raise NotImplementedError

    def HealthCheck(self, request, context):
        raise NotImplementedError

    def SendTensor(self, request, context):
        raise NotImplementedError


def add_NodeServiceServicer_to_server(servicer, server):
    """node_service_pb2_grpc.py:76-97."""
    import grpc
    handlers = {
        name: grpc.unary_unary_rpc_method_handler(
            getattr(servicer, name),
            request_deserializer=req.FromString,
            response_serializer=rep.SerializeToString)
        for name, (req, rep) in _METHODS.items()
    }
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE_NAME, handlers),))
