"""The reference's wire contract for the ORACLE side -- TEST INFRASTRUCTURE ONLY.

The CPU port of the node loop (oracle/node_oracle.py) must not serialise through the
product's ``dnn_b200.wire`` (an oracle that depends on the thing it judges proves
nothing), so this module gives it messages and a stub of its own:

  * when ``oracle/_ref/`` holds the reference's generated ``node_service_pb2*.py``
    (placed there by oracle/build_ref.py, never committed) those are used as they are;
  * otherwise the seven messages of node_service.proto:11-41 are declared once more,
    from the .proto text restated below, with nothing but the protobuf runtime.

Either way the names are the ones node.py uses (node.py:5-6, 64-68, 75-89, 172-189).
"""
from __future__ import annotations

import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref")

# node_service.proto:4-41 restated as (message, [(field, number, type, repeated, optional)])
_SCHEMA = (
    ("MessageRequest", (("sender_id", 1, "string", 0, 0), ("message_text", 2, "string", 0, 0))),
    ("MessageReply", (("confirmation_text", 1, "string", 0, 0),)),
    ("Empty", ()),
    ("HealthCheckResponse", (("is_healthy", 1, "bool", 0, 0),)),
    ("Tensor", (("tensor_data", 1, "bytes", 0, 0), ("shape", 2, "int32", 1, 0), ("dtype", 3, "string", 0, 0))),
    ("TensorRequest", (("request_id", 1, "string", 0, 0), ("tensor", 2, ".node_service.Tensor", 0, 0))),
    ("TensorResponse", (("status", 1, "string", 0, 0), ("result_tensor", 2, ".node_service.Tensor", 0, 1))),
)
_RPCS = (("SendMessage", "MessageRequest", "MessageReply"),
         ("HealthCheck", "Empty", "HealthCheckResponse"),
         ("SendTensor", "TensorRequest", "TensorResponse"))


def _from_reference_gencode():
    if not os.path.exists(os.path.join(_REF, "node_service_pb2.py")):
        return None
    if _REF not in sys.path:
        sys.path.insert(0, _REF)
    import node_service_pb2 as pb2
    import node_service_pb2_grpc as pb2_grpc
    ns = {name: getattr(pb2, name) for name, _ in _SCHEMA}
    ns["NodeServiceStub"] = pb2_grpc.NodeServiceStub
    ns["add_NodeServiceServicer_to_server"] = pb2_grpc.add_NodeServiceServicer_to_server
    ns["SOURCE"] = "reference gencode (oracle/_ref)"
    return ns


def _from_runtime():
    from google.protobuf import descriptor_pb2 as dp, descriptor_pool, message_factory
    scalar = {"string": dp.FieldDescriptorProto.TYPE_STRING, "bytes": dp.FieldDescriptorProto.TYPE_BYTES,
              "bool": dp.FieldDescriptorProto.TYPE_BOOL, "int32": dp.FieldDescriptorProto.TYPE_INT32}
    fdp = dp.FileDescriptorProto(name="oracle_node_service.proto", package="node_service", syntax="proto3")
    for mname, fields in _SCHEMA:
        msg = fdp.message_type.add(name=mname)
        for fname, num, ftype, repeated, optional in fields:
            f = msg.field.add(name=fname, number=num)
            f.label = f.LABEL_REPEATED if repeated else f.LABEL_OPTIONAL
            if ftype.startswith("."):
                f.type, f.type_name = f.TYPE_MESSAGE, ftype
            else:
                f.type = scalar[ftype]
            if optional:                           # the proto3 `optional` keyword: a synthetic one-of
                msg.oneof_decl.add(name="_" + fname)
                f.oneof_index, f.proto3_optional = len(msg.oneof_decl) - 1, True
    svc = fdp.service.add(name="NodeService")
    for rpc, req, rep in _RPCS:
        svc.method.add(name=rpc, input_type=".node_service." + req, output_type=".node_service." + rep)
    pool = descriptor_pool.DescriptorPool()        # private pool: no clash with any other declaration
    pool.AddSerializedFile(fdp.SerializeToString())
    ns = {m: message_factory.GetMessageClass(pool.FindMessageTypeByName("node_service." + m)) for m, _ in _SCHEMA}

    class NodeServiceStub:                         # node_service_pb2_grpc.py:28-51
        def __init__(self, channel):
            for rpc, req, rep in _RPCS:
                setattr(self, rpc, channel.unary_unary(f"/node_service.NodeService/{rpc}",
                                                       request_serializer=ns[req].SerializeToString,
                                                       response_deserializer=ns[rep].FromString))

    def add_NodeServiceServicer_to_server(servicer, server):     # node_service_pb2_grpc.py:76-97
        import grpc
        handlers = {rpc: grpc.unary_unary_rpc_method_handler(getattr(servicer, rpc),
                                                             request_deserializer=ns[req].FromString,
                                                             response_serializer=ns[rep].SerializeToString)
                    for rpc, req, rep in _RPCS}
        server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler("node_service.NodeService", handlers),))

    ns["NodeServiceStub"] = NodeServiceStub
    ns["add_NodeServiceServicer_to_server"] = add_NodeServiceServicer_to_server
    ns["SOURCE"] = "protobuf runtime (oracle/node_wire.py)"
    return ns


globals().update(_from_reference_gencode() or _from_runtime())
