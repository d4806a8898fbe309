"""Runtime-built protobuf contract vs bytes produced by the reference's gencode
(tests/golden/wire_vectors.npz, made by oracle/make_golden.py)."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def vec(golden_dir):
    return np.load(os.path.join(golden_dir, "wire_vectors.npz"))


def test_request_bytes_identical(engine, vec):
    w = engine.wire
    a, i = vec["a"], vec["i"]
    m = w.TensorRequest(request_id="cifar_pipe_2node_001", tensor=w.ndarray_to_tensor(a))
    assert m.SerializeToString() == vec["req_f32"].tobytes()
    m = w.TensorRequest(request_id="gpt", tensor=w.ndarray_to_tensor(i))
    assert m.SerializeToString() == vec["req_i64"].tobytes()


def test_response_presence_bit(engine, vec):
    w = engine.wire
    with_t = w.TensorResponse.FromString(vec["resp_with"].tobytes())
    assert with_t.HasField("result_tensor")
    assert with_t.status == "[node2] Processing complete. Prediction: 3"
    np.testing.assert_array_equal(w.tensor_to_ndarray(with_t.result_tensor), vec["a"])
    without = w.TensorResponse.FromString(vec["resp_without"].tobytes())
    assert not without.HasField("result_tensor")
    assert w.TensorResponse(status="[node2] Error: boom").SerializeToString() == vec["resp_without"].tobytes()
    again = w.TensorResponse(status=with_t.status, result_tensor=with_t.result_tensor)
    assert again.SerializeToString() == vec["resp_with"].tobytes()


def test_small_messages(engine, vec):
    w = engine.wire
    assert w.HealthCheckResponse(is_healthy=True).SerializeToString() == vec["health"].tobytes()
    assert w.MessageRequest(sender_id="n1", message_text="hello").SerializeToString() == vec["msg_req"].tobytes()
    assert w.MessageReply(confirmation_text="[n2] got msg 'hello'").SerializeToString() == vec["msg_rep"].tobytes()
    assert w.Empty().SerializeToString() == vec["empty"].tobytes()


def test_ndarray_roundtrip_dtypes_and_empty(engine):
    w = engine.wire
    for arr in (np.zeros((0, 4096), np.float32), np.arange(6, dtype=np.int64).reshape(2, 3),
                np.float32(3.5).reshape(()), np.asfortranarray(np.arange(12, dtype=np.float32).reshape(3, 4))):
        back = w.tensor_to_ndarray(w.Tensor.FromString(w.ndarray_to_tensor(arr).SerializeToString()))
        assert back.dtype == arr.dtype and back.shape == arr.shape
        np.testing.assert_array_equal(back, arr)


def test_grpc_loopback(engine):
    """Stub + servicer registration work end to end over a real localhost channel."""
    import grpc
    from concurrent import futures
    w = engine.wire

    class Echo(w.NodeServiceServicer):
        def SendTensor(self, request, context):
            return w.TensorResponse(status=f"echo {request.request_id}", result_tensor=request.tensor)

        def HealthCheck(self, request, context):
            return w.HealthCheckResponse(is_healthy=True)

        def SendMessage(self, request, context):
            return w.MessageReply(confirmation_text=f"got {request.message_text}")

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
    w.add_NodeServiceServicer_to_server(Echo(), server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    try:
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = w.NodeServiceStub(ch)
            x = np.arange(8, dtype=np.float32).reshape(2, 4)
            rep = stub.SendTensor(w.TensorRequest(request_id="r1", tensor=w.ndarray_to_tensor(x)))
            assert rep.status == "echo r1" and rep.HasField("result_tensor")
            np.testing.assert_array_equal(w.tensor_to_ndarray(rep.result_tensor), x)
            assert stub.HealthCheck(w.Empty()).is_healthy
            assert stub.SendMessage(w.MessageRequest(sender_id="a", message_text="hi")).confirmation_text == "got hi"
    finally:
        server.stop(0)


def test_bfloat16_payloads(engine):
    """bfloat16 travels as bit patterns under the dtype string "bfloat16"; conversion is round-to-nearest-
    even and agrees with torch's."""
    import torch
    w = engine.wire
    x = np.array([0.0, 1.0, -2.5, 3.14159265, 1e-8, 65504.0, float("inf"), 1.00390625, 1.01171875], np.float32)
    bits = w.float32_to_bfloat16_bits(x)
    want = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(bits, want)
    t = w.bfloat16_tensor(x.reshape(3, 3))
    assert t.dtype == "bfloat16" and list(t.shape) == [3, 3] and len(t.tensor_data) == 18
    back = w.bfloat16_bits_to_float32(w.tensor_to_ndarray(t))
    assert np.array_equal(back.reshape(-1), torch.from_numpy(x).to(torch.bfloat16).float().numpy())
    assert np.isnan(w.bfloat16_bits_to_float32(w.float32_to_bfloat16_bits(np.array([np.nan], np.float32))))[0]
