"""The NodeService front door.  CPU tests use a stand-in chain (host logic only: status
strings, error folding, the lifted message cap); the GPU test serves the real CIFAR chain and
replays the request recorded from the unmodified reference (tests/golden/cifar_grpc.npz)."""
import os
from importlib import import_module

import numpy as np
import pytest


def _svc():
    import dnn_b200  # noqa: F401
    return import_module("distributed-neural-networks_b200.service")


def _client(engine, port, big=False):
    import grpc
    opts = [("grpc.max_send_message_length", -1), ("grpc.max_receive_message_length", -1)] if big else []
    ch = grpc.insecure_channel(f"127.0.0.1:{port}", options=opts)
    return ch, engine.wire.NodeServiceStub(ch)


def test_status_strings_and_error_folding(engine):
    svc = _svc()
    w = engine.wire

    def chain(x):
        if x.shape[0] == 3:
            raise ValueError("boom")
        return np.tile(np.arange(10, dtype=np.float32), (x.shape[0], 1))

    servicer = svc.NodeServiceImpl(["node1", "node2"], chain)
    server, port = svc.serve(servicer, 0)
    try:
        ch, stub = _client(engine, port)
        x = np.zeros((2, 3, 32, 32), np.float32)
        rep = stub.SendTensor(w.TensorRequest(request_id="a", tensor=w.ndarray_to_tensor(x)))
        # node.py:61 takes the FLAT argmax: row 0 already holds the maximum 9
        assert rep.status == "[node1] Forwarded. Next node status: [node2] Processing complete. Prediction: 9"
        assert rep.HasField("result_tensor") and w.tensor_to_ndarray(rep.result_tensor).shape == (2, 10)
        bad = stub.SendTensor(w.TensorRequest(request_id="b", tensor=w.ndarray_to_tensor(np.zeros((3, 1), np.float32))))
        assert bad.status == "[node1] Error: boom" and not bad.HasField("result_tensor")
        garbage = stub.SendTensor(w.TensorRequest(request_id="c", tensor=w.Tensor(tensor_data=b"123", shape=[2], dtype="float32")))
        assert garbage.status.startswith("[node1] Error:") and not garbage.HasField("result_tensor")
        assert stub.HealthCheck(w.Empty()).is_healthy is True
        assert stub.SendMessage(w.MessageRequest(sender_id="c", message_text="hi")).confirmation_text == "[node1] got msg 'hi'"
        ch.close()
    finally:
        server.stop(0)
    single = svc.NodeServiceImpl(["only"], chain)
    assert single._status(4) == "[only] Processing complete. Prediction: 4"


def test_dynamic_batching_of_concurrent_requests(engine):
    """concurrent B=1 clients are merged into one chain call; every caller gets its own rows back,
    a malformed request neither joins a batch nor poisons its neighbours, and status strings stay per request."""
    import threading
    import time
    svc = _svc()
    w = engine.wire
    calls = []

    def chain(x):                                    # row i -> one-hot at int(x[i,0,0,0]) % 10; slow enough for a queue to form
        calls.append(x.shape[0])
        if x.ndim != 4:
            raise ValueError("expected [B,3,32,32]")
        if (x[:, 0, 0, 0] < 0).any():
            raise ValueError("negative tag")
        time.sleep(0.01)
        y = np.zeros((x.shape[0], 10), np.float32)
        y[np.arange(x.shape[0]), x[:, 0, 0, 0].astype(int) % 10] = 1.0
        return y

    runner = svc.BatchingRunner(chain, max_rows=16, max_wait_ms=20.0)
    servicer = svc.NodeServiceImpl(["node1", "node2"], runner)
    server, port = svc.serve(servicer, 0, workers=32)
    results = {}
    try:
        def client(tag):
            ch, stub = _client(engine, port)
            x = np.full((1, 3, 32, 32), float(tag), np.float32)
            results[tag] = stub.SendTensor(w.TensorRequest(request_id=str(tag), tensor=w.ndarray_to_tensor(x)))
            ch.close()
        tags = list(range(24)) + [-1]                # -1 makes its batch fail: neighbours must still succeed
        threads = [threading.Thread(target=client, args=(t,)) for t in tags]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=30)
        odd = _client(engine, port)
        rep = odd[1].SendTensor(w.TensorRequest(request_id="odd", tensor=w.ndarray_to_tensor(np.zeros((2, 5), np.float32))))
        assert rep.status == "[node1] Error: expected [B,3,32,32]" and not rep.HasField("result_tensor")
        odd[0].close()
    finally:
        server.stop(0)
        runner.close()
    for tag in range(24):
        rep = results[tag]
        assert rep.status.endswith(f"Prediction: {tag % 10}"), rep.status
        y = w.tensor_to_ndarray(rep.result_tensor)
        assert y.shape == (1, 10) and y[0, tag % 10] == 1.0
    assert results[-1].status == "[node1] Error: negative tag" and not results[-1].HasField("result_tensor")
    assert max(runner.batches) > 1 and max(runner.batches) <= 16, runner.batches      # requests really were merged


def test_messages_beyond_the_reference_cap(engine):
    """The reference cannot carry > 4 MiB; the front door can."""
    svc = _svc()
    w = engine.wire
    servicer = svc.NodeServiceImpl(["n"], lambda x: x[:, :2].copy())
    server, port = svc.serve(servicer, 0)
    try:
        ch, stub = _client(engine, port, big=True)
        x = np.ones((300, 4096), np.float32)                 # 4.9 MB
        rep = stub.SendTensor(w.TensorRequest(request_id="big", tensor=w.ndarray_to_tensor(x)))
        assert rep.HasField("result_tensor"), rep.status
        assert w.tensor_to_ndarray(rep.result_tensor).shape == (300, 2)
        ch.close()
    finally:
        server.stop(0)


@pytest.mark.gpu
def test_cifar_chain_replays_the_reference_recording(cuda_engine, golden_dir, tmp_path):
    import json
    import torch
    from oracle import cifar_oracle as co
    svc = _svc()
    w = cuda_engine.wire
    g = np.load(os.path.join(golden_dir, "cifar_grpc.npz"))
    sd = co.init_state_dict(int(g["weight_seed"]))
    pth = tmp_path / "w.pth"
    torch.save(sd, pth)
    cfg = {"nodes": [{"id": "node1", "address": "127.0.0.1:0", "part_index": 0, "device": 0},
                     {"id": "node2", "address": "127.0.0.1:0", "part_index": 1, "device": 0}],
           "model_weights": str(pth), "num_parts": 2, "return_to_node_id": "node1"}
    cpath = tmp_path / "config.json"
    cpath.write_text(json.dumps(cfg))
    topo = cuda_engine.load_config(str(cpath))
    servicer, pipe = svc.build_cifar_chain(topo, "node1", math="tf32x3", microbatch=128, max_batch=256)
    server, port = svc.serve(servicer, 0)
    try:
        ch, stub = _client(cuda_engine, port, big=True)
        for b in (1, 4):
            x = g[f"x{b}"]
            rep = stub.SendTensor(w.TensorRequest(request_id=f"golden_{b}", tensor=w.ndarray_to_tensor(x)))
            assert rep.status == str(g[f"status{b}"])          # byte-for-byte the reference's status line
            got = w.tensor_to_ndarray(rep.result_tensor)
            assert got.dtype == np.float32 and got.shape == (b, 10)
            np.testing.assert_allclose(got, g[f"probs{b}"], rtol=0, atol=1e-5)
        # a batch larger than the pipeline capacity and than the reference's 4 MiB cap
        x = co.make_input(600, seed=12).numpy()
        rep = stub.SendTensor(w.TensorRequest(request_id="big", tensor=w.ndarray_to_tensor(x)))
        got = w.tensor_to_ndarray(rep.result_tensor)
        want = co.full(torch.from_numpy(x), sd).numpy()
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-5)
        assert rep.status.endswith(f"Prediction: {int(np.argmax(want))}") or np.abs(want - got).max() > 0
        wrong = stub.SendTensor(w.TensorRequest(request_id="w", tensor=w.ndarray_to_tensor(np.zeros((2, 5), np.float32))))
        assert wrong.status.startswith("[node1] Error:") and not wrong.HasField("result_tensor")
        ch.close()
    finally:
        server.stop(0)


def _gpu_chain(cuda_engine, tmp_path, **servicer_kw):
    import json
    import torch
    from oracle import cifar_oracle as co
    svc = _svc()
    sd = co.init_state_dict(1234)
    pth = tmp_path / "w.pth"
    torch.save(sd, pth)
    cfg = {"nodes": [{"id": "node1", "address": "127.0.0.1:0", "part_index": 0, "device": 0},
                     {"id": "node2", "address": "127.0.0.1:0", "part_index": 1, "device": 0}],
           "model_weights": str(pth), "num_parts": 2, "return_to_node_id": "node1"}
    cpath = tmp_path / "config.json"
    cpath.write_text(json.dumps(cfg))
    topo = cuda_engine.load_config(str(cpath))
    servicer, pipe = svc.build_cifar_chain(topo, "node1", math="tf32x3", microbatch=128, max_batch=256)
    for k, v in servicer_kw.items():
        setattr(servicer, k, v)
    return svc, sd, servicer, pipe


@pytest.mark.gpu
def test_concurrent_clients_through_the_batching_front_door_on_the_gpu(cuda_engine, tmp_path):
    """SURVEY 8f-1 on the real CUDA chain: 32 client threads, each with its own images; whatever batch a request
    happened to travel in, its reply is the oracle's row for ITS image (|d prob| <= 1e-5, arg-max equal)."""
    import threading
    import torch
    from oracle import cifar_oracle as co
    svc, sd, servicer, pipe = _gpu_chain(cuda_engine, tmp_path)
    w = cuda_engine.wire
    runner = svc.BatchingRunner(svc.ChainRunner(pipe), max_rows=128, max_wait_ms=0.5)
    batched = svc.NodeServiceImpl(["node1", "node2"], runner)
    server, port = svc.serve(batched, 0, workers=32)
    n_clients, per_client = 32, 6
    xs = co.make_input(n_clients * per_client, seed=77)
    want = co.full(xs, sd).numpy()
    errs, got = [], {}

    def client(k):
        try:
            ch, stub = _client(cuda_engine, port)
            for j in range(per_client):
                i = k * per_client + j
                rep = stub.SendTensor(w.TensorRequest(request_id=f"c{k}_{j}", tensor=w.ndarray_to_tensor(xs[i:i + 1].numpy())),
                                      timeout=60)
                assert rep.HasField("result_tensor"), rep.status
                got[i] = (w.tensor_to_ndarray(rep.result_tensor).copy(), rep.status)
            ch.close()
        except Exception as e:
            errs.append(e)

    ths = [threading.Thread(target=client, args=(k,)) for k in range(n_clients)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(timeout=120)
    server.stop(0)
    runner.close()
    assert not errs, errs[:3]
    assert len(got) == n_clients * per_client
    for i, (probs, status) in got.items():
        assert probs.shape == (1, 10)
        assert np.abs(probs[0] - want[i]).max() <= 1e-5, i
        assert status.endswith(f"Prediction: {int(np.argmax(want[i]))}") or np.sort(want[i])[-1] - np.sort(want[i])[-2] < 1e-5
    assert max(runner.batches) > 1, "no two requests were ever merged: the batching path was not exercised"


@pytest.mark.gpu
def test_initiate_inference_on_a_png_through_the_gpu_chain(cuda_engine, tmp_path):
    """node.py:137-200 on the real chain: a PNG goes through the reference's preprocessing, both stages on the GPU and
    the arg-max; compared with the oracle on the very tensor the preprocessing produced."""
    import torch
    from PIL import Image
    from oracle import cifar_oracle as co
    svc, sd, servicer, pipe = _gpu_chain(cuda_engine, tmp_path)
    rng = np.random.default_rng(5)
    png = tmp_path / "img.png"
    Image.fromarray(rng.integers(0, 256, (48, 40, 3), dtype=np.uint8), "RGB").save(png)
    pred, probs, real = svc.initiate_inference(servicer, str(png))
    assert real is True and probs.shape == (1, 10)
    x, _ = svc.load_input_image(str(png))
    want = co.full(x, sd).numpy()
    assert np.abs(probs - want).max() <= 1e-5
    assert pred == int(np.argmax(want))
    # the reference's fallback when the file is missing: a random image, flagged as such (node.py:150-153)
    pred2, probs2, real2 = svc.initiate_inference(servicer, str(tmp_path / "nope.png"), seed=3)
    x2, _ = svc.load_input_image(None, seed=3)
    assert real2 is False and np.abs(probs2 - co.full(x2, sd).numpy()).max() <= 1e-5


@pytest.mark.gpu
def test_request_dtypes_and_per_row_argmax_on_the_gpu(cuda_engine, tmp_path):
    """float16 / bfloat16 / float64 payloads, replies in the request's dtype on demand, per-row arg-max."""
    import torch
    from oracle import cifar_oracle as co
    svc, sd, servicer, pipe = _gpu_chain(cuda_engine, tmp_path, per_row_argmax=True)
    w = cuda_engine.wire
    server, port = svc.serve(servicer, 0)
    try:
        ch, stub = _client(cuda_engine, port, big=True)
        x = co.make_input(5, seed=21).numpy()
        x16 = x.astype(np.float16)
        xbf = w.bfloat16_bits_to_float32(w.float32_to_bfloat16_bits(x))
        cases = [("float32", w.ndarray_to_tensor(x), x), ("float16", w.ndarray_to_tensor(x16), x16.astype(np.float32)),
                 ("float64", w.ndarray_to_tensor(x.astype(np.float64)), x), ("bfloat16", w.bfloat16_tensor(x), xbf)]
        for name, tensor, seen in cases:
            assert tensor.dtype == name
            rep = stub.SendTensor(w.TensorRequest(request_id=name, tensor=tensor))
            assert rep.HasField("result_tensor"), rep.status
            got = w.tensor_to_ndarray(rep.result_tensor)
            want = co.full(torch.from_numpy(seen), sd).numpy()      # the oracle on exactly the values the payload carries
            assert got.dtype == np.float32 and np.abs(got - want).max() <= 1e-5, name
            rows = [int(k) for k in want.argmax(1)]
            assert rep.status.endswith(f"Prediction: {int(np.argmax(want))} Per-row: {rows}"), rep.status
        servicer.reply_dtype = "same"
        rep = stub.SendTensor(w.TensorRequest(request_id="h", tensor=w.ndarray_to_tensor(x16)))
        assert rep.result_tensor.dtype == "float16" and len(rep.result_tensor.tensor_data) == 5 * 10 * 2
        rep = stub.SendTensor(w.TensorRequest(request_id="b", tensor=w.bfloat16_tensor(x)))
        assert rep.result_tensor.dtype == "bfloat16"
        back = w.bfloat16_bits_to_float32(w.tensor_to_ndarray(rep.result_tensor))
        assert np.abs(back - co.full(torch.from_numpy(xbf), sd).numpy()).max() <= 4e-3      # bf16 rounding of a probability
        # an int64 image is converted (it is numbers, not bytes to reinterpret); a string dtype is an error, reported in-band
        bad = stub.SendTensor(w.TensorRequest(request_id="s", tensor=w.Tensor(tensor_data=b"abcd", shape=[1], dtype="S4")))
        assert bad.status.startswith("[node1] Error:") and not bad.HasField("result_tensor")
        ch.close()
    finally:
        server.stop(0)
