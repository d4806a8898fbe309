#!/usr/bin/env python
"""Generate tests/golden/* by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE ONLY.  Needs /root/reference (absent on the GPU box, so the
outputs are committed).  Run:  python oracle/make_golden.py

What is produced
  cifar_b8.npz      inputs + outputs of the reference's ModelPart0_2Node /
                    ModelPart1_2Node / NeuralNetwork (cifar_model_parts.py) on
                    seeded default-init weights (cifar10_model.pth is missing
                    from the checkout, .MISSING_LARGE_BLOBS:1).
  cifar_grpc.npz    result + status string of a real two-process localhost run
                    of the unmodified node.py driven by a NodeServiceStub client
                    (SURVEY 3.4), B=1 and B=4.
  gpt_tiny.npz      outputs of the reference's ModelPart0 / ModelPartIntermediate
                    / ModelPartFinal_GPT (partitions/gpt_model_parts.py) built
                    over oracle/gpt_model.py (the stand-in for the absent
                    third-party ``model`` module).
  wire_vectors.npz  serialised protobuf messages made with the reference's
                    generated node_service_pb2.py.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(REF, "partitions"))

from oracle import cifar_oracle, gpt_model, gpt_oracle  # noqa: E402

torch.set_num_threads(1)          # fixed thread count => reproducible bits


def cifar():
    import cifar_model_parts as ref
    torch.manual_seed(1234)
    net = ref.NeuralNetwork().eval()
    sd = cifar_oracle.init_state_dict(1234)
    assert all(torch.equal(sd[k], v) for k, v in net.state_dict().items()), "init drift"
    x = cifar_oracle.make_input(8, seed=0)
    with torch.no_grad():
        p0 = ref.ModelPart0_2Node(net)(x)
        p1 = ref.ModelPart1_2Node(net)(p0)
        fu = net(x)
    assert torch.equal(p1, fu)
    np.savez_compressed(os.path.join(OUT, "cifar_b8.npz"), x=x.numpy(), part0=p0.numpy(),
                        probs=p1.numpy(), weight_seed=1234, input_seed=0,
                        digest=cifar_oracle.state_dict_digest(sd))
    return net, sd


def cifar_grpc(sd):
    """Two unmodified node.py processes on localhost + proto client into node 1."""
    import grpc
    import node_service_pb2 as pb
    import node_service_pb2_grpc as pbg
    tmp = tempfile.mkdtemp()
    pth = os.path.join(tmp, "w.pth")
    torch.save(sd, pth)
    cfg = {"nodes": [{"id": "node1", "address": "127.0.0.1:50151", "part_index": 0},
                     {"id": "node2", "address": "127.0.0.1:50152", "part_index": 1}],
           "model_weights": pth, "num_parts": 2, "return_to_node_id": "node1"}
    cpath = os.path.join(tmp, "config.json")
    json.dump(cfg, open(cpath, "w"))
    env = dict(os.environ, PYTHONPATH=REF, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, os.path.join(REF, "node.py"), "--node_id", n,
                               "--config", cpath], env=env, stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL, cwd=tmp) for n in ("node2", "node1")]
    out = {}
    try:
        chan = grpc.insecure_channel("127.0.0.1:50151")
        grpc.channel_ready_future(chan).result(timeout=120)
        stub = pbg.NodeServiceStub(chan)
        time.sleep(3)
        for b in (1, 4):
            x = cifar_oracle.make_input(b, seed=7).numpy()
            for attempt in range(30):
                rep = stub.SendTensor(pb.TensorRequest(request_id=f"golden_{b}", tensor=pb.Tensor(
                    tensor_data=x.tobytes(), shape=list(x.shape), dtype=str(x.dtype))))
                if rep.HasField("result_tensor"):
                    break
                time.sleep(1)
            res = np.frombuffer(rep.result_tensor.tensor_data,
                                np.dtype(rep.result_tensor.dtype)).reshape(tuple(rep.result_tensor.shape))
            out[f"x{b}"], out[f"probs{b}"], out[f"status{b}"] = x, res.copy(), rep.status
        hc = stub.HealthCheck(pb.Empty())
        msg = stub.SendMessage(pb.MessageRequest(sender_id="c", message_text="hi"))
        out["healthy"], out["confirmation"] = hc.is_healthy, msg.confirmation_text
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            p.wait(timeout=20)
    np.savez_compressed(os.path.join(OUT, "cifar_grpc.npz"), input_seed=7, weight_seed=1234, **out)


def gpt():
    sys.modules["model"] = gpt_model            # what gpt_model_parts.py:4 imports
    import gpt_model_parts as ref
    cfg = gpt_model.GPTConfig(block_size=64, vocab_size=256, n_layer=3, n_head=2, n_embd=128)
    torch.manual_seed(1234)
    net = gpt_model.GPT(cfg).eval()
    # make LayerNorm/bias parameters non-trivial so a dropped bias is caught
    g = torch.Generator().manual_seed(99)
    with torch.no_grad():
        for n, p in net.named_parameters():
            if n.endswith("bias") or ".ln_" in n or "ln_f" in n:
                p.add_(0.1 * torch.randn(p.shape, generator=g))
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    idx = gpt_oracle.make_tokens(2, 32, cfg.vocab_size, seed=0)
    with torch.no_grad():
        a = ref.ModelPart0(net, 0, 0)(idx)
        b = ref.ModelPartIntermediate(net, 1, 1)(a)
        c = ref.ModelPartFinal_GPT(net, 2, 2)(b)
        fu = net(idx)
    assert torch.allclose(c, fu, atol=1e-5)
    np.savez_compressed(os.path.join(OUT, "gpt_tiny.npz"), idx=idx.numpy(), stage0=a.numpy(),
                        stage1=b.numpy(), logits=c.numpy(),
                        cfg=np.array([cfg.block_size, cfg.vocab_size, cfg.n_layer, cfg.n_head, cfg.n_embd]),
                        **{"w:" + k: v.numpy() for k, v in sd.items()})


def wire():
    import node_service_pb2 as pb
    a = np.arange(24, dtype=np.float32).reshape(2, 3, 4) * 0.5
    i = np.arange(6, dtype=np.int64).reshape(2, 3)
    vec = {
        "req_f32": pb.TensorRequest(request_id="cifar_pipe_2node_001", tensor=pb.Tensor(
            tensor_data=a.tobytes(), shape=list(a.shape), dtype=str(a.dtype))).SerializeToString(),
        "req_i64": pb.TensorRequest(request_id="gpt", tensor=pb.Tensor(
            tensor_data=i.tobytes(), shape=list(i.shape), dtype=str(i.dtype))).SerializeToString(),
        "resp_with": pb.TensorResponse(status="[node2] Processing complete. Prediction: 3",
                                       result_tensor=pb.Tensor(tensor_data=a.tobytes(), shape=list(a.shape),
                                                               dtype="float32")).SerializeToString(),
        "resp_without": pb.TensorResponse(status="[node2] Error: boom").SerializeToString(),
        "health": pb.HealthCheckResponse(is_healthy=True).SerializeToString(),
        "msg_req": pb.MessageRequest(sender_id="n1", message_text="hello").SerializeToString(),
        "msg_rep": pb.MessageReply(confirmation_text="[n2] got msg 'hello'").SerializeToString(),
        "empty": pb.Empty().SerializeToString(),
    }
    np.savez(os.path.join(OUT, "wire_vectors.npz"), a=a, i=i,
             **{k: np.frombuffer(v, dtype=np.uint8) for k, v in vec.items()})


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    _, sd = cifar()
    gpt()
    wire()
    cifar_grpc(sd)
    print("golden vectors written to", OUT)
