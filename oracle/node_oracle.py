#!/usr/bin/env python
"""CPU port of the reference's node loop -- the timed CPU baseline.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  /root/reference does not exist
on the GPU box, so the "reference arm" of bench.py runs this restatement of
node.py instead (cpu_baseline.kind = "port").  It keeps every cost the reference
pays per request, in the same order:

  SendTensor (node.py:35-105)
    np.frombuffer(...).reshape -> torch.from_numpy(copy)            :45-48
    with torch.no_grad(): part(x)   (blocking, inside the event loop) :52-53
    .cpu().numpy()                                                  :55
    last node: flat np.argmax, status "Processing complete. Prediction: k" :61-68
    else: NEW grpc.aio channel per request, tobytes -> TensorRequest ->
          await stub.SendTensor, relay result_tensor upstream        :73-89
    exceptions folded into the status string, never a gRPC error     :91-100
  serve (node.py:114-133): grpc.aio server on [::]:port, default 4 MiB cap.

Checked against a recorded run of the unmodified reference
(tests/golden/cifar_grpc.npz: result bits and status strings).

Run one node:   python -m oracle.node_oracle --node_id node1 --config cfg.json
"""
from __future__ import annotations

import argparse
import asyncio
import json
import os
import socket
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import cifar_oracle  # noqa: E402

PARTS = {0: cifar_oracle.part0, 1: cifar_oracle.part1}        # node.py:29-32


def _wire():
    """The oracle's OWN messages and stub (oracle/node_wire.py): the reference's gencode when oracle/_ref/ holds
    it, else a protobuf-runtime restatement -- never the product's dnn_b200.wire."""
    from oracle import node_wire
    return node_wire


class NodeOracle:
    def __init__(self, node_id, part_index, num_parts, next_address, sd):
        self.node_id, self.part_index = node_id, part_index
        self.is_last = part_index == num_parts - 1              # node.py:260
        self.next_address, self.sd = next_address, sd
        self.wire = _wire()

    async def SendTensor(self, request, context):
        w = self.wire
        result, status = None, f"[{self.node_id}] Error processing tensor."
        try:
            x_np = np.frombuffer(request.tensor.tensor_data,
                                 dtype=np.dtype(request.tensor.dtype)).reshape(tuple(request.tensor.shape))
            x = torch.from_numpy(x_np.copy())
            with torch.no_grad():
                y = PARTS[self.part_index](x, self.sd)
            y_np = y.cpu().numpy()
            if self.is_last:
                pred = np.argmax(y_np)
                status = f"[{self.node_id}] Processing complete. Prediction: {pred}"
                result = w.Tensor(tensor_data=y_np.tobytes(), shape=list(y_np.shape), dtype=str(y_np.dtype))
            else:
                import grpc
                async with grpc.aio.insecure_channel(self.next_address) as channel:
                    stub = w.NodeServiceStub(channel)
                    nxt = w.TensorRequest(request_id=request.request_id, tensor=w.Tensor(
                        tensor_data=y_np.tobytes(), shape=list(y_np.shape), dtype=str(y_np.dtype)))
                    try:
                        rep = await stub.SendTensor(nxt)
                        status = f"[{self.node_id}] Forwarded. Next node status: {rep.status}"
                        if rep.HasField("result_tensor"):
                            result = rep.result_tensor
                    except grpc.aio.AioRpcError as e:
                        status, result = f"[{self.node_id}] Error forwarding: {e.details()}", None
        except Exception as e:                                   # node.py:96-100
            status, result = f"[{self.node_id}] Error: {e}", None
        return w.TensorResponse(status=status, result_tensor=result)

    async def HealthCheck(self, request, context):              # node.py:107-109
        return self.wire.HealthCheckResponse(is_healthy=True)

    async def SendMessage(self, request, context):              # node.py:111-113
        return self.wire.MessageReply(
            confirmation_text=f"[{self.node_id}] got msg '{request.message_text}'")


async def _serve(node: NodeOracle, port: int):
    import grpc
    server = grpc.aio.server()
    node.wire.add_NodeServiceServicer_to_server(node, server)
    server.add_insecure_port(f"[::]:{port}")
    await server.start()
    await server.wait_for_termination()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--node_id", required=True)
    ap.add_argument("--config", required=True)
    a = ap.parse_args()
    cfg = json.load(open(a.config))
    me = next(n for n in cfg["nodes"] if n["id"] == a.node_id)
    num_parts = cfg["num_parts"]
    if num_parts != 2:                                           # node.py:246-248
        print(f"ERROR: Configuration file specifies num_parts={num_parts}, but this script expects 2 parts.")
        sys.exit(1)
    nxt = next((n for n in cfg["nodes"] if n["part_index"] == me["part_index"] + 1), None)
    sd = torch.load(cfg["model_weights"], map_location=torch.device("cpu"))
    node = NodeOracle(a.node_id, me["part_index"], num_parts, nxt["address"] if nxt else None, sd)
    asyncio.run(_serve(node, int(me["address"].split(":")[-1])))


# ------------------------------------------------------------------------------------------
# harness: start a two-node localhost pipeline and drive it like SURVEY 3.4
# ------------------------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def reference_available() -> bool:
    """True when oracle/_ref/ holds the unmodified reference node program (oracle/build_ref.py)."""
    from oracle import build_ref
    return build_ref.available()


class TwoNodePipeline:
    """Context manager: a two-node localhost pipeline + a blocking stub.

    impl = "reference": two processes of the UNMODIFIED reference ``node.py`` from oracle/_ref/ (CUDA hidden from
                        them, so DEVICE resolves to "cpu", node.py:25) -- cpu_baseline.kind "reference";
    impl = "port":      two processes of this file's restatement -- cpu_baseline.kind "port";
    impl = None:        the reference when oracle/_ref/ is there, else the port.  ``self.kind`` says which ran."""

    def __init__(self, sd, threads: int | None = None, impl: str | None = None):
        self.sd, self.threads = sd, threads
        self.kind = impl or ("reference" if reference_available() else "port")
        if self.kind == "reference" and not reference_available():
            raise RuntimeError("oracle/_ref/ is not populated (run oracle/build_ref.py where /root/reference exists)")
        self.procs = []

    def __enter__(self):
        import grpc
        self.tmp = tempfile.mkdtemp(prefix="node_oracle_")
        pth = os.path.join(self.tmp, "weights.pth")
        torch.save(self.sd, pth)
        p1, p2 = _free_port(), _free_port()
        cfg = {"nodes": [{"id": "node1", "address": f"127.0.0.1:{p1}", "part_index": 0},
                         {"id": "node2", "address": f"127.0.0.1:{p2}", "part_index": 1}],
               "model_weights": pth, "num_parts": 2, "return_to_node_id": "node1"}
        cpath = os.path.join(self.tmp, "config.json")
        json.dump(cfg, open(cpath, "w"))
        env = dict(os.environ)
        if self.threads:
            env["OMP_NUM_THREADS"] = env["MKL_NUM_THREADS"] = str(self.threads)
        else:
            # torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the reference's nodes
            # run with torch's default threading (all cores), so do not let that leak in
            env.pop("OMP_NUM_THREADS", None)
            env.pop("MKL_NUM_THREADS", None)
        if self.kind == "reference":
            ref_dir = os.path.join(ROOT, "oracle", "_ref")
            env["CUDA_VISIBLE_DEVICES"] = ""          # the CPU baseline: node.py:25 picks "cuda" when it sees one
            cmd = lambda nid: [sys.executable, os.path.join(ref_dir, "node.py"), "--node_id", nid, "--config", cpath]
            cwd = ref_dir
        else:
            cmd = lambda nid: [sys.executable, "-m", "oracle.node_oracle", "--node_id", nid, "--config", cpath]
            cwd = ROOT
        for nid in ("node2", "node1"):
            self.procs.append(subprocess.Popen(cmd(nid), cwd=cwd, env=env, stdout=subprocess.DEVNULL,
                                               stderr=subprocess.DEVNULL))
        self.channel = grpc.insecure_channel(f"127.0.0.1:{p1}")
        grpc.channel_ready_future(self.channel).result(timeout=180)
        self.stub = _wire().NodeServiceStub(self.channel)
        # node2 may come up a little later than node1: retry until a result flows
        x = np.zeros((1, 3, 32, 32), np.float32)
        t0 = time.time()
        while time.time() - t0 < 180:
            if self.request(x, "warm")[0] is not None:
                break
            time.sleep(0.5)
        else:
            raise RuntimeError("two-node oracle pipeline did not come up")
        return self

    def request(self, x: np.ndarray, request_id: str = "req"):
        w = _wire()
        x = np.ascontiguousarray(x)                              # the client of node.py:172-189
        rep = self.stub.SendTensor(w.TensorRequest(request_id=request_id, tensor=w.Tensor(
            tensor_data=x.tobytes(), shape=list(x.shape), dtype=str(x.dtype))))
        if not rep.HasField("result_tensor"):
            return None, rep.status
        t = rep.result_tensor
        return np.frombuffer(t.tensor_data, dtype=np.dtype(t.dtype)).reshape(tuple(t.shape)).copy(), rep.status

    def __exit__(self, *exc):
        self.channel.close()
        for p in self.procs:
            p.terminate()
        for p in self.procs:
            try:
                p.wait(timeout=20)
            except subprocess.TimeoutExpired:
                p.kill()
        return False


if __name__ == "__main__":
    main()
