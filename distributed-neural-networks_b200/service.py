"""The NodeService front door (reference: node.py:34-133 servicer + serve, :137-200 initiator).

An unmodified ``node_service.proto`` client calls ``SendTensor`` on the first node exactly as
it would on the reference; behind the door the request's tensor goes
bytes -> pinned host buffer -> H2D -> every stage of the chain on the GPU(s) -> D2H -> bytes.
There is no per-hop gRPC, no per-request channel and no host staging between stages; the
protobuf ``Tensor`` survives only at the two ends of the pipe.

What is mirrored from the reference so that clients and log scrapers see the same thing:
  * status strings: the last node says ``[id] Processing complete. Prediction: <k>`` with k the
    flat ``np.argmax`` of the whole output array (node.py:61-63); every node in front of it
    wraps that as ``[id] Forwarded. Next node status: ...`` (node.py:86);
  * failures never surface as gRPC errors: any exception becomes ``[id] Error: ...`` with no
    ``result_tensor`` (node.py:96-105);
  * ``HealthCheck`` answers True, ``SendMessage`` echoes (node.py:107-113).
Deliberate differences: the message-size cap is lifted (the reference dies at 4 MiB),
and one process serves the whole chain (stage k on GPU k) instead of one process per stage.
"""
from __future__ import annotations

import warnings
from concurrent import futures
from typing import Callable, Sequence

import numpy as np

from . import wire


class NodeServiceImpl(wire.NodeServiceServicer):
    """``run`` maps the request ndarray to the chain's output ndarray (e.g. ``ChainRunner``).

    Wire generality beyond the reference, all off by default so that a reference client sees the
    reference's bytes:
      * request dtype strings ``float16``, ``bfloat16`` (payload = bit patterns), ``float64`` and the integer types are
        accepted and widened / narrowed to what the chain computes in (``ChainRunner``); the reference would feed such
        a tensor to its fp32 modules and fail inside torch (node.py:45-53);
      * ``reply_dtype="same"`` answers in the request's dtype (half the bytes for fp16 / bf16 clients);
      * ``per_row_argmax=True`` appends `` Per-row: [k0, k1, ...]`` to the last node's status: the reference prints the
        FLAT arg-max of the whole [B,10] array (node.py:61), which for B > 1 is not a class index."""

    def __init__(self, node_ids: Sequence[str], run: Callable[[np.ndarray], np.ndarray], per_row_argmax: bool = False,
                 reply_dtype: str = "float32"):
        self.node_ids = list(node_ids)          # this node first, then every downstream node
        self.run = run
        self.per_row_argmax, self.reply_dtype = per_row_argmax, reply_dtype

    def _status(self, prediction: int, rows=None) -> str:
        status = f"[{self.node_ids[-1]}] Processing complete. Prediction: {prediction}"
        if rows is not None:
            status += f" Per-row: {rows}"
        for nid in reversed(self.node_ids[:-1]):
            status = f"[{nid}] Forwarded. Next node status: {status}"
        return status

    def SendTensor(self, request, context):
        me = self.node_ids[0]
        try:
            x = wire.tensor_to_ndarray(request.tensor)                      # node.py:45-47
            if request.tensor.dtype in ("bfloat16", "torch.bfloat16"):
                x = wire.bfloat16_bits_to_float32(x)
            y = np.ascontiguousarray(self.run(x))
            rows = [int(k) for k in np.argmax(y.reshape(y.shape[0], -1), axis=1)] if self.per_row_argmax and y.ndim >= 2 else None
            status = self._status(int(np.argmax(y)), rows)                  # node.py:61
            if self.reply_dtype == "same" and request.tensor.dtype in ("bfloat16", "torch.bfloat16"):
                result = wire.bfloat16_tensor(y)
            elif self.reply_dtype == "same" and np.dtype(request.tensor.dtype).kind == "f":
                result = wire.ndarray_to_tensor(y.astype(np.dtype(request.tensor.dtype)))
            else:
                result = wire.ndarray_to_tensor(y)
            return wire.TensorResponse(status=status, result_tensor=result)
        except Exception as e:                                              # node.py:96-100
            return wire.TensorResponse(status=f"[{me}] Error: {e}")

    def HealthCheck(self, request, context):
        return wire.HealthCheckResponse(is_healthy=True)

    def SendMessage(self, request, context):
        return wire.MessageReply(confirmation_text=f"[{self.node_ids[0]}] got msg '{request.message_text}'")


class ChainRunner:
    """Host tensor in, host tensor out, through a ``LocalPipeline`` (pinned staging, async copies).
    Batches larger than the pipeline's capacity are fed in capacity-sized pieces."""

    def __init__(self, pipeline, in_dtype=np.float32):
        import torch
        self.pipe, self.torch = pipeline, torch
        self.in_dtype = np.dtype(in_dtype)
        self.stage_in = torch.empty((pipeline.max_batch, *pipeline.in_shape), dtype=pipeline.in_dtype).pin_memory()

    def __call__(self, x: np.ndarray) -> np.ndarray:
        torch, p = self.torch, self.pipe
        if x.dtype != self.in_dtype:
            # float16 / float64 / integer payloads: converted once, on the host, to the type the chain
            # computes in -- exact for fp16 -> fp32 and for token ids; never a silent reinterpretation of bytes
            if x.dtype.kind in "fiu" and self.in_dtype.kind in "fiu" and not (x.dtype.kind == "f" and self.in_dtype.kind != "f"):
                x = x.astype(self.in_dtype)
            else:
                raise TypeError(f"expected a {self.in_dtype} tensor, got {x.dtype}")
        if tuple(x.shape[1:]) != tuple(p.in_shape):
            raise ValueError(f"expected input [B,{','.join(map(str, p.in_shape))}], got {list(x.shape)}")
        outs = []
        for lo in range(0, x.shape[0], p.max_batch):
            n = min(p.max_batch, x.shape[0] - lo)
            with warnings.catch_warnings():                    # np.frombuffer views are read-only; they are only read here
                warnings.simplefilter("ignore", UserWarning)
                src = torch.from_numpy(x[lo:lo + n])
            self.stage_in[:n].copy_(src)                       # the one host copy (node.py:48 makes two)
            outs.append(p.run_host(self.stage_in[:n], sync=True).numpy().copy())
        return np.concatenate(outs) if len(outs) != 1 else outs[0]


class BatchingRunner:
    """Dynamic micro-batching of concurrent requests.

    The reference handles one request at a time: its forward blocks the asyncio loop (node.py:53) and every hop
    opens a fresh channel (node.py:73), so B=1 clients pay the whole pipeline latency each.  Here gRPC worker
    threads enqueue their tensors; ONE dispatcher thread collects what has arrived -- up to ``max_rows`` rows, waiting
    at most ``max_wait_ms`` after the first arrival -- concatenates along the batch dimension, runs the chain once
    and hands every caller its own rows.  No op on the path mixes samples and the K slicing is fixed, so a request's
    result does not depend on what it was batched with (bit-exact; tests/test_gpu_cifar.py).  A failing batch is
    re-run request by request so that one malformed request cannot poison its neighbours; shape / dtype mismatches
    never enter a batch (only requests with the first request's row shape and dtype are merged)."""

    def __init__(self, run: Callable[[np.ndarray], np.ndarray], max_rows: int = 128, max_wait_ms: float = 0.2):
        import queue
        import threading
        self.run, self.max_rows, self.max_wait = run, int(max_rows), max_wait_ms * 1e-3
        self._q: "queue.Queue" = queue.Queue()
        self._queue_mod = queue
        self.batches: list[int] = []            # rows per executed batch (for tests / accounting)
        self._stop = False
        self._t = threading.Thread(target=self._loop, name="b2p-batcher", daemon=True)
        self._t.start()

    def __call__(self, x: np.ndarray) -> np.ndarray:
        import threading
        item = {"x": x, "done": threading.Event(), "y": None, "err": None}
        self._q.put(item)
        item["done"].wait()
        if item["err"] is not None:
            raise item["err"]
        return item["y"]

    def close(self):
        self._stop = True
        self._q.put(None)
        self._t.join(timeout=5)

    def _run_alone(self, it):
        try:
            it["y"] = self.run(it["x"])
        except Exception as e:                  # folded into the status string by the servicer
            it["err"] = e
        it["done"].set()

    def _loop(self):
        import time
        pending = None
        while not self._stop:
            first = pending if pending is not None else self._q.get()
            pending = None
            if first is None:
                break
            group, rows = [first], first["x"].shape[0] if first["x"].ndim else 1
            key = (first["x"].shape[1:], first["x"].dtype)
            deadline = time.perf_counter() + self.max_wait
            while rows < self.max_rows:
                left = deadline - time.perf_counter()
                try:
                    nxt = self._q.get(timeout=left) if left > 0 else self._q.get_nowait()
                except self._queue_mod.Empty:
                    break
                if nxt is None:
                    self._stop = True
                    break
                ok = nxt["x"].ndim >= 1 and (nxt["x"].shape[1:], nxt["x"].dtype) == key
                if not ok or rows + nxt["x"].shape[0] > self.max_rows:
                    pending = nxt               # starts the next batch
                    break
                group.append(nxt)
                rows += nxt["x"].shape[0]
            if len(group) == 1 or first["x"].ndim < 1:
                for it in group:
                    self._run_alone(it)
                self.batches.append(rows)
                continue
            try:
                y = self.run(np.concatenate([it["x"] for it in group]))
                lo = 0
                for it in group:
                    n = it["x"].shape[0]
                    it["y"] = y[lo:lo + n]
                    lo += n
                    it["done"].set()
                self.batches.append(rows)
            except Exception:                   # find the culprit: everyone gets an individual run
                for it in group:
                    self._run_alone(it)
                    self.batches.append(it["x"].shape[0])


def build_cifar_chain(topology, node_id: str, math: str = "tf32x3", microbatch: int = 128, max_batch: int = 1024,
                      state_dict=None):
    """What node.py:294-318 does for every node from ``node_id`` to the end of the chain, in one
    process: ``MODEL_PARTS_CLASSES[k](NeuralNetwork()).to(cuda:k')``, ``load_state_dict(strict=False)``,
    ``eval()``.  Returns (servicer, pipeline)."""
    import torch
    from .pipeline import LocalPipeline
    from .stages import MODEL_PARTS_CLASSES, NeuralNetwork
    me = topology.resolve(node_id)
    chain = topology.chain()[me.part_index:]
    if state_dict is None:
        state_dict = torch.load(topology.model_weights, map_location=torch.device("cpu"))   # node.py:296
    full = NeuralNetwork()
    n_gpu = torch.cuda.device_count()
    parts = []
    for n in chain:
        cls = MODEL_PARTS_CLASSES.get(n["part_index"])                     # node.py:300
        if cls is None:
            raise KeyError(f"No model part class defined for index {n['part_index']}")
        dev = n.get("device", n["part_index"] % max(n_gpu, 1))
        part = cls(full, math=math).to(f"cuda:{dev}")
        part.load_state_dict(state_dict, strict=False)                     # node.py:306
        parts.append(part.eval())
    shapes = {0: ((3, 32, 32), (4096,)), 1: ((4096,), (10,))}
    in_shape = shapes[chain[0]["part_index"]][0]
    out_shape = shapes[chain[-1]["part_index"]][1]
    hidden = [shapes[n["part_index"]][1] for n in chain[:-1]]
    pipe = LocalPipeline(parts, microbatch, in_shape, torch.float32, out_shape, hidden, max_batch)
    return NodeServiceImpl([n["id"] for n in chain], ChainRunner(pipe)), pipe


def serve(servicer, port: int, max_message_bytes: int = -1, workers: int = 1):
    """node.py:114-133 with a blocking server; one worker keeps the reference's
    one-request-at-a-time behaviour (node.py:53 blocks its event loop).  ``workers > 1`` admits concurrent
    requests: give the servicer a ``BatchingRunner`` then, the chain itself runs one batch at a time."""
    import grpc
    opts = [("grpc.max_send_message_length", max_message_bytes), ("grpc.max_receive_message_length", max_message_bytes)]
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=workers), options=opts)
    wire.add_NodeServiceServicer_to_server(servicer, server)
    bound = server.add_insecure_port(f"[::]:{port}")
    server.start()
    return server, bound


def load_input_image(path: str | None, seed: int | None = None):
    """node.py:142-154: ``Resize((32,32)) -> ToTensor -> Normalize(0.5, 0.5)`` of an RGB image, or the
    reference's fallback ``torch.randn(1,3,32,32)`` when the file is missing or unreadable.  Host-side,
    once per request, not on the hot path (out of scope for optimisation)."""
    import torch
    try:
        from PIL import Image
        if path is None:
            raise FileNotFoundError("no input image given")
        img = Image.open(path).convert("RGB").resize((32, 32), Image.BILINEAR)
        x = torch.from_numpy(np.asarray(img, dtype=np.float32) / 255.0).permute(2, 0, 1).unsqueeze(0)
        return ((x - 0.5) / 0.5).contiguous(), True
    except Exception:
        g = torch.Generator().manual_seed(seed) if seed is not None else None
        return torch.randn(1, 3, 32, 32, generator=g), False


def initiate_inference(servicer: NodeServiceImpl, input_image_path: str | None = None, seed: int | None = None):
    """The built-in client of node.py:137-200 for the node that owns part 0: load (or fake) one image,
    push it through the chain, arg-max the reply.  In the reference part 0 runs in-process and the rest
    behind a gRPC hop; here the whole chain is one device-resident pipeline, so this is a plain call.
    Returns (predicted_class_index, probabilities ndarray, used_real_image)."""
    x, real = load_input_image(input_image_path, seed)
    probs = servicer.run(x.numpy())
    return int(np.argmax(probs)), probs, real
