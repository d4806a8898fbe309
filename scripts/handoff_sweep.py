"""BASELINE config 5: inter-stage activation handoff sweep, 64 KiB .. 256 MiB.

Run with N = 2, 4 or 8 ranks on one box:
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/handoff_sweep.py
Rank k+1 owns the ring it consumes from, rank k produces into it; all N-1 hops of the chain run at once.  For every size the payload is a counter
pattern verified by checksum on the consumer; bandwidth = bytes / device time (CUDA events on
the producer stream, max over ranks).  Three transports are compared:
  ring/copy_engine  cudaMemcpyAsync into the mapped slot
  ring/sm_copy      SM copy kernel with 128-bit peer stores
  nccl send/recv    torch.distributed (the library baseline)
plus, on rank 0's host cores, the reference's own hop (tobytes -> TensorRequest -> gRPC loopback
SendTensor -> frombuffer().copy(), node.py:73-89,45-48) with the unmodified 4 MiB cap and with
the cap raised in the harness only.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import dnn_b200 as engine
from importlib import import_module

Ring = import_module("distributed-neural-networks_b200.handoff").Ring
SIZES = [64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20]


def reference_hop(sizes, raise_cap):
    """The reference's serialisation path over a real localhost gRPC channel."""
    import grpc
    from concurrent import futures
    w = engine.wire

    class Sink(w.NodeServiceServicer):
        def SendTensor(self, request, context):
            a = np.frombuffer(request.tensor.tensor_data, dtype=np.dtype(request.tensor.dtype)).reshape(
                tuple(request.tensor.shape)).copy()
            return w.TensorResponse(status=f"ok {a.shape[0]}")

    opts = [("grpc.max_send_message_length", -1), ("grpc.max_receive_message_length", -1)] if raise_cap else []
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=1), options=opts)
    w.add_NodeServiceServicer_to_server(Sink(), server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    out = {}
    try:
        for n in sizes:
            if n > (64 << 20):
                continue
            a = np.arange(n // 4, dtype=np.float32)
            reps = 3 if n >= (16 << 20) else 10
            try:
                t0 = time.perf_counter()
                for _ in range(reps):
                    with grpc.insecure_channel(f"127.0.0.1:{port}", options=opts) as ch:     # new channel per hop
                        stub = w.NodeServiceStub(ch)
                        stub.SendTensor(w.TensorRequest(request_id="s", tensor=w.Tensor(
                            tensor_data=a.tobytes(), shape=list(a.shape), dtype=str(a.dtype))))
                out[n] = n * reps / (time.perf_counter() - t0) / 1e9
            except grpc.RpcError as e:
                out[n] = f"fails: {e.code().name}"
    finally:
        server.stop(0)
    return out


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    import_module("distributed-neural-networks_b200.rank_pipeline").quiet_nccl()
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)                                     # NCCL prints its banner on stdout
    dist.init_process_group("nccl", device_id=dev)
    assert world >= 2
    from bench import ClockSampler
    sampler = ClockSampler(local)
    sampler.start(); sampler.active.set()
    # The pipeline's traffic pattern: EVERY stage boundary active at once.  Rank k > 0 owns the ring it consumes from,
    # rank k < world-1 produces into rank k+1's ring; all world-1 hops run concurrently.
    producer, consumer = rank < world - 1, rank > 0
    results = []
    stream = torch.cuda.Stream()
    for n in SIZES:
        reps = 20 if n <= (16 << 20) else 6
        slots = 2
        ring_in = ring_out = handle = None
        if consumer:
            ring_in = Ring.create(local, slots, n)
            handle = ring_in.export()
        hs = [None] * world
        dist.all_gather_object(hs, handle)
        if producer:
            ring_out = Ring.open(hs[rank + 1], local, slots, n)
        pattern = torch.arange(n // 8, dtype=torch.int64, device=dev) * 2654435761
        payload = pattern + rank                      # every hop carries its own payload
        want = int((pattern + (rank - 1)).sum().item()) if consumer else 0
        row = {"bytes": n}
        seq = 0
        for mode in ("copy_engine", "sm_copy"):
            if producer:
                ring_out.set_mode(mode)
            ok = True
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                for i in range(reps):
                    if producer:
                        ring_out.send(payload.data_ptr(), n, seq, stream.cuda_stream)
                    if consumer:
                        ptr = ring_in.recv(seq, stream.cuda_stream)
                        if i == reps - 1:
                            class R:  # noqa
                                __cuda_array_interface__ = {"shape": (n // 8,), "typestr": "<i8", "data": (ptr, False), "version": 2}
                            got = int(torch.as_tensor(R(), device=dev).sum().item())
                            ok = got == want
                        ring_in.release(seq, stream.cuda_stream)
                    seq += 1
                e1.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1), 0.0 if ok else 1.0], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)          # the slowest rank: every hop moved n*reps bytes in that time
            row[f"ring_{mode}_GBps"] = n * reps / (float(t[0]) * 1e-3) / 1e9
            row[f"ring_{mode}_ok"] = float(t[1]) == 0.0
        # library baseline: NCCL send/recv on the same chain, all hops at once
        rbuf = torch.empty(n // 8, dtype=torch.int64, device=dev)

        def nccl_round():
            ops = []
            if producer:
                ops.append(dist.P2POp(dist.isend, payload, rank + 1))
            if consumer:
                ops.append(dist.P2POp(dist.irecv, rbuf, rank - 1))
            for r in dist.batch_isend_irecv(ops):
                r.wait()
        for _ in range(2):
            nccl_round()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            nccl_round()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        row["nccl_sendrecv_GBps"] = n * reps / (float(t[0]) * 1e-3) / 1e9
        results.append(row)
        for r in (ring_in, ring_out):
            if r is not None:
                r.close()
        dist.barrier()
    sampler.active.clear()
    clocks = sampler.finish()
    if rank == 0:
        ref = reference_hop(SIZES, raise_cap=False)
        ref_raised = reference_hop(SIZES, raise_cap=True)
        for row in results:
            row["reference_grpc_GBps"] = ref.get(row["bytes"], "not run (>64 MiB)")
            row["reference_grpc_cap_raised_GBps"] = ref_raised.get(row["bytes"], "not run (>64 MiB)")
        out = {"handoff_sweep": results, "n_gpus": world, "hops": world - 1, "host_cores": os.cpu_count(), "clocks_rank0": clocks,
               "what": "GB/s PER HOP with all world-1 stage boundaries of the chain active at once (max device time over ranks); "
                       "ring_copy_engine = plain cudaMemcpyAsync into the peer-mapped slot (the denominator), ring_sm_copy = the "
                       "engine's copy kernel with 128-bit peer stores, nccl = batch_isend_irecv on the same chain; the reference's "
                       "hop (tobytes -> TensorRequest -> gRPC -> frombuffer().copy()) on rank 0's host cores"}
        os.write(real_stdout, (json.dumps(out, indent=1) + "\n").encode())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
