"""Micro-batched pipeline schedules over the stage operators.

The reference moves ONE whole activation per request from node to node through
gRPC and overlaps nothing (node.py:52-53 blocks the event loop; node.py:73-89 is a
nested unary RPC).  Here a request batch is cut along the batch dimension into
microbatches -- exact, because no op on the path mixes samples -- and every
microbatch flows stage by stage while the next one is already being copied /
computed (1F pipelining).

``LocalPipeline``  all stages driven from one process (one or several GPUs); this is
                   what the NodeService front door and the single-GPU bench use.
``RankPipeline``   (multi-process, one rank per GPU) lives in rank_pipeline.py.
"""
from __future__ import annotations

import torch

from .stages import EngineStage


class LocalPipeline:
    """Stages on one or more GPUs of this process, activations stay in HBM.

    ``run_device``  inputs already resident on the first stage's GPU
    ``run_host``    the user-facing call: pinned host input -> H2D per microbatch on a
                    copy stream -> stages -> D2H of the result, all asynchronous; only
                    the final ``synchronize`` blocks.
    """

    def __init__(self, stages: list[EngineStage], microbatch: int, in_shape: tuple, in_dtype,
                 out_shape: tuple, hidden_shapes: list[tuple], max_batch: int):
        assert stages and all(s._handle is not None for s in stages), "stages must be on CUDA devices"
        self.stages = stages
        self.mb = int(microbatch)
        self.max_batch = int(max_batch)
        self.devs = [s._device for s in stages]
        self.in_shape, self.in_dtype, self.out_shape = tuple(in_shape), in_dtype, tuple(out_shape)
        d0, dl = self.devs[0], self.devs[-1]
        self.compute = [torch.cuda.Stream(device=d) for d in self.devs]
        self.copy_in = torch.cuda.Stream(device=d0)
        # device-resident buffers, allocated once
        self.x_dev = [torch.empty((self.max_batch, *self.in_shape), dtype=in_dtype, device=d0) for _ in range(2)]
        self.hidden = [torch.empty((self.max_batch, *shp), dtype=torch.float32, device=self.devs[k + 1])
                       for k, shp in enumerate(hidden_shapes)]
        # when consecutive stages sit on different GPUs the producer writes into a staging
        # buffer on its own GPU and a peer copy moves it (handoff); same GPU: write in place
        self.stage_out = []
        for k, shp in enumerate(hidden_shapes):
            if self.devs[k] == self.devs[k + 1]:
                self.stage_out.append(self.hidden[k])
            else:
                self.stage_out.append(torch.empty((self.max_batch, *shp), dtype=torch.float32, device=self.devs[k]))
        self.out_dev = torch.empty((self.max_batch, *self.out_shape), dtype=torch.float32, device=dl)
        self.out_host = torch.empty((self.max_batch, *self.out_shape), dtype=torch.float32).pin_memory()
        self._flip = 0
        self._x_free = [None, None]        # event: last reader of x_dev[i] has finished
        # back-pressure between steps: hidden[k] rows of microbatch i may be overwritten only
        # after stage k+1 has consumed what the previous step left there
        self._consumed = [dict() for _ in hidden_shapes]
        for s in stages:
            s._finalize()
            if getattr(s, "WHOLE_BATCH_OK", False):
                s.reserve(self.max_batch)          # no forward of this pipeline allocates (graph capture, serving)
        # CUDA graphs of the whole device-resident schedule, keyed by (input address, batch, seq)
        self._graphs: dict = {}
        self.graph_kernel_launches = 0      # kernels executed through graph replays (for accounting)
        # stages sharing a GPU run on two streams: early-started dependent grids would squat on
        # the SMs the other stream needs, so programmatic dependent launch is switched off then
        self._pdl = len(set(self.devs)) == len(self.devs)

    # -- device-resident schedule -------------------------------------------------------
    def run_device(self, x: torch.Tensor, out: torch.Tensor | None = None, seq: int = 0,
                   ready: list | None = None) -> torch.Tensor:
        """x: [B, *in_shape] on stage 0's GPU.  Enqueues everything; returns the device
        output (valid once ``self.compute[-1]`` has drained).  ``ready[i]`` (optional) is
        an event after which microbatch i of x may be read."""
        from . import _lib
        b = x.shape[0]
        assert b <= self.max_batch
        out = self.out_dev if out is None else out
        n_st = len(self.stages)
        if ready is None and len(set(self.devs)) == 1 and all(getattr(st, "WHOLE_BATCH_OK", False) for st in self.stages):
            # Every stage on ONE GPU and the batch already resident: micro-batching has nothing to overlap with (each
            # stage's kernel fills the GPU), so each stage runs once over the whole batch on one stream -- same rows,
            # same per-row arithmetic (fixed K slicing), bit-identical results, an eighth of the launches.
            s = self.compute[0]
            self._joined = []                  # no other stream took part (run_device_graph joins nothing)
            src = x
            for k, st in enumerate(self.stages):
                dst = out[:b] if k == n_st - 1 else self.hidden[k][:b]
                st.set_pdl(3)                       # per stage, not process-wide: other pipelines keep their own setting
                st.forward_ptr(src.data_ptr(), dst.data_ptr(), b, seq, s.cuda_stream)
                src = dst
            return out[:b]
        self._joined = self.compute[1:]
        for st in self.stages:
            st.set_pdl(3 if self._pdl else 0)       # shared GPU: no early starts (measured: 0.199 vs 0.238 ms/step)
        for i, lo in enumerate(range(0, b, self.mb)):
            n = min(self.mb, b - lo)
            src = x[lo:lo + n]
            for k, st in enumerate(self.stages):
                s = self.compute[k]
                if k == 0 and ready is not None:
                    s.wait_event(ready[i])
                if k > 0:
                    s.wait_event(prev_done)
                if k < n_st - 1 and i in self._consumed[k]:
                    s.wait_event(self._consumed[k][i])
                dst = out[lo:lo + n] if k == n_st - 1 else self.stage_out[k][lo:lo + n]
                st.forward_ptr(src.data_ptr(), dst.data_ptr(), n, seq, s.cuda_stream)
                if k > 0:
                    done = torch.cuda.Event()
                    done.record(s)
                    self._consumed[k - 1][i] = done
                if k < n_st - 1:
                    if self.stage_out[k] is not self.hidden[k]:
                        with torch.cuda.stream(s):         # NVLink peer copy on the producer's stream
                            self.hidden[k][lo:lo + n].copy_(dst, non_blocking=True)
                    prev_done = torch.cuda.Event()
                    prev_done.record(s)
                    src = self.hidden[k][lo:lo + n]
        return out[:b]

    # -- the same schedule as one CUDA graph ---------------------------------------------------
    def run_device_graph(self, x: torch.Tensor, seq: int = 0) -> torch.Tensor:
        """Like run_device, but the whole micro-batched schedule (all stages, all streams,
        their events) is captured once per distinct input buffer and replayed with a single
        launch: the CIFAR step is ~25 kernels of a few microseconds each, so per-kernel
        launch latency would otherwise dominate.  The result is complete when
        ``self.compute[0]`` has drained (the capture joins every stage stream back)."""
        from . import _lib
        key = (x.data_ptr(), x.shape[0], seq)
        entry = self._graphs.get(key)
        main = self.compute[0]
        if entry is None:
            self.run_device(x, seq=seq)                 # warm-up outside capture: lazy allocations
            self.synchronize()
            self._consumed = [dict() for _ in self._consumed]    # no events from outside the capture
            g = torch.cuda.CUDAGraph()
            before = _lib.launch_count()
            with torch.cuda.graph(g, stream=main):
                self.run_device(x, seq=seq)
                for st in self._joined:
                    main.wait_stream(st)
            entry = (g, _lib.launch_count() - before)
            self._graphs[key] = entry
            self._consumed = [dict() for _ in self._consumed]
        with torch.cuda.stream(main):
            entry[0].replay()
        self.graph_kernel_launches += entry[1]
        return self.out_dev[:x.shape[0]]

    # -- host-facing schedule ---------------------------------------------------------------
    def run_host(self, x_host: torch.Tensor, seq: int = 0, sync: bool = True, graph: bool = False) -> torch.Tensor:
        """x_host: pinned host tensor [B, *in_shape].  Returns the pinned host result
        [B, *out_shape] (a view of an internal buffer, overwritten by the next call).
        graph=True replays the captured schedule (one launch per step); the H2D copy of step
        s+1 still overlaps the compute of step s through the double-buffered device input."""
        b = x_host.shape[0]
        assert b <= self.max_batch
        flip = self._flip
        xd = self.x_dev[flip]
        self._flip ^= 1
        ready = []
        # the copy stream must not overwrite xd while the run before last still reads it
        # (double buffer: the copy of step s+1 overlaps the compute of step s)
        if self._x_free[flip] is not None:
            self.copy_in.wait_event(self._x_free[flip])
        with torch.cuda.stream(self.copy_in):
            if graph:                          # the graph starts when the whole batch has landed: one large copy
                xd[:b].copy_(x_host, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy_in)
                ready.append(ev)
            else:
                for lo in range(0, b, self.mb):
                    n = min(self.mb, b - lo)
                    xd[lo:lo + n].copy_(x_host[lo:lo + n], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(self.copy_in)
                    ready.append(ev)
        if graph:
            self.compute[0].wait_event(ready[-1])
            out = self.run_device_graph(xd[:b], seq=seq)
        else:
            out = self.run_device(xd[:b], seq=seq, ready=ready)
        ev = torch.cuda.Event()
        ev.record(self.compute[0])
        self._x_free[flip] = ev
        last = self.compute[0] if graph else self.compute[-1]
        with torch.cuda.stream(last):
            self.out_host[:b].copy_(out, non_blocking=True)
        self._last_stream = last
        if sync:
            last.synchronize()
        return self.out_host[:b]

    def synchronize(self):
        for s in self.compute:
            s.synchronize()


def cifar_pipeline(parts, microbatch: int = 128, max_batch: int = 1024) -> LocalPipeline:
    """BASELINE config 2: CIFAR conv | FC, batch 1024 in microbatches of 128."""
    return LocalPipeline(list(parts), microbatch, (3, 32, 32), torch.float32, (10,), [(4096,)], max_batch)
