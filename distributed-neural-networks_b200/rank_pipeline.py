"""One process per GPU: stage k of the pipeline on rank k, activations handed over NVLink.

Reference counterpart: one ``node.py`` process per machine, stage = ``part_index``
(node.py:239-260), next hop = the node with ``part_index + 1`` (node.py:262-271), one gRPC
``SendTensor`` per request per hop (node.py:73-89).  Here the hop is a peer-memory ring
(handoff.py): the producer stage's kernel stores its output straight into the consumer
GPU's receive slot.  For the CIFAR tensor-core stages the slot handshake lives inside the stage
kernels (one call and one launch per stage per step, csrc/ring_hooks.cuh); other stages order the
two streams with one-thread wait/flag kernels.  The host of neither rank ever blocks inside a step.

Launched with ``python -m torch.distributed.run --nproc-per-node N ...``; torch.distributed
is used for rendezvous, the exchange of ring handles and the max-over-ranks of the timing.
"""
from __future__ import annotations

import json
import os
import statistics
import time
from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class RankRole:
    rank: int
    replica: int        # which independent pipeline this rank belongs to
    stage: int          # 0 .. stages-1 within the replica
    stages: int
    prev_rank: int | None
    next_rank: int | None
    active: bool = True


def plan(world: int, stages: int) -> list[RankRole]:
    """Ranks [r*stages, (r+1)*stages) form replica r; left-over ranks stay idle.  For the
    CIFAR workload stages == 2 (4 and 8 GPUs run 2 and 4 independent pipelines)."""
    roles = []
    replicas = world // stages
    for rank in range(world):
        rep, st = divmod(rank, stages)
        if rep >= replicas:
            roles.append(RankRole(rank, rep, st, stages, None, None, active=False))
            continue
        roles.append(RankRole(rank, rep, st, stages,
                              rank - 1 if st > 0 else None, rank + 1 if st < stages - 1 else None))
    return roles


def quiet_nccl():
    """bench.py must print exactly one JSON line on stdout, and NCCL writes its version banner (any NCCL_DEBUG
    level from VERSION up, which some images export by default) to stdout too: send NCCL's log to stderr instead,
    unless the caller already chose a file."""
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")


def exchange_handles(my_handle: bytes | None, group=None) -> list:
    """Every rank contributes the export handle of the ring it consumes from (or None)."""
    import torch.distributed as dist
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, my_handle, group=group)
    return out


class CifarRankPipeline:
    """BASELINE config 2 across ranks: even stage = conv part, odd stage = FC part."""

    ACT_BYTES = 4096 * 4            # per sample, the reference's fp32 [B,4096] handoff tensor

    def __init__(self, role: RankRole, device: torch.device, math: str, batch: int, microbatch: int,
                 slots: int = 32, group=None, input_via_device: int | None = None):     # four steps of slots (64 MiB on the consumer GPU): the stage-0 kernels of
                                                   # consecutive steps overlap (b2p_stage_set_pdl bit 2) and the consumer
                                                   # releases a step's slots only at the end of its step; with two steps of
                                                   # slots the producer waited for them (22.5 vs 26.0 M samples/s on 2 GPUs)
        import dnn_b200 as engine
        from .handoff import Ring
        self.role, self.dev, self.batch, self.mb = role, device, batch, microbatch
        self.n_mb = (batch + microbatch - 1) // microbatch
        self.seq = 0
        self.two_streams = False
        self.stream = torch.cuda.Stream(device=device)
        self.copy_stream = torch.cuda.Stream(device=device)
        torch.manual_seed(1234)
        full = engine.NeuralNetwork()
        sd = full.state_dict()
        self.part = engine.MODEL_PARTS_CLASSES[role.stage](full, math=math).to(device)
        self.part.load_state_dict(sd, strict=False)
        self.part.eval()
        slot_bytes = microbatch * self.ACT_BYTES
        self.ring = None
        mine = None
        if role.active and role.stage == 1:
            self.ring = Ring.create(device.index, slots, slot_bytes)
            mine = self.ring.export()
        handles = exchange_handles(mine, group)
        if role.active and role.stage == 0:
            self.ring = Ring.open(handles[role.next_rank], device.index, slots, slot_bytes)
        if role.stage == 0:
            # bit 2: the ring-mode kernel of step k+1 does not wait for step k's kernel -- its input is a resident buffer
            # or an H2D copy ordered by an event, never the preceding kernel's output (b200pipe.h, b2p_stage_set_pdl)
            if os.environ.get("B2P_RING_WAIT") != "1":
                self.part.set_pdl(7)
            self.x_dev = [torch.empty((batch, 3, 32, 32), device=device) for _ in range(2)]
            self._x_free = [None, None]
            self._flip = 0
            # Host-facing path: the input link is the limit (12.6 MB per step over one PCIe x16 = 4.4 M samples/s), and
            # the stage-1 GPU of the pipeline has a host link of its own that carries 40 KB of results per step.  The
            # second half of every batch therefore enters through THAT GPU (pinned host -> its memory) and crosses to
            # this one over NVLink (6 MiB: 26 us) -- measured 110 GB/s for the two links together against 55 GB/s.
            self.via = None
            if (input_via_device is not None and input_via_device != device.index
                    and 0 <= input_via_device < torch.cuda.device_count()
                    and torch.cuda.can_device_access_peer(device.index, input_via_device)):
                self.via = torch.device("cuda", input_via_device)
                self.half = (self.n_mb // 2) * microbatch
                self.hop_stream = torch.cuda.Stream(device=device)
                self._split_events = [tuple(torch.cuda.Event() for _ in range(3)) for _ in range(2)]
                self._xd_hi = [xd[self.half:] for xd in self.x_dev]
                self._hi_views = {}
                with torch.cuda.device(self.via):
                    self.via_stream = torch.cuda.Stream(device=self.via)
                    self.x_via = [torch.empty((batch - self.half, 3, 32, 32), device=self.via) for _ in range(2)]
        else:
            self.out_dev = torch.empty((batch, 10), device=device)
            self.out_host = torch.empty((batch, 10)).pin_memory()
            # consumer GPU: the tail kernel of step s on a second stream, hidden behind fc1 of step s+1 (which spends
            # its time waiting for the producer's messages); launches stay eager here (two kernels per step)
            self.aux = torch.cuda.Stream(device=device)
            if math == "tf32x3" and os.environ.get("B2P_NO_AUX_STREAM") != "1":
                self.part.set_aux_stream(self.aux.cuda_stream)
                self.two_streams = True
        self.slots = slots
        self._graphs = {}
        self._warm = set()
        self.graph_kernel_launches = 0

    # ---- one step = one batch through this rank's stage ------------------------------------------
    def step_device(self, x: torch.Tensor | None, ready=None):
        s = self.stream.cuda_stream
        if ready is None and self.n_mb <= self.slots:
            # the whole step as ONE call: microbatch i is message seq+i; the stage-0 kernel publishes each message
            # when its last row is stored, the stage-1 kernel waits message by message (b2p_stage_forward_ring_span)
            if self.role.stage == 0:
                self.part.forward_ring(x.data_ptr(), None, self.batch, 0, None, self.ring, self.seq, s, self.mb)
            else:
                self.part.forward_ring(None, self.out_dev.data_ptr(), self.batch, 0, self.ring, None, self.seq, s, self.mb)
            self.seq += self.n_mb
            return
        for i, lo in enumerate(range(0, self.batch, self.mb)):
            n = min(self.mb, self.batch - lo)
            if self.role.stage == 0:
                if ready is not None:
                    self.stream.wait_event(ready[i])
                # epilogue stores cross NVLink into the next GPU's slot; the slot handshake is inside the kernel
                self.part.forward_ring(x[lo:].data_ptr(), None, n, 0, None, self.ring, self.seq, s)
            else:
                self.part.forward_ring(None, self.out_dev[lo:].data_ptr(), n, 0, self.ring, None, self.seq, s)
            self.seq += 1

    def step_graph(self, x: torch.Tensor | None):
        """step_device as ONE CUDA-graph launch.  Possible because the ring's sequence counters live in
        device memory: no launch carries a step-dependent argument, so a captured step replays verbatim
        whenever it starts on the slot it was captured for.  One graph per (input buffer, first slot); the
        first visit of a key runs eagerly (warm-up, lazy allocations), the second captures and replays."""
        from . import _lib
        assert self.n_mb % self.slots == 0 or self.slots % self.n_mb == 0
        if self.two_streams:                           # fork/join across replays cannot live inside one captured step
            return self.step_device(x)
        key = (x.data_ptr() if x is not None else 0, self.seq % self.slots)
        entry = self._graphs.get(key)
        if entry is None:
            if key not in self._warm:
                self._warm.add(key)
                return self.step_device(x)
            self.stream.synchronize()
            g = torch.cuda.CUDAGraph()
            before = _lib.launch_count()
            seq_before = self.seq
            with torch.cuda.graph(g, stream=self.stream):
                self.step_device(x)                    # captured, not executed
            self.seq = seq_before                      # (host seq only selects slots; nothing ran)
            entry = self._graphs[key] = (g, _lib.launch_count() - before)
        with torch.cuda.stream(self.stream):
            entry[0].replay()
        self.seq += self.n_mb
        self.graph_kernel_launches += entry[1]

    def step_host(self, x_host: torch.Tensor | None, graph: bool = False):
        """Stage 0: pinned host batch -> H2D -> stage; stage 1: stage -> D2H of the result.
        graph=True: one large H2D copy, then the captured step (the copy of step s+1 still overlaps
        the compute of step s through the double-buffered device input)."""
        if self.role.stage == 0:
            flip = self._flip
            self._flip ^= 1
            xd = self.x_dev[flip]
            if self._x_free[flip] is not None:
                self.copy_stream.wait_event(self._x_free[flip])
            ready = []
            if graph and self.via is not None and 0 < self.half < self.batch:
                arrived, own, ev = self._split_events[flip]                # (events are re-recorded every step)
                if self._x_free[flip] is not None:
                    self.via_stream.wait_event(self._x_free[flip])         # the hop that read this staging buffer is long done
                    self.hop_stream.wait_event(self._x_free[flip])
                with torch.cuda.stream(self.via_stream):                   # the stage-1 GPU's host link
                    self.x_via[flip].copy_(self._x_host_hi(x_host), non_blocking=True)
                    arrived.record(self.via_stream)
                with torch.cuda.stream(self.copy_stream):                  # this GPU's host link
                    xd[:self.half].copy_(x_host[:self.half], non_blocking=True)
                    own.record(self.copy_stream)
                with torch.cuda.stream(self.hop_stream):                   # its own stream: this GPU's host link never waits for the hop
                    self.hop_stream.wait_event(arrived)
                    self._xd_hi[flip].copy_(self.x_via[flip], non_blocking=True)   # stage-1 GPU -> this GPU over NVLink
                    self.hop_stream.wait_event(own)
                    ev.record(self.hop_stream)
                ready.append(ev)
            elif graph:
                with torch.cuda.stream(self.copy_stream):
                    xd.copy_(x_host, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(self.copy_stream)
                    ready.append(ev)
            else:
                with torch.cuda.stream(self.copy_stream):
                    for lo in range(0, self.batch, self.mb):
                        xd[lo:lo + self.mb].copy_(x_host[lo:lo + self.mb], non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(self.copy_stream)
                        ready.append(ev)
            if graph:
                self.stream.wait_event(ready[-1])
                self.step_graph(xd)
            else:
                self.step_device(xd, ready)
            ev = torch.cuda.Event()
            ev.record(self.stream)
            self._x_free[flip] = ev
        else:
            if graph:
                self.step_graph(None)
            else:
                self.step_device(None)
            # the result is complete on the stream the tail kernel ran on: copy it out from there, so that the next
            # step's fc1 (on self.stream) is not ordered behind this step's tail
            with torch.cuda.stream(self.aux if self.two_streams else self.stream):
                self.out_host.copy_(self.out_dev, non_blocking=True)

    def _x_host_hi(self, x_host: torch.Tensor) -> torch.Tensor:
        """Second half of a pinned host batch (the view is cached: host buffers recur)."""
        v = self._hi_views.get(x_host.data_ptr())
        if v is None:
            v = self._hi_views[x_host.data_ptr()] = x_host[self.half:]
        return v

    def join(self):
        """Make this rank's stream wait for tail kernels still running on the second stream."""
        if self.two_streams:
            self.part.join(self.stream.cuda_stream)


def input_over_both_links(world: int, override: str = "") -> bool:
    """Host-facing path: does a 2-GPU pipeline take its input over both of its GPUs' host links?  Yes while the links are
    the limit (up to two pipelines per node); from three pipelines up the host's memory system is (bench_main).
    override: "1" / "0" force it (B2P_SPLIT_H2D)."""
    if override in ("0", "1"):
        return override == "1"
    return world // 2 <= 2


def bench_main(args) -> int:
    """bench.py --gpus N under torchrun (N >= 2)."""
    import torch.distributed as dist
    import dnn_b200 as engine
    from bench import (BATCH, MICROBATCH, METRIC, UNIT, WORKLOAD, FLOP_PART0, BYTES_PART0, ClockSampler, _peaks)

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    quiet_nccl()
    # stdout carries exactly one JSON line: park the real stdout and point fd 1 at stderr while NCCL and torch
    # initialise and run (NCCL prints its version banner with printf, whatever NCCL_DEBUG_FILE says)
    import sys
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    dist.init_process_group("nccl", device_id=dev)
    roles = plan(world, 2)
    role = roles[rank]
    # Host-facing path: a pipeline's input may enter over both of its GPUs' host links (CifarRankPipeline).  Worth it while
    # the links are the limit: 2 pipelines (4 GPUs) 8.8 -> 13.2 M samples/s; with 4 pipelines all eight links together
    # ask the host for more than its memory system delivers (~200 GB/s measured: 15.1 M split against 17.2 M over the
    # four stage-0 links alone), so from 3 pipelines up the input stays on the stage-0 links.  B2P_SPLIT_H2D=0/1 forces it.
    via = None
    split_h2d = input_over_both_links(world, os.environ.get("B2P_SPLIT_H2D", ""))
    if role.active and role.stage == 0 and split_h2d:
        via = dev.index + (role.next_rank - rank)            # one node: the stage-1 GPU of this pipeline
    pipe = CifarRankPipeline(role, dev, args.math, BATCH, MICROBATCH, slots=int(os.environ.get("B2P_RING_SLOTS", "32")),
                             input_via_device=via)
    replicas = world // 2
    W, K = max(args.warmup, 3), args.steps
    n_pool = 16
    g = torch.Generator().manual_seed(rank)
    host_pool = dev_pool = None
    if role.active and role.stage == 0:
        host_pool = [torch.randn(BATCH, 3, 32, 32, generator=g).pin_memory() for _ in range(n_pool)]
        dev_pool = [h.to(dev) for h in host_pool]
    sampler = ClockSampler(local)
    sampler.start()

    def timed(fn):
        for i in range(W):
            fn(i)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record(pipe.stream)
        for i in range(K):
            fn(W + i)
        if role.active:
            pipe.join()
        e1.record(pipe.stream)
        torch.cuda.synchronize()
        wall = time.perf_counter() - w0
        dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1), 1e3 * wall], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)          # max over ranks
        return float(t[0]), float(t[1])

    use_graph = not getattr(args, "no_graph", False)                 # host-facing path
    dev_graph = getattr(args, "graph_device_path", False)            # device-resident path: eager launches with PDL by default
    step = pipe.step_graph if dev_graph else pipe.step_device
    if role.active and dev_graph:                      # capture one graph per input buffer (stage 0) / one (stage 1)
        for i in range(2 * n_pool + 2):
            step(dev_pool[i % n_pool] if dev_pool else None)
        torch.cuda.synchronize()
    dist.barrier()
    launches0 = engine.launch_count() + pipe.graph_kernel_launches
    sampler.active.set()
    if role.active:
        ms_dev, _ = timed(lambda i: step(dev_pool[i % n_pool] if dev_pool else None))
    else:
        ms_dev, _ = timed(lambda i: None)
    launches = engine.launch_count() + pipe.graph_kernel_launches - launches0
    if role.active:
        # the host-facing path has its own graphs (double-buffered device input x first slot): capture them outside
        # the timed region whatever W is (a key is run eagerly on its first visit and captured on its second)
        for i in range(10):
            pipe.step_host(host_pool[i % n_pool] if host_pool else None, graph=use_graph)
        torch.cuda.synchronize()
    dist.barrier()
    if role.active:
        _, ms_e2e = timed(lambda i: pipe.step_host(host_pool[i % n_pool] if host_pool else None, graph=use_graph))
    else:
        _, ms_e2e = timed(lambda i: None)
    sampler.active.clear()
    clocks = sampler.finish()
    # Who limits the 2-stage pipeline?  Each side timed WITHOUT the other in its way: the ring holds two steps of
    # messages, so the producer can run two steps while the consumer sits still (no back-pressure), and the consumer
    # can then run its two steps on messages that are all there already (no waiting).
    torch.cuda.synchronize()
    dist.barrier()
    alone = torch.zeros(2, device=dev)
    n_alone = max(2, pipe.slots // (BATCH // MICROBATCH)) if role.active else 2      # as many steps as the ring holds
    if role.active:
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if role.stage == 0:
            pe0.record(pipe.stream)
            for i in range(n_alone):
                pipe.step_device(dev_pool[i % n_pool])
            pe1.record(pipe.stream)
            torch.cuda.synchronize()
            alone[0] = 1e3 * pe0.elapsed_time(pe1) / n_alone
    dist.barrier()
    if role.active and role.stage == 1:
        pe0.record(pipe.stream)
        for i in range(n_alone):
            pipe.step_device(None)
        pipe.join()
        pe1.record(pipe.stream)
        torch.cuda.synchronize()
        alone[1] = 1e3 * pe0.elapsed_time(pe1) / n_alone
    dist.barrier()
    dist.all_reduce(alone, op=dist.ReduceOp.MAX)
    if os.environ.get("B2P_TRACE_RING") and rank == 0:       # developer aid: CTA 0's timeline of the ring-mode stage-0 kernel
        import ctypes as C
        from . import _lib
        buf = (C.c_longlong * 1024)()
        _lib.check(_lib.lib().b2p_stage_trace(pipe.part._handle, buf, 1024))      # arm
        for i in range(4):
            pipe.step_device(dev_pool[i % n_pool])
        torch.cuda.synchronize()
        _lib.check(_lib.lib().b2p_stage_trace(pipe.part._handle, buf, 1024))
        t = list(buf)
        ent = [t[256 + 3 * i] for i in range(148)]; rel = [t[257 + 3 * i] for i in range(148)]; ext = [t[258 + 3 * i] for i in range(148)]
        e0 = min(ent)
        us = lambda v: round((v - e0) / 1e3, 1)
        print(f"ring trace (last launch) entry min/max {us(min(ent))}/{us(max(ent))} released min/max {us(min(rel))}/{us(max(rel))} "
              f"exit min/median/max {us(min(ext))}/{us(sorted(ext)[74])}/{us(max(ext))} us; CTA0 {us(ent[0])} {us(rel[0])} {us(ext[0])}", file=sys.stderr)
        for n in range(7):
            row = {r: [t[(r * 8 + n) * 8 + e] - t[192] for e in range(8) if t[(r * 8 + n) * 8 + e]] for r in range(4)}
            print(f"ring trace n={n} mma={row[0]} e1={row[1]} e2={row[2]}", file=sys.stderr)
    lt = torch.tensor([float(launches)], device=dev)
    dist.all_reduce(lt)
    checksum = float(pipe.out_host.sum()) if role.active and role.stage == 1 else 0.0
    ct = torch.tensor([checksum], device=dev)
    dist.all_reduce(ct)
    # Did the ring deliver the right activations?  The FC rank regenerates its producer's last input batch (same
    # seeded generator), runs BOTH parts locally through the plain two-call path and compares with what came out
    # of the cross-GPU pipeline in the last end-to-end step: the kernels are deterministic, so it must be bit-exact.
    mismatch = 0.0
    if role.active and role.stage == 1:
        gp = torch.Generator().manual_seed(role.prev_rank)
        pool = [torch.randn(BATCH, 3, 32, 32, generator=gp) for _ in range(n_pool)]
        x_last = pool[(W + K - 1) % n_pool].to(dev)
        torch.manual_seed(1234)
        full = engine.NeuralNetwork()
        p0 = engine.MODEL_PARTS_CLASSES[0](full, math=args.math).to(dev)
        p0.load_state_dict(full.state_dict(), strict=False)
        with torch.no_grad():
            want = torch.cat([pipe.part(p0.eval()(x_last[lo:lo + MICROBATCH])) for lo in range(0, BATCH, MICROBATCH)])
        mismatch = float((want.cpu() - pipe.out_host).abs().max())
    mt = torch.tensor([mismatch], device=dev)
    dist.all_reduce(mt, op=dist.ReduceOp.MAX)

    # roofline: stage-0 kernel timed alone on rank 0's launch stream
    roofline = None
    if rank == 0:
        peaks, kind = _peaks()
        s = pipe.stream
        act = torch.empty(BATCH, 4096, device=dev)
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 0
        torch.cuda.synchronize()
        r0.record(s)
        for i in range(K):                  # the launch shape of the timed region: one launch per step
            x = dev_pool[i % n_pool]
            pipe.part.forward_ptr(x.data_ptr(), act.data_ptr(), BATCH, 0, s.cuda_stream)
            n += 1
        r1.record(s)
        torch.cuda.synchronize()
        us = 1e3 * r0.elapsed_time(r1) / n
        tf = FLOP_PART0 * BATCH / (us * 1e-6) / 1e12
        roofline = {"kernel": f"part0_tc2 (conv1+ReLU+pool+conv2+ReLU+pool, both convolutions on tcgen05), one launch per step = {BATCH} samples (output in local memory here)",
                    "bound": "tensor", "achieved": tf, "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                    "frac": tf / peaks["bf16_tflops_sustained"], "traffic": None,
                    "peak_source": f"bf16_tflops_sustained of {kind}", "us_per_launch": us,
                    "hbm_view": {"achieved_gbs": BYTES_PART0 * BATCH / (us * 1e-6) / 1e9, "peak_gbs": peaks["hbm_gbs"]}}
    dist.barrier()
    gpt = None
    if not getattr(args, "no_gpt", False):
        try:
            del pipe                                         # the CIFAR rings and stages are done
            torch.cuda.empty_cache()
            sampler2 = ClockSampler(local)
            sampler2.start(); sampler2.active.set()
            gpt = gpt_bench(world, rank, dev, _peaks()[0])
            sampler2.active.clear()
            gpt["clocks"] = sampler2.finish()
        except Exception as e:                               # never lose the CIFAR line over the extra
            gpt = {"error": f"{type(e).__name__}: {e}"}
    dist.barrier()
    if rank == 0:
        total = BATCH * K * replicas
        line = {
            "metric": METRIC, "value": total / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "f32"}.get(args.math, args.math), "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch": BATCH, "microbatch": MICROBATCH, "stages": 2,
                       "placement": f"{replicas} pipeline(s) of 2 stages, stage k on GPU 2r+k; activations stored by the "
                                    "stage-0 kernel directly into a peer-memory ring on the stage-1 GPU (NVLink)",
                       "math": args.math,
                       "launch": ("cuda_graph_per_step_per_rank" if dev_graph else "one_call_per_rank_per_step_with_programmatic_dependent_launch") +
                                 " (device path); " + ("cuda_graph_per_step_per_rank" if use_graph else "per_kernel") + " (host-facing path)",
                       "l2_policy": f"inputs rotate over {n_pool} distinct batches = 201 MB > 126 MB L2"},
            "e2e": {"value": total / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": BATCH * 3 * 32 * 32 * 4 * replicas,
                    "d2h_bytes_per_step": BATCH * 10 * 4 * replicas, "ms_per_step": ms_e2e / K, "result_checksum": float(ct[0]),
                    "input_path": ("half of every batch over the stage-0 GPU's host link, half over the stage-1 GPU's and on over NVLink"
                                   if split_h2d else "all over the stage-0 GPU's host link"),
                    "max_abs_diff_vs_single_gpu_path": float(mt[0])},
            "gpu_launches": int(lt[0]), "clocks": clocks, "roofline": roofline, "cpu_baseline": None, "gpt": gpt,
            "stage_times": {"stage0_us_ring": float(alone[0]), "stage1_us_fed": float(alone[1]), "step_us": 1e3 * ms_dev / K,
                            "stage0_busy_frac": float(alone[0]) / (1e3 * ms_dev / K), "stage1_busy_frac": float(alone[1]) / (1e3 * ms_dev / K),
                            "what": "per 1024-sample step: the stage-0 kernel storing into the peer ring with no back-pressure "
                                    "(two steps into free slots), the stage-1 kernels on messages already delivered, and the "
                                    "steady-state step of the pipeline; max over replicas"},
        }
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    dist.barrier()
    dist.destroy_process_group()
    return 0


class GptRankPipeline:
    """BASELINE configs 3/4: rank k runs the reference's part k (ModelPart0 | ModelPartIntermediate*
    | ModelPartFinal_GPT, partitions/gpt_model_parts.py) over a contiguous layer range; hidden
    states [microbatch*T, C] fp32 travel through peer-memory rings.  Each stage's last kernel
    (the mlp.c_proj GEMM epilogue with the fused residual add) stores straight into the next
    GPU's ring slot."""

    def __init__(self, role: RankRole, device: torch.device, model, math: str, mb_seqs: int, seq: int,
                 ranges=None, slots: int = 4, group=None):
        import dnn_b200 as engine
        from .handoff import Ring
        self.role, self.dev, self.mb, self.seq_len = role, device, mb_seqs, seq
        cfg = model.config
        self.cfg = cfg
        self.stream = torch.cuda.Stream(device=device)
        parts = engine.gpt_parts_for(model, role.stages, math=math, max_tokens=mb_seqs * seq, ranges=ranges)
        self.part = parts[role.stage].to(device).eval()
        # Two lanes (B2P_GPT_LANES=2): a second instance of the stage (own workspaces, own copy of the weights) on a second
        # stream takes every other microbatch, so the kernels of microbatch i+1 fill the SMs that the last wave of
        # microbatch i's kernels leaves idle (2048-token GEMMs are 1.3 to 2 waves of tiles on 148 SMs).
        self.lanes = int(os.environ.get("B2P_GPT_LANES", "2"))
        self.lanes = self.lanes if self.lanes in (1, 2, 4) and slots % self.lanes == 0 else 1    # a ring slot always sees the same lane
        self.parts_l = [self.part]
        self.streams_l = [self.stream]
        for _ in range(self.lanes - 1):
            parts2 = engine.gpt_parts_for(model, role.stages, math=math, max_tokens=mb_seqs * seq, ranges=ranges)
            self.parts_l.append(parts2[role.stage].to(device).eval())
            self.streams_l.append(torch.cuda.Stream(device=device))
        slot_bytes = mb_seqs * seq * cfg.n_embd * 4
        self.ring_in = self.ring_out = None
        mine = None
        if role.stage > 0:
            self.ring_in = Ring.create(device.index, slots, slot_bytes)
            mine = self.ring_in.export()
        handles = exchange_handles(mine, group)
        if role.next_rank is not None:
            self.ring_out = Ring.open(handles[role.next_rank], device.index, slots, slot_bytes)
        self.seq_no = 0
        self.is_last = role.next_rank is None
        if self.is_last:
            self.ids_l = [torch.empty((mb_seqs, seq), dtype=torch.int64, device=device) for _ in range(self.lanes)]
            self.ids = self.ids_l[0]          # the ids of the microbatch processed last
            self.logits = None

    def step(self, idx_batch: torch.Tensor | None, n_microbatches: int, head: str = "ids"):
        """One batch = n_microbatches microbatches of mb_seqs sequences through this rank's stage."""
        import ctypes as C
        from . import _lib
        if self.lanes > 1:
            assert head == "ids" or not self.is_last            # (the logits buffer of the final rank is not doubled)
            for st in self.streams_l[1:]:
                st.wait_stream(self.stream)                      # fork (also inside a graph capture)
        for i in range(n_microbatches):
            lane = self.seq_no % self.lanes                      # a ring slot always sees the same lane (slots are even)
            part, s = self.parts_l[lane], self.streams_l[lane].cuda_stream
            src = self.ring_in.recv(self.seq_no, s) if self.ring_in else idx_batch[i * self.mb:].data_ptr()
            if not self.is_last:
                dst = self.ring_out.acquire(self.seq_no, s)
                part.forward_ptr(src, dst, self.mb, self.seq_len, s)
                self.ring_out.publish(self.seq_no, s)
            elif head == "ids":
                self.ids = self.ids_l[lane]
                _lib.check(_lib.lib().b2p_stage_forward_argmax(part._handle, C.c_void_p(src),
                                                               C.c_void_p(self.ids.data_ptr()), self.mb, self.seq_len,
                                                               C.c_void_p(s)))
            else:
                if self.logits is None:
                    self.logits = torch.empty((self.mb, self.seq_len, self.cfg.vocab_size), device=self.dev)
                part.forward_ptr(src, self.logits.data_ptr(), self.mb, self.seq_len, s)
            if self.ring_in:
                self.ring_in.release(self.seq_no, s)
            self.seq_no += 1
        for st in self.streams_l[1:]:
            self.stream.wait_stream(st)                          # join

    def step_graph(self, idx_batch: torch.Tensor | None, n_microbatches: int, head: str = "ids"):
        """``step`` as ONE CUDA-graph launch per rank.  The ring counters live in device memory, so no launch carries a
        step-dependent argument and a captured step replays verbatim whenever it starts on the slot it was captured
        for (one graph per (token buffer, first slot); first visit of a key eager, second captures)."""
        if not hasattr(self, "_graphs"):
            self._graphs, self._warm = {}, set()
        slots = (self.ring_in or self.ring_out).slots if (self.ring_in or self.ring_out) else 1
        key = (idx_batch.data_ptr() if idx_batch is not None else 0, self.seq_no % slots, n_microbatches, head)
        g = self._graphs.get(key)
        if g is None:
            if key not in self._warm:
                self._warm.add(key)
                return self.step(idx_batch, n_microbatches, head)
            self.stream.synchronize()
            g = torch.cuda.CUDAGraph()
            seq_before = self.seq_no
            with torch.cuda.graph(g, stream=self.stream):
                self.step(idx_batch, n_microbatches, head)       # captured, not executed
            self.seq_no = seq_before
            self._graphs[key] = g
        with torch.cuda.stream(self.stream):
            g.replay()
        self.seq_no += n_microbatches


# ================================================================================================================
# BASELINE configs 3 and 4 inside bench.py's line (the `gpt` key)
# ================================================================================================================
def gpt_workload(world: int) -> dict:
    """Which GPT configuration `bench.py --gpus N` measures next to the CIFAR metric: N stages of the reference's part
    chain (gpt_model_parts.py:6-50), one stage per GPU.  N=4 is BASELINE config 3 and N=8 config 4 as written."""
    if world >= 8:
        return dict(model="medium", math="bf16", batch=32, microbatch=2, seq=1024, balance=False,
                    what="BASELINE config 4 as written: GPT-2 medium, 8 stages x 3 blocks, bf16, 16 microbatches x 2 sequences")
    if world == 4:
        return dict(model="small", math="bf16", batch=32, microbatch=8, seq=1024, balance=True,
                    what="BASELINE config 3: GPT-2 small, 4 stages (layer ranges balanced against the LM head), batch 32 as "
                         "4 microbatches x 8 sequences, bf16 throughput mode")
    return dict(model="small", math="bf16", batch=32, microbatch=8, seq=1024, balance=world > 1,
                what=f"GPT-2 small on {world} GPU(s) ({max(world, 1)} stage(s)), batch 32 as 4 microbatches x 8 sequences, bf16")


def gpt_flop_per_token(cfg, seq: int) -> float:
    """24 C^2 per block + causal attention counted as executed (2 T C) + LM head 2 C V."""
    return cfg.n_layer * (24 * cfg.n_embd ** 2 + 2 * seq * cfg.n_embd) + 2 * cfg.n_embd * cfg.vocab_size


def gpt_bench(world: int, rank: int, dev: torch.device, peaks: dict, steps: int = 8, group=None) -> dict | None:
    """Tokens/s of the GPT part chain over `world` GPUs, one CUDA-graph launch per rank per batch, timed on the device
    (max over ranks), plus a correctness bit: the greedy ids the LAST rank produced for the last microbatch, compared
    with the same tokens through a single-GPU chain of the same parts built on that rank (`ids_equal_single_gpu`)."""
    import torch.distributed as dist
    import dnn_b200 as engine
    w = gpt_workload(world)
    dims = dict(engine.GPT2_SMALL if w["model"] == "small" else engine.GPT2_MEDIUM)
    cfg = engine.GPTConfig(block_size=1024, vocab_size=50257, **dims)
    model = engine.GPT(cfg, seed=1234)
    seq, mb, batch = w["seq"], w["microbatch"], w["batch"]
    n_mb = batch // mb
    g = torch.Generator().manual_seed(0)
    idx_host = torch.randint(0, cfg.vocab_size, (batch, seq), generator=g)
    t = torch.zeros(1, device=dev)
    ids_equal = None
    if world == 1:
        import ctypes as C
        from . import _lib
        # two lanes, as in GptRankPipeline: a second instance of the chain on a second stream takes every other
        # microbatch, so its kernels fill the SMs the last wave of the other lane's kernels leaves idle
        lanes = max(1, min(2, int(os.environ.get("B2P_GPT_LANES", "2")))) if n_mb % 2 == 0 else 1
        chains = [[p.to(dev).eval() for p in engine.gpt_parts_for(model, 1, math=w["math"], max_tokens=mb * seq)]
                  for _ in range(lanes)]
        first, final = chains[0]
        idx = idx_host.to(dev)
        hiddens = [torch.empty(mb, seq, cfg.n_embd, device=dev) for _ in range(lanes)]
        ids = torch.empty(batch, seq, dtype=torch.int64, device=dev)
        stream = torch.cuda.Stream(device=dev)
        streams = [stream] + [torch.cuda.Stream(device=dev) for _ in range(lanes - 1)]

        def step():
            for st in streams[1:]:
                st.wait_stream(stream)                           # fork (also inside the graph capture)
            for i, lo in enumerate(range(0, batch, mb)):
                lane = i % lanes
                s = streams[lane].cuda_stream
                chains[lane][0].forward_ptr(idx[lo:].data_ptr(), hiddens[lane].data_ptr(), mb, seq, s)
                _lib.check(_lib.lib().b2p_stage_forward_argmax(chains[lane][1]._handle, C.c_void_p(hiddens[lane].data_ptr()),
                                                               C.c_void_p(ids[lo:].data_ptr()), mb, seq, C.c_void_p(s)))
            for st in streams[1:]:
                stream.wait_stream(st)                           # join
        step(); step()
        stream.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            graph.replay()
            e0.record(stream)
            for _ in range(steps):
                graph.replay()
            e1.record(stream)
        stream.synchronize()
        ms = e0.elapsed_time(e1) / steps
        # the fused arg-max head against the materialised logits of the same stage, last microbatch
        with torch.no_grad(), torch.cuda.stream(stream):
            logits = final(first(idx[batch - mb:]))
        stream.synchronize()
        ids_equal = bool(torch.equal(logits.argmax(-1), ids[batch - mb:]))
        ranges = "one GPU: first part (all blocks) -> final part (ln_f + head)"
    else:
        ranges = engine.balanced_ranges(cfg, world, seq) if w["balance"] else None
        role = plan(world, world)[rank]
        pipe = GptRankPipeline(role, dev, model, w["math"], mb, seq, ranges=ranges, group=group)
        idx = idx_host.to(dev) if rank == 0 else None
        for _ in range(3):                                    # eager, capture, first replay
            pipe.step_graph(idx, n_mb)
        torch.cuda.synchronize(); dist.barrier(group=group); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(pipe.stream)
        for _ in range(steps):
            pipe.step_graph(idx, n_mb)
        e1.record(pipe.stream)
        torch.cuda.synchronize(); dist.barrier(group=group)
        t[0] = e0.elapsed_time(e1)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        ms = float(t[0]) / steps
        eq = torch.ones(1, device=dev)
        if pipe.is_last:                                      # the chain's ids for the last microbatch vs one GPU
            got = pipe.ids.clone()
            parts = [p.to(dev).eval() for p in engine.gpt_parts_for(model, world, math=w["math"], max_tokens=mb * seq, ranges=ranges)]
            with torch.no_grad():
                x = idx_host[batch - mb:].to(dev)
                for p in parts[:-1]:
                    x = p(x)
                want = parts[-1].greedy_ids(x)
            eq[0] = float(torch.equal(got, want))
            del parts
        dist.all_reduce(eq, op=dist.ReduceOp.MIN, group=group)
        ids_equal = bool(eq[0] > 0)
    tokens = batch * seq
    tf = gpt_flop_per_token(cfg, seq) * tokens / (ms * 1e-3) / 1e12
    return {"what": w["what"], "model": f"gpt2-{w['model']}", "stages": world, "math": w["math"], "batch": batch, "seq_len": seq,
            "microbatch_sequences": mb, "layer_ranges": ranges or "even", "ms_per_batch": ms, "tokens_per_s": tokens / (ms * 1e-3),
            "model_tflops_total": tf, "frac_of_sustained_bf16_peak_per_gpu": tf / world / peaks["bf16_tflops_sustained"],
            "launch": "one CUDA graph per rank per batch", "lanes_per_rank": (lanes if world == 1 else pipe.lanes), "ids_equal_single_gpu": ids_equal, "steps": steps,
            "data": "synthetic tokens, random-init weights (seed 1234)"}
