"""Host-side logic of the multi-process pipeline on CPU: role planning and the ring-handle
exchange over a real world_size-2 gloo group (no GPU needed)."""
import os
import socket
from importlib import import_module

import pytest
import torch.multiprocessing as mp


def _rp():
    import dnn_b200  # noqa: F401  (puts the package on sys.path)
    return import_module("distributed-neural-networks_b200.rank_pipeline")


def test_plan_two_stage_replicas():
    rp = _rp()
    r = rp.plan(2, 2)
    assert [(x.replica, x.stage, x.prev_rank, x.next_rank) for x in r] == [(0, 0, None, 1), (0, 1, 0, None)]
    r = rp.plan(8, 2)
    assert [x.replica for x in r] == [0, 0, 1, 1, 2, 2, 3, 3]
    assert all(x.active for x in r)
    assert r[5].stage == 1 and r[5].prev_rank == 4 and r[4].next_rank == 5
    r = rp.plan(3, 2)                      # an odd rank is left idle rather than mis-wired
    assert [x.active for x in r] == [True, True, False]
    r = rp.plan(8, 8)                      # GPT: one 8-stage chain
    assert [x.stage for x in r] == list(range(8)) and r[7].next_rank is None and r[3].prev_rank == 2


def test_input_link_policy():
    rp = _rp()
    assert rp.input_over_both_links(2) and rp.input_over_both_links(4)       # one or two pipelines: both host links each
    assert not rp.input_over_both_links(8)                                   # four: the host's memory system is the limit
    assert rp.input_over_both_links(8, "1") and not rp.input_over_both_links(2, "0")


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rp = _rp()
    role = rp.plan(world, 2)[rank]
    mine = (b"ring-of-rank-%d" % rank).ljust(256, b"\0") if role.stage == 1 else None
    handles = rp.exchange_handles(mine)
    got = handles[role.next_rank] if role.stage == 0 else None
    q.put((rank, role.stage, got))
    dist.barrier()
    dist.destroy_process_group()


def test_handle_exchange_world2_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][:2] == (0, 0) and res[0][2].startswith(b"ring-of-rank-1")   # producer got its consumer's ring
    assert res[1] == (1, 1, None)
