"""World-size-2 gloo run of the data-parallel host logic (bucketed all-reduce of a flat gradient buffer)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from p2p_pytorch_b200.trainer import GradReducer
    red = GradReducer(bucket_bytes=1000 * 4)             # 1000-float buckets -> 4 buckets, ragged first one
    assert red.world == world
    flat = torch.arange(3210, dtype=torch.float32) * (rank + 1)
    red.launch(flat)
    assert len(red.pending) == 4
    red.wait()
    want = torch.arange(3210, dtype=torch.float32) * sum(r + 1 for r in range(world))
    q.put((rank, bool(torch.equal(flat, want))))
    dist.destroy_process_group()


def test_bucketed_allreduce_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]


def test_reducer_is_a_noop_without_process_group():
    from p2p_pytorch_b200.trainer import GradReducer
    red = GradReducer()
    assert red.world == 1
    t = torch.ones(10)
    red.launch(t)
    red.wait()
    assert torch.equal(t, torch.ones(10))
