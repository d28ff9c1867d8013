"""CPU (gloo, world_size 2) tests of the multi-rank host logic: the shard arithmetic and the payload all-gather
that the long-horizon path uses over NCCL on the GPUs. No compute calls (there is no GPU here)."""
import os
import socket

import numpy as np
import pytest

from rslqr_b200.dist import shard_range


def test_shard_range_partitions():
    for total in (1, 7, 4096, 4097):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from rslqr_b200.dist import agree_all, allgather_payload, shard_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pay = 904
    send = torch.full((pay,), float(rank + 1), dtype=torch.float64)
    send[0] = 100.0 + rank
    gathered = torch.zeros(pay * world, dtype=torch.float64)
    allgather_payload(send, gathered)
    lo, hi = shard_range(4096, rank, world)
    # batch sharding: ranks agree on a disjoint cover
    spans = [None] * world
    dist.all_gather_object(spans, (lo, hi))
    # transport agreement: one rank without peer access sends every rank to the fallback
    votes = (agree_all(True), agree_all(rank != 1), agree_all(False))
    q.put((rank, gathered.numpy().copy(), spans, votes))
    dist.destroy_process_group()


def test_payload_allgather_gloo_world2():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    out = [q.get(timeout=120) for _ in range(2)]
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, g, spans, votes in out:
        assert votes == (1, 0, 0)
        g = g.reshape(2, 904)
        assert g[0, 0] == 100.0 and g[1, 0] == 101.0           # rank order = shard order along the horizon
        assert (g[0, 1:] == 1.0).all() and (g[1, 1:] == 2.0).all()
        assert spans == [(0, 2048), (2048, 4096)]
