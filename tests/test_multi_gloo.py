"""world_size-2 gloo tests (CPU) of the N>1 host logic: read sharding, partition-handle exchange, max over ranks."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from charon_b200 import multi


class FakeGpu:
    """Stands in for GpuClassifier's partition plumbing (no CUDA here): records what gets attached."""

    def __init__(self, rank):
        self.rank, self.attached = rank, {}

    def export_partition(self):
        return bytes([self.rank + 1]) * 64

    def attach_ipc(self, peer, handle):
        self.attached[peer] = handle


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n = 1001
        b, e = multi.shard_range(n, rank, world)
        ids = np.arange(n)[b:e]
        gpu = FakeGpu(rank)
        handles = multi.exchange_partitions(gpu, dist, rank, world)
        t = multi.max_over_ranks(dist, 10.0 + rank)
        # gather every rank's shard on rank 0 to check the input-order merge
        sizes = [multi.shard_range(n, r, world) for r in range(world)]
        bufs = [torch.zeros(s[1] - s[0], dtype=torch.int64) for s in sizes]
        pad = torch.from_numpy(ids.astype(np.int64))
        if rank == 0:
            bufs[0] = pad
            for r in range(1, world):
                dist.recv(bufs[r], src=r)
            merged = multi.merge_in_input_order([x.numpy() for x in bufs])
        else:
            dist.send(pad, dst=0)
            merged = None
        q.put((rank, (b, e), sorted(gpu.attached.items()), [h[0] for h in handles], t, None if merged is None else merged.tolist() == list(range(n))))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding_and_handle_exchange():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, s0, a0, h0, t0, m0), (r1, s1, a1, h1, t1, m1) = res
    assert s0 == (0, 501) and s1 == (501, 1001)
    assert a0 == [(1, bytes([2]) * 64)] and a1 == [(0, bytes([1]) * 64)]      # each rank mapped exactly the other's partition
    assert h0 == [1, 2] and h1 == [1, 2]
    assert t0 == 11.0 and t1 == 11.0
    assert m0 is True


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 8, 1000003):
        for world in (1, 2, 3, 8):
            spans = [multi.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1
