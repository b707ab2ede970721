"""Multi-GPU plumbing (one process per GPU, torch.distributed): read sharding and the exchange of partition handles.

Reads are independent (reference src/dehost_main.cpp:341-376 keeps no cross-read state), so the data path has no
collective: every rank classifies its own contiguous share of the reads.  torch.distributed carries only
  * the 64-byte CUDA IPC handles of the filter partitions when the filter is split by row range, and
  * barriers / max-over-ranks of device-measured times in bench.py.
"""
import numpy as np


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous share [begin, end) of rank `rank`; shares differ by at most one item and cover [0, n_items)."""
    base, rem = divmod(n_items, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def all_gather_bytes(dist, payload: bytes, world: int, device=None):
    """all_gather of one fixed-size byte string per rank (NCCL needs device tensors, gloo takes CPU tensors)."""
    import torch
    t = torch.frombuffer(bytearray(payload), dtype=torch.uint8)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [bytes(o.cpu().numpy().tobytes()) for o in out]


def exchange_partitions(gpu, dist, rank: int, world: int, device=None):
    """Every rank exports the IPC handle of its filter partition and maps all the others (cg_partition_attach_ipc)."""
    handles = all_gather_bytes(dist, gpu.export_partition(), world, device)
    for peer, h in enumerate(handles):
        if peer != rank:
            gpu.attach_ipc(peer, h)
    dist.barrier()
    return handles


def max_over_ranks(dist, x: float, device=None) -> float:
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def merge_in_input_order(per_rank_arrays):
    """Host-side merge of per-rank outputs of contiguous shards back into input order."""
    return np.concatenate(per_rank_arrays, axis=0)
