"""Column sharding across the GPUs of one box.

Columns are independent, so every rank runs the full tree over its own contiguous column range and nothing is exchanged
while the passes run.  The single exchange is the variable-length gather of the sparse mutation records to rank 0
(`gather_records`: counts by all_gather, payloads by point-to-point send/recv -- NCCL over NVLink on GPUs, gloo in the
CPU tests).  Because ranges are rank-ordered and each rank's list is already sorted by (node,pos), ordering the
concatenation by the 64-bit key (node<<32 | pos) restores the global (node,pos) order.  The <=6-run NucMut packing must
run AFTER the gather: runs of consecutive positions can straddle shard boundaries and the greedy rule
(/root/reference/src/panman.cpp:1439-1455) is stateful from the start of each node's list.
"""
from __future__ import annotations

import numpy as np


def column_shards(n_cols: int, world: int, align: int = 32):
    """Contiguous column ranges, one per rank, every boundary a multiple of `align` (one 128-bit nibble vector)."""
    groups = (n_cols + align - 1) // align
    out = []
    for r in range(world):
        b = (groups * r // world) * align
        e = min(n_cols, (groups * (r + 1) // world) * align)
        out.append((min(b, n_cols), e if r < world - 1 else n_cols))
    return out


def records_as_tensor(ptr: int, n: int, device):
    """Wrap `n` device-resident pmfs_mut records (12 bytes each) as an int32 [n,3] torch tensor without copying."""
    import torch

    if n == 0:
        return torch.zeros((0, 3), dtype=torch.int32, device=device)

    class _Wrap:
        __cuda_array_interface__ = {"shape": (n, 3), "typestr": "<i4", "data": (ptr, False), "version": 3}

    return torch.as_tensor(_Wrap(), device=device)


def gather_records(rec, col_offset: int, rank: int, world: int, group=None):
    """rec: int32 [n,3] tensor (node, pos, packed type/code bytes), sorted by (node,pos), positions local to the shard.
    Returns on rank 0 the merged [sum n, 3] tensor sorted by (node, global pos); None elsewhere."""
    import torch
    import torch.distributed as dist

    dev = rec.device
    n = torch.tensor([rec.shape[0]], dtype=torch.int64, device=dev)
    counts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    local = rec.clone()
    local[:, 1] += col_offset
    if rank != 0:
        if rec.shape[0]:
            dist.send(local.contiguous(), dst=0, group=group)
        return None
    parts = [local]
    for r in range(1, world):
        if counts[r]:
            buf = torch.empty((counts[r], 3), dtype=torch.int32, device=dev)
            dist.recv(buf, src=r, group=group)
            parts.append(buf)
    allrec = torch.cat(parts, 0)
    key = (allrec[:, 0].to(torch.int64) << 32) | (allrec[:, 1].to(torch.int64) & 0xFFFFFFFF)
    order = torch.argsort(key, stable=True)
    return allrec.index_select(0, order)


def gather_device_records(ptr: int, n: int, rank: int, world: int, device, col_offset: int | None = None, cols_per_rank: int = 0):
    """bench.py helper: gather the records of the last pmfs_run_device call of every rank to rank 0."""
    rec = records_as_tensor(ptr, n, device)
    return gather_records(rec, col_offset if col_offset is not None else rank * cols_per_rank, rank, world)


def records_to_numpy(t) -> np.ndarray:
    from .api import MUT_DTYPE

    a = np.ascontiguousarray(t.cpu().numpy().astype("<i4"))
    return a.view(np.uint8).reshape(-1, 12).copy().view(MUT_DTYPE).reshape(-1)


def numpy_to_records(rec: np.ndarray):
    import torch

    a = np.ascontiguousarray(rec).view(np.uint8).reshape(-1, 12).copy().view("<i4").reshape(-1, 3)
    return torch.from_numpy(a.copy())
