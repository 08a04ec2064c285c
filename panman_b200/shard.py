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


def node_ranges(n_nodes: int, world: int):
    """Contiguous node-id ranges, one per rank: rank r owns the mutation lists of nodes [b, e)."""
    return [((n_nodes * r) // world, (n_nodes * (r + 1)) // world) for r in range(world)]


def exchange_by_node(rec, col_offset: int, rank: int, world: int, n_nodes: int, group=None):
    """Scalable form of the gather: an all-to-all that leaves every rank with the COMPLETE mutation lists (all column
    shards, sorted by (node, global pos)) of its own node range -- per-rank traffic and merge work stay constant as ranks
    are added, instead of piling everything on rank 0.  rec: int32 [n,3] sorted by (node,pos), positions shard-local.

    One host synchronisation (the split sizes); the merge is a scatter, not a sort: parts arrive sorted and shard
    ranges are rank-ordered, so a record's slot is  start(node) + records of that node from lower ranks + its index
    inside its own part's run of that node."""
    import torch
    import torch.distributed as dist

    dev = rec.device
    bounds = torch.tensor([b for b, _ in node_ranges(n_nodes, world)] + [n_nodes], dtype=torch.int32, device=dev)
    nodes = rec[:, 0].contiguous()
    cuts = torch.searchsorted(nodes, bounds).to(torch.int64)          # [world+1] split points of my list
    all_cuts = [torch.empty_like(cuts) for _ in range(world)]
    dist.all_gather(all_cuts, cuts, group=group)
    all_cuts = torch.stack(all_cuts).cpu()                            # the one host sync
    send = (all_cuts[rank, 1:] - all_cuts[rank, :-1]).tolist()
    recv = (all_cuts[:, rank + 1] - all_cuts[:, rank]).tolist()
    local = rec.clone()
    local[:, 1] += col_offset
    inbox = torch.empty((sum(recv), 3), dtype=torch.int32, device=dev)
    dist.all_to_all_single(inbox, local, output_split_sizes=recv, input_split_sizes=send, group=group)
    # scatter-merge of the `world` sorted parts
    b, e = node_ranges(n_nodes, world)[rank]
    nloc = max(1, e - b)
    node_l = (inbox[:, 0] - b).to(torch.int64)
    part = torch.repeat_interleave(torch.arange(world, device=dev), torch.tensor(recv, device=dev))
    counts = torch.zeros((world, nloc), dtype=torch.int64, device=dev)
    counts.view(-1).index_add_(0, part * nloc + node_l, torch.ones_like(node_l))
    per_node = counts.sum(0)
    node_start = torch.cumsum(per_node, 0) - per_node                 # first slot of each node
    before = torch.cumsum(counts, 0) - counts                         # records of that node from lower ranks
    run_start = torch.cumsum(counts, 1) - counts                      # index of the node's run inside its part
    part_start = torch.tensor([0] + recv[:-1], device=dev).cumsum(0)
    idx_in_part = torch.arange(inbox.shape[0], device=dev) - part_start[part]
    dest = node_start[node_l] + before[part, node_l] + (idx_in_part - run_start[part, node_l])
    out = torch.empty_like(inbox)
    out[dest] = inbox
    return out


def _as_tensor(ptr: int, n: int, dtype_str: str, device):
    import torch

    tdt = {"<i8": torch.int64, "<i4": torch.int32}[dtype_str]
    if n == 0:
        return torch.zeros(0, dtype=tdt, device=device)

    class _Wrap:
        __cuda_array_interface__ = {"shape": (n,), "typestr": dtype_str, "data": (ptr, False), "version": 3}

    return torch.as_tensor(_Wrap(), device=device)


_xbuf = {}


def gather_slices_by_node_range(rec, rank: int, world: int, n_nodes: int, group=None):
    """The protocol of pmfs_comm_run (panman_b200/csrc/comm.cu) restated on torch tensors, for the CPU / gloo tests of the
    multi-GPU host logic: rec is this rank's int32 [n,3] records (node, GLOBAL pos, payload) ordered by (node,pos); the
    ranks' column ranges are disjoint and ascending with the rank.
      1. noff[v] = first record of node v (the library gets it from the prefix sums that ordered the records),
         cut[r] = noff[first node of rank r]: rank r's records are one contiguous slice of every source;
      2. all-gather of the cut points; 3. all-to-all of the slices and of the matching noff ranges;
      4. stitch: node v's list is the concatenation of its runs in source order -- no sort.
    Returns this rank's node range [n_nodes*rank/world, n_nodes*(rank+1)/world) of the unsharded result."""
    import torch
    import torch.distributed as dist

    bounds = [(n_nodes * r) // world for r in range(world + 1)]
    nodes = rec[:, 0].contiguous()
    noff = torch.searchsorted(nodes, torch.arange(n_nodes + 1, dtype=nodes.dtype)).to(torch.int64)
    cuts = noff[torch.tensor(bounds)]
    all_cuts = [torch.empty_like(cuts) for _ in range(world)]
    dist.all_gather(all_cuts, cuts, group=group)
    send_n = [int(cuts[p + 1] - cuts[p]) for p in range(world)]
    recv_n = [int(all_cuts[s][rank + 1] - all_cuts[s][rank]) for s in range(world)]
    inbox = torch.empty((sum(recv_n), 3), dtype=rec.dtype)
    # (the slices are contiguous and in owner order: this rank's records ARE the send buffer)
    dist.all_to_all_single(inbox, rec[int(cuts[0]):int(cuts[world])].contiguous(), output_split_sizes=recv_n, input_split_sizes=send_n,
                           group=group)
    recv = list(torch.split(inbox, recv_n))
    n_own = bounds[rank + 1] - bounds[rank]
    send_off = torch.cat([noff[bounds[p]:bounds[p + 1] + 1] for p in range(world)])
    off_in = torch.empty(world * (n_own + 1), dtype=torch.int64)
    dist.all_to_all_single(off_in, send_off, output_split_sizes=[n_own + 1] * world,
                           input_split_sizes=[bounds[p + 1] - bounds[p] + 1 for p in range(world)], group=group)
    recv_off = list(torch.split(off_in, n_own + 1))
    runs = torch.stack([o[1:] - o[:-1] for o in recv_off])           # [source, owned node] run lengths
    total = runs.sum(0)
    start = torch.cumsum(total, 0) - total
    out = torch.empty((int(total.sum()), 3), dtype=rec.dtype)
    before = torch.cumsum(runs, 0) - runs                             # records of that node from lower ranks
    for s in range(world):
        rel = recv_off[s] - recv_off[s][0]
        node_l = torch.repeat_interleave(torch.arange(n_own), runs[s])
        idx = torch.arange(recv[s].shape[0]) - rel[node_l]
        out[start[node_l] + before[s][node_l] + idx] = recv[s]
    return out


def exchange_parts(ctx, rank: int, world: int, device, group=None, same_stream: bool = False):
    """The data path's single exchange on GPUs, after a shard-mode pmfs_run_device on every rank: all-to-all of the
    64-bit records (already grouped by owning rank when they were emitted) over NCCL, then each rank orders what it
    received with the library's radix sort.  Returns (device pointer to this rank's pmfs_mut records, count).

    (The first, radix-sorting form of the exchange, kept for the shard-mode entry points pmfs_set_shard / pmfs_get_parts /
    pmfs_sort_records; bench.py and C++ callers use pmfs_comm_run, which needs no sort.)
    same_stream: pass True only when the context was told to run on torch's current stream (Context.set_stream with that
    stream's handle): the sort is then ordered after the all-to-all on the device.  Otherwise the all-to-all is
    synchronised on the host first -- NCCL's work.wait() orders torch's stream, not the context's.  Small tensors and the
    inbox are reused across calls."""
    import torch
    import torch.distributed as dist

    base, offsets, send, col_ordered = ctx.get_parts()
    key = (str(device), world)
    if key not in _xbuf:
        _xbuf[key] = {"h": torch.empty(world + 1, dtype=torch.int64).pin_memory(),
                      "d": torch.empty(world + 1, dtype=torch.int64, device=device),
                      "t": torch.empty((world, world + 1), dtype=torch.int64, device=device),
                      "th": torch.empty((world, world + 1), dtype=torch.int64).pin_memory(),
                      "inbox": torch.empty(0, dtype=torch.int64, device=device)}
    B = _xbuf[key]
    B["h"][:world] = torch.tensor(send, dtype=torch.int64)
    B["h"][world] = int(col_ordered)
    B["d"].copy_(B["h"], non_blocking=True)
    dist.all_gather_into_tensor(B["t"].view(-1), B["d"], group=group)
    B["th"].copy_(B["t"], non_blocking=True)
    torch.cuda.current_stream(device).synchronize()  # the one host synchronisation of the exchange
    recv = B["th"][:, rank].tolist()
    all_ordered = bool(B["th"][:, world].min().item())
    total = sum(recv)
    if B["inbox"].shape[0] < total:
        B["inbox"] = torch.empty(int(total * 1.25) + 1024, dtype=torch.int64, device=device)
    inbox = B["inbox"][:total]
    if all_ordered:  # every owner's block is contiguous and the blocks follow each other: one all_to_all_single
        outbox = _as_tensor(base + 8 * offsets[0], sum(send), "<i8", device)
        dist.all_to_all_single(inbox, outbox, output_split_sizes=recv, input_split_sizes=send, group=group)
    else:
        outs = [_as_tensor(base + 8 * offsets[r], send[r], "<i8", device) for r in range(world)]
        dist.all_to_all(list(torch.split(inbox, recv)), outs, group=group)
    if not same_stream:
        torch.cuda.current_stream(device).synchronize()
    out = ctx.sort_records(inbox.data_ptr(), total, column_ordered=all_ordered)
    return out, total


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


def stitch_by_node(shards, n_nodes: int) -> np.ndarray:
    """Per-node mutation lists from column shards that kept their output (pmfs_set_shard with n_parts = 0): `shards` is
    the list, in column order, of each shard's (node,pos)-ordered record array (api.MUT_DTYPE, global positions).  A
    node's list is the concatenation of the shards' slices for that node; the result is the array an unsharded run over
    all columns returns.  Host-side pointer arithmetic on per-node counts -- this is all the "gather" a consumer of
    per-node lists needs when no rank has to own complete lists on the device."""
    counts = np.zeros((len(shards), n_nodes + 1), np.int64)
    for s, rec in enumerate(shards):
        counts[s, 1:] = np.bincount(rec["node"], minlength=n_nodes)[:n_nodes]
    per_node = counts.sum(axis=0)
    node_off = np.cumsum(per_node)                     # start of node v's list in the result: node_off[v]
    within = np.cumsum(counts[:, 1:], axis=0) - counts[:, 1:]  # [s, v]: records of v in shards before s
    total = int(node_off[-1])
    out = np.zeros(total, shards[0].dtype if shards else None)
    for s, rec in enumerate(shards):
        if rec.shape[0] == 0:
            continue
        src_off = np.cumsum(counts[s]) [:-1]            # start of node v inside shard s
        nodes = rec["node"].astype(np.int64)
        dst = node_off[nodes] + within[s, nodes] + (np.arange(rec.shape[0]) - src_off[nodes])
        out[dst] = rec
    return out
