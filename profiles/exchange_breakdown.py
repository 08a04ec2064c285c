"""torchrun helper: wall-clock breakdown of one multi-GPU step (shard-mode run + exchange), printed by rank 0."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch, torch.distributed as dist
from panman_b200 import api, trees, shard

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
N, L = 20000, 30000
tree = trees.random_binary_tree(N); V = tree.n_nodes
ctx = api.Context(tree, device=dev.index)
stride = ((L + 31) // 32) * 16
d_msa = torch.zeros((N, stride), dtype=torch.uint8, device=dev); d_cons = torch.zeros(L, dtype=torch.uint8, device=dev)
torch.cuda.synchronize()
ctx.synth_msa(api.SynthSpec(), rank * L, L, d_msa.data_ptr(), stride, d_cons.data_ptr())
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ctx.set_shard(rank * L, world * L, [b for b, _ in shard.node_ranges(V, world)] + [V])
def T():
    torch.cuda.synchronize(); return time.perf_counter()
for it in range(6):
    dist.barrier(); t0 = T()
    ctx.run_device(api.FITCH, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr()); t1 = T()
    base, offsets, send, col_ordered = ctx.get_parts(); t2 = T()
    mine = torch.tensor(send, dtype=torch.int64, device=dev); table = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(table, mine); recv = torch.stack(table)[:, rank].tolist(); t3 = T()
    total = sum(recv)
    inbox = torch.empty(total, dtype=torch.int64, device=dev)
    outbox = shard._as_tensor(base + 8 * offsets[0], sum(send), "<i8", dev)
    dist.all_to_all_single(inbox, outbox, output_split_sizes=recv, input_split_sizes=send); t4 = T()
    ctx.sort_records(inbox.data_ptr(), total, column_ordered=col_ordered); t5 = T()
    if rank == 0:
        print(f"it{it}: run {1e3*(t1-t0):.3f} wrap {1e3*(t2-t1):.3f} counts {1e3*(t3-t2):.3f} a2a {1e3*(t4-t3):.3f} sort {1e3*(t5-t4):.3f} ms; timing {ctx.timing()}")
dist.destroy_process_group()
