"""Debug: how many records of a tile shape differ from the 4x2 shape on a config-4-like input (Fitch + Sankoff)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from panman_b200 import api, shard, trees
shapes = [tuple(map(int, s.split("x"))) for s in (sys.argv[1:] or ["4x1", "2x1"])]
N, L = 1000, 5_000_000
tree = trees.random_binary_tree(N)
spec = api.SynthSpec(p_sub=2e-3, p_del=5e-3, p_ins=5e-3, p_n_block=1e-3, p_amb=1e-4, accessory_frac=0.35)
ctx = api.Context(tree)
stride = ((L + 31) // 32) * 16
d_msa = torch.zeros((N, stride), dtype=torch.uint8, device="cuda")
d_cons = torch.zeros(L, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
ctx.synth_msa(spec, 0, L, d_msa.data_ptr(), stride, d_cons.data_ptr())
def run(shape, algo):
    ctx.set_tile(*shape)
    ptr, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
    return shard.records_as_tensor(ptr, n, "cuda").clone()
def key(t):
    return (t[:, 0].to(torch.int64) << 40) | (t[:, 1].to(torch.int64) << 8) | (t[:, 2].to(torch.int64) & 255)
tag = os.environ.get("TAG", "")
for algo in (0, 1):
    A = run((4, 2), algo)
    ka = key(A)
    for shape in shapes:
        bad = []
        for rep in range(4):
            B = run(shape, algo)
            kb = key(B)
            bad.append(int((~torch.isin(kb, ka)).sum()) + int((~torch.isin(ka, kb)).sum()))
        print(f"{tag} algo {algo} shape {shape[0]}x{shape[1]}: differing records per run {bad}", flush=True)
