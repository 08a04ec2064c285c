"""Debug: records of one tile shape against the 4x2 shape on a config-4-like input; prints what differs."""
import sys, os, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np, torch
from panman_b200 import api, shard, trees

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 5_000_000
shapes = [tuple(map(int, s.split("x"))) for s in (sys.argv[3:] or ["4x1", "2x1"])]
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
    return (t[:, 0].to(torch.int64) << 40) | (t[:, 1].to(torch.int64) << 8)

for algo in (0, 1):
    A = run((4, 2), algo)
    A2 = run((4, 2), algo)
    print("algo", algo, "4x2 n", A.shape[0], "repeat equal", bool(A.shape == A2.shape and (A == A2).all()), flush=True)
    ka = key(A)
    for shape in shapes:
        for rep in range(3):
            B = run(shape, algo)
            kb = key(B)
            same = B.shape == A.shape and bool((A == B).all())
            print(f"algo {algo} shape {shape} rep {rep}: n {B.shape[0]} equal {same}", flush=True)
            if same:
                continue
            extra = B[~torch.isin(kb, ka)]
            missing = A[~torch.isin(ka, kb)]
            print("  extra (in shape, not in 4x2):", extra.shape[0], " missing:", missing.shape[0])
            common_b = B[torch.isin(kb, ka)]
            common_a = A[torch.isin(ka, kb)]
            if common_a.shape == common_b.shape:
                print("  common records with different payload:", int((common_a[:, 2] != common_b[:, 2]).sum()))
            for name, t in (("extra", extra), ("missing", missing)):
                t = t[:24].cpu().numpy()
                for r in t:
                    node, pos, pay = int(r[0]), int(r[1]), int(r[2])
                    leaf = int(tree.leaf_row[node])
                    print(f"   {name} node {node} (leaf_row {leaf}, parent {int(tree.parent[node])}) pos {pos} group {pos // 32} tile128 {pos // 32 // 128} "
                          f"grp_in_tile {(pos // 32) % 128} col_in_group {pos % 32} type {pay & 255} code {(pay >> 8) & 255} pcode {(pay >> 16) & 255}")
            if extra.shape[0]:
                e = extra.cpu().numpy()
                print("  extra: distinct nodes", len(set(e[:, 0].tolist())), "distinct groups", len(set((e[:, 1] // 32).tolist())),
                      "leaf share", float((tree.leaf_row[e[:, 0]] >= 0).mean()))
