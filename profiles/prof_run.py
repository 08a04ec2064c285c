"""Minimal driver for ncu captures: BASELINE config 2, `--steps` device-resident passes, nothing else.
usage: python profiles/prof_run.py [fitch|sankoff] [steps] [taxa] [columns] [fraction of internal edges collapsed] [split arity]
(split arity: nodes with more children are split into parts, pmfs_set_poly_split; default 128, 0 = never)"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

from panman_b200 import api, trees

algo = api.SANKOFF if len(sys.argv) > 1 and sys.argv[1] == "sankoff" else api.FITCH
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
N = int(sys.argv[3]) if len(sys.argv) > 3 else 20000
L = int(sys.argv[4]) if len(sys.argv) > 4 else 30000
tree = trees.random_binary_tree(N)
if len(sys.argv) > 5 and float(sys.argv[5]) > 0:
    tree = trees.collapse_edges(tree, float(sys.argv[5]))  # polytomies
ctx = api.Context(tree)
if len(sys.argv) > 6:
    ctx.set_poly_split(int(sys.argv[6]), 0)
stride = ((L + 31) // 32) * 16
d_msa = torch.zeros((N, stride), dtype=torch.uint8, device="cuda")
d_cons = torch.zeros(L, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
ctx.synth_msa(api.SynthSpec(), 0, L, d_msa.data_ptr(), stride, d_cons.data_ptr())
for _ in range(steps):
    _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
print("records", n, ctx.timing(), ctx.plan_info())
