"""Minimal driver for ncu captures and wait profiles: one BASELINE shape, a few device-resident passes, nothing else.
usage: python profiles/prof_cfg.py <c2|c2poly|c3|c4|c5> <fitch|sankoff> [columns] [taxa] [steps]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

from panman_b200 import api, trees

name = sys.argv[1]
algo = api.SANKOFF if sys.argv[2] == "sankoff" else api.FITCH
SARS = api.SynthSpec()
CFG = {"c2": (lambda n: trees.random_binary_tree(n), 20000, 30000, SARS),
       "c2poly": (lambda n: trees.collapse_edges(trees.random_binary_tree(n), 0.6), 20000, 30000, SARS),
       "c3": (lambda n: trees.random_binary_tree(n), 8_000_000, 2048, SARS),
       "c4": (lambda n: trees.random_binary_tree(n), 1000, 5_000_000,
              api.SynthSpec(p_sub=2e-3, p_del=5e-3, p_ins=5e-3, p_n_block=1e-3, p_amb=1e-4, accessory_frac=0.35)),
       "c5": (lambda n: trees.caterpillar_tree(n), 100_000, 100_000,
              api.SynthSpec(p_sub=1e-4, p_del=2e-6, p_ins=1e-6, p_n_block=1e-2, p_amb=1e-4))}
build, N, L, spec = CFG[name]
if len(sys.argv) > 3 and int(sys.argv[3]) > 0:
    L = int(sys.argv[3])
if len(sys.argv) > 4 and int(sys.argv[4]) > 0:
    N = int(sys.argv[4])
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
tree = build(N)
ctx = api.Context(tree)
stride = ((L + 31) // 32) * 16
d_msa = torch.zeros((N, stride), dtype=torch.uint8, device="cuda")
d_cons = torch.zeros(L, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
ctx.synth_msa(spec, 0, L, d_msa.data_ptr(), stride, d_cons.data_ptr())
for _ in range(steps):
    _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
print("records", n, ctx.timing(), ctx.plan_info())
