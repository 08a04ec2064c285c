"""Debug: a small run of one tile shape for compute-sanitizer (racecheck / memcheck)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np
import synth
from oracle import oracle_api as O
from panman_b200 import api, trees
nw, ng = map(int, sys.argv[1].split("x"))
algo = int(sys.argv[2])
tree = trees.random_binary_tree(300, seed=3)
rng = np.random.default_rng(1)
L = 4096 * 3
msa = synth.simulate_msa(tree, L, rng, p_sub=0.02, p_del=0.05, p_ins=0.05, p_n=0.01, p_amb=0.01, root_gap_frac=0.35)
cons = synth.consensus_first_nongap(msa)
with api.Context(tree) as ctx:
    ctx.set_tile(nw, ng)
    ctx.set_watchdog(0)
    got = ctx.run(algo, api.pack_nibbles(msa), L, cons)
want = O.run_port(algo, tree, msa, cons, None, threads=4)
print("records", got.shape[0], "equal", got.shape == want.shape and all((got[k] == want[k]).all() for k in ("node", "pos", "type", "code")))
