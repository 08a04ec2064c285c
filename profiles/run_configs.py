"""Runs the remaining BASELINE.json configs once each on one B200 (C4 bacterial-shaped, C5 ladder, a column slice of C3)
for both algorithms: device-side synthetic MSA, full pipeline timing (pmfs_run_device, CUDA events inside the library)
and a bit-exactness check against the plain-C oracle on a column subsample.  One JSON line per (config, algo).
usage: python profiles/run_configs.py [c4] [c5] [c3]"""
import json
import os
import sys
import time

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import numpy as np
import torch

from oracle import oracle_api as O
from panman_b200 import api, trees

CONFIGS = {
    # name: (tree builder, n_leaves, n_cols, spec, oracle columns)
    "c4": (lambda: trees.random_binary_tree(1000), 1000, 5_000_000,
           api.SynthSpec(p_sub=2e-3, p_del=5e-3, p_ins=5e-3, p_n_block=1e-3, p_amb=1e-4, accessory_frac=0.35), 4096),
    "c5": (lambda: trees.caterpillar_tree(100_000), 100_000, 100_000,
           api.SynthSpec(p_sub=1e-4, p_del=2e-6, p_ins=1e-6, p_n_block=1e-2, p_amb=1e-4), 64),
    "c3": (lambda: trees.random_binary_tree(8_000_000), 8_000_000, 2048,
           api.SynthSpec(), 8),
}


def run(name):
    build, N, L, spec, ncheck = CONFIGS[name]
    t0 = time.time()
    tree = build()
    t_tree = time.time() - t0
    V = tree.n_nodes
    t0 = time.time()
    ctx = api.Context(tree)
    t_plan = time.time() - t0
    stride = ((L + 31) // 32) * 16
    d_msa = torch.zeros((N, stride), dtype=torch.uint8, device="cuda")
    d_cons = torch.zeros(L, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    t0 = time.time()
    ctx.synth_msa(spec, 0, L, d_msa.data_ptr(), stride, d_cons.data_ptr())
    t_synth = time.time() - t0
    cols = np.sort(np.random.default_rng(1).choice(L, size=min(ncheck, L), replace=False))
    byte_cols = torch.from_numpy(cols // 2).cuda()
    sub_bytes = d_msa[:, byte_cols].cpu().numpy()
    sub = np.where(cols % 2 == 0, sub_bytes & 15, sub_bytes >> 4).astype(np.uint8)
    cons = d_cons.cpu().numpy()
    for algo_name, algo in (("fitch", api.FITCH), ("sankoff", api.SANKOFF)):
        best = None
        for it in range(3):
            _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
            t = ctx.timing()
            if best is None or t["wall_ms"] < best["wall_ms"]:
                best = t
        rec = ctx.download(n)
        t0 = time.time()
        want = O.run_port(algo, tree, sub, cons[cols], None, threads=os.cpu_count() or 1)
        t_cpu = time.time() - t0
        sel = np.isin(rec["pos"], cols)
        mine = rec[sel].copy()
        mine["pos"] = np.searchsorted(cols, mine["pos"])
        mine = mine[np.lexsort((mine["pos"], mine["node"]))]
        exact = mine.shape == want.shape and all((mine[k] == want[k]).all() for k in ("node", "pos", "type", "code"))
        key = rec["node"].astype(np.int64) << 32 | rec["pos"].astype(np.int64)
        line = {"config": name, "algo": algo_name, "taxa": N, "nodes": V, "columns": L, "records": int(n),
                "updates_per_s": V * L / (best["wall_ms"] * 1e-3), "ms": {k: best[k] for k in ("forward_ms", "backward_ms", "sort_ms", "wall_ms")},
                "plan": ctx.plan_info(), "bit_exact_on_sample": bool(exact), "sample_columns": int(cols.size),
                "sorted_unique": bool((np.diff(key) > 0).all()),
                "cpu_port_updates_per_s": V * cols.size / t_cpu, "cpu_threads": os.cpu_count(),
                "setup_s": {"tree": t_tree, "plan": t_plan, "synth": t_synth}}
        print(json.dumps(line), flush=True)
        del rec, key
    ctx.close()


if __name__ == "__main__":
    for name in (sys.argv[1:] or ["c4", "c5", "c3"]):
        run(name)
        torch.cuda.empty_cache()
