"""Column-tile shape sweep: every built shape of the walk kernels (walking warps x column groups per thread) on the
BASELINE shapes, device-resident inputs, timings from CUDA events inside the library.  One JSON line per
(config, algo, shape); `auto` is what the library picks on its own.
usage: python profiles/sweep_shapes.py [c2] [c3] [c4] [c5] [c2poly]"""
import json
import os
import sys
import time

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import torch

from panman_b200 import api, trees

SHAPES = [(0, 0), (4, 2), (2, 2), (4, 1), (1, 2), (2, 1), (1, 1)]
SARS = api.SynthSpec()
CONFIGS = {
    # name: (tree builder, n_leaves, [column counts], spec, algos)
    "c2": (lambda: trees.random_binary_tree(20000), 20000, [30000, 3750], SARS, ("fitch", "sankoff")),
    "c2poly": (lambda: trees.collapse_edges(trees.random_binary_tree(20000), 0.6), 20000, [30000], SARS, ("fitch", "sankoff")),
    "c4": (lambda: trees.random_binary_tree(1000), 1000, [5_000_000],
           api.SynthSpec(p_sub=2e-3, p_del=5e-3, p_ins=5e-3, p_n_block=1e-3, p_amb=1e-4, accessory_frac=0.35), ("fitch", "sankoff")),
    "c5": (lambda: trees.caterpillar_tree(100_000), 100_000, [100_000],
           api.SynthSpec(p_sub=1e-4, p_del=2e-6, p_ins=1e-6, p_n_block=1e-2, p_amb=1e-4), ("fitch", "sankoff")),
    # config 3: the widths a run meets -- 2 048 columns (memory-limited batch on one GPU), 3 776 = 30 000 / 8 GPUs
    "c3": (lambda: trees.random_binary_tree(8_000_000), 8_000_000, [2048, 3776, 1024], SARS, ("fitch", "sankoff")),
}


def run(name):
    build, N, widths, spec, algos = CONFIGS[name]
    t0 = time.time()
    tree = build()
    V = tree.n_nodes
    ctx = api.Context(tree)
    if name == "c3":
        ctx.set_tuning(chunk_nodes=1126)  # (what the library picks for one 256-group tile; pinned: a re-plan of 16 M nodes takes 5 s)
    setup = time.time() - t0
    for L in widths:
        stride = ((L + 31) // 32) * 16
        d_msa = torch.zeros((N, stride), dtype=torch.uint8, device="cuda")
        d_cons = torch.zeros(L, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        ctx.synth_msa(spec, 0, L, d_msa.data_ptr(), stride, d_cons.data_ptr())
        for algo_name in algos:
            algo = api.FITCH if algo_name == "fitch" else api.SANKOFF
            if name == "c3" and algo_name == "sankoff" and L > 2048:
                continue  # S + D1 of 16 M nodes: 1 GB per column group
            for nw, ng in SHAPES:
                ctx.set_tile(nw, ng)
                best = None
                try:
                    for it in range(4):
                        _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
                        t = ctx.timing()
                        if it and (best is None or t["wall_ms"] < best["wall_ms"]):
                            best = t
                except api.PmfsError as e:
                    print(json.dumps({"config": name, "algo": algo_name, "columns": L, "shape": f"{nw}x{ng}", "error": str(e)}), flush=True)
                    continue
                design = best["forward_bytes"] + best["backward_bytes"]
                line = {"config": name, "algo": algo_name, "nodes": V, "columns": L, "shape": f"{nw}x{ng}" if nw else "auto",
                        "records": int(n), "updates_per_s": V * L / (best["wall_ms"] * 1e-3),
                        "ms": {k: round(best[k], 4) for k in ("forward_ms", "backward_ms", "sort_ms", "wall_ms")},
                        "design_gbs_passes": design / ((best["forward_ms"] + best["backward_ms"]) * 1e-3) / 1e9,
                        "chunks": ctx.plan_info()["n_chunks"], "setup_s": round(setup, 1)}
                print(json.dumps(line), flush=True)
        del d_msa, d_cons
        torch.cuda.empty_cache()
    ctx.close()


if __name__ == "__main__":
    for name in (sys.argv[1:] or ["c2", "c5", "c4", "c3"]):
        run(name)
        torch.cuda.empty_cache()
