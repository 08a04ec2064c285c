#!/usr/bin/env python
"""bench.py -- column.node updates/s of the Fitch / Sankoff pass on BASELINE config 2, one JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--algo fitch|sankoff] [--impl ours|reference]

A step = one full pass of the hot path (bottom-up sets, top-down assignment, mutation emission, ordering by (node,pos))
over one synthetic SARS-CoV-2-shaped alignment of 20 000 taxa x 30 000 columns on a random binary tree (39 999 nodes),
already resident in HBM.  With N GPUs every rank owns its own 30 000-column shard of a 20 000 x (30 000*N) alignment
(weak scaling; columns are independent: every rank runs the whole step on its shard and keeps its (node,pos)-ordered
records with global positions -- no data-path collective; a node's complete list is the concatenation of the ranks'
slices, panman_b200/shard.py stitch_by_node).  The variant in which an NCCL all-to-all leaves every rank with the
complete lists of a node range ON THE DEVICE is timed after the main loop and reported as config.gathered_by_node.
`value` counts all ranks' updates over the max-over-ranks device time; `e2e` is the same metric through the
host-buffer C-ABI call (pmfs_run: H2D of the nibble-packed MSA and D2H of the records inside the timed region).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# NCCL_DEBUG=VERSION makes NCCL print its version banner on STDOUT, next to the one JSON line the driver parses
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
    del os.environ["NCCL_DEBUG"]

N_TAXA = 20000
N_COLS = 30000
B_CONTRACT = {"fitch": 4.25, "sankoff": 6.25}  # level-sweep bytes per column.node update


def ncu_traffic(algo):
    """DRAM bytes per step of fwd_kernel + bwd_kernel from the committed `ncu --set full` capture (profiles/)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic_r1.json")))
        return float(t[algo]["traffic_bytes_per_step"])
    except Exception:
        return None


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=lambda: [self.rows.append(l) for l in self.proc.stdout], daemon=True).start()
        except Exception:
            self.proc = None

    def mark(self):
        """Samples taken before this call (warm-up) are dropped, unless nothing else arrives."""
        self.skip = len(self.rows)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        rows = self.rows[getattr(self, "skip", 0):] or self.rows
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_sample(tree, host_codes, cons, algo, threads):
    """Times the reference's CPU path on a column sample.  oracle/_ref (the reference's own fitchSankoff.cpp) when its
    prebuilt library is present, else the plain-C port."""
    from oracle import oracle_api as O
    a = 0 if algo == "fitch" else 1
    if O.ref_available() and os.path.exists(O.REF_SO):
        kind = "reference"
        chars, cchars = O.codes_to_chars(host_codes), O.codes_to_chars(cons)
        t0 = time.perf_counter()
        rec = O.run_ref_columns(a, tree, chars, cchars, None, threads=threads)
        dt = time.perf_counter() - t0
    else:
        kind = "port"
        t0 = time.perf_counter()
        rec = O.run_port(a, tree, host_codes, cons, None, threads=threads)
        dt = time.perf_counter() - t0
    return kind, dt, rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--algo", choices=["fitch", "sankoff"], default="fitch")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--cpu-cols", type=int, default=0, help="columns of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--chunk-nodes", type=int, default=0, help="plan tuning: nodes per chunk (0 = automatic)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (tuning sweeps)")
    ap.add_argument("--no-exchange", action="store_true", help="N > 1: skip the extra gathered-by-node leg")
    ap.add_argument("--p-n-block", type=float, default=None, help="experiment: override the N-block rate of the synthetic MSA")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0

    import numpy as np
    import torch

    from panman_b200 import api, trees

    algo = api.FITCH if args.algo == "fitch" else api.SANKOFF
    tree = trees.random_binary_tree(N_TAXA)
    V = tree.n_nodes
    ncpu = os.cpu_count() or 1
    workload = (f"c2: synthetic SARS-CoV-2-shaped MSA, {N_TAXA} taxa x {N_COLS} columns per GPU, "
                f"random binary tree ({V} nodes), algo={args.algo}")
    hbm_peak, peak_src = peaks()

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (the synthetic MSA is generated on the GPU)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    ctx = api.Context(tree, device=local_rank)
    if args.chunk_nodes:
        ctx.set_tuning(chunk_nodes=args.chunk_nodes)
    stride = ((N_COLS + 31) // 32) * 16
    d_msa = torch.zeros((N_TAXA, stride), dtype=torch.uint8, device=dev)
    d_cons = torch.zeros(N_COLS, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    spec = api.SynthSpec()
    if args.p_n_block is not None:
        spec.p_n_block = args.p_n_block
    ctx.synth_msa(spec, rank * N_COLS, N_COLS, d_msa.data_ptr(), stride, d_cons.data_ptr())

    # ------------------------------------------------------------------ reference arm: the CPU path on the host cores
    if args.impl == "reference":
        cols_per_step = args.cpu_cols or max(64, min(2048, 32 * ncpu))
        rng = np.random.default_rng(0)
        cons_h = d_cons.cpu().numpy()
        times = []
        full = api.unpack_nibbles(d_msa.cpu().numpy(), N_COLS)
        for it in range(args.warmup + args.steps):
            cols = np.sort(rng.choice(N_COLS, size=cols_per_step, replace=False))
            sub = full[:, cols]
            kind, dt, _ = cpu_reference_sample(tree, np.ascontiguousarray(sub), cons_h[cols], args.algo, ncpu)
            if it >= args.warmup:
                times.append(dt)
        total = sum(times)
        val = V * cols_per_step * len(times) / total
        line = {"metric": "column.node updates/sec", "value": val, "unit": "updates/s", "impl": "reference",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16 bitsets / int32 costs",
                "data": "synthetic", "config": {"workload": workload},
                "cpu_baseline": {"value": val, "unit": "updates/s", "cores": ncpu, "kind": kind,
                                 "sample": f"{cols_per_step} random columns of the workload per step, all {V} nodes, "
                                           f"{ncpu} threads over columns, g++ -O2"},
                "e2e": {"value": val, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------ our arm
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    if world > 1:
        ctx.set_shard(rank * N_COLS, world * N_COLS, None)  # a column shard that keeps its output: global positions

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from panman_b200 import shard

    def step(exchange=False):
        ptr, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N_TAXA, N_COLS, d_cons.data_ptr())
        if exchange:  # every rank ends up owning the complete, ordered lists of its node range
            shard.exchange_parts(ctx, rank, world, dev)
        return n

    sampler = ClockSampler(local_rank)
    sampler.start()  # started before the warm-up so that nvidia-smi is already sampling when the timed region begins
    for _ in range(args.warmup):
        n_rec = step()
    barrier()
    sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pass_ms, launches = 0.0, 0
    e0.record()
    for _ in range(args.steps):
        n_rec = step()
        t = ctx.timing()
        pass_ms += t["forward_ms"] + t["backward_ms"]
        launches += t["n_launches"]
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1), pass_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms, pass_ms = float(ms[0]), float(ms[1])
    updates_per_step = V * N_COLS * world
    value = updates_per_step * args.steps / (total_ms * 1e-3)
    tinfo = ctx.timing()
    plan = ctx.plan_info()

    # the same step followed by the all-to-all by node range (lists complete on the device of the owning rank)
    gathered = None
    if world > 1 and not args.no_exchange:
        ctx.set_shard(rank * N_COLS, world * N_COLS, [b for b, _ in shard.node_ranges(V, world)] + [V])
        xsteps = max(1, min(args.steps, 50))
        for _ in range(3):
            step(True)
        barrier()
        x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        x0.record()
        for _ in range(xsteps):
            step(True)
        x1.record()
        barrier()
        xms = torch.tensor([x0.elapsed_time(x1)], device=dev, dtype=torch.float64)
        dist.all_reduce(xms, op=dist.ReduceOp.MAX)
        gathered = {"ms_per_step": float(xms[0]) / xsteps, "value": updates_per_step * xsteps / (float(xms[0]) * 1e-3),
                    "steps": xsteps, "what": "step + NCCL all-to-all of the 8-byte records by node range + owner-side ordering"}
        ctx.set_shard(0, 0, None)

    # end to end: pinned host buffers through pmfs_run (H2D + passes + sort + D2H of the records), every step
    h_msa = torch.empty((N_TAXA, stride), dtype=torch.uint8, pin_memory=True)
    h_msa.copy_(d_msa)
    h_cons = torch.empty(N_COLS, dtype=torch.uint8, pin_memory=True)
    h_cons.copy_(d_cons)
    torch.cuda.synchronize()
    h_msa_np, h_cons_np = h_msa.numpy(), h_cons.numpy()
    ctx.set_stream(0)
    if world > 1:
        ctx.set_shard(rank * N_COLS, world * N_COLS, None)
    for _ in range(max(1, min(args.warmup, 2))):
        rec = ctx.run(algo, h_msa_np, N_COLS, h_cons_np, copy=False)
    e2e_steps = 0 if args.no_e2e else max(1, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        rec = ctx.run(algo, h_msa_np, N_COLS, h_cons_np, copy=False)  # view of the context's pinned result buffer
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_val = updates_per_step * e2e_steps / float(e2e_s[0]) if e2e_steps else None
    e2e_phases = ctx.timing()
    h2d = int(N_TAXA * stride + N_COLS)
    d2h = int(rec.shape[0] * 12)

    if rank == 0:
        b = B_CONTRACT[args.algo]
        per_step_pass_s = pass_ms * 1e-3 / args.steps
        achieved = V * N_COLS * b / per_step_pass_s / 1e9  # per GPU
        design_bytes = tinfo["forward_bytes"] + tinfo["backward_bytes"]
        line = {
            "metric": "column.node updates/sec", "value": value, "unit": "updates/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 bit-planes (16-state sets, 4-bit codes)", "data": "synthetic",
            "config": {"workload": workload, "l2": "inputs larger than L2 (MSA 300 MB + set array 1.2 GB touched per step)",
                       "parallelism": (f"column shards x{world}, no data-path collective: every rank keeps the (node,pos)-ordered "
                                       "records of its own columns (global positions); a node's list is the concatenation "
                                       "over ranks") if world > 1 else "1 GPU",
                       "gathered_by_node": gathered,
                       "records_per_step": int(n_rec), "plan": plan},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "traffic": ncu_traffic(args.algo), "peak_source": peak_src,
                         "traffic_note": "dram__bytes_read+write of fwd_kernel+bwd_kernel per step, ncu --set full capture of "
                                         "this workload (profiles/traffic_r1.json); algorithmic bytes at the contract "
                                         "figure: %.3e" % (V * N_COLS * b),
                         "kernel": "fwd_kernel + bwd_kernel (all chunk levels of one step)",
                         "bytes_per_update_contract": b, "pass_ms_per_step": 1e3 * per_step_pass_s,
                         "frac_whole_step": V * N_COLS * b / (total_ms * 1e-3 / args.steps) / 1e9 / hbm_peak / world,
                         "bytes_per_update_design": design_bytes / (V * N_COLS),
                         "achieved_design_gbs": design_bytes / per_step_pass_s / 1e9},
            "e2e": {"value": e2e_val, "unit": "updates/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "phases_ms_last_step": {k: e2e_phases[k] for k in
                                                                ("h2d_ms", "forward_ms", "backward_ms", "sort_ms", "d2h_ms")}},
            "gpu_launches": int(launches), "clocks": clocks,
            "phases_ms_last_step": {k: tinfo[k] for k in ("forward_ms", "backward_ms", "sort_ms", "wall_ms")},
        }
        if world == 1 and not args.no_cpu_baseline:
            cols_n = args.cpu_cols or 8192  # about 10 s (Fitch) to 30 s (Sankoff) of CPU work on 16 cores
            cols = np.sort(np.random.default_rng(0).choice(N_COLS, size=cols_n, replace=False))
            full = api.unpack_nibbles(h_msa_np, N_COLS)
            sub = np.ascontiguousarray(full[:, cols])
            kind, dt, want = cpu_reference_sample(tree, sub, h_cons_np[cols], args.algo, ncpu)
            # the sample doubles as a parity check of the benchmarked run
            sel = np.isin(rec["pos"], cols)
            mine = rec[sel].copy()
            mine["pos"] = np.searchsorted(cols, mine["pos"])
            mine = mine[np.lexsort((mine["pos"], mine["node"]))]
            exact = mine.shape == want.shape and all((mine[k] == want[k]).all() for k in ("node", "pos", "type", "code"))
            line["cpu_baseline"] = {"value": V * cols_n / dt, "unit": "updates/s", "cores": ncpu, "kind": kind,
                                    "sample": f"{cols_n} random columns of the workload, all {V} nodes, {ncpu} threads over "
                                              f"columns, g++ -O2, {dt:.2f} s", "bit_exact_on_sample": bool(exact)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
