#!/usr/bin/env python
"""bench.py -- column.node updates/s of the Fitch / Sankoff pass, one JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--algo fitch|sankoff] [--impl ours|reference] [--workload c3|c2]

Workload (default, every N): BASELINE config 3 -- the 8 000 000-taxon x 30 000-column SARS-CoV-2-shaped alignment
(random binary tree, 15 999 999 nodes), COLUMN-SHARDED over the N GPUs, strong scaling: the same 30 000 columns whatever
N.  A step = one full pass of the hot path over all 30 000 columns: bottom-up sets, top-down assignment, mutation
emission, ordering by (node,pos) and -- for N > 1 -- the gather of the per-node lists by node range over NVLink
(pmfs_comm_run: every rank ends up owning the complete (node,pos)-ordered lists of its node range, in HBM), all inside
the timed region.  The alignment is resident in HBM when the timed region starts (120 GB on one GPU, 15 GB per GPU on
eight).  The per-node set arrays of 8 M internal nodes take 0.5 GB per 32 columns, so the columns go through in
ascending windows (one window = N x the columns that fit per GPU); each window's result is complete and ordered, and a
node's list is the concatenation over windows -- the shape of the reference's own --low-mem-mode batch loop
(src/panman.cpp:1503-1514).  `--workload c2` is BASELINE config 2 (20 000 taxa x 30 000 columns per GPU); on one GPU it
is also measured next to the main workload and reported as `c2_single_gpu`.

`value` counts all ranks' updates over the max-over-ranks device time; `e2e` is the same metric through host buffers
(N = 1: pmfs_run; N > 1: H2D of the rank's columns + pmfs_comm_run + D2H of the rank's node range) on a column batch of
the workload; `cpu_baseline` / `--impl reference` time the reference's own fitchSankoff.cpp (oracle/_ref) on the host
cores on a column sample of the same workload.
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

B_CONTRACT = {"fitch": 4.25, "sankoff": 6.25}  # level-sweep bytes per column.node update
WORKLOADS = {
    # name: (taxa, columns, tree, description)
    "c3": (8_000_000, 30_000, "random binary tree", "c3: synthetic SARS-CoV-2-shaped MSA, 8000000 taxa x 30000 columns"),
    "c2": (20_000, 30_000, "random binary tree", "c2: synthetic SARS-CoV-2-shaped MSA, 20000 taxa x 30000 columns per GPU"),
}


def ncu_traffic(workload, algo):
    """DRAM bytes of fwd_kernel + bwd_kernel from the committed `ncu --set full` capture (profiles/traffic_r2.json),
    with the commit and tile shape it was taken at; None when there is no capture for this workload."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic_r2.json")))
        return t[{"c3": "c3_8m_taxa_2048col_batch"}.get(workload, workload)][algo]
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
                "reasons": sorted(reasons), "samples": len(sm), "sm_min_mhz": sm[0] if sm else None}


def ref_threads(n_nodes, ncpu):
    """Threads for the reference's CPU code: each holds one column's node-name-keyed hash maps (~200 B per node)."""
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 32 << 30
    return max(1, min(ncpu, int(0.4 * avail / (n_nodes * 200 + 1))))


class CpuReference:
    """The reference's CPU path on column samples: oracle/_ref (the reference's own fitchSankoff.cpp) when its prebuilt
    library is present, else the plain-C port.  The Node graph is built once; only the column loop is timed."""

    def __init__(self, tree, ncpu):
        from oracle import oracle_api as O
        self.O, self.tree = O, tree
        self.threads = ref_threads(tree.n_nodes, ncpu)
        self.kind = "reference" if (O.ref_available() and os.path.exists(O.REF_SO)) else "port"
        self.graph = O.RefGraph(tree) if self.kind == "reference" else None

    def run(self, algo_name, codes, cons):
        """-> (records, seconds of the column loop)"""
        O, a = self.O, 0 if algo_name == "fitch" else 1
        if self.graph is not None:
            rec = self.graph.columns(a, O.codes_to_chars(codes), O.codes_to_chars(cons), None, threads=self.threads)
            return rec, O.last_ref_loop_seconds()
        t0 = time.perf_counter()
        rec = O.run_port(a, self.tree, codes, cons, None, threads=self.threads)
        return rec, time.perf_counter() - t0


def consensus_of(codes):
    """First non-gap code per column in row order (the rule the device generator applies, src/panman.cpp:1326-1340)."""
    import numpy as np
    nz = codes != 0
    return codes[nz.argmax(0), np.arange(codes.shape[1])]


def subtree_sample(tree, max_nodes):
    """A subtree of at most max_nodes nodes (ids are pre-order, so a subtree is the id range [v, v + size)), as a tree of
    its own with its leaves' MSA rows renumbered -> (FlatTree, rows of the full alignment its leaves use)."""
    import numpy as np
    from panman_b200.trees import FlatTree
    V = tree.n_nodes
    size = np.ones(V, np.int64)
    par = tree.parent
    for v in range(V - 1, 0, -1):
        size[par[v]] += size[v]
    cand = np.nonzero((size <= max_nodes) & (size >= max_nodes // 4) & (tree.leaf_row == -1))[0]
    v = int(cand[np.argmax(size[cand])]) if cand.size else 0
    n = int(size[v])
    ids = np.arange(v, v + n)
    parent = (par[ids] - v).astype(np.int32)
    parent[0] = -1
    off = tree.child_off[v:v + n + 1].astype(np.int64)
    child_idx = (tree.child_idx[off[0]:off[-1]] - v).astype(np.int32)
    child_off = (off - off[0]).astype(np.int32)
    rows_full = tree.leaf_row[ids].copy()
    used = rows_full[rows_full >= 0]
    remap = -np.ones(int(used.max()) + 1 if used.size else 1, np.int64)
    remap[used] = np.arange(used.size)
    leaf_row = np.where(rows_full >= 0, remap[np.maximum(rows_full, 0)], rows_full).astype(np.int32)
    return FlatTree(parent, child_off, child_idx, leaf_row), used


def reference_arm(args, tree, spec, workload, n_taxa, n_cols, ncpu):
    """bench.py --impl reference: the reference's CPU code alone -- no GPU, no product library.  Input: the CPU twin of
    the synthetic alignment generator (oracle/synth_twin.c), a fresh random column sample per step.  On the 16 M-node tree
    one column costs the reference's node-name-keyed hash maps about half a minute of one core, so a step of K-step runs
    would take hours: there the sample of column.node updates is (random columns) x (one subtree of ~1 M nodes) -- the
    reference's own code on a real part of the workload; its smaller hash maps make it, if anything, faster per update
    than on the whole tree (the cpu_baseline of the main arm times whole columns once)."""
    import numpy as np
    from oracle import oracle_api as O
    V_full = tree.n_nodes
    sub, sub_rows = (tree, None) if V_full <= 2_000_000 else subtree_sample(tree, 1_000_000)
    V = sub.n_nodes
    ref = CpuReference(sub, ncpu)
    cols_per_step = args.cpu_cols or max(ref.threads, min(2048, int(4e7 / V)))  # a few seconds of CPU work per step
    rng = np.random.default_rng(0)
    times = []
    for it in range(args.warmup + args.steps):
        cols = np.sort(rng.choice(n_cols, size=cols_per_step, replace=False))
        codes = O.synth_columns(tree, spec, cols, n_taxa, threads=ncpu)
        if sub_rows is not None:
            codes = np.ascontiguousarray(codes[sub_rows])
        _, dt = ref.run(args.algo, codes, consensus_of(codes))
        if it >= args.warmup:
            times.append(dt)
    total = sum(times)
    val = V * cols_per_step * len(times) / total
    what = (f"{cols_per_step} random columns of the workload per step x all {V} nodes" if sub_rows is None else
            f"{cols_per_step} random columns of the workload per step x one subtree of {V} of the {V_full} nodes")
    line = {"metric": "column.node updates/sec", "value": val, "unit": "updates/s", "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "strong" if args.workload == "c3" else "weak", "vs_baseline": None,
            "dtype": "u16 bitsets / int32 costs", "data": "synthetic", "config": {"workload": workload},
            "cpu_baseline": {"value": val, "unit": "updates/s", "cores": ref.threads, "kind": ref.kind,
                             "sample": what + f", {ref.threads} threads over columns, g++ -O2; the Node graph is built once, the "
                                              "column loop alone is timed"},
            "e2e": {"value": val, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--algo", choices=["fitch", "sankoff"], default="fitch")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=["c3", "c2"], default="c3")
    ap.add_argument("--cpu-cols", type=int, default=0, help="columns of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (tuning sweeps)")
    ap.add_argument("--no-c2", action="store_true", help="N = 1: skip the extra config-2 measurement")
    ap.add_argument("--chunk-nodes", type=int, default=0, help="plan tuning: nodes per chunk (0 = automatic)")
    ap.add_argument("--window-cols", type=int, default=0, help="columns per GPU and window (0 = as many as fit)")
    ap.add_argument("--taxa", type=int, default=0, help="experiment: override the number of taxa")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0

    import numpy as np

    from panman_b200 import trees
    n_taxa, n_cols, _, wl_text = WORKLOADS[args.workload]
    if args.taxa:
        n_taxa = args.taxa
        wl_text = wl_text.replace(str(WORKLOADS[args.workload][0]), str(n_taxa))
    ncpu = os.cpu_count() or 1
    t_setup = time.time()
    tree = trees.random_binary_tree(n_taxa)
    V = tree.n_nodes
    strong = args.workload == "c3"
    workload = f"{wl_text}, random binary tree ({V} nodes), algo={args.algo}" + (f", column-sharded x{world}" if strong else "")

    if args.impl == "reference":
        from panman_b200.api import SynthSpec  # (a dataclass: the product library is NOT loaded by this arm)
        reference_arm(args, tree, SynthSpec(), workload, n_taxa, n_cols, ncpu)
        return 0

    import torch

    from panman_b200 import api, shard
    algo = api.FITCH if args.algo == "fitch" else api.SANKOFF
    hbm_peak, peak_src = peaks()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor(x, device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t]

    ctx = api.Context(tree, device=local_rank)
    if args.chunk_nodes:
        ctx.set_tuning(chunk_nodes=args.chunk_nodes)
    # one stream for everything that is timed: the library's kernels, its NCCL exchange and torch's events and copies
    bench_stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(bench_stream)
    ctx.set_stream(bench_stream.cuda_stream)
    comm = None
    if world > 1 and strong:
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(api.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        comm = api.Comm(ctx, idt.cpu().numpy().tobytes(), rank, world)
    spec = api.SynthSpec()

    # ---- column layout.  c3: the G column groups go through in windows; in window w rank r owns the groups
    # [w*world*gb + r*gb, +gb): its local alignment is the concatenation of its shares, window after window.
    # c2: every rank has its own 30 000 columns of a 20 000 x (30 000*world) alignment (weak scaling, no gather).
    G = (n_cols + 31) // 32
    if strong:
        share = (G + world - 1) // world                   # groups per rank if everything went in one window
        local_groups_max = share
    else:
        share = local_groups_max = G
    stride = local_groups_max * 16
    # (allocate the alignment first: what fits per window is measured against what is free afterwards)
    d_msa = torch.zeros((n_taxa, stride), dtype=torch.uint8, device=dev)
    d_cons = torch.zeros(local_groups_max * 32, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    if not strong:
        gb, n_win = G, 1
    else:
        if args.window_cols:
            gb = max(1, min(share, (args.window_cols + 31) // 32))
        elif comm is not None:
            gb = max(1, (comm.batch_cols(algo, share * 32) + 31) // 32)
        else:
            # one GPU: the widest power-of-two batch whose set arrays (0.5 GB per column group on the 8 M-taxon tree, twice
            # that for Sankoff) take at most 0.62 of what the resident alignment leaves free -- the record buffers need
            # the rest.  Config 3: 64 groups (2 048 columns) for Fitch, 32 for Sankoff.
            free_b = torch.cuda.mem_get_info(dev)[0]
            per_group = (V // 2) * 64 * (2 if args.algo == "sankoff" else 1)
            gb = 128
            while gb > 1 and gb * per_group > 0.62 * free_b:
                gb //= 2
            gb = min(share, gb)
        n_win = (share + gb - 1) // gb

    def window(w):
        """(local group offset, global first column, columns) of this rank's share of window w.  Windows are ascending
        column ranges; inside a window the ranks' shares are ascending with the rank; every rank's shares, window after
        window, tile its local alignment [0, share) (the last window is narrower)."""
        if not strong:
            return 0, rank * n_cols, n_cols
        width = gb if w < n_win - 1 else share - (n_win - 1) * gb
        g_lo = min(G, w * world * gb + rank * width)
        g_hi = min(G, g_lo + width)
        c0 = min(n_cols, g_lo * 32)
        return w * gb, c0, max(0, min(n_cols, g_hi * 32) - c0)

    # the synthetic alignment, generated on the device window by window (a pure function of (seed, node, column))
    for w in range(n_win):
        lg, c0, nc = window(w)
        for o in range(0, nc, 2048):
            m = min(2048, nc - o)
            ctx.synth_msa(spec, c0 + o, m, d_msa.data_ptr() + lg * 16 + o // 2, stride, d_cons.data_ptr() + lg * 32 + o)
    torch.cuda.synchronize()
    setup_s = time.time() - t_setup

    totals = {"fwd": 0.0, "bwd": 0.0, "sort": 0.0, "xchg": 0.0, "stitch": 0.0, "launches": 0, "design_bytes": 0, "records": 0, "batches": 0}

    def account(t, n):
        totals["fwd"] += t["forward_ms"]; totals["bwd"] += t["backward_ms"]; totals["sort"] += t["sort_ms"]
        totals["xchg"] += t["exchange_ms"]; totals["stitch"] += t["stitch_ms"]; totals["launches"] += t["n_launches"]
        totals["design_bytes"] += t["forward_bytes"] + t["backward_bytes"]
        totals["records"] += n

    def step():
        if not strong:
            _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, n_taxa, n_cols, d_cons.data_ptr())
            account(ctx.timing(), n)
            totals["batches"] += 1
        elif comm is None:
            n, nb = ctx.run_batched(algo, d_msa.data_ptr(), stride, n_taxa, n_cols, d_cons.data_ptr(), batch_cols=gb * 32)
            account(ctx.timing(), n)
            totals["batches"] += nb
        else:
            for w in range(n_win):
                lg, c0, nc = window(w)
                _, n = comm.run(algo, d_msa.data_ptr() + lg * 16, stride, n_taxa, nc, c0, n_cols, d_cons.data_ptr() + lg * 32)
                account(ctx.timing(), n)
                totals["batches"] += 1

    if not strong and world > 1:
        ctx.set_shard(rank * n_cols, world * n_cols, None)  # a column shard that keeps its output: global positions
    sampler = ClockSampler(local_rank)
    sampler.start()  # started before the warm-up so that nvidia-smi is already sampling when the timed region begins
    for _ in range(args.warmup):
        step()
    barrier()
    sampler.mark()
    for k in totals:
        totals[k] = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    mine_ms = [e0.elapsed_time(e1), totals["fwd"] + totals["bwd"], totals["xchg"], totals["stitch"], totals["sort"]]
    total_ms, pass_ms, xchg_ms, stitch_ms, sort_ms = allmax(mine_ms)
    per_rank = None
    if world > 1:  # every rank's own phase times (the maxima above hide how much of `exchange` is waiting for the slowest rank)
        g = torch.zeros((world, 5), device=dev, dtype=torch.float64)
        dist.all_gather_into_tensor(g.view(-1), torch.tensor(mine_ms, device=dev, dtype=torch.float64))
        per_rank = [[round(float(v) / args.steps, 3) for v in row] for row in g.cpu()]
        # every rank samples its own GPU: median / lowest SM clock and a bit mask of the throttle reasons it saw
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        ck = torch.tensor([clocks["sm_mhz"] or 0, clocks.get("sm_min_mhz") or 0,
                           sum(1 << i for i, nm in enumerate(names) if nm in clocks["reasons"])], device=dev, dtype=torch.float64)
        gk = torch.zeros((world, 3), device=dev, dtype=torch.float64)
        dist.all_gather_into_tensor(gk.view(-1), ck)
        clocks["per_rank_median_min_reasons"] = [[float(a), float(b), [nm for i, nm in enumerate(names) if int(c) >> i & 1]]
                                                 for a, b, c in gk.cpu().tolist()]
        clocks["reasons"] = sorted({nm for row in clocks["per_rank_median_min_reasons"] for nm in row[2]})
    updates_per_step = V * n_cols * (1 if strong else world)
    value = updates_per_step * args.steps / (total_ms * 1e-3)
    plan = ctx.plan_info()
    rec_all = torch.tensor([totals["records"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rec_all)
    records_per_step = int(float(rec_all[0]) / args.steps)
    mine = dict(totals)

    # ---- N > 1, untimed: the gathered result of a narrow window against the same columns run unsharded on this GPU,
    # restricted to the rank's node range -- record for record (every rank generates all the window's columns itself)
    gather_ok = None
    if comm is not None:
        try:
            from panman_b200 import shard
            vc = 128
            tot = vc * world
            v_msa = torch.zeros((n_taxa, tot // 2), dtype=torch.uint8, device=dev)
            v_cons = torch.zeros(tot, dtype=torch.uint8, device=dev)
            ctx.synth_msa(spec, 0, tot, v_msa.data_ptr(), tot // 2, v_cons.data_ptr())
            torch.cuda.synchronize()
            ptr, n = comm.run(algo, v_msa.data_ptr() + rank * (vc // 2), tot // 2, n_taxa, vc, rank * vc, tot, v_cons.data_ptr() + rank * vc)
            got = shard.records_as_tensor(ptr, n, dev).clone()
            ptr2, n2 = ctx.run_device(algo, v_msa.data_ptr(), tot // 2, n_taxa, tot, v_cons.data_ptr())
            whole = shard.records_as_tensor(ptr2, n2, dev)
            nb_, ne_ = comm.node_range()
            sel = (whole[:, 0] >= nb_) & (whole[:, 0] < ne_)
            ok = torch.tensor([1 if (n > 0 and torch.equal(got, whole[sel])) else 0], device=dev, dtype=torch.int32)
            del v_msa, got, whole
        except Exception as e:  # (the check must not cost the measurement its line)
            ok = torch.tensor([0], device=dev, dtype=torch.int32)
            print(f"rank {rank}: gather check failed to run: {e}", file=sys.stderr)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        gather_ok = bool(int(ok[0]))

    # ---- end to end through host buffers, on a column batch of the workload
    e2e = None
    if not args.no_e2e:
        lg, c0, nc = window(0)
        e_cols = min(nc, 1024 if n_taxa > 1_000_000 else n_cols)
        e_groups = (e_cols + 31) // 32
        h_msa = torch.empty((n_taxa, e_groups * 16), dtype=torch.uint8, pin_memory=True)
        h_msa.copy_(d_msa[:, lg * 16: lg * 16 + e_groups * 16])
        h_cons = torch.empty(e_cols, dtype=torch.uint8, pin_memory=True)
        h_cons.copy_(d_cons[lg * 32: lg * 32 + e_cols])
        torch.cuda.synchronize()
        e_steps = max(1, min(args.steps, 3 if n_taxa > 1_000_000 else 10))
        if comm is None:
            if world > 1:
                ctx.set_shard(rank * n_cols, world * n_cols, None)
            h_msa_np, h_cons_np = h_msa.numpy(), h_cons.numpy()

            def e_step():
                return ctx.run(algo, h_msa_np, e_cols, h_cons_np, copy=False).shape[0]
        else:
            dd_msa = torch.empty((n_taxa, e_groups * 16), dtype=torch.uint8, device=dev)
            dd_cons = torch.empty(e_cols, dtype=torch.uint8, device=dev)
            h_out = torch.empty(0, dtype=torch.uint8, pin_memory=True)
            e_total = e_cols * world  # every rank brings e_cols columns: one window of the job

            ev_h = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

            def e_step():
                nonlocal h_out
                ev_h[0].record()
                dd_msa.copy_(h_msa, non_blocking=True)
                dd_cons.copy_(h_cons, non_blocking=True)
                ev_h[1].record()
                _, n = comm.run(algo, dd_msa.data_ptr(), e_groups * 16, n_taxa, e_cols, rank * e_cols, e_total, dd_cons.data_ptr())
                if h_out.numel() < n * 12:
                    h_out = torch.empty(int(n * 12 * 1.25) + 1024, dtype=torch.uint8, pin_memory=True)
                ctx.download_into(h_out.data_ptr(), n)
                return n
        for _ in range(2):
            n_e = e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            n_e = e_step()
        torch.cuda.synchronize()
        (e_s,) = allmax([time.perf_counter() - t0])
        e_updates = V * e_cols * world
        e2e = {"value": e_updates * e_steps / e_s, "unit": "updates/s", "h2d_bytes_per_step": int(n_taxa * e_groups * 16 + e_cols),
               "d2h_bytes_per_step": int(n_e * 12), "steps": e_steps,
               "what": (f"host buffers in, host records out on a {e_cols}-column batch per GPU of the workload "
                        + ("(pmfs_run, pinned buffers)" if comm is None else
                           "(H2D of the rank's columns, pmfs_comm_run with the gather by node range, D2H of the rank's node range)")
                        + "; bytes are per rank"),
               "phases_ms_last_step": {k: ctx.timing()[k] for k in ("h2d_ms", "forward_ms", "backward_ms", "sort_ms", "d2h_ms",
                                                                     "exchange_ms", "stitch_ms")}}
        if comm is not None:
            # what the host side delivers with all ranks uploading at once (they share the host's memory and PCIe
            # root complexes): this, not the GPUs, bounds the end-to-end figure at N > 1
            torch.cuda.synchronize()
            gbs = torch.tensor([h_msa.numel() / (ev_h[0].elapsed_time(ev_h[1]) * 1e-3) / 1e9], device=dev, dtype=torch.float64)
            allg = torch.zeros(world, device=dev, dtype=torch.float64)
            dist.all_gather_into_tensor(allg, gbs)
            e2e["h2d_gbs_per_rank_last_step"] = [round(float(x), 1) for x in allg.cpu()]
            e2e["h2d_gbs_all_ranks"] = round(float(allg.sum()), 1)
        del h_msa
        if world > 1 and not strong:
            ctx.set_shard(0, 0, None)

    if rank == 0:
        b = B_CONTRACT[args.algo]
        pass_s = pass_ms * 1e-3 / args.steps                 # fwd + bwd kernel time of one step on the slowest rank
        design = mine["design_bytes"] / args.steps + 8.0 * mine["records"] / args.steps   # this rank's bytes per step
        per_gpu_updates = V * n_cols / (world if strong else 1)
        achieved = design / pass_s / 1e9
        step_s = total_ms * 1e-3 / args.steps
        traffic = ncu_traffic(args.workload, args.algo)
        line = {
            "metric": "column.node updates/sec", "value": value, "unit": "updates/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "u32 bit-planes (16-state sets, 4-bit codes)", "data": "synthetic",
            "config": {"workload": workload,
                       "l2": "inputs larger than L2 (alignment and set arrays are tens of GB per GPU)" if strong else
                             "inputs larger than L2 (MSA 300 MB + set array 1.2 GB touched per step)",
                       "parallelism": (f"column shards x{world}; per window every rank runs the whole tree over its columns, then the "
                                       f"per-node lists are gathered by node range (pmfs_comm_run, transport: {comm.transport}) inside the timed step"
                                       if comm is not None else
                                       (f"column shards x{world}, no data-path collective (weak scaling)" if world > 1 else "1 GPU")),
                       "windows_per_step": mine["batches"] // args.steps, "columns_per_gpu_and_window": gb * 32,
                       "records_per_step": records_per_step, "plan": plan, "setup_s": round(setup_s, 1)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "kernel": "fwd_kernel + bwd_kernel (every work item of one step, per GPU)",
                         "bytes": "design bytes: what the walk kernels must move (leaf nibbles once per pass, every internal "
                                  "node's set written once and read once, D1 for Sankoff, 8 B per record), DESIGN.md §5",
                         "bytes_per_update_design": design / per_gpu_updates, "pass_ms_per_step": 1e3 * pass_s,
                         "peak_source": peak_src, "traffic": None if traffic is None else traffic.get("traffic_bytes_per_step"),
                         "traffic_source": None if traffic is None else traffic.get("source"),
                         "frac_contract": per_gpu_updates * b / pass_s / 1e9 / hbm_peak, "bytes_per_update_contract": b,
                         "frac_whole_step": design / step_s / 1e9 / hbm_peak},
            "e2e": e2e, "gpu_launches": int(mine["launches"]), "clocks": clocks,
            "gather_bit_exact_vs_unsharded": gather_ok,
            "phases_ms_per_step": {"forward+backward": pass_ms / args.steps, "ordering": sort_ms / args.steps,
                                   "exchange": xchg_ms / args.steps, "stitch": stitch_ms / args.steps,
                                   "note": "maxima over ranks; exchange includes waiting for the slowest rank",
                                   "per_rank_step_pass_exchange_stitch_order": per_rank},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, ctx, tree, algo, d_msa, d_cons, stride, n_taxa, min(n_cols, gb * 32), ncpu)
        if world == 1 and strong and not args.no_c2:
            line["c2_single_gpu"] = c2_leg(args, algo, hbm_peak)
        print(json.dumps(line))
    if comm is not None:
        comm.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


def cpu_baseline(args, ctx, tree, algo, d_msa, d_cons, stride, n_taxa, n_cols, ncpu):
    """The reference's CPU code on a random column sample of the benchmarked alignment (rank 0, N = 1), doubling as a
    bit-exactness check of the benchmarked path."""
    import numpy as np
    import torch

    from panman_b200 import shard
    V = tree.n_nodes
    span = min(n_cols, 2048)  # sample inside the first columns: one device run of `span` columns gives the CUDA records
    # (on the 16 M-node tree one column costs the reference ~30 s of one core: one column per host thread there)
    cols_n = args.cpu_cols or max(min(ncpu, 16), min(8192, int(3e8 / V)))
    cols = np.sort(np.random.default_rng(0).choice(span, size=min(cols_n, span), replace=False))
    byte_cols = torch.from_numpy(cols // 2).to(d_msa.device)
    sub_bytes = d_msa[:, byte_cols].cpu().numpy()
    sub = np.ascontiguousarray(np.where(cols % 2 == 0, sub_bytes & 15, sub_bytes >> 4).astype(np.uint8))
    cons = d_cons[:span].cpu().numpy()[cols]
    ref = CpuReference(tree, ncpu)
    want, dt = ref.run(args.algo, sub, cons)
    ptr, n = ctx.run_device(algo, d_msa.data_ptr(), stride, n_taxa, span, d_cons.data_ptr())
    t = shard.records_as_tensor(ptr, n, d_msa.device)
    sel = torch.isin(t[:, 1], torch.from_numpy(cols.astype(np.int32)).to(d_msa.device))
    mine = shard.records_to_numpy(t[sel])
    mine["pos"] = np.searchsorted(cols, mine["pos"])
    mine = mine[np.lexsort((mine["pos"], mine["node"]))]
    exact = mine.shape == want.shape and all((mine[k] == want[k]).all() for k in ("node", "pos", "type", "code"))
    return {"value": V * cols.size / dt, "unit": "updates/s", "cores": ref.threads, "kind": ref.kind,
            "sample": f"{cols.size} random columns of the workload, all {V} nodes, {ref.threads} threads over columns, "
                      f"g++ -O2, {dt:.2f} s (column loop; the Node graph is built once, untimed)",
            "bit_exact_on_sample": bool(exact)}


def c2_leg(args, algo, hbm_peak):
    """BASELINE config 2 (20 000 taxa x 30 000 columns, one GPU, everything in one pass) measured in the same run."""
    import torch

    from panman_b200 import api, trees
    n_taxa, n_cols = WORKLOADS["c2"][:2]
    tree = trees.random_binary_tree(n_taxa)
    V = tree.n_nodes
    stride = ((n_cols + 31) // 32) * 16
    with api.Context(tree) as ctx:
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)
        d_msa = torch.zeros((n_taxa, stride), dtype=torch.uint8, device="cuda")
        d_cons = torch.zeros(n_cols, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        ctx.synth_msa(api.SynthSpec(), 0, n_cols, d_msa.data_ptr(), stride, d_cons.data_ptr())
        for _ in range(5):
            ctx.run_device(algo, d_msa.data_ptr(), stride, n_taxa, n_cols, d_cons.data_ptr())
        torch.cuda.synchronize()
        steps, pass_ms, design, n = 50, 0.0, 0, 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, n_taxa, n_cols, d_cons.data_ptr())
            t = ctx.timing()
            pass_ms += t["forward_ms"] + t["backward_ms"]
            design = t["forward_bytes"] + t["backward_bytes"] + 8 * n
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        return {"workload": "c2: 20000 taxa x 30000 columns, random binary tree (39999 nodes), one pass",
                "value": V * n_cols / (ms * 1e-3), "ms_per_step": ms, "steps": steps, "records_per_step": int(n),
                "roofline_frac": design / (pass_ms / steps * 1e-3) / 1e9 / hbm_peak,
                "frac_whole_step": design / (ms * 1e-3) / 1e9 / hbm_peak, "pass_ms_per_step": pass_ms / steps}


if __name__ == "__main__":
    sys.exit(main())
