"""Per-kernel metrics out of `ncu --set full` reports -> profiles/<tag>_ncu_full_<workload>_<algo>.csv and
profiles/traffic_<round>.json (DRAM bytes per launch of the walk kernels, with the commit they were captured at).
usage: python profiles/ncu_extract.py <tag> <workload>:<algo>=<rep> ...   (tag e.g. r2_c)"""
import csv
import io
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
METRICS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
           "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__occupancy_limit_warps",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio"]
tag = sys.argv[1]
traffic = {}
commit = os.environ.get("CAPTURE_COMMIT") or subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=HERE).stdout.strip()
for arg in sys.argv[2:]:
    wa, rep = arg.split("=")
    workload, algo = wa.split(":")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [hdr.index("Kernel Name")] + [hdr.index(m) for m in METRICS if m in hdr]
    with open(os.path.join(HERE, f"{tag}_ncu_full_{workload}_{algo}.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in cols])
        w.writerow([units[i] for i in cols])
        for r in data:
            w.writerow([r[i] for i in cols])
    def val(r, m):
        i = hdr.index(m)
        x = float(r[i])
        u = units[i].lower()
        return x * {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1}.get(u, 1)
    ks, per_kernel = [], {}
    for r in data:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        k = {"kernel": name, "dram_read_bytes": val(r, "dram__bytes_read.sum"), "dram_write_bytes": val(r, "dram__bytes_write.sum"),
             "duration_us_under_ncu": float(r[hdr.index("gpu__time_duration.sum")]),
             "inst_executed": float(r[hdr.index("smsp__inst_executed.sum")]),
             "registers": int(float(r[hdr.index("launch__registers_per_thread")])), "grid": r[hdr.index("launch__grid_size")],
             "issue_active_pct": float(r[hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")]),
             "warps_active_pct": float(r[hdr.index("sm__warps_active.avg.pct_of_peak_sustained_active")])}
        ks.append(k)
        if "fwd_kernel" in name or "bwd_kernel" in name:
            per_kernel.setdefault(name, []).append(k["dram_read_bytes"] + k["dram_write_bytes"])
    # one step launches each walk kernel once: average over the captured launches of each (a capture may hold several steps)
    total = sum(sum(v) / len(v) for v in per_kernel.values())
    traffic.setdefault(workload, {})[algo] = {"kernels": ks, "traffic_bytes_per_step": total,
                                              "source": f"ncu --set full --clock-control none, {os.path.basename(rep)}, capture {tag} at commit {commit}"}
out_path = os.path.join(HERE, "traffic_" + tag.split("_")[0] + ".json")
old = json.load(open(out_path)) if os.path.exists(out_path) else {}
for w, d in traffic.items():
    old.setdefault(w, {}).update(d)
json.dump(old, open(out_path, "w"), indent=1)
for w, d in traffic.items():
    for a, t in d.items():
        print(w, a, t["traffic_bytes_per_step"])
        for k in t["kernels"]:
            print("  ", k)
