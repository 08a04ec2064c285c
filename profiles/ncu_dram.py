"""DRAM bytes of the walk kernels on ONE 2 048-column batch of the real 16 M-node config-3 tree, from
`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --csv` (a `--set full` capture of these launches would have
ncu save and restore 40 GB around each of ~40 replays; the two DRAM counters need one pass).  Writes the entry
bench.py reads as `roofline.traffic` for the c3 workload: a step of 30 000 columns is 30 000 / 2 048 such batches.
usage: python profiles/ncu_dram.py <csv> <tag>"""
import csv
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
iname, imetric, iunit, ival = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
per = {}
for r in rows[1:]:
    if r[imetric] not in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        continue
    scale = {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1}[r[iunit].lower()]
    name = r[iname].split("(")[0]
    per.setdefault(name, {}).setdefault(r[imetric], []).append(float(r[ival].replace(",", "")) * scale)
kernels, batch = [], 0.0
for name, m in per.items():
    rd = sum(m["dram__bytes_read.sum"]) / len(m["dram__bytes_read.sum"])
    wr = sum(m["dram__bytes_write.sum"]) / len(m["dram__bytes_write.sum"])
    kernels.append({"kernel": name, "dram_read_bytes": rd, "dram_write_bytes": wr, "launches_averaged": len(m["dram__bytes_read.sum"])})
    if "fwd_kernel" in name or "bwd_kernel" in name:
        batch += rd + wr
commit = os.environ.get("CAPTURE_COMMIT") or subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=HERE).stdout.strip()
path = os.path.join(HERE, "traffic_r2.json")
t = json.load(open(path))
t.setdefault("c3_8m_taxa_2048col_batch", {})["fitch"] = {
    "kernels": kernels, "traffic_bytes_per_batch": batch, "traffic_bytes_per_step": batch * 30000 / 2048,
    "source": f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none on one 2 048-column batch of the "
              f"8 M-taxon tree (fwd_kernel + bwd_kernel, one launch each), x 30 000 / 2 048 batches per step; capture {sys.argv[2]} at commit {commit}"}
json.dump(t, open(path, "w"), indent=1)
print("batch", batch, "step", batch * 30000 / 2048)
