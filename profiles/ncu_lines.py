"""Per-source-line stall samples of one kernel from an .ncu-rep captured with --import-source on.
usage: python profiles/ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
agg, fn, cols = {}, None, {}
stall_tot = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fn = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        cols = {h: i for i, h in enumerate(r)}
        continue
    if r[0] in ("Function Name",):
        continue
    if r[0] != "":
        try:
            s, ins = float(r[cols["# Samples"]]), float(r[cols["Instructions Executed"]])
        except (ValueError, KeyError, IndexError):
            continue
        a = agg.setdefault((fn, int(r[0]), r[1][:100]), [0, 0, {}])
        a[0] += s
        a[1] += ins
        for h, i in cols.items():
            if h.startswith("stall_") and "Not Issued" not in h:
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                a[2][h] = a[2].get(h, 0) + v
                stall_tot[h] = stall_tot.get(h, 0) + v
tot = sum(v[0] for v in agg.values())
print("samples", tot, "warp instructions", sum(v[1] for v in agg.values()))
print({k: round(100 * v / tot, 1) for k, v in sorted(stall_tot.items(), key=lambda kv: -kv[1])[:10]})
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    main = max(v[2].items(), key=lambda kv: kv[1])[0] if v[2] else ""
    print(f"{k[0]}:{k[1]:5d} {100 * v[0] / tot:5.1f}% inst {v[1]:10.0f} {main:18s} {k[2]}")
