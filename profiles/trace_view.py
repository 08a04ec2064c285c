"""Reads the per-item timelines written by a -DPMFS_WAITPROF build (PMFS_TRACE_DIR) and prints how many work items
were executing over time.  usage: python profiles/trace_view.py gpurun_out/trace/trace_fwd.bin [bins]"""
import sys

import numpy as np

t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, 4)
bins = int(sys.argv[2]) if len(sys.argv) > 2 else 25
t0, t1 = t[:, 0].astype(np.int64), t[:, 1].astype(np.int64)
base = t0.min()
t0 -= base
t1 -= base
span = t1.max()
dur = t1 - t0
print(f"items {len(t)}  span {span / 1e3:.1f} us  item duration: mean {dur.mean() / 1e3:.2f} us, median {np.median(dur) / 1e3:.2f}, "
      f"p95 {np.percentile(dur, 95) / 1e3:.2f}, max {dur.max() / 1e3:.2f}")
print(f"busy item-time {dur.sum() / 1e3:.0f} us = {dur.sum() / span:.1f} items executing on average")
edges = np.linspace(0, span, bins + 1)
for a, b in zip(edges[:-1], edges[1:]):
    ov = np.clip(np.minimum(t1, b) - np.maximum(t0, a), 0, None).sum() / (b - a)
    started = ((t0 >= a) & (t0 < b)).sum()
    print(f"{a / 1e3:7.1f}-{b / 1e3:7.1f} us: {ov:6.1f} executing, {started:5d} started  " + "#" * int(ov / 8))
