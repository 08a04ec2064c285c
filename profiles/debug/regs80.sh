#!/bin/bash
# A/B: top-down Fitch kernel of the one-group-per-thread shapes compiled for 80 registers (6 warps per SM sub-partition)
# instead of 96 (5): config-3 slices of 2 048 columns (2x1 tiles) and 1 024 (1x1).  usage: regs80.sh <outdir>
out=gpurun_out/${1:-r2v}; mkdir -p $out
cd "$(dirname "$0")/../.."
for lib in libpmfs.so libpmfs_r80.so; do
  for cols in 2048 1024; do
    echo "== $lib c3 fitch $cols"
    PMFS_LIB=$PWD/panman_b200/$lib timeout 200 python profiles/prof_cfg.py c3 fitch $cols 0 4 2>&1 | tail -1 | cut -c 1-330
  done
done > $out/regs80.log 2>&1
