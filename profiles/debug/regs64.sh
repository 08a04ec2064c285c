#!/bin/bash
# A/B: 2x1 top-down Fitch kernel at 64 registers (8 warps per SM sub-partition) with shallower set rings so that shared
# memory lets the extra CTAs in.  usage: regs64.sh <outdir>
out=gpurun_out/${1:-r2x}; mkdir -p $out
cd "$(dirname "$0")/../.."
run() { echo "== $1 rings=${2:-default}"; PMFS_LIB=$PWD/panman_b200/$1 PMFS_RINGS=$2 timeout 200 python profiles/prof_cfg.py c3 fitch 2048 0 4 2>&1 | tail -1 | cut -c 1-330; }
run64() { echo "== $1 rings=default"; PMFS_LIB=$PWD/panman_b200/$1 timeout 200 python profiles/prof_cfg.py c3 fitch 2048 0 4 2>&1 | tail -1 | cut -c 1-330; }
{ run64 libpmfs_r64.so; run libpmfs_r64.so 3,1,2,1; run libpmfs_r64.so 3,1,1,1; } > $out/regs64.log 2>&1
