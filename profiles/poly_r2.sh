#!/bin/bash
# Polytomies, round 2: the 20 000-taxon tree with a fraction of its internal edges collapsed (1.0 = a star), whole
# nodes (split arity 0) against split ones (default: more than 128 children, 63 per part), both algorithms.
out=gpurun_out/${1:-r2p}; mkdir -p $out
cd "$(dirname "$0")/.."
for algo in fitch sankoff; do
  for frac in 0.6 0.9 1.0; do
    for split in 0 128; do
      echo "== $algo collapsed $frac split_arity $split"
      timeout 120 python profiles/prof_run.py $algo 4 20000 30000 $frac $split 2>&1 | tail -1
    done
  done
done > $out/poly_split.log 2>&1
