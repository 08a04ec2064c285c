#!/bin/bash
# C5 (ladder) and C3-slice with different operand ring depths (PMFS_RINGS = log2 of fwd leaf, fwd set, bwd leaf, bwd set)
cd /root/repo
for r in "3,2,3,2" "4,0,3,3" "5,0,4,4" "5,2,5,4"; do
  echo "== c5 rings $r"; PMFS_RINGS=$r PMFS_TILE_SHAPE=1x1 timeout 120 python profiles/sweep_shapes.py c5 2>&1 | grep auto
done
for r in "2,0,1,1" "3,1,2,2" "4,2,3,3"; do
  echo "== c3 rings $r"; PMFS_RINGS=$r timeout 300 python profiles/sweep_shapes.py c3 2>&1 | grep -E "auto|2x1"
done
