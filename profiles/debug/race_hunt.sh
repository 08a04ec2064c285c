#!/bin/bash
cd /root/repo
L=panman_b200
run() { echo "== $1"; TAG="$1" timeout 200 python profiles/debug/count_diff.py 4x1 2x1 2>&1 | tail -4; }
run baseline
PMFS_RINGS="0,0,1,1" run "fwd leaf ring 1"
PMFS_RINGS="4,1,2,1" run "fwd leaf ring 16"
PMFS_RINGS="2,1,0,0" run "bwd rings 1+1"
PMFS_RINGS="2,1,4,3" run "bwd rings 16+8"
PMFS_LIB=$PWD/$L/libpmfs_NO_CHERRY.so run "no cherry fast path"
PMFS_LIB=$PWD/$L/libpmfs_PROXY_FENCE.so run "proxy fence before every TMA refill"
