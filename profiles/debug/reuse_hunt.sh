#!/bin/bash
# correctness (all records of config 4, narrow shapes vs 4x2) and speed of ring hand-back variants: libpmfs.so (default
# mode) against libpmfs_reuse<N>.so builds
cd /root/repo
L=$PWD/panman_b200
for v in "" $*; do
  name=${v:-default}
  lib=$L/libpmfs${v:+_$v}.so
  echo "===== $name"
  PMFS_LIB=$lib TAG=$name timeout 200 python profiles/debug/count_diff.py 4x1 2x1 1x1 2>&1 | tail -6
  PMFS_LIB=$lib timeout 120 python bench.py --workload c2 --steps 50 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c2 ms/step', d['ms_per_step'], 'pass', d['roofline']['pass_ms_per_step'])"
  PMFS_LIB=$lib timeout 200 python profiles/sweep_shapes.py c3 2>&1 | grep auto | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('c3', d['algo'], d['columns'], d['ms'])"
  PMFS_LIB=$lib timeout 100 python profiles/sweep_shapes.py c4 c5 2>&1 | grep auto | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'], d['algo'], d['columns'], d['ms'])"
done
