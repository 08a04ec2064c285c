# usage: bash profiles/run_variants.sh A B C ...   (libraries built into panman_b200/variants/ beforehand)
cp panman_b200/libpmfs.so /tmp/libpmfs_keep.so
for v in "$@"; do
  cp panman_b200/variants/libpmfs_$v.so panman_b200/libpmfs.so; touch panman_b200/libpmfs.so
  for A in ${ALGOS:-fitch}; do
    echo "variant $v $A"
    timeout 120 python bench.py --algo $A --no-cpu-baseline --no-e2e 2>&1 | grep "^{" | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['phases_ms_last_step'])"
  done
done
cp /tmp/libpmfs_keep.so panman_b200/libpmfs.so
