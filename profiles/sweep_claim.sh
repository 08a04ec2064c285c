for L in ${LEADS:-4}; do
  for A in fitch sankoff; do
  echo "lead $L $A"
  PMFS_CLAIM_LEAD=$L timeout 120 python bench.py --algo $A --no-cpu-baseline --no-e2e 2>&1 | grep "^{" | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['phases_ms_last_step'])"
  done
done
