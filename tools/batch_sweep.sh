B="python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e"
for v in ${BATCHES:-1 2 4}; do
  echo "batch=$v: $(EDL_DIRECT_BATCH=$v timeout 200 $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('update_ms', round(d['phases']['update_ms'],4), 'step_ms', round(d['ms_per_step'],4))")"
done
