# timing experiments for k_seg_tile: EDL_SEG_COPY = producer copy engine, EDL_SEG_DEBUG = 1 skips the reduce
B="python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e"
for copy in ${COPIES:-ldgsts}; do for dbg in ${DEBUGS:-0 1}; do
  echo "copy=$copy debug=$dbg: $(EDL_SEG_COPY=$copy EDL_SEG_DEBUG=$dbg timeout 200 $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('update_ms', round(d['phases']['update_ms'],4))")"
done; done
