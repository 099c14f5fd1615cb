#!/bin/bash
# hot-id threshold with the forked update: tools/thr_fork_sweep.sh 8 12 16 24 32
for t in "$@"; do
  EDL_LONG_THR=$t python bench.py --no-e2e --no-cpu-baseline --steps 100 --warmup 5 2>/dev/null | tail -1 > /tmp/thr_line.json
  python - $t <<'PY'
import json, sys
d = json.loads(open('/tmp/thr_line.json').read())
print("thr", sys.argv[1], round(d["ms_per_step"], 4), "ms/step", "update", round(d["phases"]["update_ms"], 4))
PY
done
