#!/bin/bash
# A/B of an environment switch on the N-GPU bench: tools/ab_sharded.sh N VAR val1 val2 ... [-- extra bench args]
N=$1; VAR=$2; shift 2
vals=(); while [ $# -gt 0 ] && [ "$1" != "--" ]; do vals+=("$1"); shift; done; shift
port=29600
for v in "${vals[@]}"; do
  port=$((port + 1))
  env $VAR=$v python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $N --no-e2e --steps 50 --warmup 5 "$@" 2>/dev/null | tail -1 > /tmp/ab_line.json
  python - "$VAR=$v" <<'PY'
import json, sys
d = json.loads(open('/tmp/ab_line.json').read())
print(sys.argv[1], round(d["value"] / 1e6, 1), "M samples/s", round(d["ms_per_step"], 3), "ms/step", d["config"].get("exchange"), d["config"].get("gradient_wire"))
PY
done
