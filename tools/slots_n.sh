#!/bin/bash
# parity + bench of the slot exchange on N real ranks: tools/slots_n.sh N
N=${1:-2}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29611 tools/mgpu_check.py slots adam 2>&1 | grep -E "parity|Error|error|Traceback" | head -5
run 29612 tools/mgpu_check.py slots-graph adagrad 2>&1 | grep -E "parity|graph replay|Error|error|Traceback" | head -5
port=29620
for ex in slots peer; do
  port=$((port + 1))
  run $port bench.py --gpus $N --exchange $ex --no-e2e --steps 50 --warmup 5 2>/tmp/err_$ex.log | tail -1 > /tmp/line_$ex.json
  python - $ex <<'PY'
import json, sys
try:
    d = json.loads(open(f'/tmp/line_{sys.argv[1]}.json').read())
    print(sys.argv[1], round(d["value"] / 1e6, 1), "M samples/s", round(d["ms_per_step"], 3), "ms/step", d.get("launch_mode"))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open(f'/tmp/err_{sys.argv[1]}.log').read()[-1500:])
PY
done
