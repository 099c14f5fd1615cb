#!/bin/bash
# lean N-rank confirmation of the slot exchange: parity (Adam, eager) then the graphed bench
N=${1:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 200 bash -c "$(declare -f run); N=$N; run 29641 tools/mgpu_check.py slots adam" 2>&1 | grep -E "parity|Error|Traceback" | head -5
timeout 300 bash -c "$(declare -f run); N=$N; run 29642 bench.py --gpus $N --exchange slots --no-e2e --steps 50 --warmup 5" 2>/tmp/err_slots.log | tail -1 > gpurun_out/slots_n$N.json
python - $N <<'PY'
import json, sys
try:
    d = json.loads(open(f'gpurun_out/slots_n{sys.argv[1]}.json').read())
    print("slots", d["n_gpus"], "GPUs", round(d["value"] / 1e6, 1), "M samples/s", round(d["ms_per_step"], 3), "ms/step", d.get("launch_mode"))
except Exception as e:
    print("FAILED", e); print(open('/tmp/err_slots.log').read()[-2500:])
PY
