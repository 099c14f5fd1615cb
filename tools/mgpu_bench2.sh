#!/bin/bash
# Only the two bench lines of tools/mgpu_big.sh (cfg2 shape, configs[2]) on N GPUs: tools/mgpu_bench2.sh N [tag]
N=${1:-8}; T=${2:-b}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
mkdir -p gpurun_out
P=gpurun_out/${T}${N}
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    r = d.get("roofline", {})
    print(sys.argv[2], "|", round(d["value"] / 1e6, 1), "M samples/s", round(d["ms_per_step"], 3), "ms/step", d.get("launch_mode"),
          "| parity", d.get("parity", {}).get("ok"), "| roofline", r.get("bound"), round(r.get("frac", 0), 3))
    if "phases" in d: print("   phases", {k: round(v, 3) for k, v in d["phases"].items()})
except Exception as e:
    print(sys.argv[2], "FAILED", e); print(open(sys.argv[1].replace(".json", ".err")).read()[-2500:])
PY
}
(timeout 500 bash -c "$(declare -f run); N=$N; run 29811 bench.py --gpus $N --steps 100 --warmup 10 --no-cpu-baseline --no-e2e" 2> ${P}_cfg2.err | tail -1) > ${P}_cfg2.json
show ${P}_cfg2.json "cfg2 shape"
(timeout 600 bash -c "$(declare -f run); N=$N; run 29812 bench.py --gpus $N --rows 40000000 --dim 128 --batch 32768 --optimizer adam --lazy-rows 24000000 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e" 2> ${P}_cfg3.err | tail -1) > ${P}_cfg3.json
show ${P}_cfg3.json "configs[2]"
