#!/bin/bash
# The measurements that need N real GPUs (N = 8 for the headline): tools/mgpu_big.sh N [quick]
#   default cfg2-shape bench, configs[2] (DLRM 26 x 40M x dim 128, Adam, lazy shards), host copy ceiling,
#   real-rank tests (parity of every exchange, elastic re-shard).  JSON lines kept under gpurun_out/.
N=${1:-8}; Q=${2:-}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
mkdir -p gpurun_out
P=gpurun_out/big${N}
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    r = d.get("roofline", {})
    print(sys.argv[2], "|", round(d["value"] / 1e6, 1), "M samples/s", round(d["ms_per_step"], 3), "ms/step", d.get("launch_mode"),
          "| parity", d.get("parity", {}).get("ok"), d.get("parity", {}).get("max_rel_err"),
          "| roofline", r.get("bound"), round(r.get("frac", 0), 3), "| e2e ms", round(d.get("e2e", {}).get("ms_per_step", 0), 2))
    if "phases" in d: print("   phases", {k: round(v, 3) for k, v in d["phases"].items()})
except Exception as e:
    print(sys.argv[2], "FAILED", e); print(open(sys.argv[1].replace(".json", ".err")).read()[-2500:])
PY
}
steps=100; [ -n "$Q" ] && steps=30
(timeout 500 bash -c "$(declare -f run); N=$N; run 29801 bench.py --gpus $N --steps $steps --warmup 10 --no-cpu-baseline --e2e-steps 5" 2> ${P}_cfg2.err | tail -1) > ${P}_cfg2.json
show ${P}_cfg2.json "cfg2 shape, slots, mixed wire"
lazy=$((24000000)); [ -n "$Q" ] && lazy=6000000
(timeout 600 bash -c "$(declare -f run); N=$N; run 29802 bench.py --gpus $N --rows 40000000 --dim 128 --batch 32768 --optimizer adam --lazy-rows $lazy --steps $steps --warmup 10 --no-cpu-baseline --e2e-steps 5" 2> ${P}_cfg3.err | tail -1) > ${P}_cfg3.json
show ${P}_cfg3.json "configs[2]: 26 x 40M x dim 128, batch 32K/GPU, Adam, lazy shards"
(timeout 200 bash -c "$(declare -f run); N=$N; run 29803 tools/pcie_probe.py" 2> ${P}_pcie.err | tail -1) > ${P}_pcie.json
cat ${P}_pcie.json
(timeout 900 python -m pytest tests/test_gpu_peer.py -m gpu -q 2>&1 | tail -15) > ${P}_tests.log 2>&1
cat ${P}_tests.log
