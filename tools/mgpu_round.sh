#!/bin/bash
# One multi-GPU measurement round: tools/mgpu_round.sh N [out_prefix]
# real-rank parity tests, elastic re-shard, then bench lines for the exchange variants (JSON lines kept).
N=${1:-2}; P=${2:-gpurun_out/mg${N}}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_peer.py -m gpu -q 2>&1 | tail -15) > ${P}_tests.log 2>&1
port=29700
for cfg in "--exchange slots" "--exchange slots --wire f64" "--exchange slots-pipe --lanes 2" "--exchange slots-pipe --lanes 3"; do
  port=$((port + 1))
  tag=$(echo $cfg | tr -d ' -' )
  (timeout 400 bash -c "$(declare -f run); N=$N; run $port bench.py --gpus $N $cfg --steps 100 --warmup 10 --no-cpu-baseline --e2e-steps 5" 2> ${P}_${tag}.err | tail -1) > ${P}_${tag}.json
  python - ${P}_${tag}.json "$cfg" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    r = d.get("roofline", {})
    print(sys.argv[2], "|", round(d["value"] / 1e6, 1), "M samples/s", round(d["ms_per_step"], 3), "ms/step", d.get("launch_mode"),
          "| parity", d.get("parity", {}).get("ok"), d.get("parity", {}).get("max_rel_err"),
          "| roofline", r.get("bound"), round(r.get("frac", 0), 3), "| e2e ms", round(d.get("e2e", {}).get("ms_per_step", 0), 2))
    if "phases" in d: print("   phases", {k: round(v, 3) for k, v in d["phases"].items()})
except Exception as e:
    print(sys.argv[2], "FAILED", e); print(open(sys.argv[1].replace(".json", ".err")).read()[-1500:])
PY
done
cat ${P}_tests.log
