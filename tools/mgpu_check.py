"""Multi-GPU parity of the sharded step on real ranks (torchrun, one rank per GPU): tools/parity.py's
check for one exchange -- K steps against the oracle's global float64 pass (Adagrad and Adam), and for
the slot exchanges graph replay == eager.  Prints one JSON line on rank 0.
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/mgpu_check.py [peer|nccl|slots|slots-pipe] [lanes]"""
import json, os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.parity import check_exchange

kind = sys.argv[1] if len(sys.argv) > 1 else "slots"
kind = {"pipe": "slots-pipe", "slots-graph": "slots", "pipe-graph": "slots-pipe"}.get(kind, kind)
lanes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
rank, lr = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lr}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
res = check_exchange(kind, dev, dist.group.WORLD, torch.float64, lanes)
if rank == 0:
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
sys.exit(0 if res["ok"] else 1)
