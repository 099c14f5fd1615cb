"""GPU time of every operation of a sharded step on a real multi-GPU run (CUDA events around each
ops / exchange call, median over 10 steps, rank 0 printed).
torchrun --nproc-per-node N tools/sharded_ops_trace.py [peer|nccl] [f32|f64]"""
import os, statistics, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from easydl_b200.criteo import CriteoSynthetic
from easydl_b200.exchange import PeerExchange, TorchExchange
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.sharded_step import ShardedSparseStep

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lr}"); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "peer"
wire = torch.float32 if (len(sys.argv) > 2 and sys.argv[2] == "f32") else torch.float64
rows, D, B = [10_000_000] * 26, 64, 65536
P = ShardParameters(rows, D, OptimizerSpec.adagrad(), world, rank, device=dev)
ex = PeerExchange(dist.group.WORLD, dev, int(1.5 * B * 26), D, wire) if mode == "peer" else TorchExchange(dist.group.WORLD, dev)
st = ShardedSparseStep(P, B, ex, wire_dtype=wire)
src = CriteoSynthetic(rows, B)
ids = torch.from_numpy(src.ids(rank)).to(dev); grad = torch.from_numpy(src.grad(D, rank)).to(dev)

log = []
def wrap(obj, name, label=None):
    fn = getattr(obj, name)
    def w(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = fn(*a, **k); e1.record()
        log.append((label or name, e0, e1))
        return out
    setattr(obj, name, w)
for n in ("unique", "partition", "lookup", "lookup_into", "expand", "reduce_for_send", "reduce_into_peers", "apply"):
    wrap(st.ops, n, "ops." + n)
if mode == "peer":
    wrap(ex, "_fence", "fence"); wrap(ex, "dst_table", "dst_table (H2D)")
wrap(ex, "handle", "exchange.handle")

def step():
    log.clear()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); st.run(ids, grad); e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e), [(n, a.elapsed_time(b)) for n, a, b in log]

for _ in range(5): step()
dist.barrier(); torch.cuda.synchronize()
runs = [step() for _ in range(10)]
if rank == 0:
    names = [n for n, _ in runs[0][1]]
    print(f"{mode} wire={wire} world={world}")
    for i, n in enumerate(names):
        print(f"{n:28s} {statistics.median(r[1][i][1] for r in runs) * 1e3:8.1f} us")
    print(f"{'whole step':28s} {statistics.median(r[0] for r in runs) * 1e3:8.1f} us")
dist.destroy_process_group()
