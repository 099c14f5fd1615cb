"""Where a sharded step's time goes on a real multi-GPU run: host wall time and GPU time (CUDA
events) between the exchange points.  torchrun --nproc-per-node N tools/sharded_trace.py"""
import os, statistics, sys, time, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from easydl_b200.criteo import CriteoSynthetic
from easydl_b200.exchange import PeerExchange, TorchExchange
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.sharded_step import ShardedSparseStep

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lr}"); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
rows, D, B = [10_000_000] * 26, 64, 65536
P = ShardParameters(rows, D, OptimizerSpec.adagrad(), world, rank, device=dev)
ex = PeerExchange(dist.group.WORLD, dev, int(1.5 * B * 26), D) if (len(sys.argv) > 1 and sys.argv[1] == 'peer') else TorchExchange(dist.group.WORLD, dev)
st = ShardedSparseStep(P, B, ex)
src = CriteoSynthetic(rows, B)
ids = torch.from_numpy(src.ids(rank)).to(dev); grad = torch.from_numpy(src.grad(D, rank)).to(dev)

marks = []
def mark(name):
    e = torch.cuda.Event(enable_timing=True); e.record(); marks.append((name, time.perf_counter(), e))

def traced_step():
    marks.clear(); mark("start")
    gen = st.gen_step(ids, grad)
    req = next(gen); mark("dedup+partition issued")
    try:
        while True:
            rep = ex.handle(req); mark(f"{req[0]} done")
            req = gen.send(rep); mark("compute issued")
    except StopIteration:
        pass
    mark("end"); torch.cuda.synchronize()

for _ in range(5): traced_step()
dist.barrier(); torch.cuda.synchronize()
acc = {}
for it in range(10):
    traced_step()
    for (n0, t0, e0), (n1, t1, e1) in zip(marks[:-1], marks[1:]):
        acc.setdefault((len(acc) if it == 0 else None, n1), None)
    rowsum = [(n1, (t1 - t0) * 1e3, e0.elapsed_time(e1)) for (n0, t0, e0), (n1, t1, e1) in zip(marks[:-1], marks[1:])]
    if it == 9 and rank == 0:
        print(f"{'segment':28s} host_ms  gpu_ms")
        for n, h, g in rowsum: print(f"{n:28s} {h:7.3f} {g:7.3f}")
        print(f"{'total':28s} {sum(h for _, h, _ in rowsum):7.3f} {marks[0][2].elapsed_time(marks[-1][2]):7.3f}")
dist.destroy_process_group()
