"""Multi-GPU parity + timing of the sharded step on real ranks (torchrun, one rank per GPU):
every rank runs K steps through NCCL (TorchExchange) or peer writes (PeerExchange) and the final
shard state is compared with the oracle's global float64 pass.  Small tables so the oracle is quick.
  torchrun --nproc-per-node N tools/mgpu_check.py [peer|nccl|slots|slots-graph|pipe|pipe-graph] [adam|adagrad]"""
import os, sys, time, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from easydl_b200.exchange import PeerExchange, TorchExchange
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.sharded_step import ShardedSparseStep
from oracle import edl_oracle as O

mode = sys.argv[1] if len(sys.argv) > 1 else "peer"
opt_name = sys.argv[2] if len(sys.argv) > 2 else "adam"
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lr}"); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
rows, D, B, K = [50_000, 301, 700_001, 11, 6, 90_000], 32, 4096, 3
po = OptimizerSpec.adam() if opt_name == "adam" else OptimizerSpec.adagrad()
oo = O.OptimizerSpec.adam() if opt_name == "adam" else O.OptimizerSpec.adagrad()
P = ShardParameters(rows, D, po, world, rank, device=dev)
if mode.startswith("slots") or mode.startswith("pipe"):
    from easydl_b200.slot_step import GraphedSlotStep, PipelinedSlotStep, SlotShardedStep
    ex = None
    Step = PipelinedSlotStep if mode.startswith("pipe") else SlotShardedStep
    st = Step(P, B, group=dist.group.WORLD)
else:
    ex = PeerExchange(dist.group.WORLD, dev, 2 * B * len(rows), D) if mode == "peer" else TorchExchange(dist.group.WORLD, dev)
    st = ShardedSparseStep(P, B, ex)
bases = O.key_bases(rows)
embs = []
for k in range(K):
    _, ids, _ = O.criteo_batch(rows, B, 1.05, step=k * world + rank)
    g = O.synthetic_grad(ids.size, D, step=k * world + rank)
    embs.append(st.run(torch.from_numpy(ids).to(dev), torch.from_numpy(g).to(dev)).cpu().numpy())
torch.cuda.synchronize()
if mode.startswith("slots") or mode.startswith("pipe"):
    st.check()
# oracle (every rank recomputes the global pass; cheap at this size)
osh = [O.OracleShard(rows, D, oo, world, s) for s in range(world)]
ok = True
for k in range(K):
    kw, gw = [], []
    for r in range(world):
        _, ids, _ = O.criteo_batch(rows, B, 1.05, step=k * world + r)
        kw.append(O.fuse_keys(ids, bases)); gw.append(O.synthetic_grad(ids.size, D, step=k * world + r))
    want = O.sharded_step(osh, kw, gw)[rank]
    err = np.abs(embs[k].astype(np.float64) - want)
    ok &= bool(np.all(err <= 1e-9 + 1e-6 * np.abs(want)))
rec = P.export_records().cpu()
ok &= np.array_equal(rec.keys.numpy(), osh[rank].keys)
for got, want in [(rec.rows.numpy(), osh[rank].w)] + [(rec.slots[i].numpy(), osh[rank].slots[i]) for i in range(oo.n_slots)]:
    err = np.abs(got.astype(np.float64) - want)
    ok &= bool(np.all(err <= 1e-9 + 1e-6 * np.abs(want)))
flag = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"[{mode}/{opt_name}] world={world} parity vs oracle: {'OK' if int(flag.item()) else 'FAIL'}")
if mode in ("slots-graph", "pipe-graph"):
    # graph replay on top of the checked state: K more steps replayed == K more steps issued eagerly
    ids_d, g_d = torch.from_numpy(ids).to(dev), torch.from_numpy(g).to(dev)
    P2 = ShardParameters(rows, D, po, world, rank, device=dev)
    st2 = Step(P2, B, group=dist.group.WORLD)
    for _ in range(4):
        st2.run(ids_d, g_d)
    st2.check()
    P3 = ShardParameters(rows, D, po, world, rank, device=dev)
    st3 = Step(P3, B, group=dist.group.WORLD)
    gs = GraphedSlotStep(st3, ids_d, g_d)          # two eager warm-up steps inside
    gs.run(); gs.run()
    st3.check()
    same = torch.equal(P2.slab, P3.slab) and torch.equal(P2.slots[0], P3.slots[0]) and P2.step == P3.step
    f2 = torch.tensor([1 if same else 0], device=dev); dist.all_reduce(f2, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"[{mode}] graph replay == eager: {'OK' if int(f2.item()) else 'FAIL'}")
    ok &= bool(int(f2.item()))
if mode == "peer":
    ex.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
