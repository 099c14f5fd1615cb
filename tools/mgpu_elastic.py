"""BASELINE.json configs[3] on real ranks: elastic re-shard S0 -> S1 -> S2 mid-training, one process per
GPU, world W = number of GPUs (8: 4 -> 8 -> 6; 4: 2 -> 4 -> 3; 2: 1 -> 2 -> 1).

    torchrun --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 tools/mgpu_elastic.py [adagrad|adam] [slots|nccl]

All W processes and their communicator stay alive through the run; a re-shard only changes which ranks
hold a shard.  Training between the re-shards runs among the active ranks (their own process group) and
is checked against the oracle's global float64 pass; each re-shard is the all-to-all-v of
easydl_b200.elastic.gen_reshard over NCCL among ALL W ranks (ranks >= S send nothing, ranks >= S' end
with no shard) and is checked two ways: the multiset of (key, row bits, slot bits) over all shards is
unchanged (a 64-bit checksum of checksums, bit-exact) and every new shard equals the oracle's re-shard.
Prints one JSON line on rank 0; exit code 0 only if everything held.
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from easydl_b200.elastic import gen_reshard
from easydl_b200.exchange import SingleExchange, TorchExchange
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.sharded_step import ShardedSparseStep
from oracle import edl_oracle as O

opt_name = sys.argv[1] if len(sys.argv) > 1 else "adagrad"
xchg = sys.argv[2] if len(sys.argv) > 2 else "slots"
rank, W, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lrank}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
SEQ = {8: (4, 8, 6), 4: (2, 4, 3), 2: (1, 2, 1)}.get(W, (max(1, W // 2), W, max(1, W - 1)))
groups = {S: dist.new_group(list(range(S))) for S in sorted(set(SEQ))}     # every rank creates every group
rows, D, B, K = [50_000, 301, 700_001, 11, 6, 90_000], 32, 2048, 2
RTOL, ATOL = 1e-6, 1e-9
po = OptimizerSpec.adam() if opt_name == "adam" else OptimizerSpec.adagrad()
oo = O.OptimizerSpec.adam() if opt_name == "adam" else O.OptimizerSpec.adagrad()
bases = O.key_bases(rows)
ok, worst, step0 = True, 0.0, 0
log = []


def close(a, b):
    global worst
    err = np.abs(a.astype(np.float64) - b)
    if err.size:
        worst = max(worst, float(np.max(err / (np.abs(b) + ATOL / RTOL))))
    return bool(np.all(err <= ATOL + RTOL * np.abs(b)))


def checksum(P):
    """64-bit wrap-around sum over this shard's records of a hash of (key, row bits, slot bits)."""
    if P is None:
        return 0
    rec = P.export_records().cpu()
    if len(rec) == 0:
        return 0
    h = rec.keys.numpy().astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)
    for m, arr in enumerate([rec.rows] + list(rec.slots)):
        w = arr.numpy().view(np.uint32).astype(np.uint64)
        mult = (np.arange(w.shape[1], dtype=np.uint64) * np.uint64(2654435761) + np.uint64(97 + m)) | np.uint64(1)
        with np.errstate(over="ignore"):
            h = h ^ O.splitmix64((w * mult).sum(axis=1, dtype=np.uint64) + np.uint64(m))
    with np.errstate(over="ignore"):
        return int(O.splitmix64(h).sum(dtype=np.uint64).astype(np.int64))


def train(P, S, osh, steps):
    """K steps among ranks 0..S-1 (idle ranks wait at the barrier); embeddings against the oracle."""
    global ok, step0
    good = True
    if rank < S:
        if S == 1:
            st = ShardedSparseStep(P, B, SingleExchange())
        elif xchg == "slots":
            from easydl_b200.slot_step import SlotShardedStep
            st = SlotShardedStep(P, B, group=groups[S])
        else:
            st = ShardedSparseStep(P, B, TorchExchange(groups[S], dev))
        embs = []
        for k in range(steps):
            _, ids, _ = O.criteo_batch(rows, B, 1.05, step=(step0 + k) * 8 + rank)
            g = O.synthetic_grad(ids.size, D, step=(step0 + k) * 8 + rank)
            embs.append(st.run(torch.from_numpy(ids).to(dev), torch.from_numpy(g).to(dev)).cpu().numpy())
        torch.cuda.synchronize()
        if hasattr(st, "check"):
            st.check()
        if hasattr(st, "buf"):
            st.buf.close()
    for k in range(steps):      # every rank advances the oracle (cheap at this size)
        kw, gw = [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, B, 1.05, step=(step0 + k) * 8 + r)
            kw.append(O.fuse_keys(ids, bases))
            gw.append(O.synthetic_grad(ids.size, D, step=(step0 + k) * 8 + r))
        want = O.sharded_step(osh, kw, gw)
        if rank < S:
            good &= close(embs[k], want[rank])
    step0 += steps
    ok &= good
    return good


def state_matches(P, S, osh):
    if rank >= S:
        return P is None
    rec = P.export_records().cpu()
    o = osh[rank]
    good = np.array_equal(rec.keys.numpy(), o.keys) and rec.step == o.step
    if good:
        good &= close(rec.rows.numpy(), o.w)
        for i in range(oo.n_slots):
            good &= close(rec.slots[i].numpy(), o.slots[i])
    return bool(good)


S = SEQ[0]
P = ShardParameters(rows, D, po, S, rank, device=dev) if rank < S else None
osh = [O.OracleShard(rows, D, oo, S, s) for s in range(S)]
log.append({"phase": f"train S={S}", "ok": train(P, S, osh, K)})
ex_world = TorchExchange(dist.group.WORLD, dev)
for new_S in SEQ[1:]:
    before = torch.tensor([checksum(P)], dtype=torch.int64, device=dev)
    dist.all_reduce(before)
    t0 = time.perf_counter()
    P_new = ex_world.run(gen_reshard(P, new_S, rank, W, rows=rows, dim=D, opt=po, seed=O.DEFAULT_SEED, device=dev))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    after = torch.tensor([checksum(P_new)], dtype=torch.int64, device=dev)
    dist.all_reduce(after)
    osh = O.reshard(osh, new_S)
    moved_exact = int(before.item()) == int(after.item())
    same = state_matches(P_new, new_S, osh)
    ok &= moved_exact and same
    log.append({"phase": f"re-shard {S} -> {new_S}", "bit_exact_move": moved_exact, "matches_oracle_reshard": same,
                "idle_ranks": list(range(new_S, W)), "seconds": round(dt, 4)})
    P, S = P_new, new_S
    log.append({"phase": f"train S={S}", "ok": train(P, S, osh, K)})
    ok &= state_matches(P, S, osh)
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
w = torch.tensor([worst], dtype=torch.float64, device=dev)
dist.all_reduce(w, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"check": "elastic re-shard on real ranks", "world": W, "sequence": list(SEQ), "optimizer": opt_name,
                      "exchange": xchg, "ok": bool(int(flag.item())), "max_rel_err": float(w.item()),
                      "tolerance": RTOL, "log": log}), flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) else 1)
