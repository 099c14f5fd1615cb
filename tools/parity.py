"""Real-rank parity of the sharded sparse step against the oracle (one process per GPU, NCCL rendezvous).

Used by bench.py (N > 1: before the warm-up, the result goes into the JSON line as "parity") and by
tools/mgpu_check.py.  Small tables, so every rank can afford the oracle's global float64 pass:
K steps through the exchange under test, then the embeddings of every step and the final rows /
optimizer slots of this rank's shard are compared with the oracle at the north_star tolerance
(1e-6 relative, 1e-9 absolute); for the slot exchanges additionally K steps replayed from CUDA graphs
must equal K steps issued eagerly bit for bit.  The oracle is the checker here, never the thing timed.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ROWS = [50_000, 301, 700_001, 11, 6, 90_000]
DIM, BATCH, STEPS = 32, 4096, 3
RTOL, ATOL = 1e-6, 1e-9


def _make_step(kind, P, group, dev, wire_dtype, lanes):
    from easydl_b200.exchange import PeerExchange, TorchExchange
    from easydl_b200.sharded_step import ShardedSparseStep
    if kind in ("slots", "slots-pipe"):
        from easydl_b200.slot_step import PipelinedSlotStep, SlotShardedStep
        if kind == "slots-pipe":
            return PipelinedSlotStep(P, BATCH, group=group, wire_dtype=wire_dtype, n_lanes=lanes), None
        return SlotShardedStep(P, BATCH, group=group, wire_dtype=wire_dtype), None
    if kind == "peer":
        ex = PeerExchange(group, dev, 2 * BATCH * len(ROWS), DIM, wire_dtype)
    else:
        ex = TorchExchange(group, dev)
    return ShardedSparseStep(P, BATCH, ex, wire_dtype=wire_dtype), ex


def _close(step, ex):
    if ex is not None and hasattr(ex, "close"):
        ex.close()
    for lane in getattr(step, "lanes", [step]):
        buf = getattr(lane, "buf", None)
        if buf is not None:
            buf.close()


def _rel_err(got, want):
    got = got.astype(np.float64)
    err = np.abs(got - want)
    ok = bool(np.all(err <= ATOL + RTOL * np.abs(want)))
    rel = float(np.max(err / (np.abs(want) + ATOL / RTOL))) if err.size else 0.0
    return ok, rel


def check_exchange(kind: str, dev, group=None, wire_dtype=torch.float64, lanes: int = 2,
                   optimizers=("adagrad", "adam"), graph: bool = True, params_factory=None):
    """-> dict for the bench line.  Every rank must call it; the verdict is all-reduced (MIN)."""
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from oracle import edl_oracle as O
    group = group if group is not None else dist.group.WORLD
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    bases = O.key_bases(ROWS)
    ok, worst, graph_equal = True, 0.0, None
    for opt_name in optimizers:
        po = OptimizerSpec.adam() if opt_name == "adam" else OptimizerSpec.adagrad()
        oo = O.OptimizerSpec.adam() if opt_name == "adam" else O.OptimizerSpec.adagrad()
        P = params_factory(po) if params_factory else ShardParameters(ROWS, DIM, po, world, rank, device=dev)
        st, ex = _make_step(kind, P, group, dev, wire_dtype, lanes)
        embs, last, prev_ids = [], None, None
        for k in range(STEPS):
            _, ids, _ = O.criteo_batch(ROWS, BATCH, 1.05, step=k * world + rank)
            g = O.synthetic_grad(ids.size, DIM, step=k * world + rank)
            if last is not None:
                prev_ids = last[0]
            last = (torch.from_numpy(ids).to(dev), torch.from_numpy(g).to(dev))
            embs.append(st.run(*last).cpu().numpy())
        torch.cuda.synchronize()
        if hasattr(st, "check"):
            st.check()
        osh = [O.OracleShard(ROWS, DIM, oo, world, s) for s in range(world)]
        for k in range(STEPS):
            kw, gw = [], []
            for r in range(world):
                _, ids, _ = O.criteo_batch(ROWS, BATCH, 1.05, step=k * world + r)
                kw.append(O.fuse_keys(ids, bases))
                gw.append(O.synthetic_grad(ids.size, DIM, step=k * world + r))
            want = O.sharded_step(osh, kw, gw)[rank]
            o, r_ = _rel_err(embs[k], want)
            ok &= o
            worst = max(worst, r_)
        rec = P.export_records().cpu()
        ok &= bool(np.array_equal(rec.keys.numpy(), osh[rank].keys))
        if ok:
            pairs = [(rec.rows.numpy(), osh[rank].w)] + [(rec.slots[i].numpy(), osh[rank].slots[i])
                                                         for i in range(oo.n_slots)]
            for got, want in pairs:
                o, r_ = _rel_err(got, want)
                ok &= o
                worst = max(worst, r_)
        _close(st, ex)
        del st, ex, P
        if graph and kind in ("slots", "slots-pipe"):
            # K steps replayed from CUDA graphs == K steps issued eagerly (same buffers, same batch)
            from easydl_b200.slot_step import GraphedSlotStep
            ids_d, g_d = last
            states = []
            for replay in (False, True):
                P2 = params_factory(po) if params_factory else ShardParameters(ROWS, DIM, po, world, rank, device=dev)
                st2, _ = _make_step(kind, P2, group, dev, wire_dtype, lanes)
                if replay:
                    gs = GraphedSlotStep(st2, ids_d, g_d)      # two eager steps inside
                    gs.run()
                    gs.run()
                else:
                    for _ in range(4):
                        st2.run(ids_d, g_d)
                st2.check()
                r2 = P2.export_records()     # dense and lazy shards alike: the materialised rows, ascending by key
                states.append((r2.keys.clone(), r2.rows.clone(), [s.clone() for s in r2.slots], P2.step))
                _close(st2, None)
                del st2, P2
            same = torch.equal(states[0][0], states[1][0]) and torch.equal(states[0][1], states[1][1]) and \
                states[0][3] == states[1][3] and all(torch.equal(a, b) for a, b in zip(states[0][2], states[1][2]))
            if kind == "slots":
                # and the prefetching ring (next batch's phase A beside this step; two eager + two replayed steps
                # over two different batches) == four plain eager steps over the same batch sequence
                from easydl_b200.slot_step import GraphedSlotRing
                ring_ids = [prev_ids, ids_d]
                rstates = []
                for use_ring in (False, True):
                    P3 = params_factory(po) if params_factory else ShardParameters(ROWS, DIM, po, world, rank, device=dev)
                    st3, _ = _make_step(kind, P3, group, dev, wire_dtype, lanes)
                    if use_ring:
                        ring = GraphedSlotRing(st3, ring_ids, g_d)
                        ring.run()
                        ring.run()
                    else:
                        for k in range(4):
                            st3.run(ring_ids[k % 2], g_d)
                    st3.check()
                    r3 = P3.export_records()
                    rstates.append((r3.keys.clone(), r3.rows.clone(), [x.clone() for x in r3.slots], P3.step))
                    _close(st3, None)
                    del st3, P3
                same = same and torch.equal(rstates[0][0], rstates[1][0]) and torch.equal(rstates[0][1], rstates[1][1]) \
                    and rstates[0][3] == rstates[1][3] and all(torch.equal(a, b) for a, b in zip(rstates[0][2], rstates[1][2]))
            graph_equal = same if graph_equal is None else (graph_equal and same)
            ok &= same
    flag = torch.tensor([1 if ok else 0], device=dev)
    w = torch.tensor([worst], dtype=torch.float64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(w, op=dist.ReduceOp.MAX, group=group)
    torch.cuda.empty_cache()
    out = {"checked": True, "ok": bool(int(flag.item())), "max_rel_err": float(w.item()), "tolerance": RTOL,
           "exchange": kind, "steps": STEPS, "optimizers": list(optimizers), "world": world,
           "against": "oracle/edl_oracle.py global float64 pass; embeddings of every step + final rows and slots"}
    if graph_equal is not None:
        out["graph_replay_equals_eager"] = bool(graph_equal)
    return out
