"""World-size-2 (and 3) gloo tests on CPU of the multi-rank host logic: the exchange layer
(TorchExchange over torch.distributed) and the sharded step's choreography -- which buffer goes to
which rank with which split sizes -- bound to a NumPy test double instead of the CUDA kernels
(there is no GPU here; the kernels themselves are covered by the -m gpu tests, and the very same
generator runs there on virtual ranks).  The expected values come from the oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ROWS, DIM, BATCH = [500, 37, 3001, 11], 8, 96


class _Uq:
    pass


class FakeOps:
    """NumPy stand-in for CudaOps with the same interface (test double, CPU tensors)."""

    def __init__(self, shard):
        from oracle import edl_oracle as O
        self.O, self.shard = O, shard
        self.bases = O.key_bases(ROWS)

    def unique(self, flat_ids):
        O = self.O
        ids = flat_ids.numpy().reshape(-1, len(ROWS))
        keys = O.fuse_keys(ids, self.bases)
        u, inv, cnt, order, seg = O.dedup(keys)
        q = _Uq()
        q.uniq, q.inverse, q.order, q.seg = u, inv, order, seg
        q.n_unique = torch.tensor([len(u)], dtype=torch.int32)
        q.n = len(keys)
        return q

    def partition(self, uq, S):
        pk, off, perm = self.O.partition_mod(uq.uniq, self.bases, S)
        p = _Uq()
        p.keys = torch.from_numpy(pk)
        p.offsets = torch.from_numpy(off)
        p.perm = perm
        pos = np.empty_like(perm)
        pos[perm] = np.arange(len(perm), dtype=perm.dtype)
        p.pos = pos
        return p

    def lookup(self, keys):
        return torch.from_numpy(self.shard.lookup(keys.numpy()))

    def expand(self, rows_in, uq, part):
        return rows_in[torch.from_numpy(part.pos[uq.inverse].astype(np.int64))]

    def reduce_for_send(self, grad, uq, part, n_send, wire_dtype):
        g = np.add.reduceat(grad.numpy()[uq.order].astype(np.float64), uq.seg[:-1].astype(np.int64), axis=0)
        out = np.empty_like(g)
        out[part.pos] = g
        return torch.from_numpy(out).to(wire_dtype)

    def apply(self, recv_keys, g_recv):
        O = self.O
        u, _, _, order, seg = O.dedup(recv_keys.numpy())
        if len(u):
            g = np.add.reduceat(g_recv.numpy()[order].astype(np.float64), seg[:-1].astype(np.int64), axis=0)
            self.shard.apply(u, g.astype(np.float32))
        else:
            self.shard.step += 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from easydl_b200.exchange import TorchExchange
        from easydl_b200.sharded_step import ShardedSparseStep
        from oracle import edl_oracle as O
        ex = TorchExchange(None, torch.device("cpu"))
        assert ex.world == world and ex.rank == rank
        # raw exchange layer
        c = torch.arange(world * 2, dtype=torch.int64).reshape(world, 2) + 100 * rank
        got = ex.handle(("counts", c))
        assert got.tolist() == [[100 * r + 2 * rank, 100 * r + 2 * rank + 1] for r in range(world)]
        send = [rank + 1 + d for d in range(world)]
        recv = [s + 1 + rank for s in range(world)]
        x = torch.arange(sum(send), dtype=torch.float32).reshape(-1, 1) + 1000 * rank
        y = ex.handle(("a2av", x, send, recv))
        want = []
        for s in range(world):
            off = sum(s + 1 + d for d in range(rank))
            want += [1000 * s + off + i for i in range(s + 1 + rank)]
        assert y.flatten().tolist() == want
        # the sharded step on a NumPy shard
        shard = O.OracleShard(ROWS, DIM, O.OptimizerSpec.adagrad(), world, rank)
        step = ShardedSparseStep(None, BATCH, ex, n_fields=len(ROWS), num_shards=world, _ops=FakeOps(shard))
        embs = []
        for k in range(3):
            _, ids, _ = O.criteo_batch(ROWS, BATCH, 1.05, step=k * world + rank)
            g = O.synthetic_grad(ids.size, DIM, step=k * world + rank)
            emb = step.run(torch.from_numpy(ids), torch.from_numpy(g))
            embs.append(emb.numpy())
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), embs=np.stack(embs), keys=shard.keys, w=shard.w,
                 acc=shard.slots[0], step=shard.step)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_step_over_gloo_matches_oracle(world, tmp_path):
    from oracle import edl_oracle as O
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    shards = [O.OracleShard(ROWS, DIM, O.OptimizerSpec.adagrad(), world, r) for r in range(world)]
    bases = O.key_bases(ROWS)
    want_embs = [[] for _ in range(world)]
    for k in range(3):
        kw, gw = [], []
        for r in range(world):
            _, ids, _ = O.criteo_batch(ROWS, BATCH, 1.05, step=k * world + r)
            kw.append(O.fuse_keys(ids, bases))
            gw.append(O.synthetic_grad(ids.size, DIM, step=k * world + r))
        for r, e in enumerate(O.sharded_step(shards, kw, gw)):
            want_embs[r].append(e)
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), f"r{r}.npz"))
        assert np.array_equal(got["keys"], shards[r].keys)
        assert int(got["step"]) == shards[r].step == 3
        np.testing.assert_allclose(got["embs"], np.stack(want_embs[r]), rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(got["w"], shards[r].w, rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(got["acc"], shards[r].slots[0], rtol=1e-6, atol=1e-9)


def test_lockstep_driver_equals_real_ranks():
    """run_lockstep (virtual ranks in one process, used by the -m gpu sharded tests) serves the
    same requests the way a real process group does."""
    from easydl_b200.exchange import run_lockstep

    def gen(rank, world):
        c = torch.arange(world * 2, dtype=torch.int64).reshape(world, 2) + 100 * rank
        got = yield ("counts", c)
        send = [rank + 1 + d for d in range(world)]
        recv = [s + 1 + rank for s in range(world)]
        x = torch.arange(sum(send), dtype=torch.float32).reshape(-1, 1) + 1000 * rank
        y = yield ("a2av", x, send, recv)
        return got, y

    world = 3
    res = run_lockstep([gen(r, world) for r in range(world)])
    for rank, (got, y) in enumerate(res):
        assert got.tolist() == [[100 * r + 2 * rank, 100 * r + 2 * rank + 1] for r in range(world)]
        want = []
        for s in range(world):
            off = sum(s + 1 + d for d in range(rank))
            want += [1000 * s + off + i for i in range(s + 1 + rank)]
        assert y.flatten().tolist() == want
