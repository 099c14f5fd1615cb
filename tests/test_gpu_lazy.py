"""Lazily materialised shards against the oracle, which is lazy by construction:
configs[2] at its FULL id space (26 tables x 40M rows x dim 128, Adam, 8 shards) on one GPU."""
import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-6, 1e-9


def dev(a, cuda):
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _close(got, want, what):
    err = np.abs(got.astype(np.float64) - want)
    assert np.all(err <= ATOL + RTOL * np.abs(want)), f"{what}: max err {err.max():.3e}"


@pytest.mark.parametrize("opt", ["adagrad", "adam"])
def test_lazy_single_shard_steps(cuda, opt):
    from easydl_b200 import kernels as K
    from easydl_b200.lazy_parameters import LazyShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    rows, dim, B = [10_000_000, 301, 40_000_000, 11], 32, 2048
    po = OptimizerSpec.adagrad() if opt == "adagrad" else OptimizerSpec.adam()
    oo = O.OptimizerSpec.adagrad() if opt == "adagrad" else O.OptimizerSpec.adam()
    P = LazyShardParameters(rows, dim, po, max_rows=20000, device=cuda)
    ora = O.OracleShard(rows, dim, oo)
    bases = O.key_bases(rows)
    fb = dev(bases[:len(rows)], cuda)
    for step in range(3):
        _, ids, _ = O.criteo_batch(rows, B, 1.05, step=step)
        keys = O.fuse_keys(ids, bases)
        grad = O.synthetic_grad(len(keys), dim, step=step)
        emb = P.lookup(dev(keys, cuda)).cpu().numpy()
        want, _, _ = O.sparse_step(ora, keys, grad)
        assert np.array_equal(bits(emb), bits(want)), f"lookup step {step}"   # single shard: exact
        uq = K.unique(dev(ids, cuda).view(-1), fb, P.layout.key_bits, field_bits=P.layout.field_bits)
        P.apply_gradients(uq, dev(grad, cuda))
    assert P.materialised() == len(ora.keys)
    rec = P.export_records().cpu()
    assert np.array_equal(rec.keys.numpy(), ora.keys) and rec.step == ora.step
    _close(rec.rows.numpy(), ora.w, "rows")
    for k in range(oo.n_slots):
        _close(rec.slots[k].numpy(), ora.slots[k], f"slot{k}")
    assert np.mean(bits(rec.rows.numpy()) == bits(ora.w)) > 0.999


def test_lazy_overflow_is_reported(cuda):
    from easydl_b200 import kernels as K
    from easydl_b200.lazy_parameters import LazyShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    P = LazyShardParameters([100000], 8, OptimizerSpec.adagrad(), max_rows=64, device=cuda)
    keys = torch.arange(500, dtype=torch.int64, device=cuda)
    uq = K.unique(keys, None, 17)
    P.apply_gradients(uq, torch.ones((500, 8), device=cuda))
    with pytest.raises(RuntimeError, match="max_rows"):
        P.materialised()


def test_cfg3_full_id_space_8_shards_and_reshard(cuda):
    """configs[2] shape at its real table size: 26 x 40M rows x dim 128, Adam, 8 shards (virtual ranks
    on one GPU, small per-rank batch), then the elastic re-shard 8 -> 6 of the lazy shards."""
    from easydl_b200.elastic import gen_reshard
    from easydl_b200.exchange import run_lockstep
    from easydl_b200.lazy_parameters import LazyShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.sharded_step import ShardedSparseStep
    rows, dim, S, B = [40_000_000] * 26, 128, 8, 512
    po, oo = OptimizerSpec.adam(), O.OptimizerSpec.adam()
    params = [LazyShardParameters(rows, dim, po, S, s, max_rows=65536, device=cuda) for s in range(S)]
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    bases = O.key_bases(rows)

    def train(params, osh, k0, n):
        st = [ShardedSparseStep(p, B) for p in params]
        for k in range(k0, k0 + n):
            kw, gw, gens = [], [], []
            for r in range(len(params)):
                _, ids, _ = O.criteo_batch(rows, B, 1.05, step=k * 8 + r)
                g = O.synthetic_grad(ids.size, dim, step=k * 8 + r)
                kw.append(O.fuse_keys(ids, bases)); gw.append(g)
                gens.append(st[r].gen_step(dev(ids, cuda), dev(g, cuda)))
            embs = run_lockstep(gens)
            want = O.sharded_step(osh, kw, gw)
            for r in range(len(params)):
                _close(embs[r].cpu().numpy(), want[r], f"emb step {k} rank {r}")

    train(params, osh, 0, 2)
    for p, o in zip(params, osh):
        rec = p.export_records().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys)
        _close(rec.rows.numpy(), o.w, "rows")
        _close(rec.slots[1].numpy(), o.slots[1], "v")
    # elastic: 8 -> 6 lazy shards
    before = {}
    for p in params:
        rec = p.export_records().cpu()
        for i, k in enumerate(rec.keys.tolist()):
            before[k] = rec.rows[i].numpy().copy()
    gens = [gen_reshard(params[r], 6, r, 8, rows=rows, dim=dim, opt=po, seed=O.DEFAULT_SEED, device=cuda,
                        factory=lambda s: LazyShardParameters(rows, dim, po, 6, s, max_rows=65536, device=cuda))
            for r in range(8)]
    new = run_lockstep(gens)[:6]
    osh = O.reshard(osh, 6)
    for p, o in zip(new, osh):
        rec = p.export_records().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys) and rec.step == o.step
        for i, k in enumerate(rec.keys.tolist()):
            assert np.array_equal(bits(rec.rows[i].numpy()), bits(before[k]))
    train(new, osh, 2, 1)


@pytest.mark.parametrize("opt,wire", [("adam", "mixed"), ("adagrad", torch.float64)])
def test_slot_step_on_lazy_shards(cuda, opt, wire):
    """The host-sync-free slot step on lazily materialised shards (what configs[2] runs on real ranks):
    4 virtual ranks, full 40M-row id space, against the oracle; then the slot-limit / row-limit check."""
    from easydl_b200.lazy_parameters import LazyShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.slot_step import SlotBuffers, SlotShardedStep, run_virtual
    rows, dim, S, B = [40_000_000] * 6, 32, 4, 1024
    po = OptimizerSpec.adam() if opt == "adam" else OptimizerSpec.adagrad()
    oo = O.OptimizerSpec.adam() if opt == "adam" else O.OptimizerSpec.adagrad()
    params = [LazyShardParameters(rows, dim, po, S, s, max_rows=1 << 15, device=cuda) for s in range(S)]
    N = B * len(rows)
    stride = min(N, int(1.5 * N / S) + 1024)
    bufs = [SlotBuffers(S, s, cuda, stride, dim, wire) for s in range(S)]
    for b in bufs:
        b.connect_local(bufs)
    st = [SlotShardedStep(p, B, b, wire_dtype=wire) for p, b in zip(params, bufs)]
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    bases = O.key_bases(rows)
    for k in range(3):
        kw, gw, di, dg = [], [], [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, B, 1.05, step=k * 8 + r)
            g = O.synthetic_grad(ids.size, dim, step=k * 8 + r)
            kw.append(O.fuse_keys(ids, bases)); gw.append(g)
            di.append(dev(ids, cuda)); dg.append(dev(g, cuda))
        embs = run_virtual(st, di, dg)
        want = O.sharded_step(osh, kw, gw)
        for r in range(S):
            _close(embs[r].cpu().numpy(), want[r], f"emb step {k} rank {r}")
    for s in st:
        s.check()
    for p, o in zip(params, osh):
        rec = p.export_records().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys)
        _close(rec.rows.numpy(), o.w, "rows")
        for i in range(oo.n_slots):
            _close(rec.slots[i].numpy(), o.slots[i], f"slot {i}")
    for b in bufs:
        b.close()
