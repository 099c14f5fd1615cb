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


def _lazy_trained(cuda, max_rows, steps=3, opt="adam"):
    from easydl_b200.lazy_parameters import LazyShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200 import kernels as K
    rows, dim, B = [3_000_000] * 4, 16, 512
    po = OptimizerSpec.adam() if opt == "adam" else OptimizerSpec.adagrad()
    P = LazyShardParameters(rows, dim, po, max_rows=max_rows, device=cuda)
    bases = O.key_bases(rows)
    batches = []
    for k in range(steps):
        _, ids, _ = O.criteo_batch(rows, B, 1.05, step=k)
        g = O.synthetic_grad(ids.size, dim, step=k)
        batches.append((dev(O.fuse_keys(ids, bases).reshape(-1), cuda), dev(g, cuda)))
    return P, batches, K


def _apply(P, K, keys, g):
    uq = K.unique(keys, None, P.layout.key_bits, want_inverse=False, want_counts=False)
    P.apply_gradients(uq, g)


def test_lazy_shard_grows_without_changing_results(cuda):
    """8(f).3 growth: a shard that runs out of room is re-housed in a larger slab between steps; rows, slots,
    stamps and the step move, and training continues bit-identically to a shard that was large from the start."""
    big, batches, K = _lazy_trained(cuda, 1 << 15)
    small, _, _ = _lazy_trained(cuda, 1536)
    for keys, g in batches[:2]:
        _apply(big, K, keys, g)
        small.ensure_room(keys.numel())           # grows when the next batch might not fit
        _apply(small, K, keys, g)
    assert small.max_rows > 1536
    small.grow(small.max_rows * 2)
    for keys, g in batches[2:]:
        _apply(big, K, keys, g)
        _apply(small, K, keys, g)
    a, b = big.export_records(), small.export_records()
    assert torch.equal(a.keys, b.keys) and torch.equal(a.rows, b.rows) and a.step == b.step
    for x, y in zip(a.slots, b.slots):
        assert torch.equal(x, y)


def test_lazy_shard_eviction_and_return(cuda):
    """8(f).3 eviction: rows not updated since a cut-off step leave the slab and are handed back; the rest
    stays put; re-importing the evicted records restores the shard bit for bit."""
    P, batches, K = _lazy_trained(cuda, 1 << 15)
    for keys, g in batches:
        _apply(P, K, keys, g)
    before = P.export_records()
    n0 = P.materialised()
    gone = P.evict_older_than(3)                  # keep only what step 3 touched
    assert 0 < len(gone) < n0 and P.materialised() == n0 - len(gone) and gone.step == before.step
    kept = P.export_records()
    both = torch.cat([kept.keys, gone.keys]).sort().values
    assert torch.equal(both, before.keys)
    # an evicted key reads as its initial value again (lossy unless re-imported)
    fresh = P.lookup(gone.keys[:8].contiguous())
    t, ids = O.split_key(gone.keys[:8].cpu().numpy(), O.key_bases([3_000_000] * 4))
    want = np.stack([O.init_rows(O.DEFAULT_SEED, int(tt), np.array([ii]), 16)[0] for tt, ii in zip(t, ids)])
    assert np.array_equal(bits(fresh.cpu().numpy()), bits(want))
    P.import_records(gone)
    after = P.export_records()
    assert torch.equal(after.keys, before.keys) and torch.equal(after.rows, before.rows)
    for x, y in zip(after.slots, before.slots):
        assert torch.equal(x, y)


def test_lazy_shard_raw_checkpoint_restores_index_and_slab(cuda, tmp_path):
    """8(f).3 index checkpoint: save_raw / load_raw bring back the hash index, the slab prefix, the stamps and
    the step as they lay -- no key is re-inserted -- and training continues bit-identically."""
    from easydl_b200.lazy_parameters import LazyShardParameters
    P, batches, K = _lazy_trained(cuda, 1 << 14)
    for keys, g in batches[:2]:
        _apply(P, K, keys, g)
    path = str(tmp_path / "shard0.raw")
    P.save_raw(path)
    Q = LazyShardParameters([3_000_000] * 4, 16, P.opt, max_rows=1 << 12, device=cuda)   # other size: re-allocated
    Q.load_raw(path)
    assert Q.step == P.step and Q.materialised() == P.materialised()
    assert torch.equal(Q.index, P.index) and torch.equal(Q.stamp[:P.materialised()], P.stamp[:P.materialised()])
    keys, g = batches[2]
    _apply(P, K, keys, g)
    _apply(Q, K, keys, g)
    a, b = P.export_records(), Q.export_records()
    assert torch.equal(a.keys, b.keys) and torch.equal(a.rows, b.rows)
    for x, y in zip(a.slots, b.slots):
        assert torch.equal(x, y)
    assert torch.equal(P.lookup(keys), Q.lookup(keys))
