"""Multi-shard parity on ONE GPU: S virtual ranks run in lock-step in one process
(easydl_b200.exchange.run_lockstep), each with its own shard slab, against the oracle's
single float64 pass over the global batch.  Also the elastic re-shard 4 -> 8 -> 6."""
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


def _make(cuda, rows, dim, S, opt_name):
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.sharded_step import ShardedSparseStep
    po = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    oo = O.OptimizerSpec.adagrad() if opt_name == "adagrad" else O.OptimizerSpec.adam()
    params = [ShardParameters(rows, dim, po, S, s, device=cuda) for s in range(S)]
    return params, po, oo


def _train(cuda, params, oshards, rows, dim, batch, steps, alpha, step0=0, fresh=None, wire=torch.float64):
    from easydl_b200.exchange import run_lockstep
    from easydl_b200.sharded_step import ShardedSparseStep
    S = len(params)
    fresh = (step0 == 0) if fresh is None else fresh
    bases = O.key_bases(rows)
    st = [ShardedSparseStep(p, batch, wire_dtype=wire) for p in params]
    for k in range(steps):
        kw, gw, gens = [], [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, batch, alpha, step=(step0 + k) * 8 + r)
            g = O.synthetic_grad(ids.size, dim, step=(step0 + k) * 8 + r)
            kw.append(O.fuse_keys(ids, bases))
            gw.append(g)
            gens.append(st[r].gen_step(dev(ids, cuda), dev(g, cuda)))
        embs = run_lockstep(gens)
        want = O.sharded_step(oshards, kw, gw)
        for r in range(S):
            got = embs[r].cpu().numpy()
            if fresh and k == 0:   # nothing updated yet: the lookup is pure data movement
                assert np.array_equal(bits(got), bits(want[r])), f"emb step {k} rank {r}"
            # later lookups read rows the GPU updated itself: per-worker fp32 partial sums are
            # merged by the owner, the oracle does one float64 pass (SPEC A4) -> tolerance
            err = np.abs(got.astype(np.float64) - want[r])
            assert np.all(err <= ATOL + RTOL * np.abs(want[r])), f"emb step {k} rank {r}: {err.max():.3e}"


def _compare_state(cuda, params, oshards, n_slots, exact_floor=0.98):
    for p, o in zip(params, oshards):
        rec = p.export_records().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys), "materialised key sets differ"
        assert rec.step == o.step
        for got, want, nm in [(rec.rows.numpy(), o.w, "rows")] + [
                (rec.slots[k].numpy(), o.slots[k], f"slot{k}") for k in range(n_slots)]:
            err = np.abs(got.astype(np.float64) - want)
            assert np.all(err <= ATOL + RTOL * np.abs(want)), f"{nm}: max err {err.max():.3e}"


@pytest.mark.parametrize("S,opt", [(2, "adagrad"), (4, "adam"), (6, "adagrad"), (8, "adam")])
def test_sharded_step_matches_global_oracle(cuda, S, opt):
    rows, dim = [5000, 301, 70001, 11, 6], 32
    params, po, oo = _make(cuda, rows, dim, S, opt)
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    _train(cuda, params, osh, rows, dim, 512, 3, 1.05)
    _compare_state(cuda, params, osh, oo.n_slots)


@pytest.mark.parametrize("S", [2, 8])
def test_fp32_gradient_wire_error_bound(cuda, S):
    """The optional fp32 wire (per-worker partial sums rounded to fp32 before the exchange, half the
    gradient bytes) is NOT inside the 1e-6 tolerance for ids that every worker hits hundreds of times
    (tables of 11 and 6 rows here): measured up to ~2e-6 relative with 8 workers.  This test pins the
    looser bound it does keep (1e-5), which is why float64 is the default wire."""
    rows, dim = [5000, 301, 70001, 11, 6], 64
    params, po, oo = _make(cuda, rows, dim, S, "adagrad")
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    from easydl_b200.exchange import run_lockstep
    from easydl_b200.sharded_step import ShardedSparseStep
    bases = O.key_bases(rows)
    st = [ShardedSparseStep(p, 1024, wire_dtype=torch.float32) for p in params]
    for k in range(3):
        kw, gw, gens = [], [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, 1024, 1.2, step=k * 8 + r)
            g = O.synthetic_grad(ids.size, dim, step=k * 8 + r)
            kw.append(O.fuse_keys(ids, bases)); gw.append(g)
            gens.append(st[r].gen_step(dev(ids, cuda), dev(g, cuda)))
        run_lockstep(gens)
        O.sharded_step(osh, kw, gw)
    for p, o in zip(params, osh):
        rec = p.export_records().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys)
        np.testing.assert_allclose(rec.rows.numpy(), o.w, rtol=1e-5, atol=1e-8)
        np.testing.assert_allclose(rec.slots[0].numpy(), o.slots[0], rtol=1e-5, atol=1e-8)


def test_cfg3_shape_reduced_rows(cuda):
    """BASELINE.json configs[2] shape (26 tables, dim 128, Adam, 8 shards) with table rows
    reduced from 40M to 40K so 8 virtual shards fit beside each other on one GPU."""
    rows, dim, S = [40_000] * 26, 128, 8
    params, po, oo = _make(cuda, rows, dim, S, "adam")
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    _train(cuda, params, osh, rows, dim, 256, 2, 1.05)
    _compare_state(cuda, params, osh, 2)


@pytest.mark.parametrize("opt", ["adagrad", "adam"])
def test_elastic_reshard_4_8_6_bit_exact(cuda, opt):
    """configs[3]: train at S=4, re-shard to 8, train, re-shard to 6, train.  The re-shard must
    move every row, slot and the step bit-exactly (checked against the GPU's own pre-move
    state), and the whole trajectory must stay within tolerance of the oracle's."""
    from easydl_b200.elastic import gen_reshard
    from easydl_b200.exchange import run_lockstep
    rows, dim, W = [5000, 301, 70001, 11, 6], 16, 8
    params, po, oo = _make(cuda, rows, dim, 4, opt)
    osh = [O.OracleShard(rows, dim, oo, 4, s) for s in range(4)]
    _train(cuda, params, osh, rows, dim, 256, 2, 1.05, step0=0)
    cur, step0 = params, 2
    for new_S in (8, 6):
        before = {}
        for p in cur:
            rec = p.export_records().cpu()
            for k, r, *sl in zip(rec.keys.tolist(), rec.rows.numpy(), *[s.numpy() for s in rec.slots]):
                before[k] = (r.copy(), [x.copy() for x in sl])
        gens = [gen_reshard(cur[r] if r < len(cur) else None, new_S, r, W, rows=rows, dim=dim, opt=po,
                            seed=O.DEFAULT_SEED, device=cuda) for r in range(W)]
        out = run_lockstep(gens)
        assert all(o is None for o in out[new_S:])
        cur = out[:new_S]
        osh = O.reshard(osh, new_S)
        seen = 0
        for p, o in zip(cur, osh):
            rec = p.export_records().cpu()
            assert np.array_equal(rec.keys.numpy(), o.keys)
            assert rec.step == o.step == step0
            for i, k in enumerate(rec.keys.tolist()):
                r0, s0 = before[k]
                assert np.array_equal(bits(rec.rows[i].numpy()), bits(r0))
                for kk in range(oo.n_slots):
                    assert np.array_equal(bits(rec.slots[kk][i].numpy()), bits(s0[kk]))
            seen += len(rec)
        assert seen == len(before)
        _train(cuda, cur, osh, rows, dim, 256, 2, 1.05, step0=step0)
        step0 += 2
    _compare_state(cuda, cur, osh, oo.n_slots)


@pytest.mark.parametrize("opt", ["adagrad", "adam"])
def test_checkpoint_reshard_files(cuda, opt, tmp_path):
    """File form of A6: save 4 shard files, restore into 6 and into 3 shards, compare with the
    oracle's restore of the oracle's files and with the pre-save GPU state bit for bit."""
    from easydl_b200 import checkpoint as ck
    from easydl_b200.parameters import ShardParameters
    rows, dim = [5000, 301, 70001, 11, 6], 16
    params, po, oo = _make(cuda, rows, dim, 4, opt)
    osh = [O.OracleShard(rows, dim, oo, 4, s) for s in range(4)]
    _train(cuda, params, osh, rows, dim, 256, 2, 1.2)
    saver = ck.CheckpointSaver(str(tmp_path))
    files = [saver.save(p) for p in params]
    assert ck.list_versions(str(tmp_path)) == [2]
    before = {}
    for p in params:
        rec = p.export_records().cpu()
        for i, k in enumerate(rec.keys.tolist()):
            before[k] = (rec.rows[i].numpy().copy(), [s[i].numpy().copy() for s in rec.slots])
    # the files parse with the oracle's reader and hold the same key sets
    for f, o in zip(files, osh):
        meta = O.read_checkpoint(f)
        assert meta["S"] == 4 and meta["step"] == 2
        assert sum(len(t[2]) for t in meta["tables"]) == len(o.keys)
    for new_S in (6, 3):
        tot = 0
        for s in range(new_S):
            q = ShardParameters(rows, dim, po, new_S, s, device=cuda)
            assert ck.restore_params_from_checkpoint(str(tmp_path), q) == 2
            rec = q.export_records().cpu()
            want = O.reshard(osh, new_S)[s]
            assert np.array_equal(rec.keys.numpy(), want.keys)
            for i, k in enumerate(rec.keys.tolist()):
                assert np.array_equal(bits(rec.rows[i].numpy()), bits(before[k][0]))
                for kk in range(oo.n_slots):
                    assert np.array_equal(bits(rec.slots[kk][i].numpy()), bits(before[k][1][kk]))
            tot += len(rec)
        assert tot == len(before)


@pytest.mark.parametrize("kind,opt", [("dense", "adagrad"), ("dense", "adam"), ("lazy", "adam")])
def test_failed_ps_shard_recovers_from_its_checkpoint(cuda, tmp_path, kind, opt):
    """ref:README.md:27 "recover failed parameter servers ... and resume the training": three PS shards
    train, checkpoint at step 2, train on; shard 1 then dies (its device state is gone).  A fresh shard 1
    restores from the newest COMPLETE version, the survivors roll back to the same version, training
    resumes from there -- and the whole trajectory still matches the oracle, which took no failure."""
    import copy
    from easydl_b200 import checkpoint as ck
    from easydl_b200.lazy_parameters import LazyShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    rows, dim, S, B = [5000, 301, 70001, 11, 6], 16, 3, 256
    po = OptimizerSpec.adagrad() if opt == "adagrad" else OptimizerSpec.adam()
    oo = O.OptimizerSpec.adagrad() if opt == "adagrad" else O.OptimizerSpec.adam()

    def make(s):
        if kind == "lazy":
            return LazyShardParameters(rows, dim, po, S, s, max_rows=1 << 14, device=cuda)
        return ShardParameters(rows, dim, po, S, s, device=cuda)
    params = [make(s) for s in range(S)]
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    _train(cuda, params, osh, rows, dim, B, 2, 1.05, step0=0)
    saver = ck.CheckpointSaver(str(tmp_path), keep_max=2)
    for p in params:
        saver.save(p)
    osh_at_ckpt = copy.deepcopy(osh)
    _train(cuda, params, osh, rows, dim, B, 1, 1.05, step0=2, fresh=False)          # progress that will be lost
    # a version that only shard 0 managed to write before the failure: must be skipped by the restore
    ck.write_shard_file(ck.shard_file(str(tmp_path), 3, 0, S), params[0].layout, dim, po.kind, params[0].export_records().cpu())
    params[1] = None                                                                  # the PS Pod is gone
    params[1] = make(1)
    assert ck.latest_complete_version(str(tmp_path)) == 2
    for p in params:
        assert ck.restore_params_from_checkpoint(str(tmp_path), p) == 2
    osh = osh_at_ckpt
    _compare_state(cuda, params, osh, oo.n_slots)
    _train(cuda, params, osh, rows, dim, B, 2, 1.05, step0=2, fresh=False)
    _compare_state(cuda, params, osh, oo.n_slots)
