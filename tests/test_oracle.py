"""CPU tests (no GPU): the oracle against published known answers, independent
implementations (torch.unique, torch.optim.Adagrad, float64 re-evaluation), its own golden
fixtures, and size-independent properties."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    with open(os.path.join(G, name)) as f:
        return json.load(f)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_splitmix64_published_vectors():
    # First three outputs of SplitMix64 seeded with 0 (Vigna's reference splitmix64.c).
    g = 0x9E3779B97F4A7C15
    xs = np.array([0, g, (2 * g) & (2 ** 64 - 1)], dtype=np.uint64)
    want = [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]
    assert [int(v) for v in O.splitmix64(xs)] == want


def test_init_rows_range_and_golden():
    w = O.init_rows(O.DEFAULT_SEED, 3, np.arange(1000), 16)
    assert w.dtype == np.float32 and w.min() >= -0.05 and w.max() < 0.05
    assert abs(float(w.mean())) < 2e-3
    for c in load("init_rows.json"):
        got = O.init_rows(c["seed"], c["t"], np.array(c["ids"]), c["dim"])
        assert bits(got).tolist() == np.array(c["bits"], dtype=np.uint32).tolist()


def test_init_is_pure_and_shard_independent():
    rows = [100, 33]
    a = O.OracleShard(rows, 8, O.OptimizerSpec.adagrad(), 1, 0)
    keys = np.arange(133)
    full = a.lookup(keys)
    for S in (2, 3, 6, 8):
        for s in range(S):
            sh = O.OracleShard(rows, 8, O.OptimizerSpec.adagrad(), S, s)
            mine = keys[sh.owns(keys)]
            assert np.array_equal(bits(sh.lookup(mine)), bits(full[mine]))


def test_dedup_matches_torch_unique():
    rng = np.random.default_rng(0)
    for n, hi in [(1, 5), (1000, 10), (5000, 1 << 40), (4096, 300)]:
        keys = rng.integers(0, hi, n)
        u, inv, cnt, order, seg = O.dedup(keys)
        tu, tinv, tcnt = torch.unique(torch.from_numpy(keys), sorted=True, return_inverse=True, return_counts=True)
        assert np.array_equal(u, tu.numpy()) and np.array_equal(inv, tinv.numpy()) and np.array_equal(cnt, tcnt.numpy())
        assert np.array_equal(keys[order], np.sort(keys, kind="stable"))
        assert np.array_equal(u[inv], keys)
        assert seg[-1] == n and np.array_equal(np.diff(seg), cnt)
        for k in range(len(u)):  # stable: positions ascending inside a segment
            assert np.all(np.diff(order[seg[k]:seg[k + 1]]) > 0)


def test_dedup_partition_golden():
    g = load("dedup_partition.json")
    u, inv, cnt, order, seg = O.dedup(np.array(g["keys"]))
    assert u.tolist() == g["uniq"] and inv.tolist() == g["inverse"] and cnt.tolist() == g["counts"]
    assert order.tolist() == g["order"] and seg.tolist() == g["seg_offsets"]
    bases = O.key_bases(g["rows"])
    for S, p in g["partition"].items():
        pk, off, perm = O.partition_mod(u, bases, int(S))
        assert pk.tolist() == p["keys"] and off.tolist() == p["offsets"] and perm.tolist() == p["perm"]


def test_partition_uses_per_table_id_not_fused_key():
    rows = [7, 10]  # key_base = [0, 7]: 7 mod 6 != 0, so fused-key mod and per-table mod differ
    bases = O.key_bases(rows)
    keys = np.array([7, 8, 13])  # table 1 ids 0, 1, 6
    pk, off, perm = O.partition_mod(keys, bases, 6)
    assert off.tolist() == [0, 2, 3, 3, 3, 3, 3] and pk.tolist() == [7, 13, 8]


def test_segment_sum_forms_agree_and_beat_fp32_order():
    rng = np.random.default_rng(1)
    keys = rng.integers(0, 40, 20000)
    grad = O.synthetic_grad(20000, 16)
    _, _, _, order, seg = O.dedup(keys)
    a = O.segment_sum_f64(grad, order, seg)
    b = O.segment_sum_f64_reduceat(grad, order, seg)
    assert np.array_equal(a, b)
    exact = np.stack([grad[order[seg[k]:seg[k + 1]]].astype(np.float64).sum(0) for k in range(len(seg) - 1)])
    assert np.array_equal(a, exact.astype(np.float32))


def test_adagrad_matches_torch_optim():
    rng = np.random.default_rng(2)
    p0 = rng.standard_normal((50, 8)).astype(np.float32) * 0.05
    gs = [rng.standard_normal((50, 8)).astype(np.float32) * 0.01 for _ in range(4)]
    p, acc = p0.copy(), np.full_like(p0, 0.1)
    for g in gs:
        p, acc = O.adagrad_rows(p, acc, g, 0.01, 1e-7)
    tp = torch.nn.Parameter(torch.from_numpy(p0.copy()))
    opt = torch.optim.Adagrad([tp], lr=0.01, eps=1e-7, initial_accumulator_value=0.1)
    for g in gs:
        tp.grad = torch.from_numpy(g.copy())
        opt.step()
    np.testing.assert_allclose(p, tp.detach().numpy(), rtol=1e-6, atol=1e-9)


def test_adam_matches_float64_evaluation():
    rng = np.random.default_rng(3)
    p0 = rng.standard_normal((40, 4)).astype(np.float32) * 0.05
    p, m, v = p0.copy(), np.zeros_like(p0), np.zeros_like(p0)
    p64, m64, v64 = p0.astype(np.float64), np.zeros((40, 4)), np.zeros((40, 4))
    for t in range(1, 6):
        g = rng.standard_normal((40, 4)).astype(np.float32) * 0.01
        lr_t = O.adam_lr_t(1e-3, 0.9, 0.999, t)
        p, m, v = O.adam_rows(p, m, v, g, lr_t, 0.9, 0.999, 1e-8)
        m64 = 0.9 * m64 + 0.1 * g
        v64 = 0.999 * v64 + 0.001 * g.astype(np.float64) ** 2
        p64 = p64 - 1e-3 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t) * m64 / (np.sqrt(v64) + 1e-8)
    np.testing.assert_allclose(p, p64, rtol=1e-5, atol=1e-7)  # fp32 vs float64 evaluation: loose


@pytest.mark.parametrize("name", ["adagrad", "adam"])
def test_steps_and_reshard_golden(name, tmp_path):
    g = load(f"steps_{name}.json")
    opt = O.OptimizerSpec.adagrad() if name == "adagrad" else O.OptimizerSpec.adam()
    rows, dim = g["rows"], g["dim"]
    bases = O.key_bases(rows)
    sh = O.OracleShard(rows, dim, opt)
    for step in range(g["n_steps"]):
        _, ids, _ = O.criteo_batch(rows, g["batch"], g["alpha"], step=step)
        k = O.fuse_keys(ids, bases)
        emb, _, _ = O.sparse_step(sh, k, O.synthetic_grad(len(k), dim, step=step))
        assert hashlib.sha256(emb.tobytes()).hexdigest() == g["emb_sha256"][step]
    assert sh.keys.tolist() == g["keys"] and sh.step == g["step"]
    assert bits(sh.w).tolist() == np.array(g["w_bits"], dtype=np.uint32).tolist()
    for k in range(opt.n_slots):
        assert bits(sh.slots[k]).tolist() == np.array(g["slot_bits"][k], dtype=np.uint32).tolist()
    s4 = O.reshard([sh], 4)
    s8 = O.reshard(s4, 8)
    s6 = O.reshard(s8, 6)
    assert [len(s.keys) for s in s4] == g["shard_sizes"]["4"]
    assert [len(s.keys) for s in s8] == g["shard_sizes"]["8"]
    assert [s.keys.tolist() for s in s6] == g["shard6_keys"]
    back = O.reshard(s6, 1)[0]  # 1 -> 4 -> 8 -> 6 -> 1 is the identity, bit for bit
    assert np.array_equal(back.keys, sh.keys) and np.array_equal(bits(back.w), bits(sh.w))
    for k in range(opt.n_slots):
        assert np.array_equal(bits(back.slots[k]), bits(sh.slots[k]))
    ck = load(f"checkpoint_{name}.json")
    p = str(tmp_path / "s.ckpt")
    O.write_checkpoint(p, s6[1])
    blob = open(p, "rb").read()
    assert len(blob) == ck["n_bytes"] and hashlib.sha256(blob).hexdigest() == ck["sha256"]
    # file form of the re-shard: 6 files -> shard 2 of 4
    paths = []
    for s in s6:
        q = str(tmp_path / f"v-{s.shard}.ckpt")
        O.write_checkpoint(q, s)
        paths.append(q)
    r = O.restore_from_checkpoints(paths, rows, dim, opt, 4, 2)
    assert np.array_equal(r.keys, s4[2].keys) and np.array_equal(bits(r.w), bits(s4[2].w)) and r.step == sh.step


def test_sharded_update_equals_single_shard_within_tolerance():
    rows, dim = [500, 40, 3000], 8
    bases = O.key_bases(rows)
    one = O.OracleShard(rows, dim, O.OptimizerSpec.adagrad())
    S = 3
    many = [O.OracleShard(rows, dim, O.OptimizerSpec.adagrad(), S, s) for s in range(S)]
    for step in range(3):
        kw, gw = [], []
        for w in range(2):
            _, ids, _ = O.criteo_batch(rows, 128, 1.05, step=step * 2 + w)
            kw.append(O.fuse_keys(ids, bases))
            gw.append(O.synthetic_grad(len(kw[-1]), dim, step=step * 2 + w))
        embs = O.sharded_step(many, kw, gw)
        e1, _, _ = O.sparse_step(one, np.concatenate(kw), np.concatenate(gw))
        assert np.array_equal(bits(np.concatenate(embs)), bits(e1))  # lookups bit-exact, any S
    merged = O.reshard(many, 1)[0]
    assert np.array_equal(merged.keys, one.keys)
    assert np.array_equal(bits(merged.w), bits(one.w))  # one float64 pass per id: independent of S


def test_per_table_step_equals_fused_step():
    rows, dim = [500, 40, 3000], 8
    bases = O.key_bases(rows)
    a = O.OracleShard(rows, dim, O.OptimizerSpec.adam())
    b = O.OracleShard(rows, dim, O.OptimizerSpec.adam())
    for step in range(2):
        _, ids, _ = O.criteo_batch(rows, 200, 1.2, step=step)
        g = O.synthetic_grad(ids.size, dim, step=step)
        e1, _, _ = O.sparse_step(a, O.fuse_keys(ids, bases), g)
        e2, _ = O.sparse_step_per_table(b, ids, g)
        assert np.array_equal(bits(e1), bits(e2))
    assert np.array_equal(a.keys, b.keys) and np.array_equal(bits(a.w), bits(b.w))


def test_zipf_generator_golden_and_bounds():
    g = load("criteo_ids.json")
    ids = O.criteo_batch(g["rows"], g["batch"], g["alpha"], g["seed"], g["step"])[1]
    assert ids.tolist() == g["ids"]
    assert O.criteo_batch(g["rows"], g["batch"], 1.0, g["seed"], 0)[1].tolist() == g["ids_alpha1"]
    big = O.criteo_batch([1000, 7], 5000, 0.5)[1]
    assert big.min() >= 0 and big[:, 0].max() < 1000 and big[:, 1].max() < 7
