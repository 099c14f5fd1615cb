"""The two CPU restatements of SPEC.md -- NumPy (oracle/edl_oracle.py) and plain C (oracle/edl_oracle_c.c) --
must agree bit for bit.  They share no code (np.unique / argsort vs a hand-written radix sort, sorted arrays +
searchsorted vs a hash map, ufuncs vs scalar loops), so agreement is evidence that both read the SPEC the same
way; it does not pin either to the reference, which holds nothing to pin to."""
import json
import os
import subprocess

import numpy as np
import pytest

from oracle import edl_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def OC():
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    from oracle import edl_oracle_c
    edl_oracle_c.load()
    return edl_oracle_c


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_splitmix64_and_init_rows(OC):
    # published SplitMix64 outputs for the seed-0 stream are mix(k * golden gamma): same values as tests/test_oracle.py
    for x in (0, 1, 0x9E3779B97F4A7C15, 2 ** 64 - 1, 1234567):
        assert OC.splitmix64(x) == int(O.splitmix64(np.array([x], dtype=np.uint64))[0])
    rng = np.random.default_rng(1)
    for dim in (4, 16, 64, 128):
        for t in (0, 3, 25):
            ids = rng.integers(0, 40_000_000, size=200, dtype=np.int64)
            assert np.array_equal(bits(OC.init_rows(O.DEFAULT_SEED, t, ids, dim)), bits(O.init_rows(O.DEFAULT_SEED, t, ids, dim)))
    for c in json.load(open(os.path.join(ROOT, "tests", "golden", "init_rows.json"))):
        got = OC.init_rows(c["seed"], c["t"], np.array(c["ids"], dtype=np.int64), c["dim"])
        assert np.array_equal(bits(got), np.array(c["bits"], dtype=np.uint32))


@pytest.mark.parametrize("case", ["random", "all_equal", "empty", "one", "wide", "negative", "zipf"])
def test_dedup_matches_numpy(OC, case):
    rng = np.random.default_rng(7)
    keys = {
        "random": rng.integers(0, 5000, size=20000, dtype=np.int64),
        "all_equal": np.full(777, 42, dtype=np.int64),
        "empty": np.zeros(0, dtype=np.int64),
        "one": np.array([9], dtype=np.int64),
        "wide": rng.integers(0, 2 ** 62, size=5000, dtype=np.int64),
        "negative": rng.integers(-2 ** 40, 2 ** 40, size=5000, dtype=np.int64),
        "zipf": O.fuse_keys(O.criteo_batch([100000] * 26, 512, 1.05)[1], O.key_bases([100000] * 26)),
    }[case]
    want, got = O.dedup(keys), OC.dedup(keys)
    for w, g in zip(want, got):
        assert w.dtype == g.dtype and np.array_equal(w, g)


@pytest.mark.parametrize("S", [1, 2, 3, 6, 8, 255])
def test_partition_matches_numpy(OC, S):
    rows = [10, 7, 40_000_000, 1, 12345]
    bases = O.key_bases(rows)
    rng = np.random.default_rng(S)
    keys = np.unique(rng.integers(0, bases[-1], size=4000, dtype=np.int64))
    keys = np.concatenate([keys, bases[:-1]])           # every table's id 0
    keys = np.unique(keys)
    for w, g in zip(O.partition_mod(keys, bases, S), OC.partition_mod(keys, bases, S)):
        assert np.array_equal(w, g)
    e = OC.partition_mod(np.zeros(0, dtype=np.int64), bases, S)
    assert len(e[0]) == 0 and np.all(e[1] == 0)


def test_segment_sum_matches_numpy(OC):
    rng = np.random.default_rng(3)
    keys = np.concatenate([rng.integers(0, 300, size=4000), np.full(2500, 5)]).astype(np.int64)   # one hot id
    grad = (rng.standard_normal((len(keys), 16)) * 10.0 ** rng.integers(-6, 3, size=(len(keys), 1))).astype(np.float32)
    _, _, _, order, seg = O.dedup(keys)
    assert np.array_equal(bits(OC.segment_sum_f64(grad, order, seg)), bits(O.segment_sum_f64_reduceat(grad, order, seg)))
    assert np.array_equal(bits(OC.segment_sum_f64(grad, order, seg)), bits(O.segment_sum_f64(grad, order, seg)))


def test_optimizer_rows_match_numpy(OC):
    rng = np.random.default_rng(5)
    n = 20000
    g = (rng.standard_normal(n) * 10.0 ** rng.integers(-8, 2, size=n)).astype(np.float32)
    g[:8] = [0.0, -0.0, 1e-38, -1e-38, 1e-45, 3e38, -3e38, 1.0]
    p = (rng.standard_normal(n) * 0.05).astype(np.float32)
    acc = np.abs(rng.standard_normal(n)).astype(np.float32) * np.float32(0.1)
    with np.errstate(all="ignore"):
        wp, wa = O.adagrad_rows(p, acc, g, 0.01, 1e-7)
        gp, ga = OC.adagrad_rows(p, acc, g, 0.01, 1e-7)
        assert np.array_equal(bits(wp), bits(gp)) and np.array_equal(bits(wa), bits(ga))
        m = (rng.standard_normal(n) * 1e-3).astype(np.float32)
        v = np.abs(rng.standard_normal(n) * 1e-5).astype(np.float32)
        for step in (1, 2, 10, 1000, 100000):
            lr_t = O.adam_lr_t(1e-3, 0.9, 0.999, step)
            assert bits(lr_t) == bits(OC.adam_lr_t(1e-3, 0.9, 0.999, step))
            want = O.adam_rows(p, m, v, g, lr_t, 0.9, 0.999, 1e-8)
            got = OC.adam_rows(p, m, v, g, lr_t, 0.9, 0.999, 1e-8)
            for w_, g_ in zip(want, got):
                assert np.array_equal(bits(w_), bits(g_))


@pytest.mark.parametrize("opt_name,threads", [("adagrad", 1), ("adagrad", 4), ("adam", 1), ("adam", 3)])
def test_step_trajectory_matches_numpy(OC, opt_name, threads):
    """Four steps of the whole path (A1 -> A3 -> A4 -> A5) on skewed ids that repeat across steps: every
    embedding of every step and the final materialised rows + slots + step counter are bit-identical."""
    rows, dim, B = [5000, 300, 70000, 1, 9999, 64], 16, 700
    opt = getattr(O.OptimizerSpec, opt_name)()
    ns, cs = O.OracleShard(rows, dim, opt), OC.CShard(rows, dim, opt)
    for s in range(4):
        ids = O.criteo_batch(rows, B, 1.05, step=s % 3)[1]
        grad = O.synthetic_grad(B * len(rows), dim, step=s)
        emb_n, _, _ = O.sparse_step(ns, O.fuse_keys(ids, ns.bases), grad)
        emb_c, n_u = cs.sparse_step(ids, grad, threads=threads)
        assert np.array_equal(bits(emb_n), bits(emb_c)), f"step {s}"
        assert n_u == len(np.unique(O.fuse_keys(ids, ns.bases)))
    kn, wn, sn, stn = ns.export()
    kc, wc, sc, stc = cs.export()
    assert stn == stc == 4
    assert np.array_equal(kn, kc) and np.array_equal(bits(wn), bits(wc))
    assert len(sn) == len(sc)
    for a, b in zip(sn, sc):
        assert np.array_equal(bits(a), bits(b))
    cs.close()


def test_c_steps_match_the_golden_trajectories(OC):
    """tests/golden/steps_*.json were written by the NumPy oracle (make_golden.py); the C restatement, given the
    same generator inputs, must land on the same rows."""
    import hashlib
    for name in ("adagrad", "adam"):
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", f"steps_{name}.json")))
        rows, dim, batch = gold["rows"], gold["dim"], gold["batch"]
        cs = OC.CShard(rows, dim, getattr(O.OptimizerSpec, name)())
        for s in range(gold["n_steps"]):
            ids = O.criteo_batch(rows, batch, gold["alpha"], step=s)[1]
            emb, _ = cs.sparse_step(ids, O.synthetic_grad(batch * len(rows), dim, step=s))
            assert hashlib.sha256(emb.tobytes()).hexdigest() == gold["emb_sha256"][s]
        keys, w, slots, step = cs.export()
        assert step == gold["step"] and keys.tolist() == gold["keys"]
        assert np.array_equal(bits(w), np.array(gold["w_bits"], dtype=np.uint32))
        for a, b in zip(slots, gold["slot_bits"]):
            assert np.array_equal(bits(a), np.array(b, dtype=np.uint32))
        cs.close()


@pytest.mark.parametrize("name", ["adagrad", "adam"])
def test_c_reshard_checkpoint_bytes_match_numpy_and_golden(OC, name, tmp_path):
    """SPEC A6 + 'Checkpoint file': the file of shard 1 of 6 written by the C restatement straight from the single
    shard is byte-identical to the NumPy oracle's after its 1 -> 4 -> 8 -> 6 re-shards (the committed sha256), and
    to the NumPy writer for every shard of S' = 1, 3, 6."""
    import hashlib
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", f"steps_{name}.json")))
    ck = json.load(open(os.path.join(ROOT, "tests", "golden", f"checkpoint_{name}.json")))
    rows, dim, batch = gold["rows"], gold["dim"], gold["batch"]
    opt = getattr(O.OptimizerSpec, name)()
    ns, cs = O.OracleShard(rows, dim, opt), OC.CShard(rows, dim, opt)
    for s in range(gold["n_steps"]):
        ids = O.criteo_batch(rows, batch, gold["alpha"], step=s)[1]
        g = O.synthetic_grad(batch * len(rows), dim, step=s)
        O.sparse_step(ns, O.fuse_keys(ids, ns.bases), g)
        cs.sparse_step(ids, g, want_emb=False)
    p = str(tmp_path / "c.ckpt")
    cs.write_checkpoint(p, ck["S"], ck["shard"])
    blob = open(p, "rb").read()
    assert len(blob) == ck["n_bytes"] and hashlib.sha256(blob).hexdigest() == ck["sha256"]
    for S in (1, 3, 6):
        for sh in O.reshard([ns], S):
            a, b = str(tmp_path / "n.ckpt"), str(tmp_path / "c2.ckpt")
            O.write_checkpoint(a, sh)
            cs.write_checkpoint(b, S, sh.shard)
            assert open(a, "rb").read() == open(b, "rb").read(), (S, sh.shard)
    cs.close()
