"""Property tests of the oracle: with no reference code or vectors to pin it,
the restatement is held to the identities its definitions imply, over generated inputs."""
import numpy as np
from hypothesis import given, settings, strategies as st

from oracle import edl_oracle as O

SET = settings(max_examples=40, deadline=None)

tables = st.lists(st.integers(min_value=1, max_value=300), min_size=1, max_size=5)


@st.composite
def batches(draw):
    rows = draw(tables)
    bases = O.key_bases(rows)
    n = draw(st.integers(min_value=0, max_value=200))
    t = np.array(draw(st.lists(st.integers(0, len(rows) - 1), min_size=n, max_size=n)), dtype=np.int64)
    u = np.array(draw(st.lists(st.floats(0, 1, exclude_max=True), min_size=n, max_size=n)), dtype=np.float64)
    ids = np.floor(u * np.array(rows, dtype=np.float64)[t]).astype(np.int64) if n else np.zeros(0, np.int64)
    keys = (bases[t] + ids).astype(np.int64) if n else np.zeros(0, np.int64)
    return rows, bases, keys


@SET
@given(batches())
def test_dedup_identities(b):
    rows, bases, keys = b
    uniq, inverse, counts, order, seg = O.dedup(keys)
    assert np.all(np.diff(uniq) > 0)                                   # ascending, no repeats
    assert np.array_equal(uniq[inverse], keys)                         # inverse reconstructs the input
    assert counts.sum() == keys.size and seg[-1] == keys.size and seg[0] == 0
    assert np.array_equal(np.diff(seg), counts)
    assert np.array_equal(np.sort(order), np.arange(keys.size))        # order is a permutation ...
    assert np.array_equal(keys[order], np.sort(keys, kind="stable"))   # ... that sorts
    for u in range(len(uniq)):                                         # ... stably: positions ascend per id
        seg_pos = order[seg[u]:seg[u + 1]]
        assert np.all(np.diff(seg_pos) > 0) and np.all(keys[seg_pos] == uniq[u])


@SET
@given(batches(), st.integers(min_value=1, max_value=9))
def test_partition_identities(b, S):
    rows, bases, keys = b
    uniq = O.dedup(keys)[0]
    grouped, offsets, perm = O.partition_mod(uniq, bases, S)
    assert offsets[0] == 0 and offsets[-1] == uniq.size and np.all(np.diff(offsets) >= 0)
    assert np.array_equal(np.sort(perm), np.arange(uniq.size)) and np.array_equal(grouped, uniq[perm])
    _, ids = O.split_key(grouped, bases)
    for s in range(S):
        g = grouped[offsets[s]:offsets[s + 1]]
        assert np.all(ids[offsets[s]:offsets[s + 1]] % S == s)          # per-TABLE id mod S, not key mod S
        assert np.all(np.diff(g) > 0)                                   # stable: ascending inside a group


@SET
@given(batches(), st.integers(1, 6), st.integers(1, 6), st.sampled_from(["adagrad", "adam"]))
def test_reshard_round_trip_and_shard_independence(b, S, S2, opt_name):
    rows, bases, keys = b
    dim = 4
    opt = O.OptimizerSpec.adagrad() if opt_name == "adagrad" else O.OptimizerSpec.adam()
    grad = O.synthetic_grad(keys.size, dim, step=1)
    one = [O.OracleShard(rows, dim, opt, 1, 0)]
    many = [O.OracleShard(rows, dim, opt, S, s) for s in range(S)]
    e1 = O.sharded_step(one, [keys], [grad])[0]
    eS = O.sharded_step(many, [keys], [grad])[0]
    assert np.array_equal(e1.view(np.uint32), eS.view(np.uint32))      # lookup does not depend on S
    e1 = O.sharded_step(one, [keys], [grad])[0]                         # second step reads updated rows
    eS = O.sharded_step(many, [keys], [grad])[0]
    assert np.array_equal(e1.view(np.uint32), eS.view(np.uint32))      # one global f64 pass: update neither
    there = O.reshard(many, S2)
    back = O.reshard(there, S)
    for a, c in zip(many, back):                                        # S -> S' -> S is the identity, bit for bit
        assert np.array_equal(a.keys, c.keys) and a.step == c.step
        assert np.array_equal(a.w.view(np.uint32), c.w.view(np.uint32))
        for x, y in zip(a.slots, c.slots):
            assert np.array_equal(x.view(np.uint32), y.view(np.uint32))
    _, ids = O.split_key(np.concatenate([s.keys for s in there]) if there else np.zeros(0, np.int64), bases)
    own = np.concatenate([np.full(len(s.keys), i) for i, s in enumerate(there)]) if there else np.zeros(0)
    assert np.array_equal(ids % S2, own)


@SET
@given(st.integers(0, 2 ** 63 - 1), st.integers(0, 40), st.integers(0, 10 ** 9), st.integers(1, 16))
def test_init_rows_is_a_pure_function_in_range(seed, t, row, dim4):
    ids = np.array([row, row + 1, row], dtype=np.int64)
    a = O.init_rows(seed, t, ids, dim4 * 4)
    assert a.dtype == np.float32 and a.shape == (3, dim4 * 4)
    assert np.all(a >= np.float32(-0.05)) and np.all(a < np.float32(0.05))
    assert np.array_equal(a[0].view(np.uint32), a[2].view(np.uint32))   # same (seed, t, id) -> same bits
    b = O.init_rows(seed, t, ids[:1], dim4 * 4)
    assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32))   # independent of the batch around it
