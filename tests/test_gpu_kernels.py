"""GPU parity tests, kernel by kernel, through the C ABI (ctypes) against the NumPy oracle.

Bars: bit-exact for dedup / partition / gather / expand / init / re-shard; optimizer-updated
rows within 1e-6 relative (absolute floor 1e-9) as BASELINE.json's north_star states -- the
tests also report bit-exactness, which the explicit-rounding kernels are expected to reach.
"""
import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-6, 1e-9


def dev(a, cuda):
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


def assert_close(got, want, what):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    err = np.abs(got - want)
    bound = ATOL + RTOL * np.abs(want)
    bad = err > bound
    assert not bad.any(), f"{what}: {bad.sum()} of {bad.size} beyond 1e-6 rel; max err {err.max():.3e}"


# ------------------------------------------------------------------------------- init
@pytest.mark.parametrize("rows,dim,S", [([50, 7, 33], 16, 1), ([1000, 1, 257, 64], 64, 3),
                                        ([97, 101], 8, 6), ([300], 128, 8), ([40, 9], 256, 2)])
def test_init_rows_bit_exact(cuda, rows, dim, S):
    from easydl_b200 import kernels as K
    from easydl_b200.layout import ShardLayout
    for shard in range(S):
        lay = ShardLayout(rows, S, shard)
        slab = torch.empty((max(lay.local_rows, 1), dim), dtype=torch.float32, device=cuda)
        K.init_rows(slab, lay.to_device(cuda), O.DEFAULT_SEED, 0, lay.local_rows)
        got = slab.cpu().numpy()
        for t, r in enumerate(rows):
            ids = np.arange(shard, r, S, dtype=np.int64)
            want = O.init_rows(O.DEFAULT_SEED, t, ids, dim)
            b = int(lay.local_base[t])
            assert np.array_equal(got[b:b + len(ids)].view(np.uint32), want.view(np.uint32)), (t, shard)


# ------------------------------------------------------------------------------- A1
def _check_unique(cuda, keys_np, field_base=None, key_bits=64, ids_np=None):
    from easydl_b200 import kernels as K
    src = keys_np if ids_np is None else ids_np
    fb = None if field_base is None else dev(field_base, cuda)
    r = K.unique(dev(src, cuda).view(-1), fb, key_bits)
    U = r.count()
    uniq, inv, cnt, order, seg = O.dedup(keys_np)
    assert U == len(uniq)
    assert np.array_equal(r.uniq[:U].cpu().numpy(), uniq)
    assert np.array_equal(r.inverse.cpu().numpy(), inv)
    assert np.array_equal(r.counts[:U].cpu().numpy(), cnt)
    assert np.array_equal(r.order.cpu().numpy(), order)
    assert np.array_equal(r.seg_offsets[:U + 1].cpu().numpy(), seg)
    assert np.array_equal(r.seg_first[:U].cpu().numpy(), order[seg[:-1]])


@pytest.mark.parametrize("n,hi,bits", [(1, 10, 8), (2, 2, 1), (31, 5, 3), (2048, 300, 9), (2049, 1 << 20, 20),
                                       (100000, 1 << 27, 28), (300001, 1000, 32), (50000, 1 << 40, 41),
                                       (70000, (1 << 62), 63)])
def test_unique_matches_np_unique(cuda, n, hi, bits):
    rng = np.random.default_rng(n + bits)
    keys = rng.integers(0, hi, size=n, dtype=np.int64)
    _check_unique(cuda, keys, key_bits=bits)


def test_unique_edge_cases(cuda):
    from easydl_b200 import kernels as K
    # empty input
    r = K.unique(torch.empty(0, dtype=torch.int64, device=cuda), None, 16)
    assert r.count() == 0 and int(r.seg_offsets[0].item()) == 0
    # all equal, already sorted, reverse sorted, max 32-bit key
    _check_unique(cuda, np.full(5000, 77, dtype=np.int64), key_bits=7)
    _check_unique(cuda, np.arange(9000, dtype=np.int64), key_bits=14)
    _check_unique(cuda, np.arange(9000, dtype=np.int64)[::-1].copy(), key_bits=14)
    _check_unique(cuda, np.array([2 ** 32 - 1, 0, 2 ** 32 - 1, 5], dtype=np.int64), key_bits=32)
    _check_unique(cuda, np.array([2 ** 63 - 1, 0, 2 ** 63 - 1, 5], dtype=np.int64), key_bits=64)


def test_unique_fused_fields_equals_per_table(cuda):
    rows = [1000, 50, 7777, 3]
    src = __import__("easydl_b200.criteo", fromlist=["CriteoSynthetic"]).CriteoSynthetic(rows, 4096, 1.05)
    ids = src.ids(0)
    bases = O.key_bases(rows)
    keys = O.fuse_keys(ids, bases)
    _check_unique(cuda, keys, field_base=bases[:len(rows)], key_bits=int(bases[-1] - 1).bit_length(), ids_np=ids)
    # per-table dedups concatenated == fused dedup (SPEC A1)
    uniq = O.dedup(keys)[0]
    cat = np.concatenate([np.unique(ids[:, t]) + bases[t] for t in range(len(rows))])
    assert np.array_equal(uniq, cat)


@pytest.mark.parametrize("impl,shape", [(2, 0), (2, 1), (1, 0), (1, 1), (0, 0)])
@pytest.mark.parametrize("B,rows", [(1, [5, 9]), (37, [1000, 50, 7777, 3]), (2048, [300, 300, 70000]),
                                    (2049, [1 << 20, 11]), (5000, [10_000_000] * 26), (4096, [40_000_000] * 5),
                                    (8192, [60000] * 7), (65536, [100, 100]), (65536, [10_000_000] * 26),
                                    (32768, [40_000_000] * 26), (65537, [1 << 20] * 3), (65535, [1 << 24, 3]),
                                    (3072, [1 << 31, 1 << 31]), (65536, [1 << 20] * 40),   # 40 clusters: two waves
                                    (300, [5000] * 100)])      # too many fields for the transposing tile: ids read in place
def test_unique_field_major_mode(cuda, B, rows, impl, shape):
    """field_bits > 0: per-field sort on the per-field id bits -- identical outputs to the flat sort.
    impl 1 / 2: fields of at most 65536 samples are sorted by k_cluster_sort (one cluster per field, 1..4 passes
    in shared memory); 2 lets that kernel write the unique step's outputs as well (fields handed out by ticket,
    per-field totals passed on through global memory), 1 runs the unique kernels after it (head count fused
    when the tiles line up); impl 0 and larger fields: the global passes.  shape: 4 CTAs x 1024 threads or
    8 CTAs x 512 threads per field."""
    from easydl_b200 import _lib
    from easydl_b200 import kernels as K
    from easydl_b200.criteo import CriteoSynthetic
    ids = CriteoSynthetic(rows, B, 1.05).ids(1)
    bases = O.key_bases(rows)
    keys = O.fuse_keys(ids, bases)
    key_bits = int(bases[-1] - 1).bit_length()
    field_bits = max(1, int(max(rows) - 1).bit_length())
    saved = _lib.tuning_get("dedup_impl"), _lib.tuning_get("dedup_shape")
    _lib.tuning_set("dedup_impl", impl)
    _lib.tuning_set("dedup_shape", shape)
    try:
        r = K.unique(dev(ids, cuda).view(-1), dev(bases[:len(rows)], cuda), key_bits, field_bits=field_bits)
        U = r.count()
    finally:
        _lib.tuning_set("dedup_impl", saved[0])
        _lib.tuning_set("dedup_shape", saved[1])
    uniq, inv, cnt, order, seg = O.dedup(keys)
    assert U == len(uniq)
    assert np.array_equal(r.uniq[:U].cpu().numpy(), uniq)
    assert np.array_equal(r.inverse.cpu().numpy(), inv)
    assert np.array_equal(r.counts[:U].cpu().numpy(), cnt)
    assert np.array_equal(r.order.cpu().numpy(), order)
    assert np.array_equal(r.seg_offsets[:U + 1].cpu().numpy(), seg)
    assert np.array_equal(r.seg_first[:U].cpu().numpy(), order[seg[:-1]])


def test_unique_n_dev(cuda):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(5)
    keys = rng.integers(0, 500, size=10000, dtype=np.int64)
    m = 6001
    n_dev = torch.tensor([m], dtype=torch.int32, device=cuda)
    r = K.unique(dev(keys, cuda), None, 9, n_dev=n_dev)
    U = r.count()
    uniq, inv, cnt, order, seg = O.dedup(keys[:m])
    assert U == len(uniq)
    assert np.array_equal(r.uniq[:U].cpu().numpy(), uniq)
    assert np.array_equal(r.inverse[:m].cpu().numpy(), inv)
    assert np.array_equal(r.order[:m].cpu().numpy(), order)
    assert np.array_equal(r.seg_offsets[:U + 1].cpu().numpy(), seg)


# ------------------------------------------------------------------------------- A2
@pytest.mark.parametrize("S", [1, 2, 3, 6, 8, 64, 255])
def test_partition_mod(cuda, S):
    from easydl_b200 import kernels as K
    from easydl_b200.layout import ShardLayout
    rows = [1000, 37, 40001]
    bases = O.key_bases(rows)
    rng = np.random.default_rng(S)
    keys = np.unique(rng.integers(0, bases[-1], size=20000, dtype=np.int64))
    lay = ShardLayout(rows, 1, 0).to_device(cuda)
    r = K.partition_mod(dev(keys, cuda), lay, S)
    want_keys, want_off, want_perm = O.partition_mod(keys, bases, S)
    assert np.array_equal(r.offsets.cpu().numpy(), want_off)
    assert np.array_equal(r.keys.cpu().numpy(), want_keys)
    assert np.array_equal(r.perm.cpu().numpy(), want_perm)
    pos = r.pos.cpu().numpy()
    assert np.array_equal(pos[want_perm], np.arange(len(keys)))
    # ascending within each group
    for s in range(S):
        g = want_keys[want_off[s]:want_off[s + 1]]
        assert np.all(np.diff(g) > 0)


def test_partition_mod_empty_and_ndev(cuda):
    from easydl_b200 import kernels as K
    r = K.partition_mod(torch.empty(0, dtype=torch.int64, device=cuda), None, 4)
    assert r.offsets.cpu().tolist() == [0, 0, 0, 0, 0]
    keys = np.arange(5000, dtype=np.int64) * 3
    n_dev = torch.tensor([1234], dtype=torch.int32, device=cuda)
    r = K.partition_mod(dev(keys, cuda), None, 6, n_dev=n_dev)
    wk, wo, wp = O.partition_mod(keys[:1234], np.array([0, 1 << 40], dtype=np.int64), 6)
    assert np.array_equal(r.offsets.cpu().numpy(), wo)
    assert np.array_equal(r.keys[:1234].cpu().numpy(), wk)


# ------------------------------------------------------------------------------- A3
@pytest.mark.parametrize("dim", [4, 16, 32, 64, 128, 256, 24])
def test_gather_and_expand_bit_exact(cuda, dim):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(dim)
    table = rng.standard_normal((5000, dim)).astype(np.float32)
    idx = rng.integers(0, 5000, size=12345, dtype=np.int64)
    t = dev(table, cuda)
    got = K.gather_keys(t, dev(idx, cuda), None).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), table[idx].view(np.uint32))
    got = K.expand(t, dev(idx.astype(np.int32), cuda)).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), table[idx].view(np.uint32))
    # [B, F] ids with field bases
    fb = np.array([0, 1000, 3000], dtype=np.int64)
    ids = np.stack([rng.integers(0, 1000, 999), rng.integers(0, 2000, 999), rng.integers(0, 2000, 999)], 1)
    got = K.gather_ids(t, dev(ids, cuda).view(-1), dev(fb, cuda)).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), table[(ids + fb[None]).reshape(-1)].view(np.uint32))
    # empty
    assert K.gather_keys(t, torch.empty(0, dtype=torch.int64, device=cuda), None).shape == (0, dim)


@pytest.mark.parametrize("S", [2, 6, 8])
def test_gather_through_shard_layout(cuda, S):
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.optimizer import OptimizerSpec
    rows = [997, 64, 5003]
    bases = O.key_bases(rows)
    rng = np.random.default_rng(S)
    keys = rng.integers(0, bases[-1], size=7000, dtype=np.int64)
    for shard in range(S):
        P = ShardParameters(rows, 32, OptimizerSpec.adagrad(), S, shard, device=cuda)
        _, ids = O.split_key(keys, bases)
        mine = keys[ids % S == shard]
        got = P.lookup(dev(mine, cuda)).cpu().numpy()
        ora = O.OracleShard(rows, 32, O.OptimizerSpec.adagrad(), S, shard)
        assert np.array_equal(got.view(np.uint32), ora.lookup(mine).view(np.uint32))


def test_scatter_rows_and_compose(cuda):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(0)
    rows = rng.standard_normal((3000, 64)).astype(np.float32)
    pos = rng.permutation(3000).astype(np.int32)
    out = torch.zeros((3000, 64), dtype=torch.float32, device=cuda)
    K.scatter_rows(dev(rows, cuda), dev(pos, cuda), out)
    want = np.zeros_like(rows)
    want[pos] = rows
    assert np.array_equal(out.cpu().numpy(), want)
    idx = rng.integers(0, 3000, 10000).astype(np.int32)
    got = K.compose_index(dev(idx, cuda), dev(pos, cuda)).cpu().numpy()
    assert np.array_equal(got, pos[idx])


# ------------------------------------------------------------------------------- A4
@pytest.mark.parametrize("dim,n,hi", [(16, 5000, 300), (64, 30000, 2000), (128, 9000, 50), (256, 4000, 10),
                                      (64, 40000, 3), (8, 1000, 1000), (40, 3000, 100)])
def test_segreduce_f64_accumulate(cuda, dim, n, hi):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(dim + n)
    keys = rng.integers(0, hi, size=n, dtype=np.int64)
    grad = (rng.standard_normal((n, dim)) * 1e-2).astype(np.float32)
    r = K.unique(dev(keys, cuda), None, 20)
    U = r.count()
    g = K.segreduce(dev(grad, cuda), r.order, r.seg_offsets, r.n_unique, n)[:U].cpu().numpy()
    _, _, _, order, seg = O.dedup(keys)
    want = O.segment_sum_f64(grad, order, seg)
    assert_close(g, want, "g_uniq")
    exact = np.mean(g.view(np.uint32) == want.view(np.uint32))
    assert exact > 0.999, f"only {exact:.4f} of g_uniq bit-exact vs float64 oracle"


def test_segreduce_out_pos(cuda):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(1)
    keys = rng.integers(0, 700, size=9000, dtype=np.int64)
    grad = rng.standard_normal((9000, 32)).astype(np.float32)
    r = K.unique(dev(keys, cuda), None, 10)
    U = r.count()
    pos = rng.permutation(U).astype(np.int32)
    posd = torch.zeros(9000, dtype=torch.int32, device=cuda)
    posd[:U] = dev(pos, cuda)
    g = K.segreduce(dev(grad, cuda), r.order, r.seg_offsets, r.n_unique, 9000, out_pos=posd)[:U].cpu().numpy()
    _, _, _, order, seg = O.dedup(keys)
    want = O.segment_sum_f64(grad, order, seg)
    assert_close(g[pos], want, "g_uniq scattered")


# ------------------------------------------------------------------------------- A4 + A5
def _run_steps(cuda, rows, dim, opt_o, opt_p, n_steps, batch, alpha, S=1, shard=0):
    from easydl_b200 import kernels as K
    from easydl_b200.parameters import ShardParameters
    P = ShardParameters(rows, dim, opt_p, 1, 0, device=cuda)
    ora = O.OracleShard(rows, dim, opt_o)
    bases = O.key_bases(rows)
    fb = dev(bases[:len(rows)], cuda)
    touched = []
    for step in range(n_steps):
        _, ids, _ = O.criteo_batch(rows, batch, alpha, step=step)
        keys = O.fuse_keys(ids, bases)
        grad = O.synthetic_grad(len(keys), dim, step=step)
        emb_o, (uniq, *_), _ = O.sparse_step(ora, keys, grad)
        emb = K.gather_ids(P.slab, dev(ids, cuda).view(-1), fb)
        assert np.array_equal(emb.cpu().numpy().view(np.uint32), emb_o.view(np.uint32)), f"emb step {step}"
        uq = K.unique(dev(ids, cuda).view(-1), fb, P.layout.key_bits)
        P.apply_gradients(uq, dev(grad, cuda))
        touched.append(uniq)
    allk = np.unique(np.concatenate(touched))
    got = [x.cpu().numpy() for x in P.read_rows(dev(allk, cuda))]
    want = [ora.lookup(allk)] + [ora.slot_lookup(k, allk) for k in range(opt_o.n_slots)]
    names = ["rows", "slot0", "slot1"]
    exact = []
    for gk, wk, nm in zip(got, want, names):
        assert_close(gk, wk, nm)
        exact.append(float(np.mean(gk.view(np.uint32) == wk.view(np.uint32))))
    return P, ora, exact


@pytest.mark.parametrize("dim", [16, 64, 128])
def test_adagrad_steps_match_oracle(cuda, dim):
    from easydl_b200.optimizer import OptimizerSpec
    rows = [5000, 300, 70000, 11]
    _, _, exact = _run_steps(cuda, rows, dim, O.OptimizerSpec.adagrad(), OptimizerSpec.adagrad(), 3, 2048, 1.05)
    assert min(exact) > 0.999, exact


@pytest.mark.parametrize("dim", [16, 128])
def test_adam_steps_match_oracle(cuda, dim):
    from easydl_b200.optimizer import OptimizerSpec
    rows = [5000, 300, 70000, 11]
    _, _, exact = _run_steps(cuda, rows, dim, O.OptimizerSpec.adam(), OptimizerSpec.adam(), 3, 2048, 1.2)
    assert min(exact) > 0.999, exact


def test_long_segments_hot_ids(cuda):
    """Zipf 1.2 on a tiny table: segments far above the long-segment threshold."""
    from easydl_b200.optimizer import OptimizerSpec
    rows = [50, 3]
    _run_steps(cuda, rows, 64, O.OptimizerSpec.adagrad(), OptimizerSpec.adagrad(), 2, 16384, 1.2)


# ------------------------------------------------------------------------------- A6
@pytest.mark.parametrize("opt", ["adagrad", "adam"])
def test_export_import_bit_exact(cuda, opt):
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    rows = [5000, 300, 70000, 11]
    dim = 32
    oo = O.OptimizerSpec.adagrad() if opt == "adagrad" else O.OptimizerSpec.adam()
    po = OptimizerSpec.adagrad() if opt == "adagrad" else OptimizerSpec.adam()
    P, ora, _ = _run_steps(cuda, rows, dim, oo, po, 2, 1024, 1.05)
    rec = P.export_records().cpu()
    keys, w, slots, step = ora.export()
    assert np.array_equal(rec.keys.numpy(), keys)
    assert rec.step == step
    # the GPU rows are the GPU's own (within 1e-6 of the oracle's); export must move them bit-exactly
    got = [x.cpu().numpy() for x in P.read_rows(dev(keys, cuda))]
    assert np.array_equal(rec.rows.numpy().view(np.uint32), got[0].view(np.uint32))
    for k in range(oo.n_slots):
        assert np.array_equal(rec.slots[k].numpy().view(np.uint32), got[1 + k].view(np.uint32))
    # import into S'=3 shards and read everything back
    bases = O.key_bases(rows)
    _, ids = O.split_key(keys, bases)
    for sp in range(3):
        Q = ShardParameters(rows, dim, po, 3, sp, device=cuda)
        Q.import_records(rec.to(cuda))
        mine = keys[ids % 3 == sp]
        back = Q.export_records().cpu()
        assert np.array_equal(back.keys.numpy(), mine)
        sel = np.nonzero(ids % 3 == sp)[0]
        assert np.array_equal(back.rows.numpy().view(np.uint32), rec.rows.numpy()[sel].view(np.uint32))
        for k in range(oo.n_slots):
            assert np.array_equal(back.slots[k].numpy().view(np.uint32),
                                  rec.slots[k].numpy()[sel].view(np.uint32))
        assert back.step == step


# ------------------------------------------------------------------------------- errors
def test_errors_are_reported_not_thrown(cuda):
    from easydl_b200 import _lib, kernels as K
    with pytest.raises(_lib.EdlError, match="multiple of 4"):
        K.gather_keys(torch.zeros((4, 6), device=cuda), torch.zeros(2, dtype=torch.int64, device=cuda), None)
    with pytest.raises(_lib.EdlError, match="key_bits"):
        K.unique(torch.zeros(4, dtype=torch.int64, device=cuda), None, 0)
    with pytest.raises(_lib.EdlError, match="CUDA tensor"):
        K.unique(torch.zeros(4, dtype=torch.int64), None, 8)


def _three_steps_digest(dev, opt, dim=64, alpha=1.2, interleaved=True):
    import hashlib
    from easydl_b200.criteo import CriteoSynthetic
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.step import SparseStep
    rows = [50_000] * 26
    P = ShardParameters(rows, dim, opt, device=dev, interleaved=interleaved)
    src = CriteoSynthetic(rows, 8192, alpha=alpha)
    st = SparseStep(P, 8192)
    h = hashlib.sha256()
    for k in range(3):
        ids = torch.from_numpy(src.ids(k)).to(dev)
        grad = torch.from_numpy(src.grad(dim, k)).to(dev)
        emb = st.run(ids, grad)
        torch.cuda.synchronize()
        h.update(emb.cpu().numpy().tobytes())
    h.update(P.slab.cpu().numpy().tobytes())
    for s in P.slots:
        h.update(s.cpu().numpy().tobytes())
    return h.hexdigest()


@pytest.mark.parametrize("opt_name,dim", [("adagrad", 64), ("adam", 64), ("adagrad", 16), ("adam", 128), ("adagrad", 256)])
def test_update_launch_variants_agree(cuda, opt_name, dim):
    """Every launch shape of the update gives the same bits: the streamed main kernel (default) and
    the register-path one, hot-id kernels on the helper stream or in a row, the hot-id listing on
    either stream, different hot-id thresholds and ring occupancies, the prefetching gather and the
    plain one.  The knobs are switched in-process (edl_tuning_set)."""
    from easydl_b200 import _lib
    from easydl_b200.optimizer import OptimizerSpec
    knobs = ("seg_impl", "seg_fork", "classify_side", "long_thr", "stream_ctas", "stream_stages", "direct_batch", "gather_impl")
    saved = {k: _lib.tuning_get(k) for k in knobs}
    variants = [
        dict(),                                               # defaults
        dict(seg_impl=0, gather_impl=0),                      # round-1 kernels
        dict(seg_impl=0, seg_fork=0, direct_batch=1),
        dict(seg_impl=1, seg_fork=0),
        dict(seg_impl=1, classify_side=0, long_thr=4, stream_ctas=1, stream_stages=2),
        dict(seg_impl=1, long_thr=100, stream_ctas=2, stream_stages=3),
        dict(seg_impl=0, long_thr=3),
    ]
    opt = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    digests = []
    try:
        for v in variants:
            for k in knobs:
                _lib.tuning_set(k, v.get(k, saved[k]))
            digests.append(_three_steps_digest(cuda, opt, dim))
        for v in (dict(seg_impl=0), dict(seg_impl=1)):        # separate dense arrays instead of [p | slots] records
            for k in knobs:
                _lib.tuning_set(k, v.get(k, saved[k]))
            digests.append(_three_steps_digest(cuda, opt, dim, interleaved=False))
    finally:
        for k in knobs:
            _lib.tuning_set(k, saved[k])
    assert len(set(digests)) == 1, digests


@pytest.mark.parametrize("opt_name", ["adagrad", "adam"])
def test_prefetched_dedup_gives_the_same_bits(cuda, opt_name):
    """SparseStep.run_prefetch (the next batch's dedup beside this batch's lookup and update), eagerly and
    replayed from the graph ring, against the plain step: identical embeddings, rows and slots."""
    from easydl_b200.criteo import CriteoSynthetic
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.step import GraphedPrefetchRing, SparseStep
    rows, B, dim, n, steps = [50_000] * 26, 4096, 32, 4, 9
    opt = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    src = CriteoSynthetic(rows, B, alpha=1.1)
    ids = [torch.from_numpy(src.ids(k)).to(cuda) for k in range(n)]
    grad = torch.from_numpy(src.grad(dim, 0)).to(cuda)
    outs = []
    for mode in ("plain", "prefetch", "graph"):
        P = ShardParameters(rows, dim, opt, device=cuda)
        st = SparseStep(P, B)
        ring = GraphedPrefetchRing(st, ids, grad) if mode == "graph" else None
        if ring is not None:                      # the ring's two warm-up steps touched the state: start over
            P.reset()
        embs = []
        for k in range(steps):
            if mode == "plain":
                e = st.run(ids[k % n], grad)
            elif mode == "prefetch":
                e = st.run_prefetch(ids[k % n], grad, ids[(k + 1) % n])
            else:
                e = ring.run()
            embs.append(e.clone())
        torch.cuda.synchronize()
        outs.append((embs, [t.clone() for t in P.state_tensors()], P.step))
    for other in outs[1:]:
        assert other[2] == outs[0][2]
        for a, b in zip(outs[0][0], other[0]):
            assert torch.equal(a, b)
        for a, b in zip(outs[0][1], other[1]):
            assert torch.equal(a, b)


def test_tuning_knobs_reject_bad_input(cuda):
    from easydl_b200 import _lib
    with pytest.raises(_lib.EdlError, match="unknown knob"):
        _lib.tuning_set("no_such_knob", 1)
    with pytest.raises(_lib.EdlError, match="outside"):
        _lib.tuning_set("seg_impl", 7)
    assert _lib.tuning_get("long_thr") >= 2


def test_c_host_program(cuda, tmp_path):
    """examples/c_host.c: dedup + segmented reduce + Adagrad from a plain C program (cudaMalloc'ed
    buffers, no Python in the process), checked inside the program against host arithmetic."""
    import os, subprocess
    from test_host import _c_host_cmd
    exe = str(tmp_path / "c_host")
    root, cmd = _c_host_cmd(exe)
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    env = dict(os.environ, LD_LIBRARY_PATH=os.pathsep.join(
        [os.path.join(root, "easydl_b200"), os.path.join(cuda_home, "lib64"), os.environ.get("LD_LIBRARY_PATH", "")]))
    r = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "mismatching elements: 0" in r.stdout


def test_four_wide_sqrt_div_are_ieee_exact(cuda):
    """The update kernels' side-by-side sqrt / divide sequences against __fsqrt_rn / __fdiv_rn on 2^27
    random operand triples (raw bit patterns and training-like magnitudes), element-wise and composed as
    Adagrad / Adam steps: not one differing bit."""
    from easydl_b200 import _lib
    lib = _lib.load()
    bad = torch.zeros(1, dtype=torch.int64, device=cuda)
    for seed in (1, 0xDEADBEEF):
        _lib.check(lib.edl_selftest_rounding(1 << 26, seed, bad.data_ptr(), torch.cuda.current_stream().cuda_stream),
                   "edl_selftest_rounding")
    assert int(bad.item()) == 0
