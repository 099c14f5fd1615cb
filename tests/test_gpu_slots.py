"""The host-sync-free sharded step (easydl_b200/slot_step.py, csrc/slots.cu) on ONE GPU: the slot
kernels one by one, then S virtual ranks phase by phase against the oracle's global float64 pass
and against ShardedSparseStep (same arithmetic, same merge order -> same bits)."""
import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-6, 1e-9


def dev(a, cuda):
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


def test_slot_send_positions_compact_remap(cuda):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(11)
    S, stride, limit, dim = 3, 64, 40, 4
    offs = np.array([0, 17, 17, 50], dtype=np.int32)            # middle group empty
    rows = rng.standard_normal((50, dim)).astype(np.float64)
    keys = rng.integers(0, 1 << 40, size=50, dtype=np.int64)
    bufs_r = [torch.zeros((stride, dim), dtype=torch.float64, device=cuda) for _ in range(S)]
    keyslots = torch.full((S * stride,), -1, dtype=torch.int64, device=cuda)
    counts = torch.full((S,), -7, dtype=torch.int32, device=cuda)
    err = torch.zeros(1, dtype=torch.int32, device=cuda)
    d_offs = dev(offs, cuda)
    K.slot_send(dev(rows, cuda), d_offs, torch.tensor([b.data_ptr() for b in bufs_r], dtype=torch.int64, device=cuda),
                limit, None, err, 64)
    kb = torch.tensor([keyslots.data_ptr() + s * stride * 8 for s in range(S)], dtype=torch.int64, device=cuda)
    cp = torch.tensor([counts.data_ptr() + 4 * s for s in range(S)], dtype=torch.int64, device=cuda)
    K.slot_send(dev(keys, cuda), d_offs, kb, limit, cp, err, 64)
    assert int(err.item()) == 0
    for s in range(S):
        c = offs[s + 1] - offs[s]
        assert np.array_equal(bufs_r[s][:c].cpu().numpy(), rows[offs[s]:offs[s + 1]])
        assert float(bufs_r[s][c:].abs().sum()) == 0.0
        assert np.array_equal(keyslots[s * stride:s * stride + c].cpu().numpy(), keys[offs[s]:offs[s + 1]])
    assert counts.cpu().tolist() == [17, 0, 33]
    # positions: partition position -> slot-space row
    pos = dev(rng.permutation(50).astype(np.int32), cuda)
    n_dev = torch.tensor([45], dtype=torch.int32, device=cuda)
    sp = K.slot_positions(pos, n_dev, d_offs, stride, out=torch.full((50,), -1, dtype=torch.int32, device=cuda))
    p = pos.cpu().numpy()
    d = np.searchsorted(offs[1:], p, side="right")
    want = d * stride + (p - offs[d])
    assert np.array_equal(sp.cpu().numpy()[:45], want[:45]) and np.all(sp.cpu().numpy()[45:] == -1)
    # compact: slot order, counts clipped at the limit
    keys_c = torch.empty(S * limit, dtype=torch.int64, device=cuda)
    map_c = torch.empty(S * limit, dtype=torch.int32, device=cuda)
    n_out = torch.zeros(1, dtype=torch.int32, device=cuda)
    K.slot_compact(keyslots, counts, stride, limit, keys_c, map_c, n_out)
    assert int(n_out.item()) == 50
    assert np.array_equal(keys_c[:50].cpu().numpy(), keys)
    assert np.array_equal(map_c[:50].cpu().numpy(), np.concatenate([np.arange(17), 2 * stride + np.arange(33)]))
    # remap with a device count
    idx = dev(rng.integers(0, 50, size=80).astype(np.int32), cuda)
    out = K.remap_index(idx, map_c, torch.tensor([60], dtype=torch.int32, device=cuda),
                        out=torch.full((80,), -5, dtype=torch.int32, device=cuda))
    assert np.array_equal(out[:60].cpu().numpy(), map_c.cpu().numpy()[idx.cpu().numpy()[:60]])
    assert np.all(out[60:].cpu().numpy() == -5)
    # overflow: a group longer than the limit is cut and flagged
    K.slot_send(dev(keys, cuda), d_offs, kb, 20, cp, err, 64)
    assert int(err.item()) == 1 and counts.cpu().tolist() == [17, 0, 20]


@pytest.mark.parametrize("S,limit,hi", [(1, 100, 50), (2, 5000, 20000), (3, 64, 40), (8, 3000, 9000), (8, 20000, 1 << 27),
                                        (64, 300, 5000)])
def test_slot_merge_unique_equals_compact_sort_remap(cuda, S, limit, hi):
    """edl_slot_merge_unique (the owner's S sorted key slots merged) against the calls it replaces --
    edl_slot_compact -> edl_unique -> edl_remap_index x 2 -- and against the oracle's dedup of the
    concatenated keys: empty slots, a count beyond the limit (cut there), keys marked with bit 62."""
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(S * 1000 + limit)
    stride = limit + 7
    slots = np.zeros(S * stride, dtype=np.int64)
    counts = np.zeros(S, dtype=np.int32)
    runs = []
    for s in range(S):
        n = 0 if (s == 1 and S > 2) else int(rng.integers(1, min(limit, hi) + 1))
        run = np.unique(rng.integers(0, hi, size=n, dtype=np.int64))
        n = len(run)
        flags = rng.integers(0, 2, size=n).astype(np.int64) << 62
        slots[s * stride:s * stride + n] = run | flags
        counts[s] = n
        runs.append(run)
    if S > 1:                       # a sender that reports more than the slot holds: cut at the limit
        full = np.unique(rng.integers(0, max(hi, 4 * limit), size=8 * limit, dtype=np.int64))[:limit]
        assert len(full) == limit
        slots[(S - 1) * stride:(S - 1) * stride + limit] = full
        counts[S - 1] = limit + 5
        runs[S - 1] = full
    R = S * limit
    d_slots, d_counts = dev(slots, cuda), dev(counts, cuda)
    i32 = dict(dtype=torch.int32, device=cuda)

    def result():
        return K.UniqueResult(uniq=torch.empty(R, dtype=torch.int64, device=cuda), inverse=None, counts=None,
                              order=torch.empty(R, **i32), seg_offsets=torch.empty(R + 1, **i32),
                              n_unique=torch.zeros(1, **i32), n=R, seg_first=torch.empty(R, **i32))
    # the calls it replaces
    keys_c, map_c, n_recv = torch.empty(R, dtype=torch.int64, device=cuda), torch.empty(R, **i32), torch.zeros(1, **i32)
    K.slot_compact(d_slots, d_counts, stride, limit, keys_c, map_c, n_recv)
    ref = K.unique(keys_c, None, 63, want_inverse=False, want_counts=False, n_dev=n_recv, out=result())
    order_ref = K.remap_index(ref.order, map_c, n_recv, out=torch.empty(R, **i32))
    first_ref = K.remap_index(ref.seg_first, map_c, ref.n_unique, out=torch.empty(R, **i32))
    # merged
    got, n_out = result(), torch.zeros(1, **i32)
    K.slot_merge_unique(d_slots, d_counts, stride, limit, got, n_out)
    n, U = int(n_recv.item()), ref.count()
    assert int(n_out.item()) == n == sum(len(r) for r in runs)
    assert got.count() == U
    assert torch.equal(got.uniq[:U], ref.uniq[:U])
    assert torch.equal(got.seg_offsets[:U + 1], ref.seg_offsets[:U + 1])
    assert torch.equal(got.order[:n], order_ref[:n])
    assert torch.equal(got.seg_first[:U], first_ref[:U])
    # and the oracle on the concatenation
    uniq, _, _, order, seg = O.dedup(np.concatenate(runs) if n else np.zeros(0, dtype=np.int64))
    assert np.array_equal(got.uniq[:U].cpu().numpy(), uniq)
    assert np.array_equal(got.seg_offsets[:U + 1].cpu().numpy(), seg)
    starts = np.concatenate([[0], np.cumsum([len(r) for r in runs])])
    src = np.concatenate([s * stride + np.arange(len(r)) for s, r in enumerate(runs)]) if n else np.zeros(0, dtype=np.int64)
    assert np.array_equal(got.order[:n].cpu().numpy() & ((1 << 30) - 1), src[order])


def test_fence_with_device_counter_world_of_one(cuda):
    import ctypes as C
    from easydl_b200 import _lib
    lib = _lib.load()
    flags = torch.zeros(2, dtype=torch.int64, device=cuda)
    counter = torch.zeros(1, dtype=torch.int64, device=cuda)
    ptrs = torch.tensor([flags.data_ptr()], dtype=torch.int64, device=cuda)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        assert lib.edl_peer_signal_wait_counter(ptrs.data_ptr(), flags.data_ptr(), 0, 1, counter.data_ptr(), st) == 0
    torch.cuda.synchronize()
    assert flags.cpu().tolist() == [3, 0] and int(counter.item()) == 3


def _virtual(cuda, rows, dim, S, opt_name, batch, wire=torch.float64, slack=1.5, tight=True):
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.slot_step import SlotBuffers, SlotShardedStep
    po = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    params = [ShardParameters(rows, dim, po, S, s, device=cuda) for s in range(S)]
    N = batch * len(rows)
    # tight: slots exactly as long as the step's slot limit (what a real rank allocates); else worst case
    stride = min(N, int(slack * N / S) + 1024) if tight else N
    bufs = [SlotBuffers(S, s, cuda, stride, dim, wire) for s in range(S)]
    for b in bufs:
        b.connect_local(bufs)
    steps = [SlotShardedStep(p, batch, b, wire_dtype=wire, slot_slack=slack) for p, b in zip(params, bufs)]
    return params, bufs, steps


@pytest.mark.parametrize("S,opt,wire", [(1, "adagrad", torch.float64), (2, "adagrad", "mixed"), (4, "adam", torch.float64),
                                        (6, "adagrad", torch.float64), (8, "adam", "mixed"), (3, "adam", "mixed")])
def test_slot_step_matches_global_oracle(cuda, S, opt, wire):
    from easydl_b200.slot_step import run_virtual
    from test_gpu_sharded import _compare_state, bits
    rows, dim, batch, steps = [5000, 301, 70001, 11, 6], 32, 512, 3
    params, bufs, st = _virtual(cuda, rows, dim, S, opt, batch, wire=wire)
    oo = O.OptimizerSpec.adagrad() if opt == "adagrad" else O.OptimizerSpec.adam()
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    bases = O.key_bases(rows)
    for k in range(steps):
        kw, gw, di, dg = [], [], [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, batch, 1.05, step=k * 8 + r)
            g = O.synthetic_grad(ids.size, dim, step=k * 8 + r)
            kw.append(O.fuse_keys(ids, bases)); gw.append(g)
            di.append(dev(ids, cuda)); dg.append(dev(g, cuda))
        embs = run_virtual(st, di, dg)
        want = O.sharded_step(osh, kw, gw)
        for r in range(S):
            got = embs[r].cpu().numpy()
            if k == 0:
                assert np.array_equal(bits(got), bits(want[r])), f"emb step {k} rank {r}"
            err = np.abs(got.astype(np.float64) - want[r])
            assert np.all(err <= ATOL + RTOL * np.abs(want[r])), f"emb step {k} rank {r}: {err.max():.3e}"
    for s in st:
        s.check()
    _compare_state(cuda, params, osh, oo.n_slots)
    for b in bufs:
        b.close()


def test_slot_step_same_bits_as_packed_exchange(cuda):
    """Same kernels, same merge order: identical shard state to ShardedSparseStep (lock-step ranks)."""
    from easydl_b200.exchange import run_lockstep
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.sharded_step import ShardedSparseStep
    from easydl_b200.slot_step import run_virtual
    rows, dim, batch, S = [20000, 3000, 17], 64, 2048, 4
    params, bufs, st = _virtual(cuda, rows, dim, S, "adagrad", batch)
    ref_p = [ShardParameters(rows, dim, OptimizerSpec.adagrad(), S, s, device=cuda) for s in range(S)]
    ref_st = [ShardedSparseStep(p, batch) for p in ref_p]
    for k in range(3):
        di, dg = [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, batch, 1.2, step=k * 8 + r)
            di.append(dev(ids, cuda)); dg.append(dev(O.synthetic_grad(ids.size, dim, step=k * 8 + r), cuda))
        embs = run_virtual(st, di, dg)
        ref = run_lockstep([s.gen_step(i, g) for s, i, g in zip(ref_st, di, dg)])
        for a, b in zip(embs, ref):
            assert torch.equal(a, b)
    for p, q in zip(params, ref_p):
        assert torch.equal(p.slab, q.slab) and torch.equal(p.slots[0], q.slots[0]) and p.step == q.step
    for b in bufs:
        b.close()


def test_mixed_wire_gives_the_bits_of_the_float64_wire(cuda):
    """fp32 rows for ids with one local gradient row + float64 partial sums for the rest: the owner's
    float64 total, and with it every updated row and slot, is bit-identical to the all-float64 wire."""
    from easydl_b200.slot_step import run_virtual
    rows, dim, batch, S = [20000, 3000, 17, 500000], 64, 2048, 4
    for opt in ("adagrad", "adam"):
        pa, ba, sa = _virtual(cuda, rows, dim, S, opt, batch, wire=torch.float64, tight=False)
        pb, bb, sb = _virtual(cuda, rows, dim, S, opt, batch, wire="mixed")
        for k in range(3):
            di, dg = [], []
            for r in range(S):
                _, ids, _ = O.criteo_batch(rows, batch, 1.1, step=k * 8 + r)
                di.append(dev(ids, cuda)); dg.append(dev(O.synthetic_grad(ids.size, dim, step=k * 8 + r), cuda))
            ea = run_virtual(sa, di, dg)
            eb = run_virtual(sb, di, dg)
            for x, y in zip(ea, eb):
                assert torch.equal(x, y)
        for s in sa + sb:
            s.check()
        for p, q in zip(pa, pb):
            assert torch.equal(p.slab, q.slab) and p.step == q.step
            for u, v in zip(p.slots, q.slots):
                assert torch.equal(u, v)
        for b in ba + bb:
            b.close()


def test_slot_overflow_is_reported(cuda):
    from easydl_b200.slot_step import run_virtual
    rows, dim, batch, S = [100000], 16, 4096, 2
    params, bufs, st = _virtual(cuda, rows, dim, S, "adagrad", batch, slack=0.0)   # limit = 1024 rows per pair
    _, ids, _ = O.criteo_batch(rows, batch, 0.5, step=0)
    g = O.synthetic_grad(ids.size, dim, step=0)
    run_virtual(st, [dev(ids, cuda)] * S, [dev(g, cuda)] * S)
    with pytest.raises(RuntimeError, match="slot_limit"):
        st[0].check()
    for b in bufs:
        b.close()


def test_expand_window(cuda):
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(2)
    B, F, Fw, f0, dim, R = 37, 7, 3, 2, 16, 50
    rows = rng.standard_normal((R, dim)).astype(np.float32)
    idx = rng.integers(0, R, size=B * Fw).astype(np.int32)
    out = torch.full((B * F, dim), -1.0, dtype=torch.float32, device=cuda)
    K.expand_window(dev(rows, cuda), dev(idx, cuda), Fw, F, f0, out)
    want = np.full((B, F, dim), -1.0, dtype=np.float32)
    want[:, f0:f0 + Fw, :] = rows[idx].reshape(B, Fw, dim)
    assert np.array_equal(out.cpu().numpy().reshape(B, F, dim), want)


def _virtual_pipelined(cuda, rows, dim, S, opt_name, batch, n_lanes):
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.slot_step import PipelinedSlotStep, SlotBuffers
    po = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    params = [ShardParameters(rows, dim, po, S, s, device=cuda) for s in range(S)]
    F = len(rows)
    cuts = [F * g // n_lanes for g in range(n_lanes + 1)]
    bufs = []                                     # bufs[lane][rank]
    for g in range(n_lanes):
        lane = [SlotBuffers(S, s, cuda, batch * (cuts[g + 1] - cuts[g]), dim) for s in range(S)]
        for b in lane:
            b.connect_local(lane)
        bufs.append(lane)
    steps = [PipelinedSlotStep(params[s], batch, buffers=[bufs[g][s] for g in range(n_lanes)], n_lanes=n_lanes)
             for s in range(S)]
    return params, bufs, steps


@pytest.mark.parametrize("S,opt,n_lanes", [(2, "adagrad", 2), (4, "adam", 3), (3, "adagrad", 5)])
def test_pipelined_slot_step_same_bits_as_unpipelined_and_oracle(cuda, S, opt, n_lanes):
    from easydl_b200.slot_step import run_virtual, run_virtual_pipelined
    from test_gpu_sharded import _compare_state
    rows, dim, batch, steps = [5000, 301, 70001, 11, 6], 32, 512, 3
    params, bufs, st = _virtual_pipelined(cuda, rows, dim, S, opt, batch, n_lanes)
    ref_p, ref_b, ref_st = _virtual(cuda, rows, dim, S, opt, batch)
    oo = O.OptimizerSpec.adagrad() if opt == "adagrad" else O.OptimizerSpec.adam()
    osh = [O.OracleShard(rows, dim, oo, S, s) for s in range(S)]
    bases = O.key_bases(rows)
    for k in range(steps):
        kw, gw, di, dg = [], [], [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, batch, 1.05, step=k * 8 + r)
            g = O.synthetic_grad(ids.size, dim, step=k * 8 + r)
            kw.append(O.fuse_keys(ids, bases)); gw.append(g)
            di.append(dev(ids, cuda)); dg.append(dev(g, cuda))
        embs = run_virtual_pipelined(st, di, dg)
        ref = run_virtual(ref_st, di, dg)
        O.sharded_step(osh, kw, gw)
        for a, b in zip(embs, ref):
            assert torch.equal(a, b), f"emb step {k}"
    for s in st:
        s.check()
    for p, q in zip(params, ref_p):
        assert p.step == q.step and torch.equal(p.slab, q.slab) and torch.equal(p.slots[0], q.slots[0])
    _compare_state(cuda, params, osh, oo.n_slots)
    for lane in bufs:
        for b in lane:
            b.close()
    for b in ref_b:
        b.close()


@pytest.mark.parametrize("n_lanes", [1, 2])
def test_real_drivers_world_of_one_eager_equals_graph(cuda, n_lanes):
    """The real-rank drivers (helper streams, fences, CUDA-graph replay with alternating buffer
    generations) with a world of one: graph replay == eager issue == the phase-by-phase virtual run."""
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.slot_step import (GraphedSlotStep, PipelinedSlotStep, SlotBuffers, SlotShardedStep,
                                       run_virtual)
    rows, dim, batch = [20000, 3000, 17, 900], 32, 1024
    F = len(rows)
    _, ids, _ = O.criteo_batch(rows, batch, 1.1, step=3)
    d_ids, d_g = dev(ids, cuda), dev(O.synthetic_grad(ids.size, dim, step=3), cuda)

    def make():
        p = ShardParameters(rows, dim, OptimizerSpec.adagrad(), 1, 0, device=cuda)
        if n_lanes == 1:
            b = SlotBuffers(1, 0, cuda, batch * F, dim)
            b.connect_local([b])
            return p, [b], SlotShardedStep(p, batch, b)
        cuts = [F * g // n_lanes for g in range(n_lanes + 1)]
        bs = [SlotBuffers(1, 0, cuda, batch * (cuts[g + 1] - cuts[g]), dim) for g in range(n_lanes)]
        for b in bs:
            b.connect_local([b])
        return p, bs, PipelinedSlotStep(p, batch, buffers=bs, n_lanes=n_lanes)

    p1, b1, eager = make()
    for _ in range(4):
        emb_e = eager.run(d_ids, d_g).clone()
    eager.check()
    p2, b2, st2 = make()
    graphed = GraphedSlotStep(st2, d_ids, d_g)       # two eager steps inside
    graphed.run()
    emb_g = graphed.run().clone()
    st2.check()
    assert p1.step == p2.step == 4
    assert torch.equal(emb_e, emb_g)
    assert torch.equal(p1.slab, p2.slab) and torch.equal(p1.slots[0], p2.slots[0])
    # the unpipelined phase-by-phase run gives the same bits
    p3 = ShardParameters(rows, dim, OptimizerSpec.adagrad(), 1, 0, device=cuda)
    b3 = SlotBuffers(1, 0, cuda, batch * F, dim)
    b3.connect_local([b3])
    st3 = SlotShardedStep(p3, batch, b3)
    for _ in range(4):
        run_virtual([st3], [d_ids], [d_g])
    assert torch.equal(p1.slab, p3.slab) and torch.equal(p1.slots[0], p3.slots[0])
    for b in b1 + b2 + [b3]:
        b.close()


@pytest.mark.parametrize("opt_name,wire", [("adagrad", "mixed"), ("adam", torch.float64)])
def test_prefetching_slot_step_world_of_one(cuda, opt_name, wire):
    """SlotShardedStep.run_prefetch (the next batch's phase A on a helper stream, sent into the other buffer
    generation) eagerly and replayed from GraphedSlotRing, against plain eager steps over the same batch
    sequence: identical embeddings, rows and slots.  World of one: the fences are real kernels."""
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.slot_step import GraphedSlotRing, SlotBuffers, SlotShardedStep
    rows, dim, batch, n, steps = [20000, 3000, 17, 900], 32, 1024, 4, 10
    F = len(rows)
    ids = [dev(O.criteo_batch(rows, batch, 1.1, step=k)[1], cuda) for k in range(n)]
    g = dev(O.synthetic_grad(batch * F, dim, step=0), cuda)
    po = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    outs = []
    for mode in ("plain", "prefetch", "ring"):
        p = ShardParameters(rows, dim, po, 1, 0, device=cuda)
        b = SlotBuffers(1, 0, cuda, batch * F, dim, wire)
        b.connect_local([b])
        st = SlotShardedStep(p, batch, b, wire_dtype=wire)
        embs = []
        ring = None
        for k in range(steps):
            if mode == "plain":
                e = st.run(ids[k % n], g)
            elif mode == "prefetch":
                e = st.run_prefetch(ids[k % n], g, ids[(k + 1) % n])
            else:
                if ring is None:
                    ring = GraphedSlotRing(st, ids, g)          # runs steps 0 and 1 eagerly
                    embs += [None, None]
                if k < 2:
                    continue
                e = ring.run()
            embs.append(e.clone())
        st.check()
        outs.append((embs, [t.clone() for t in p.state_tensors()], p.step))
        b.close()
    for other in outs[1:]:
        assert other[2] == outs[0][2] == steps
        for a, c in zip(outs[0][0], other[0]):
            assert c is None or torch.equal(a, c)
        for a, c in zip(outs[0][1], other[1]):
            assert torch.equal(a, c)
