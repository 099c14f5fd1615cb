"""The PS / worker API surface (lineage names) against the oracle, through host arrays."""
import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

pytestmark = pytest.mark.gpu
ROWS, DIM = [5000, 301, 70001, 11, 6], 32
NAMES = [f"emb_{i}" for i in range(len(ROWS))]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _servers(cuda, S, opt="adagrad"):
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.ps import ParameterServer
    po = OptimizerSpec.adagrad() if opt == "adagrad" else OptimizerSpec.adam()
    return [ParameterServer(NAMES, ROWS, DIM, po, S, s, device=cuda) for s in range(S)]


@pytest.mark.parametrize("S", [1, 3, 8])
def test_psclient_batch_pull_push_matches_oracle(cuda, S):
    from easydl_b200.ps_client import PSClient
    client = PSClient(_servers(cuda, S))
    osh = [O.OracleShard(ROWS, DIM, O.OptimizerSpec.adagrad(), S, s) for s in range(S)]
    bases = O.key_bases(ROWS)
    for step in range(3):
        _, ids, _ = O.criteo_batch(ROWS, 300, 1.05, step=step)
        grad = O.synthetic_grad(ids.size, DIM, step=step)
        emb = client.pull_embedding_batch(ids)
        client.push_gradient_batch(grad.reshape(300, len(ROWS), DIM))
        want = O.sharded_step(osh, [O.fuse_keys(ids, bases)], [grad])[0]
        got = emb.reshape(-1, DIM)
        if step == 0:
            assert np.array_equal(bits(got), bits(want))
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-9)
    for ps, o in zip(client.servers, osh):
        rec = ps.export_shard().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys) and ps.version == o.step == 3
        np.testing.assert_allclose(rec.rows.numpy(), o.w, rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(rec.slots[0].numpy(), o.slots[0], rtol=1e-6, atol=1e-9)


def test_per_table_calls_and_embedding_table_get_set(cuda):
    from easydl_b200.hash_utils import int_to_id, string_to_id
    from easydl_b200.ps_client import PSClient
    S = 4
    servers = _servers(cuda, S, "adam")
    client = PSClient(servers)
    rng = np.random.default_rng(0)
    ids = rng.integers(0, ROWS[2], 777)
    got = client.pull_embedding_vectors("emb_2", ids)
    want = O.init_rows(O.DEFAULT_SEED, 2, ids, DIM)
    assert np.array_equal(bits(got), bits(want))
    # IndexedSlices push with duplicates == oracle's dedup + apply on the owning shards
    g = (rng.standard_normal((777, DIM)) * 1e-2).astype(np.float32)
    client.push_gradients("emb_2", ids, g)
    osh = [O.OracleShard(ROWS, DIM, O.OptimizerSpec.adam(), S, s) for s in range(S)]
    O.sharded_step(osh, [ids + O.key_bases(ROWS)[2]], [g])
    after = client.pull_embedding_vectors("emb_2", ids)
    keys = ids + O.key_bases(ROWS)[2]
    want_after = np.empty_like(after)
    for s in range(S):
        sel = ids % S == s
        want_after[sel] = osh[s].lookup(keys[sel])
    np.testing.assert_allclose(after, want_after, rtol=1e-6, atol=1e-9)
    # EmbeddingTable.get / set on one shard
    tab = servers[1].tables["emb_0"]
    mine = torch.arange(1, ROWS[0], S)[:50]
    vals = torch.from_numpy(rng.standard_normal((50, DIM)).astype(np.float32))
    tab.set(mine, vals)
    assert torch.equal(tab.get(mine).cpu(), vals)
    assert int_to_id(np.array([5, 9]), 4).tolist() == [1, 1]
    assert 0 <= string_to_id("dense/kernel:0", 4) < 4
    parts = client.partition_dense_parameters(["a", "b", "c", "d", "e"])
    assert sorted(sum(parts.values(), [])) == ["a", "b", "c", "d", "e"]


def test_deduplicate_indexed_slices(cuda):
    from easydl_b200.tensor_utils import deduplicate_indexed_slices, merge_indexed_slices
    rng = np.random.default_rng(1)
    idx = rng.integers(0, 100, 5000)
    vals = rng.standard_normal((5000, 16)).astype(np.float32)
    v, i = merge_indexed_slices([(torch.from_numpy(vals[:3000]).to(cuda), torch.from_numpy(idx[:3000]).to(cuda)),
                                 (torch.from_numpy(vals[3000:]).to(cuda), torch.from_numpy(idx[3000:]).to(cuda))])
    g, u = deduplicate_indexed_slices(v, i, 8)
    uo, _, _, order, seg = O.dedup(idx)
    assert np.array_equal(u.cpu().numpy(), uo)
    np.testing.assert_allclose(g.cpu().numpy(), O.segment_sum_f64(vals, order, seg), rtol=1e-6, atol=1e-9)
