"""BASELINE.json configs as parity cases.

configs[0]  Wide&Deep shape, 26 tables x 1M rows x dim 16, 1 PS + 2 workers: through the PSClient
            host-array API, pushes applied in worker order.
configs[1]  DeepFM shape at FULL size (26 x 10M rows x dim 64, batch 64K, Adagrad) on one B200:
            the oracle keeps only touched rows (the initialiser is a pure function), so the whole
            step is checked -- every looked-up embedding bit for bit and every updated row and
            accumulator -- not just size-independent properties.
"""
import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_cfg1_one_ps_two_workers(cuda):
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.ps import ParameterServer
    from easydl_b200.ps_client import PSClient
    rows, dim, B = [1_000_000] * 26, 16, 2048        # 2 workers x 2048 = the 4096-sample batch
    names = [f"C{i + 1}" for i in range(26)]
    ps = ParameterServer(names, rows, dim, OptimizerSpec.adagrad(), 1, 0, device=cuda)
    workers = [PSClient([ps]), PSClient([ps])]
    ora = O.OracleShard(rows, dim, O.OptimizerSpec.adagrad())
    bases = O.key_bases(rows)
    for step in range(2):
        for w, client in enumerate(workers):        # asynchronous PS: each push is its own update
            _, ids, _ = O.criteo_batch(rows, B, 1.05, step=step * 2 + w)
            grad = O.synthetic_grad(ids.size, dim, step=step * 2 + w)
            emb = client.pull_embedding_batch(ids)
            client.push_gradient_batch(grad.reshape(B, 26, dim))
            want, _, _ = O.sparse_step(ora, O.fuse_keys(ids, bases), grad)
            got = emb.reshape(-1, dim)
            np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-9)
            if step == 0 and w == 0:
                assert np.array_equal(bits(got), bits(want))
    rec = ps.export_shard().cpu()
    assert np.array_equal(rec.keys.numpy(), ora.keys) and ps.version == ora.step == 4
    np.testing.assert_allclose(rec.rows.numpy(), ora.w, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(rec.slots[0].numpy(), ora.slots[0], rtol=1e-6, atol=1e-9)
    exact = np.mean(bits(rec.rows.numpy()) == bits(ora.w))
    assert exact > 0.999, exact


def test_cfg2_full_size_step(cuda):
    from easydl_b200.criteo import CriteoSynthetic
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.step import GraphedStep, SparseStep
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    if free < 150e9:
        pytest.skip(f"needs ~140 GB of free HBM, {free / 1e9:.0f} GB available")
    rows, dim, B = [10_000_000] * 26, 64, 65536
    P = ShardParameters(rows, dim, OptimizerSpec.adagrad(), device=cuda)
    step = SparseStep(P, B)
    src = CriteoSynthetic(rows, B, 1.05)
    ora = O.OracleShard(rows, dim, O.OptimizerSpec.adagrad())
    bases = O.key_bases(rows)
    for k in range(2):
        ids, grad = src.ids(k), src.grad(dim, k)
        ids_d, grad_d = torch.from_numpy(ids).to(cuda), torch.from_numpy(grad).to(cuda)
        if k == 0:
            emb = step.run(ids_d, grad_d).cpu().numpy()
        else:                                   # the CUDA-graph replay path bench.py times
            host_step = P.step
            g = GraphedStep(step, ids_d, grad_d)            # its warm-up applies the batch once...
            emb = step.emb.cpu().numpy()
            assert P.step == host_step + 1
        want, (uniq, *_), _ = O.sparse_step(ora, O.fuse_keys(ids, bases), grad)
        assert np.array_equal(bits(emb), bits(want)), f"step {k}: lookup differs"
    got = [x.cpu().numpy() for x in P.read_rows(torch.from_numpy(ora.keys).to(cuda))]
    for a, b, nm in [(got[0], ora.w, "rows"), (got[1], ora.slots[0], "accumulators")]:
        err = np.abs(a.astype(np.float64) - b)
        assert np.all(err <= 1e-9 + 1e-6 * np.abs(b)), f"{nm}: max err {err.max():.3e}"
        assert np.mean(bits(a) == bits(b)) > 0.9999, nm
    # one graph replay == one more eager step on the same batch
    g.run()
    O.sparse_step(ora, O.fuse_keys(ids, bases), grad)
    got = P.read_rows(torch.from_numpy(ora.keys).to(cuda))[0].cpu().numpy()
    np.testing.assert_allclose(got, ora.w, rtol=1e-6, atol=1e-9)
    del P, step, g
    torch.cuda.empty_cache()
