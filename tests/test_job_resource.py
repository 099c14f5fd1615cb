"""The JobResource shim: the reference's own example YAML (ref:docs/design/elastic-training-operator.md:57-95,
transcribed below including its `//` comment) parses, and a change of parameter_server.replicas
drives the bit-exact re-shard (GPU part marked gpu)."""
import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

REFERENCE_EXAMPLE = """
apiVersion: "elastic.easydl.org/v1alpha1"
kind: "JobResource"
metadata:
  name: "elastic-training-resource"
spec:
  selector:
    name: elastic-deepctr-job  // Job Name
  parameter_server:
    replicas: 4
    resource:
      cpu: 4
      memory: 4096
      disk: 8192
      gpu: 1
  worker:
    replicas: 4
    resource:
      cpu: 4
      memory: 4096
      disk: 8192
      gpu: 1
  evaluator:
    replicas: 1
    resource:
      cpu: 4
      memory: 4096
      disk: 8192
      gpu: 1
  resource_updation:
    - name: "elastic-deepctr-job-ps-0"
      resource:
        cpu: 8
        memory: 8192
    - name: "elastic-deepctr-job-ps-1"
      resource:
        cpu: 16
        memory: 8192
"""


def test_reference_example_parses():
    from easydl_b200.job_resource import parse_job_resource
    jr = parse_job_resource(REFERENCE_EXAMPLE)
    assert jr.name == "elastic-training-resource" and jr.job == "elastic-deepctr-job"
    assert jr.ps_replicas == 4 and jr.worker_replicas == 4 and jr.ps_gpu == 1
    assert [u["name"] for u in jr.resource_updation] == ["elastic-deepctr-job-ps-0", "elastic-deepctr-job-ps-1"]
    with pytest.raises(ValueError):
        parse_job_resource(REFERENCE_EXAMPLE.replace('"JobResource"', '"ElasticJob"'))


@pytest.mark.gpu
def test_replica_change_drives_reshard(cuda):
    from easydl_b200.exchange import run_lockstep
    from easydl_b200.job_resource import ElasticShardSet, parse_job_resource
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.sharded_step import ShardedSparseStep
    rows, dim = [5000, 301, 70001, 11], 16
    po = OptimizerSpec.adagrad()
    shards = [ShardParameters(rows, dim, po, 4, s, device=cuda) for s in range(4)]
    osh = [O.OracleShard(rows, dim, O.OptimizerSpec.adagrad(), 4, s) for s in range(4)]
    bases = O.key_bases(rows)

    def train(params, oshards, k0):
        S = len(params)
        st = [ShardedSparseStep(p, 128) for p in params]
        kw, gw, gens = [], [], []
        for r in range(S):
            _, ids, _ = O.criteo_batch(rows, 128, 1.05, step=k0 * 8 + r)
            g = O.synthetic_grad(ids.size, dim, step=k0 * 8 + r)
            kw.append(O.fuse_keys(ids, bases)); gw.append(g)
            gens.append(st[r].gen_step(torch.from_numpy(ids).to(cuda), torch.from_numpy(g).to(cuda)))
        run_lockstep(gens)
        O.sharded_step(oshards, kw, gw)

    es = ElasticShardSet(shards, rows=rows, dim=dim, opt=po, seed=O.DEFAULT_SEED, device=cuda, world=8)
    train(es.shards, osh, 0)
    assert es.reconcile(parse_job_resource(REFERENCE_EXAMPLE)) is False          # replicas: 4 == current
    for k, new_s in enumerate((8, 6), start=1):
        jr = parse_job_resource(REFERENCE_EXAMPLE.replace("parameter_server:\n    replicas: 4",
                                                          f"parameter_server:\n    replicas: {new_s}"))
        assert es.reconcile(jr) is True and es.replicas == new_s
        osh = O.reshard(osh, new_s)
        train(es.shards, osh, k)
    assert es.history == [4, 8, 6]
    for p, o in zip(es.shards, osh):
        rec = p.export_records().cpu()
        assert np.array_equal(rec.keys.numpy(), o.keys) and rec.step == o.step
        np.testing.assert_allclose(rec.rows.numpy(), o.w, rtol=1e-6, atol=1e-9)
