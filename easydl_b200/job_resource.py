"""Control-plane shim: read the `JobResource` the ElasticTrainer publishes
(ref:docs/design/elastic-training-operator.md:57-95) and turn a change of
`spec.parameter_server.replicas` into the A6 re-shard of the embedding state.  Nothing of the
operator itself is implemented here -- no CRDs, no Pods; only the one field that drives the hot path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional

import yaml


@dataclass
class JobResource:
    name: str
    job: str
    ps_replicas: int
    worker_replicas: int
    ps_gpu: int = 0
    resource_updation: Optional[List[Dict]] = None   # ref :86-95: Pods to replace with new resources


def parse_job_resource(text: str) -> JobResource:
    """Parse the YAML of ref:docs/design/elastic-training-operator.md:57-95 (comments after `//`
    in the reference's example are tolerated)."""
    cleaned = "\n".join(line.split("//")[0].rstrip() for line in text.splitlines())
    doc = yaml.safe_load(cleaned)
    if doc.get("kind") != "JobResource":
        raise ValueError(f"not a JobResource: kind={doc.get('kind')!r}")
    spec = doc["spec"]
    ps = spec.get("parameter_server", {})
    wk = spec.get("worker", {})
    return JobResource(
        name=doc.get("metadata", {}).get("name", ""),
        job=spec.get("selector", {}).get("name", ""),
        ps_replicas=int(ps.get("replicas", 0)),
        worker_replicas=int(wk.get("replicas", 0)),
        ps_gpu=int(ps.get("resource", {}).get("gpu", 0)),
        resource_updation=spec.get("resource_updation"),
    )


class ElasticShardSet:
    """The active PS shards of one process group (real ranks) or of one process (virtual shards
    on a single GPU, which is what the tests drive).  `reconcile(job_resource)` re-shards the
    embedding state when the requested replica count differs from the current one."""

    def __init__(self, shards, *, rows, dim, opt, seed, device, world: Optional[int] = None):
        self.shards = list(shards)               # ShardParameters, index = shard id
        self.rows, self.dim, self.opt, self.seed, self.device = rows, dim, opt, seed, device
        self.world = world if world is not None else max(len(self.shards), 1)
        self.history = [len(self.shards)]

    @property
    def replicas(self) -> int:
        return len(self.shards)

    def reconcile(self, jr: JobResource) -> bool:
        """Returns True when a re-shard happened."""
        new_s = jr.ps_replicas
        if new_s <= 0:
            raise ValueError("parameter_server.replicas must be positive")
        if new_s == self.replicas:
            return False
        from .elastic import gen_reshard
        from .exchange import run_lockstep
        world = max(self.world, new_s, self.replicas)
        gens = [gen_reshard(self.shards[r] if r < self.replicas else None, new_s, r, world, rows=self.rows,
                            dim=self.dim, opt=self.opt, seed=self.seed, device=self.device)
                for r in range(world)]
        out = run_lockstep(gens)
        self.shards = out[:new_s]
        self.world = world
        self.history.append(new_s)
        return True
