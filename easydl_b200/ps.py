"""One parameter-server shard (lineage: `Parameters` + the PS servicer).

The embedding tables of the shard live in a ShardParameters slab on the GPU; dense (non
embedding) variables are kept as plain tensors and only stored / served -- their optimizer is
outside the hot path this repository covers.

Table handles follow the lineage: `EmbeddingTable.get(ids)` / `.set(ids, values)` by per-table
id, `pull_embedding_vectors(name, ids)`, `push_gradients(...)`.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import kernels as K
from .criteo import DEFAULT_SEED
from .optimizer import OptimizerSpec
from .parameters import ShardParameters, ShardRecords


class EmbeddingTable:
    """Handle on one table of a shard.  Rows exist from the start (the initialiser is a pure
    function of (seed, table, id), so the lineage's lazy initialisation and eager storage are
    indistinguishable)."""

    def __init__(self, params: ShardParameters, index: int, name: str):
        self.params, self.index, self.name = params, index, name
        self.dim = params.dim
        self.key_base = int(params.layout.key_base[index])
        self.rows = int(params.layout.rows[index])

    def _keys(self, ids: torch.Tensor) -> torch.Tensor:
        return ids.to(device=self.params.device, dtype=torch.int64).reshape(-1) + self.key_base

    def get(self, ids: torch.Tensor) -> torch.Tensor:
        """Rows of per-table ids owned by this shard -> [n, dim] on the device."""
        return self.params.lookup(self._keys(ids))

    def set(self, ids: torch.Tensor, values: torch.Tensor):
        """Overwrite rows (ids owned by this shard; others are ignored)."""
        keys = self._keys(ids)
        K.import_shard(self.params.slab, None, None, self.params.touched, self.params._layout_arg,
                       keys, values.to(self.params.device).contiguous(), None, None)


class ParameterServer:
    """Shard `shard` of `num_shards`.  Ids passed to pull / push must be owned by this shard
    (the client partitions with int_to_id)."""

    def __init__(self, table_names: Sequence[str], rows: Sequence[int], dim: int, opt: OptimizerSpec,
                 num_shards: int = 1, shard: int = 0, seed: int = DEFAULT_SEED, device="cuda:0", params=None):
        # params: serve an existing ShardParameters / LazyShardParameters instead of allocating tables
        self.names = list(table_names)
        self.embedding_params = params if params is not None else \
            ShardParameters(rows, dim, opt, num_shards, shard, seed=seed, device=device)
        self.tables: Dict[str, EmbeddingTable] = {
            n: EmbeddingTable(self.embedding_params, i, n) for i, n in enumerate(self.names)}
        self.non_embedding_params: Dict[str, torch.Tensor] = {}
        self.key_bits = self.embedding_params.layout.key_bits

    # ---- lineage servicer calls
    @property
    def version(self) -> int:
        return self.embedding_params.step

    def push_model(self, dense: Dict[str, torch.Tensor]):
        for k, v in dense.items():
            self.non_embedding_params[k] = v.to(self.embedding_params.device).clone()

    def pull_dense_parameters(self) -> Dict[str, torch.Tensor]:
        return dict(self.non_embedding_params)

    def pull_embedding_vectors(self, name: str, ids: torch.Tensor) -> torch.Tensor:
        return self.tables[name].get(ids)

    def pull_embedding_keys(self, keys: torch.Tensor) -> torch.Tensor:
        """Fused form: rows of fused keys (any table) owned by this shard."""
        return self.embedding_params.lookup(keys)

    def push_gradients(self, keys: torch.Tensor, grads: torch.Tensor):
        """Apply one batch of embedding gradients given as (fused keys, rows) with duplicates:
        same-key rows are summed in float64 in the order given, then the optimizer runs once
        per touched row (A1 + A4 + A5).  grads may be float32 or float64 rows."""
        p = self.embedding_params
        n = keys.numel()
        if n == 0:
            p.step += 1
            return
        uq = K.unique(keys, None, self.key_bits, want_inverse=False, want_counts=False)
        p.apply_gradients(uq, grads, n)

    def push_table_gradients(self, name: str, ids: torch.Tensor, grads: torch.Tensor):
        self.push_gradients(self.tables[name]._keys(ids), grads)

    # ---- elastic
    def export_shard(self) -> ShardRecords:
        return self.embedding_params.export_records()

    def import_shard(self, rec: ShardRecords):
        self.embedding_params.import_records(rec)


Parameters = ShardParameters  # lineage name for the PS-side parameter container
