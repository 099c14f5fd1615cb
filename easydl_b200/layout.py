"""Shard layout math on the host (SPEC "Shard function"): which shard owns an id, where its
row lives in the shard's slab, and the small int64 array the kernels decode keys with."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Sequence

import numpy as np


def shard_rows(rows_t: int, shard: int, num_shards: int) -> int:
    """Shard s of a table with rows_t buckets holds ceil((rows_t - s) / S) local rows."""
    if shard >= rows_t:
        return 0
    return (rows_t - shard + num_shards - 1) // num_shards


@dataclass
class ShardLayout:
    rows: Sequence[int]      # rows_t per table
    num_shards: int = 1
    shard: int = 0
    key_base: np.ndarray = field(init=False)    # [T+1] fused-key bases (same on every shard)
    local_base: np.ndarray = field(init=False)  # [T+1] slab-row bases of this shard

    def __post_init__(self):
        self.rows = [int(r) for r in self.rows]
        if not (0 <= self.shard < self.num_shards):
            raise ValueError("shard index out of range")
        if len(self.rows) > 64:
            raise ValueError("at most 64 tables per layout (EDL_MAX_TABLES)")
        T = len(self.rows)
        self.key_base = np.zeros(T + 1, dtype=np.int64)
        self.key_base[1:] = np.cumsum(self.rows)
        loc = [shard_rows(r, self.shard, self.num_shards) for r in self.rows]
        self.local_base = np.zeros(T + 1, dtype=np.int64)
        self.local_base[1:] = np.cumsum(loc)

    @property
    def n_tables(self) -> int:
        return len(self.rows)

    @property
    def total_rows(self) -> int:
        return int(self.key_base[-1])

    @property
    def local_rows(self) -> int:
        return int(self.local_base[-1])

    @property
    def key_bits(self) -> int:
        return max(1, int(self.total_rows - 1).bit_length())

    @property
    def field_bits(self) -> int:
        """Bits of the largest per-table id (the radix sort of a fused batch runs on these)."""
        return max(1, int(max(self.rows) - 1).bit_length())

    def to_array(self) -> np.ndarray:
        """[T, S, shard, 0, key_base[T+1], local_base[T+1]] as the kernels expect."""
        T = self.n_tables
        a = np.zeros(4 + 2 * (T + 1), dtype=np.int64)
        a[0], a[1], a[2] = T, self.num_shards, self.shard
        a[4:4 + T + 1] = self.key_base
        a[4 + T + 1:] = self.local_base
        return a

    def to_device(self, device):
        import torch
        return torch.from_numpy(self.to_array()).to(device)

    def resharded(self, num_shards: int, shard: int) -> "ShardLayout":
        return ShardLayout(self.rows, num_shards, shard)
