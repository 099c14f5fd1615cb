"""Shard-assignment helpers with the lineage's names: which PS shard owns an
embedding id (`int_to_id`) or a dense variable (`string_to_id`), and the device-side grouping of
ids / rows by owner (`scatter_embedding_vector`, A2)."""
from __future__ import annotations

import hashlib
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import kernels as K


def int_to_id(number, bucket_num: int):
    """Owner shard of an embedding id: id mod S (SPEC "Shard function")."""
    if isinstance(number, torch.Tensor):
        return number % bucket_num
    return np.asarray(number, dtype=np.int64) % np.int64(bucket_num)


def string_to_id(name: str, bucket_num: int) -> int:
    """Owner shard of a dense variable: sha256 digest of its name, as an integer, mod S."""
    return int(hashlib.sha256(name.encode("utf-8")).hexdigest(), 16) % bucket_num


def scatter_embedding_vector(values: Optional[torch.Tensor], keys: torch.Tensor, bucket_num: int,
                             layout: Optional[torch.Tensor] = None):
    """Group `keys` (and the matching rows of `values`) by owner shard on the GPU.

    Returns (keys_by_shard, offsets[S+1] on the host as a list, rows_by_shard or None, perm).
    Stable: ascending keys stay ascending inside every shard's group."""
    part = K.partition_mod(keys, layout, bucket_num)
    rows = None
    if values is not None and keys.numel():
        rows = K.expand(values, part.perm)
    return part.keys, [int(x) for x in part.offsets.cpu()], rows, part.perm
