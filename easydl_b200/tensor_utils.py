"""IndexedSlices helpers with the lineage's names on the GPU kernels."""
from __future__ import annotations

from typing import Sequence, Tuple

import torch

from . import kernels as K


def deduplicate_indexed_slices(values: torch.Tensor, indices: torch.Tensor, key_bits: int = 64
                               ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Sum the rows of `values` that share an index (A1 + A4): returns (summed_values[U, D],
    unique_indices[U] ascending).  Sums are accumulated in float64 and rounded once."""
    uq = K.unique(indices.reshape(-1), None, key_bits, want_inverse=False, want_counts=False)
    n = indices.numel()
    g = K.segreduce(values, uq.order, uq.seg_offsets, uq.n_unique, n, seg_first=uq.seg_first)
    u = uq.count()
    return g[:u], uq.uniq[:u]


def merge_indexed_slices(slices: Sequence[Tuple[torch.Tensor, torch.Tensor]]
                         ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Concatenate several (values, indices) pairs (duplicates are kept, as in the lineage)."""
    return torch.cat([v for v, _ in slices]), torch.cat([i for _, i in slices])
