"""Synthetic Criteo-shape input (SPEC "Synthetic Criteo-shape input"): 13 dense features,
26 categorical ids, 1 label per sample.  Generated on the host with PCG64 so the GPU path and
the CPU oracle can be handed byte-identical batches; no cross-device RNG matching needed."""
from __future__ import annotations

from typing import Sequence

import numpy as np

DEFAULT_SEED = 20240607
N_DENSE = 13
N_SPARSE = 26
_GOLDEN = 2654435761  # Knuth multiplicative constant: scatters hot Zipf ranks over the table


def zipf_field(rows: int, alpha: float, field: int, batch: int, seed: int) -> np.ndarray:
    """Bounded-Zipf ids for one categorical field, int64 in [0, rows)."""
    u = np.random.Generator(np.random.PCG64(seed + field)).random(batch)
    if abs(alpha - 1.0) < 1e-12:
        x = np.power(float(rows + 1), u)
    else:
        one_m = 1.0 - alpha
        x = np.power((np.power(float(rows + 1), one_m) - 1.0) * u + 1.0, 1.0 / one_m)
    rank = np.clip(np.floor(x).astype(np.int64) - 1, 0, rows - 1)
    return (rank * np.int64(_GOLDEN) + np.int64(field)) % np.int64(rows)


class CriteoSynthetic:
    """Batch source.  `batch(step)` is deterministic in (seed, step)."""

    def __init__(self, rows: Sequence[int], batch_size: int, alpha: float = 1.05,
                 seed: int = DEFAULT_SEED):
        self.rows = [int(r) for r in rows]
        self.batch_size = int(batch_size)
        self.alpha = float(alpha)
        self.seed = int(seed)

    def ids(self, step: int = 0) -> np.ndarray:
        s = self.seed + 7919 * step
        cols = [zipf_field(r, self.alpha, t, self.batch_size, s) for t, r in enumerate(self.rows)]
        return np.ascontiguousarray(np.stack(cols, axis=1))

    def batch(self, step: int = 0):
        s = self.seed + 7919 * step
        g = np.random.Generator(np.random.PCG64(s + 500))
        dense = np.log1p(g.lognormal(0.0, 1.0, size=(self.batch_size, N_DENSE))).astype(np.float32)
        labels = (g.random(self.batch_size) < 0.25).astype(np.float32)
        return dense, self.ids(step), labels

    def grad(self, dim: int, step: int = 0) -> np.ndarray:
        """Synthetic upstream gradient [B*F, dim] ~ N(0, sigma=1e-2), fp32."""
        g = np.random.Generator(np.random.PCG64(self.seed + 1000 + 7919 * step))
        n = self.batch_size * len(self.rows)
        return g.standard_normal((n, dim), dtype=np.float32) * np.float32(1e-2)
