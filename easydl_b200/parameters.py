"""One parameter-server shard's device-resident state.

Lineage name: `Parameters` (embedding_params of every table + optimizer slot tables + the
model version), SURVEY.md 2d.  Here all tables of one embedding dimension live in ONE slab
[local_rows, dim] in HBM (table-major, local row = id div S), with the optimizer slots in
slabs of the same shape and a 1-bit-per-row "touched" bitmap that records which rows ever
received an update -- exactly the rows a checkpoint / re-shard has to move (SPEC A6).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import kernels as K
from .criteo import DEFAULT_SEED
from .layout import ShardLayout
from .optimizer import OptimizerSpec


@dataclass
class ShardRecords:
    """Materialised rows of a shard, ascending by fused key (host or device tensors)."""
    keys: torch.Tensor            # int64 [n]
    rows: torch.Tensor            # f32 [n, dim]
    slots: List[torch.Tensor]     # n_slots x f32 [n, dim]
    step: int

    def __len__(self):
        return int(self.keys.numel())

    def cpu(self) -> "ShardRecords":
        return ShardRecords(self.keys.cpu(), self.rows.cpu(), [s.cpu() for s in self.slots], self.step)

    def to(self, device) -> "ShardRecords":
        return ShardRecords(self.keys.to(device), self.rows.to(device),
                            [s.to(device) for s in self.slots], self.step)


class ShardParameters:
    def __init__(self, rows: Sequence[int], dim: int, opt: OptimizerSpec, num_shards: int = 1,
                 shard: int = 0, seed: int = DEFAULT_SEED, device="cuda:0", init: bool = True,
                 interleaved: bool = False):
        if dim <= 0 or dim % 4:
            raise ValueError("dim must be a positive multiple of 4")
        self.layout = ShardLayout(rows, num_shards, shard)
        self.dim = int(dim)
        self.opt = opt
        self.seed = int(seed)
        self.device = torch.device(device)
        self.step = 0                      # number of applied batches (Adam's t)
        n = max(self.layout.local_rows, 1)
        self.layout_dev = self.layout.to_device(self.device)
        # interleaved: one record [p | slot0 | slot1] per row -- an update touches one DRAM page and one TLB
        # entry per id instead of one per array; slab / slots are then column windows of the records and the
        # kernels take the row stride.  Measured on B200 at cfg2 (profiles/r2_experiments.md): the bare access
        # pattern gains 27 % (tools/rowbw_probe), the update kernels as they are gain nothing (they are bound
        # by dependent-load latency, not by row activations) and the lookup loses 15 % (it reads every other
        # 256-byte block), so separate dense arrays stay the default.
        self.interleaved = bool(interleaved)
        if self.interleaved:
            self.records = torch.empty((n, dim * (1 + opt.n_slots)), dtype=torch.float32, device=self.device)
            self.slab = self.records[:, :dim]
            self.slots = [self.records[:, dim * (k + 1):dim * (k + 2)] for k in range(opt.n_slots)]
        else:
            self.records = None
            self.slab = torch.empty((n, dim), dtype=torch.float32, device=self.device)
            self.slots = [torch.empty((n, dim), dtype=torch.float32, device=self.device)
                          for _ in range(opt.n_slots)]
        self.touched = torch.zeros((n + 31) // 32 + 1, dtype=torch.int32, device=self.device)
        # the optimizer rate as a device scalar, for launch sequences that are captured once and
        # replayed (Adam's lr_t changes every step): stage_rate() copies rate(step) into it
        self.lr_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._lr_ring = None
        if init:
            self.reset()

    # ------------------------------------------------------------------ init
    def reset(self):
        """(Re)initialise every local row with the pure initialiser and clear the slots."""
        K.init_rows(self.slab, self.layout_dev, self.seed, 0, self.layout.local_rows)
        for k, s in enumerate(self.slots):
            K.fill(s, self.opt.slot_init(k))
        self.touched.zero_()
        self.step = 0

    @property
    def num_shards(self):
        return self.layout.num_shards

    @property
    def shard(self):
        return self.layout.shard

    @property
    def _layout_arg(self):
        # S == 1: slab row == fused key, the kernels skip the table search entirely
        return None if self.layout.num_shards == 1 else self.layout_dev

    # ------------------------------------------------------------------ A3
    def lookup(self, keys: torch.Tensor, *, n_dev=None, out=None) -> torch.Tensor:
        """rows[i] = table row of fused key keys[i] (keys must be owned by this shard)."""
        return K.gather_keys(self.slab, keys, self._layout_arg, n_dev=n_dev, out=out)

    # ------------------------------------------------------------------ A4 + A5
    _RING = 1024

    @property
    def rate_changes(self) -> bool:
        """True if the optimizer rate differs from step to step (Adam's bias correction): a captured launch
        sequence must then read it from lr_dev; a constant rate (Adagrad) is captured by value."""
        from ._lib import EDL_OPT_ADAM
        return self.opt.kind == EDL_OPT_ADAM

    def stage_rate(self, step: int):
        """Enqueue rate(step) -> lr_dev on the current stream (from a ring of pinned host slots, so the
        host never waits).  Call before replaying a launch sequence captured with rate_on_device."""
        if self._lr_ring is None:
            self._lr_ring = torch.zeros(self._RING, dtype=torch.float32).pin_memory()
        k = step % self._RING
        self._lr_ring[k] = self.opt.rate(step)
        self.lr_dev.copy_(self._lr_ring[k:k + 1], non_blocking=True)

    def apply_gradients(self, uq: K.UniqueResult, grad: torch.Tensor, max_unique: Optional[int] = None,
                        workspace=None, n_rows: Optional[int] = None, bump_step: bool = True,
                        rate_on_device: bool = False, grad32: Optional[torch.Tensor] = None):
        """Segment-sum grad rows per unique key and run the optimizer on the touched rows.
        bump_step=False: a further part of the same training step (disjoint ids), same step number.
        rate_on_device: the kernels read the rate from lr_dev (see stage_rate) instead of taking it by value."""
        if bump_step:
            self.step += 1
        slot1 = self.slots[1] if len(self.slots) > 1 else None
        opt = self.opt.params(self.step, self.lr_dev.data_ptr() if rate_on_device else 0)
        K.segreduce_update(self.slab, self.slots[0], slot1, uq.uniq, uq.order, uq.seg_offsets,
                           uq.n_unique, uq.n if max_unique is None else max_unique, grad,
                           self._layout_arg, opt, self.touched, workspace,
                           seg_first=uq.seg_first, n_rows=n_rows, grad32=grad32)

    # ------------------------------------------------------------------ A6
    def export_records(self) -> ShardRecords:
        """Materialised rows, ascending by key (device tensors).  One small D2H sync for n."""
        n_rows = self.layout.local_rows
        n = int(K.export_count(self.touched, n_rows).item())
        s0 = self.slots[0]
        s1 = self.slots[1] if len(self.slots) > 1 else None
        keys, rows, o0, o1, _ = K.export_shard(self.slab, s0, s1, self.touched, n_rows,
                                               self._layout_arg, max(n, 1))
        slots = [o0[:n]] + ([o1[:n]] if o1 is not None else [])
        return ShardRecords(keys[:n], rows[:n], slots, self.step)

    def import_records(self, rec: ShardRecords):
        """Write records this shard owns (id mod S == shard) into the slab; others are skipped."""
        if len(rec) == 0:
            self.step = max(self.step, rec.step)
            return
        s1 = self.slots[1] if len(self.slots) > 1 else None
        K.import_shard(self.slab, self.slots[0], s1, self.touched, self._layout_arg,
                       rec.keys.contiguous(), rec.rows.contiguous(), rec.slots[0].contiguous(),
                       rec.slots[1].contiguous() if len(rec.slots) > 1 else None)
        self.step = max(self.step, rec.step)

    # ------------------------------------------------------------------ test / debug helpers
    def read_rows(self, keys: torch.Tensor):
        """(rows, slots...) for fused keys owned by this shard -- through the gather kernel."""
        out = [K.gather_keys(self.slab, keys, self._layout_arg)]
        for s in self.slots:
            out.append(K.gather_keys(s, keys, self._layout_arg))
        return out

    def nbytes(self) -> int:
        return self.slab.numel() * 4 * (1 + len(self.slots)) + self.touched.numel() * 4

    def state_tensors(self):
        """[parameter rows, slot rows...] as dense copies (tests, digests)."""
        return [self.slab.contiguous()] + [s.contiguous() for s in self.slots]
