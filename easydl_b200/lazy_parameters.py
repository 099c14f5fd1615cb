"""A PS shard that stores only the rows it has updated.

Same surface as ShardParameters (lookup / apply_gradients / export_records / import_records), so
the sharded step and the elastic re-shard run on it unchanged.  Needed when the dense slab does not
fit: BASELINE.json configs[2] (26 x 40M rows x dim 128, Adam) is 1.6 TB dense, 200 GB per GPU on
eight 180 GB B200s.  `max_rows` bounds the rows this shard can materialise; running out raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib
from . import kernels as K
from ._lib import check
from .criteo import DEFAULT_SEED
from .layout import ShardLayout
from .optimizer import OptimizerSpec
from .parameters import ShardRecords


def _pow2_at_least(x: int) -> int:
    p = 1
    while p < x:
        p <<= 1
    return p


class LazyShardParameters:
    def __init__(self, rows: Sequence[int], dim: int, opt: OptimizerSpec, num_shards: int = 1,
                 shard: int = 0, max_rows: int = 1 << 20, seed: int = DEFAULT_SEED, device="cuda:0"):
        if dim <= 0 or dim % 4:
            raise ValueError("dim must be a positive multiple of 4")
        self.layout = ShardLayout(rows, num_shards, shard)
        self.dim, self.opt, self.seed = int(dim), opt, int(seed)
        self.device = torch.device(device)
        self.step = 0
        self.layout_dev = self.layout.to_device(self.device)
        self.lr_dev = torch.zeros(1, dtype=torch.float32, device=self.device)   # see ShardParameters.stage_rate
        self._lr_ring = None
        self.touched = None                                          # every stored row is a touched row
        self._allocate(int(max_rows))

    def _allocate(self, max_rows: int):
        """Slab of records, the stamp array and an empty hash index for `max_rows` materialised rows."""
        dim, opt = self.dim, self.opt
        self.max_rows = int(max_rows)
        # + 1: the spare row that absorbs writes after an overflow (reported by materialised()).
        # One record [p | slot0 | slot1] per row, as in ShardParameters (one DRAM page per updated id).
        self.records = torch.empty((self.max_rows + 1, dim * (1 + opt.n_slots)), dtype=torch.float32, device=self.device)
        self.slab = self.records[:, :dim]
        self.slots = [self.records[:, dim * (k + 1):dim * (k + 2)] for k in range(opt.n_slots)]
        self.row_stride = dim * (1 + opt.n_slots)
        self.stamp = torch.zeros(self.max_rows + 1, dtype=torch.int32, device=self.device)   # step of the last update
        self.index_capacity = _pow2_at_least(2 * self.max_rows)
        lib = _lib.load()
        self.index = torch.empty(lib.edl_lazy_index_bytes(self.index_capacity), dtype=torch.uint8,
                                 device=self.device)
        check(lib.edl_lazy_clear(self.index.data_ptr(), self.index_capacity, K._stream()), "edl_lazy_clear")
        off = self.index_capacity * 12
        self.status = self.index[off:off + 8].view(torch.int32)     # [rows materialised, overflow]
        self._slot_buf = None

    num_shards = property(lambda self: self.layout.num_shards)
    shard = property(lambda self: self.layout.shard)
    _layout_arg = property(lambda self: self.layout_dev)

    def reset(self):
        check(_lib.load().edl_lazy_clear(self.index.data_ptr(), self.index_capacity, K._stream()), "edl_lazy_clear")
        self.step = 0

    def materialised(self) -> int:
        n, err = (int(x) for x in self.status.cpu())
        if err:
            raise RuntimeError(f"lazy shard {self.shard}: more than max_rows={self.max_rows} rows were updated")
        return n

    # ------------------------------------------------------------------ A3
    def lookup(self, keys: torch.Tensor, *, n_dev=None, out=None) -> torch.Tensor:
        lib = _lib.load()
        n = keys.numel()
        if out is None:
            out = torch.empty((n, self.dim), dtype=torch.float32, device=self.device)
        check(lib.edl_lazy_gather(self.index.data_ptr(), self.index_capacity, self.slab.data_ptr(), self.dim,
                                  self.row_stride, K._ptr(keys), n, K._ptr(n_dev), self.layout_dev.data_ptr(),
                                  C.c_uint64(self.seed & (2 ** 64 - 1)), K._ptr(out), K._stream()),
              "edl_lazy_gather")
        return out

    # ------------------------------------------------------------------ A4 + A5
    def _slots_for(self, keys: torch.Tensor, n_dev=None, reuse: bool = False) -> torch.Tensor:
        lib = _lib.load()
        n = keys.numel()
        if reuse:   # the step's own scratch: one buffer, grown on demand, so a captured step allocates nothing
            if self._slot_buf is None or self._slot_buf.numel() < max(n, 1):
                self._slot_buf = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
            slots = self._slot_buf
        else:
            slots = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        s1 = self.slots[1] if len(self.slots) > 1 else None
        check(lib.edl_lazy_insert(self.index.data_ptr(), self.index_capacity, self.slab.data_ptr(),
                                  self.slots[0].data_ptr(), K._ptr(s1), self.dim, self.row_stride, self.max_rows,
                                  K._ptr(keys), n,
                                  K._ptr(n_dev), self.layout_dev.data_ptr(), C.c_uint64(self.seed & (2 ** 64 - 1)),
                                  C.c_float(self.opt.slot_init(0)), C.c_float(self.opt.slot_init(1)),
                                  slots.data_ptr(), K._stream()), "edl_lazy_insert")
        return slots

    _RING = 1024

    @property
    def rate_changes(self) -> bool:
        from ._lib import EDL_OPT_ADAM
        return self.opt.kind == EDL_OPT_ADAM

    def stage_rate(self, step: int):
        """rate(step) -> lr_dev on the current stream (same contract as ShardParameters.stage_rate)."""
        if self._lr_ring is None:
            self._lr_ring = torch.zeros(self._RING, dtype=torch.float32).pin_memory()
        k = step % self._RING
        self._lr_ring[k] = self.opt.rate(step)
        self.lr_dev.copy_(self._lr_ring[k:k + 1], non_blocking=True)

    def apply_gradients(self, uq: K.UniqueResult, grad: torch.Tensor, max_unique: Optional[int] = None,
                        workspace=None, n_rows: Optional[int] = None, bump_step: bool = True,
                        rate_on_device: bool = False, grad32: Optional[torch.Tensor] = None):
        """Same signature as ShardParameters.apply_gradients: insert-if-absent for the unique keys, then the
        fused reduce + optimizer on their slab rows."""
        if bump_step:
            self.step += 1
        slots = self._slots_for(uq.uniq, uq.n_unique, reuse=True)
        check(_lib.load().edl_lazy_stamp(self.stamp.data_ptr(), slots.data_ptr(), uq.n, K._ptr(uq.n_unique),
                                         int(self.step), K._stream()), "edl_lazy_stamp")
        s1 = self.slots[1] if len(self.slots) > 1 else None
        opt = self.opt.params(self.step, self.lr_dev.data_ptr() if rate_on_device else 0)
        K.segreduce_update(self.slab, self.slots[0], s1, slots, uq.order, uq.seg_offsets, uq.n_unique,
                           uq.n if max_unique is None else max_unique, grad, None,
                           opt, None, workspace, seg_first=uq.seg_first, n_rows=n_rows, grad32=grad32)

    # ------------------------------------------------------------------ A6
    def export_records(self) -> ShardRecords:
        return self._export()[0]

    def _export(self):
        """(records ascending by key, the last-update stamp of each)."""
        lib = _lib.load()
        n = self.materialised()
        keys = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        slots = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        if n:
            check(lib.edl_lazy_dump(self.index.data_ptr(), self.index_capacity, keys.data_ptr(), slots.data_ptr(),
                                    K._stream()), "edl_lazy_dump")
            uq = K.unique(keys[:n].contiguous(), None, self.layout.key_bits, want_inverse=False, want_counts=False)
            order = uq.order[:n].to(torch.int64)               # ascending-key order of the slots
            keys_sorted = uq.uniq[:n]
            rows = K.gather_keys(self.slab, order, None)
            sl = [K.gather_keys(s, order, None) for s in self.slots]
            stamps = self.stamp[order]
        else:
            keys_sorted = keys[:0]
            rows = torch.empty((0, self.dim), dtype=torch.float32, device=self.device)
            sl = [rows.clone() for _ in self.slots]
            stamps = torch.empty(0, dtype=torch.int32, device=self.device)
        return ShardRecords(keys_sorted, rows, sl, self.step), stamps

    def import_records(self, rec: ShardRecords, stamps: Optional[torch.Tensor] = None):
        self.step = max(self.step, rec.step)
        if len(rec) == 0:
            return
        # keep the records this shard owns (id mod S == shard), then insert + overwrite
        part = K.partition_mod(rec.keys.contiguous(), self.layout_dev, self.num_shards, want_pos=False)
        offs = [int(x) for x in part.offsets.cpu()]
        a, b = offs[self.shard], offs[self.shard + 1]
        if b == a:
            return
        keys = part.keys[a:b].contiguous()
        perm = part.perm[a:b].contiguous()
        slots = self._slots_for(keys)[:b - a].to(torch.int32)
        K.scatter_rows(K.expand(rec.rows.contiguous(), perm), slots, self.slab)
        for k, s in enumerate(self.slots):
            K.scatter_rows(K.expand(rec.slots[k].contiguous(), perm), slots, s)
        self.stamp[slots.long()] = rec.step if stamps is None else stamps[perm.long()]

    # ------------------------------------------------------------------ growth, eviction, raw checkpoint (8(f).3)
    def grow(self, new_max_rows: int):
        """Re-house the shard in a slab of `new_max_rows` rows (host-synchronising; between steps): every
        materialised row, its slots, its stamp and the step move; results are unchanged."""
        if new_max_rows <= self.max_rows:
            raise ValueError("grow() needs a larger max_rows")
        rec, stamps = self._export()
        step = self.step
        self._allocate(new_max_rows)
        self.import_records(rec, stamps)
        self.step = step

    def ensure_room(self, n_new: int, factor: float = 2.0) -> bool:
        """Grow (by `factor`) if fewer than n_new rows are free; call between steps with the number of ids
        the next step may materialise (at most its unique ids).  -> True if the shard grew."""
        if self.materialised() + int(n_new) <= self.max_rows:
            return False
        self.grow(max(int(self.max_rows * factor), self.materialised() + int(n_new)))
        return True

    def evict_older_than(self, step_cutoff: int) -> ShardRecords:
        """Drop the rows whose last update is older than `step_cutoff` and return them (keys ascending).  The
        slab keeps the rest.  LOSSY unless the caller keeps the returned records (a checkpoint file, host
        memory) and import_records() them again before the ids come back: a dropped row reads as its initial
        value and restarts its optimizer state."""
        rec, stamps = self._export()
        keep = stamps >= int(step_cutoff)
        step = self.step

        def pick(mask):
            return ShardRecords(rec.keys[mask].contiguous(), rec.rows[mask].contiguous(),
                                [s[mask].contiguous() for s in rec.slots], step)
        gone, kept, kept_stamps = pick(~keep), pick(keep), stamps[keep].contiguous()
        self._allocate(self.max_rows)
        self.import_records(kept, kept_stamps)
        self.step = step
        return gone

    def save_raw(self, path: str):
        """Checkpoint the shard AS IT LIES: the hash index, the materialised prefix of the slab and the
        stamps.  load_raw() restores it without re-inserting a single key (same slots, same probes)."""
        n = self.materialised()
        torch.save({"index": self.index.cpu(), "records": self.records[:n].cpu(), "stamp": self.stamp[:n].cpu(),
                    "step": self.step, "max_rows": self.max_rows, "dim": self.dim, "opt_kind": self.opt.kind,
                    "S": self.num_shards, "shard": self.shard, "seed": self.seed}, path)

    def load_raw(self, path: str):
        blob = torch.load(path, map_location="cpu")
        for k, v in (("dim", self.dim), ("opt_kind", self.opt.kind), ("S", self.num_shards), ("shard", self.shard),
                     ("seed", self.seed)):
            if blob[k] != v:
                raise ValueError(f"{path}: {k} = {blob[k]} in the file, {v} here")
        if blob["max_rows"] != self.max_rows:
            self._allocate(blob["max_rows"])
        n = blob["records"].shape[0]
        self.index.copy_(blob["index"])
        self.records[:n].copy_(blob["records"])
        self.stamp[:n].copy_(blob["stamp"])
        self.step = int(blob["step"])

    # ------------------------------------------------------------------ test helper
    def read_rows(self, keys: torch.Tensor):
        """(rows, slots...) of materialised keys."""
        slots = self._slots_for(keys.contiguous())[:keys.numel()]
        return [K.gather_keys(self.slab, slots, None)] + [K.gather_keys(s, slots, None) for s in self.slots]

    def nbytes(self) -> int:
        return self.slab.numel() * 4 * (1 + len(self.slots)) + self.index.numel()
