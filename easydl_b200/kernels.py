"""Thin torch-tensor wrappers over the C ABI (include/edlsparse.h).

torch is plumbing here: it owns device memory and streams.  Every function enqueues on
torch's current CUDA stream and returns without synchronising; counts produced on the device
stay on the device (`n_unique` tensors) unless the caller asks for a trimmed result.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import OptParams, check


def _ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    return t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _need(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise _lib.EdlError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


def _rows(t: torch.Tensor, name: str):
    """A [rows, dim] fp32 table or slot array whose rows may lie apart (a column window of interleaved
    records): -> row stride in floats for the C ABI."""
    if not t.is_cuda:
        raise _lib.EdlError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if t.dtype != torch.float32 or t.dim() != 2:
        raise TypeError(f"{name} must be float32 [rows, dim], got {t.dtype} {tuple(t.shape)}")
    if t.shape[0] > 1 and t.shape[1] > 1 and t.stride(1) != 1:
        raise ValueError(f"{name}: the elements of a row must be contiguous")
    return int(t.stride(0)) if t.shape[0] > 1 else int(t.shape[1])


def _same_stride(slab, *slots):
    rs = _rows(slab, "slab")
    for k, sl in enumerate(slots):
        if sl is not None and _rows(sl, f"slot{k}") != rs:
            raise ValueError("slab and optimizer slots must share one row stride")
    return rs


def _ws(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


def key_bits_for(total_rows: int) -> int:
    """Number of significant bits of a fused key in [0, total_rows)."""
    return max(1, int(total_rows - 1).bit_length())


# ------------------------------------------------------------------------------- init / fill
def init_rows(slab: torch.Tensor, layout: torch.Tensor, seed: int, row_begin: int = 0,
              n_rows: Optional[int] = None):
    lib = _lib.load()
    rs = _rows(slab, "slab")
    n_rows = slab.shape[0] - row_begin if n_rows is None else n_rows
    check(lib.edl_init_rows(_ptr(slab), slab.shape[1], rs, _ptr(layout), row_begin, n_rows,
                            C.c_uint64(seed & (2 ** 64 - 1)), _stream()), "edl_init_rows")


def fill(t: torch.Tensor, value: float):
    lib = _lib.load()
    if t.dim() == 2 and not t.is_contiguous():     # a slot inside interleaved records
        rs = _rows(t, "t")
        check(lib.edl_fill_rows_f32(_ptr(t), t.shape[1], rs, t.shape[0], C.c_float(value), _stream()),
              "edl_fill_rows_f32")
        return
    _need(t, torch.float32, "t")
    check(lib.edl_fill_f32(_ptr(t), t.numel(), C.c_float(value), _stream()), "edl_fill_f32")


# ------------------------------------------------------------------------------- A1
@dataclass
class UniqueResult:
    """Device-resident result of A1.  Arrays have capacity n; the first U entries are valid."""
    uniq: torch.Tensor          # int64 [n]
    inverse: Optional[torch.Tensor]  # int32 [n]
    counts: Optional[torch.Tensor]   # int32 [n]
    order: torch.Tensor         # int32 [n]
    seg_offsets: torch.Tensor   # int32 [n+1]
    n_unique: torch.Tensor      # int32 [1] on device
    n: int
    seg_first: Optional[torch.Tensor] = None  # int32 [n]: order[seg_offsets[u]]

    def count(self) -> int:
        """Host value of U (synchronises)."""
        return int(self.n_unique.item())

    def trimmed(self):
        u = self.count()
        return (self.uniq[:u], self.inverse, None if self.counts is None else self.counts[:u],
                self.order, self.seg_offsets[:u + 1])


def unique(ids: torch.Tensor, field_base: Optional[torch.Tensor] = None, key_bits: int = 64, *,
           field_bits: int = 0, want_inverse: bool = True, want_counts: bool = True,
           n_dev: Optional[torch.Tensor] = None, out: Optional[UniqueResult] = None,
           workspace: Optional[torch.Tensor] = None) -> UniqueResult:
    lib = _lib.load()
    _need(ids, torch.int64, "ids")
    n = ids.numel()
    dev = ids.device
    n_fields = 0
    if field_base is not None:
        _need(field_base, torch.int64, "field_base")
        n_fields = field_base.numel()
    if out is None:
        out = UniqueResult(
            uniq=torch.empty(n, dtype=torch.int64, device=dev),
            inverse=torch.empty(n, dtype=torch.int32, device=dev) if want_inverse else None,
            counts=torch.empty(n, dtype=torch.int32, device=dev) if want_counts else None,
            order=torch.empty(n, dtype=torch.int32, device=dev),
            seg_offsets=torch.empty(n + 1, dtype=torch.int32, device=dev),
            n_unique=torch.empty(1, dtype=torch.int32, device=dev), n=n,
            seg_first=torch.empty(max(n, 1), dtype=torch.int32, device=dev))
    need = lib.edl_unique_workspace_bytes(n, key_bits, n_fields, field_bits)
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, dev)
    check(lib.edl_unique(_ptr(ids), n, _ptr(n_dev), _ptr(field_base), n_fields, key_bits, field_bits,
                         _ptr(out.uniq), _ptr(out.inverse), _ptr(out.counts), _ptr(out.order),
                         _ptr(out.seg_offsets), _ptr(out.seg_first), _ptr(out.n_unique),
                         _ptr(workspace), workspace.numel(), _stream()), "edl_unique")
    return out


def unique_launches(n: int, n_fields: int, key_bits: int, field_bits: int = 0, *, n_dev: bool = False,
                    want_counts: bool = False) -> int:
    """Kernel launches one edl_unique call makes (sort.cu, unique_impl) -- for the launch counts bench.py
    reports.  Mirrors the path choice there: field-major batches of at most 65536 samples per field go
    through k_field_major_keys + k_cluster_sort (dedup_impl 1 / 2), everything else through three
    kernels per 8-bit radix pass plus three unique kernels."""
    if n == 0:
        return 1
    extra = 1 if want_counts else 0
    passes_flat = (key_bits + 7) // 8
    fm_ok = field_bits > 0 and n_fields > 1 and not n_dev and n % n_fields == 0
    if fm_ok:
        impl = _lib.tuning_get("dedup_impl")
        B = n // n_fields
        in_cluster = field_bits <= 32 and B <= 65536 and impl >= 1
        passes_fm = (field_bits + 7) // 8
        if in_cluster:
            transposed = 1 if n_fields * 129 * 4 <= 48 * 1024 else 0
            if impl == 2:
                return transposed + 1 + extra
            cs, ct = (8, 512) if _lib.tuning_get("dedup_shape") == 1 else (4, 1024)
            chunk = ((B + cs - 1) // cs + ct - 1) // ct * ct
            heads_fused = B % 2048 == 0 and chunk % 2048 == 0
            return transposed + 1 + (0 if heads_fused else 1) + 2 + extra
        if passes_fm < passes_flat:
            return 3 * passes_fm + 3 + extra
    return 3 * passes_flat + 3 + extra


# ------------------------------------------------------------------------------- A2
@dataclass
class PartitionResult:
    keys: torch.Tensor      # int64 [n] grouped by shard
    offsets: torch.Tensor   # int32 [S+1]
    perm: torch.Tensor      # int32 [n] original position of each output
    pos: Optional[torch.Tensor]  # int32 [n] output position of each input


def partition_buffers(n: int, num_shards: int, device, want_pos: bool = True) -> PartitionResult:
    return PartitionResult(torch.empty(n, dtype=torch.int64, device=device),
                           torch.empty(num_shards + 1, dtype=torch.int32, device=device),
                           torch.empty(n, dtype=torch.int32, device=device),
                           torch.empty(n, dtype=torch.int32, device=device) if want_pos else None)


def partition_mod(keys: torch.Tensor, layout: Optional[torch.Tensor], num_shards: int, *,
                  n_dev: Optional[torch.Tensor] = None, want_pos: bool = True,
                  workspace: Optional[torch.Tensor] = None, out: Optional[PartitionResult] = None) -> PartitionResult:
    lib = _lib.load()
    _need(keys, torch.int64, "keys")
    n = keys.numel()
    dev = keys.device
    res = out if out is not None else partition_buffers(n, num_shards, dev, want_pos)
    need = lib.edl_partition_mod_workspace_bytes(n, num_shards)
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, dev)
    check(lib.edl_partition_mod(_ptr(keys), n, _ptr(n_dev), _ptr(layout), num_shards,
                                _ptr(res.keys), _ptr(res.offsets), _ptr(res.perm), _ptr(res.pos),
                                _ptr(workspace), workspace.numel(), _stream()), "edl_partition_mod")
    return res


# ------------------------------------------------------------------------------- A3
def gather_keys(slab: torch.Tensor, keys: torch.Tensor, layout: Optional[torch.Tensor] = None, *,
                n_dev: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None):
    lib = _lib.load()
    rs = _rows(slab, "slab")
    _need(keys, torch.int64, "keys")
    n, dim = keys.numel(), slab.shape[1]
    if out is None:
        out = torch.empty((n, dim), dtype=torch.float32, device=slab.device)
    check(lib.edl_gather_keys(_ptr(slab), dim, rs, _ptr(keys), n, _ptr(n_dev), _ptr(layout), _ptr(out),
                              _stream()), "edl_gather_keys")
    return out


def gather_ids(slab: torch.Tensor, ids: torch.Tensor, field_base: Optional[torch.Tensor], *,
               out: Optional[torch.Tensor] = None):
    lib = _lib.load()
    rs = _rows(slab, "slab")
    _need(ids, torch.int64, "ids")
    n, dim = ids.numel(), slab.shape[1]
    if out is None:
        out = torch.empty((n, dim), dtype=torch.float32, device=slab.device)
    nf = 0 if field_base is None else field_base.numel()
    check(lib.edl_gather_ids(_ptr(slab), dim, rs, _ptr(ids), n, _ptr(field_base), nf, _ptr(out),
                             _stream()), "edl_gather_ids")
    return out


def expand(rows: torch.Tensor, index: torch.Tensor, *, out: Optional[torch.Tensor] = None):
    lib = _lib.load()
    _need(rows, torch.float32, "rows")
    _need(index, torch.int32, "index")
    n, dim = index.numel(), rows.shape[1]
    if out is None:
        out = torch.empty((n, dim), dtype=torch.float32, device=rows.device)
    check(lib.edl_expand(_ptr(rows), dim, _ptr(index), n, _ptr(out), _stream()), "edl_expand")
    return out


def expand_window(rows: torch.Tensor, index: torch.Tensor, fields_in_window: int, fields_total: int, field0: int,
                  out: torch.Tensor):
    """out[(i // Fw) * F + field0 + i % Fw] = rows[index[i]]: one field group of a [B, F] batch."""
    lib = _lib.load()
    _need(rows, torch.float32, "rows")
    _need(index, torch.int32, "index")
    _need(out, torch.float32, "out")
    n, dim = index.numel(), rows.shape[1]
    if out.numel() < (n // max(fields_in_window, 1)) * fields_total * dim:
        raise ValueError("expand_window: out is smaller than [n / fields_in_window * fields_total, dim]")
    if rows.numel() // 4 >= 1 << 32:
        raise ValueError("expand_window: rows larger than 2^32 float4")
    check(lib.edl_expand_window(_ptr(rows), dim, _ptr(index), n, fields_in_window, fields_total, field0, _ptr(out),
                                _stream()), "edl_expand_window")
    return out


def compose_index(index: torch.Tensor, mapping: torch.Tensor):
    lib = _lib.load()
    _need(index, torch.int32, "index")
    _need(mapping, torch.int32, "mapping")
    out = torch.empty_like(index)
    check(lib.edl_compose_index(_ptr(index), _ptr(mapping), index.numel(), _ptr(out), _stream()),
          "edl_compose_index")
    return out


# ------------------------------------------------------------------------------- A7 (slot rounds)
def slot_send(rows: torch.Tensor, offsets: torch.Tensor, dst_bases: torch.Tensor, slot_limit: int,
              count_ptrs: Optional[torch.Tensor], err: torch.Tensor, n_max: int,
              flag_seg: Optional[torch.Tensor] = None, flag_perm: Optional[torch.Tensor] = None):
    """rows (partition order, any dtype with 8-byte-multiple rows) -> this rank's slot at every peer.
    flag_seg / flag_perm: mark the keys of ids with one local gradient row (bit 62; the narrow wire)."""
    lib = _lib.load()
    _need(offsets, torch.int32, "offsets")
    _need(dst_bases, torch.int64, "dst_bases")
    _need(err, torch.int32, "err")
    if not rows.is_contiguous():
        raise ValueError("rows must be contiguous")
    row_bytes = rows.element_size() * (rows.numel() // max(rows.shape[0], 1)) if rows.dim() > 1 else rows.element_size()
    check(lib.edl_slot_send(_ptr(rows), row_bytes, _ptr(offsets), offsets.numel() - 1, _ptr(dst_bases),
                            slot_limit, _ptr(count_ptrs), _ptr(err), n_max, _ptr(flag_seg), _ptr(flag_perm),
                            _stream()), "edl_slot_send")


def slot_positions(pos: torch.Tensor, n_dev: Optional[torch.Tensor], offsets: torch.Tensor, slot_stride: int,
                   out: Optional[torch.Tensor] = None):
    lib = _lib.load()
    _need(pos, torch.int32, "pos")
    _need(offsets, torch.int32, "offsets")
    if out is None:
        out = torch.empty_like(pos)
    check(lib.edl_slot_positions(_ptr(pos), pos.numel(), _ptr(n_dev), _ptr(offsets), offsets.numel() - 1,
                                 slot_stride, _ptr(out), _stream()), "edl_slot_positions")
    return out


def slot_compact(keys_slots: torch.Tensor, counts: torch.Tensor, slot_stride: int, slot_limit: int,
                 keys_out: torch.Tensor, map_out: torch.Tensor, n_out: torch.Tensor):
    lib = _lib.load()
    _need(keys_slots, torch.int64, "keys_slots")
    _need(counts, torch.int32, "counts")
    _need(keys_out, torch.int64, "keys_out")
    _need(map_out, torch.int32, "map_out")
    _need(n_out, torch.int32, "n_out")
    S = counts.numel()
    if keys_out.numel() < S * slot_limit or map_out.numel() < S * slot_limit:
        raise ValueError("slot_compact: outputs smaller than num_slots * slot_limit")
    check(lib.edl_slot_compact(_ptr(keys_slots), _ptr(counts), S, slot_stride, slot_limit, _ptr(keys_out),
                               _ptr(map_out), _ptr(n_out), _stream()), "edl_slot_compact")


def slot_merge_unique(keys_slots: torch.Tensor, counts: torch.Tensor, slot_stride: int, slot_limit: int,
                      out: UniqueResult, n_out: torch.Tensor, workspace: Optional[torch.Tensor] = None) -> UniqueResult:
    """Owner side of the slot exchange in one call: the S sorted key slots merged into the dedup result
    `out` (uniq, seg_offsets, n_unique; order and seg_first as slot-space rows with bit 30 = fp32 gradient
    row) -- slot_compact + unique + 2 x remap_index without the sort."""
    lib = _lib.load()
    _need(keys_slots, torch.int64, "keys_slots")
    _need(counts, torch.int32, "counts")
    _need(n_out, torch.int32, "n_out")
    S = counts.numel()
    cap = S * slot_limit
    if out.uniq.numel() < cap or out.order.numel() < cap or out.seg_offsets.numel() < cap + 1:
        raise ValueError("slot_merge_unique: outputs smaller than num_slots * slot_limit")
    if out.seg_first is not None and out.seg_first.numel() < cap:
        raise ValueError("slot_merge_unique: seg_first smaller than num_slots * slot_limit")
    need = lib.edl_slot_merge_unique_workspace_bytes(S, slot_limit)
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, keys_slots.device)
    check(lib.edl_slot_merge_unique(_ptr(keys_slots), _ptr(counts), S, slot_stride, slot_limit, _ptr(out.uniq),
                                    _ptr(out.order), _ptr(out.seg_offsets), _ptr(out.seg_first), _ptr(out.n_unique),
                                    _ptr(n_out), _ptr(workspace), workspace.numel(), _stream()), "edl_slot_merge_unique")
    return out


def remap_index(index: torch.Tensor, mapping: torch.Tensor, n_dev: Optional[torch.Tensor],
                out: Optional[torch.Tensor] = None):
    lib = _lib.load()
    _need(index, torch.int32, "index")
    _need(mapping, torch.int32, "mapping")
    if out is None:
        out = torch.empty_like(index)
    check(lib.edl_remap_index(_ptr(index), _ptr(mapping), index.numel(), _ptr(n_dev), _ptr(out), _stream()),
          "edl_remap_index")
    return out


def scatter_rows(rows: torch.Tensor, pos: torch.Tensor, out: torch.Tensor, *,
                 n_dev: Optional[torch.Tensor] = None):
    lib = _lib.load()
    _need(rows, torch.float32, "rows")
    _need(pos, torch.int32, "pos")
    check(lib.edl_scatter_rows(_ptr(rows), rows.shape[1], _ptr(pos), pos.numel(), _ptr(n_dev),
                               _ptr(out), _rows(out, "out"), _stream()), "edl_scatter_rows")
    return out


# ------------------------------------------------------------------------------- A4 / A5
def segreduce(grad: torch.Tensor, order: torch.Tensor, seg_offsets: torch.Tensor,
              n_unique: torch.Tensor, max_unique: int, *, seg_first: Optional[torch.Tensor] = None,
              out_pos: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None, out_dtype=torch.float32,
              dst_table: Optional[torch.Tensor] = None,
              workspace: Optional[torch.Tensor] = None):
    """g[dst(u)] = float64 segment sum; out_dtype float32 (rounded once) or float64 (exact wire).
    dst_table: fused reduce -> send, rows are written to per-group base addresses (see the header)."""
    lib = _lib.load()
    _need(grad, torch.float32, "grad")
    _need(order, torch.int32, "order")
    _need(seg_offsets, torch.int32, "seg_offsets")
    dim = grad.shape[1]
    if out is None and dst_table is None:
        out = torch.empty((max_unique, dim), dtype=out_dtype, device=grad.device)
    odt = out.dtype if out is not None else out_dtype
    mixed = odt == "mixed"
    if not mixed and odt not in (torch.float32, torch.float64):
        raise TypeError("segreduce output must be float32, float64 or 'mixed'")
    if mixed and dst_table is None:
        raise ValueError("the mixed wire needs a dst_table")
    n_rows = grad.shape[0]
    need = lib.edl_segreduce_workspace_bytes(max_unique, n_rows, dim)
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, grad.device)
    check(lib.edl_segreduce(_ptr(grad), dim, _ptr(order), _ptr(seg_offsets), _ptr(seg_first),
                            _ptr(n_unique), max_unique, n_rows, _ptr(out_pos), _ptr(out),
                            2 if mixed else int(odt == torch.float64),
                            _ptr(dst_table), _ptr(workspace), workspace.numel(), _stream()), "edl_segreduce")
    return out


def segreduce_update(slab: torch.Tensor, slot0: torch.Tensor, slot1: Optional[torch.Tensor],
                     uniq: torch.Tensor, order: torch.Tensor, seg_offsets: torch.Tensor,
                     n_unique: torch.Tensor, max_unique: int, grad: torch.Tensor,
                     layout: Optional[torch.Tensor], opt: OptParams,
                     touched: Optional[torch.Tensor] = None,
                     workspace: Optional[torch.Tensor] = None,
                     seg_first: Optional[torch.Tensor] = None, n_rows: Optional[int] = None,
                     grad32: Optional[torch.Tensor] = None):
    """grad rows may be float32 (a worker's raw gradient) or float64 (exchanged partial sums); with
    grad32 given they are MIXED: row indices with bit 30 set name fp32 rows of grad32, the others float64
    rows of grad.  n_rows: bound on the number of gradient rows referenced (default: all rows of `grad`)."""
    lib = _lib.load()
    rs = _same_stride(slab, slot0, slot1)
    if grad.dtype not in (torch.float32, torch.float64):
        raise TypeError("grad must be float32 or float64")
    _need(grad, grad.dtype, "grad")
    _need(uniq, torch.int64, "uniq")
    dim = slab.shape[1]
    if grad.shape[1] != dim:
        raise ValueError("grad and slab dims differ")
    n_rows = grad.shape[0] if n_rows is None else int(n_rows)
    need = lib.edl_segreduce_workspace_bytes(max_unique, n_rows, dim)
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, slab.device)
    check(lib.edl_segreduce_update(_ptr(slab), _ptr(slot0), _ptr(slot1), dim, rs, _ptr(uniq),
                                   _ptr(order), _ptr(seg_offsets), _ptr(seg_first), _ptr(n_unique),
                                   max_unique, n_rows, _ptr(grad),
                                   2 if grad32 is not None else int(grad.dtype == torch.float64), _ptr(grad32), _ptr(layout),
                                   C.byref(opt), _ptr(touched), _ptr(workspace), workspace.numel(),
                                   _stream()), "edl_segreduce_update")


# ------------------------------------------------------------------------------- A6
def export_count(touched: torch.Tensor, n_rows: int) -> torch.Tensor:
    lib = _lib.load()
    n_out = torch.empty(1, dtype=torch.int32, device=touched.device)
    ws = _ws(lib.edl_export_workspace_bytes(n_rows), touched.device)
    check(lib.edl_export_count(_ptr(touched), n_rows, _ptr(n_out), _ptr(ws), ws.numel(), _stream()),
          "edl_export_count")
    return n_out


def export_shard(slab, slot0, slot1, touched, n_rows: int, layout, capacity: int):
    """-> keys[capacity] i64, rows, slot0, slot1 [capacity, dim], n_out (device int32)."""
    lib = _lib.load()
    dev, dim = slab.device, slab.shape[1]
    keys = torch.empty(capacity, dtype=torch.int64, device=dev)
    rows = torch.empty((capacity, dim), dtype=torch.float32, device=dev)
    o0 = torch.empty_like(rows) if slot0 is not None else None
    o1 = torch.empty_like(rows) if slot1 is not None else None
    n_out = torch.empty(1, dtype=torch.int32, device=dev)
    ws = _ws(lib.edl_export_workspace_bytes(n_rows), dev)
    check(lib.edl_export_shard(_ptr(slab), _ptr(slot0), _ptr(slot1), dim, _same_stride(slab, slot0, slot1), _ptr(touched), n_rows,
                               _ptr(layout), capacity, _ptr(keys), _ptr(rows), _ptr(o0), _ptr(o1),
                               _ptr(n_out), _ptr(ws), ws.numel(), _stream()), "edl_export_shard")
    return keys, rows, o0, o1, n_out


def import_shard(slab, slot0, slot1, touched, layout, keys, rows, slot0_in, slot1_in):
    lib = _lib.load()
    _need(keys, torch.int64, "keys")
    _need(rows, torch.float32, "rows")
    check(lib.edl_import_shard(_ptr(slab), _ptr(slot0), _ptr(slot1), slab.shape[1], _same_stride(slab, slot0, slot1), _ptr(touched),
                               _ptr(layout), _ptr(keys), _ptr(rows), _ptr(slot0_in), _ptr(slot1_in),
                               keys.numel(), _stream()), "edl_import_shard")
