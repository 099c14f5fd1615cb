"""The sharded sparse step without a host round trip.

ShardedSparseStep needs the count matrix of every all-to-all-v on the host, so the GPU drains once
per step and the host issues ~45 launches behind it.  Here each receive buffer is cut into one
fixed slot per sending rank (csrc/slots.cu): senders need only their own partition offsets, which
stay on the device, and row counts travel with the data.  The step is then a fixed launch sequence

    A  dedup(local ids) -> partition by owner -> keys into my slot at every owner        | fence
    B  owner: gather the rows of every requester's slot straight into that requester     | fence
    C  expand my rows to emb;  reduce grad per unique id straight into the owners' slots | fence
    D  owner: compact the key slots (peer-rank order) -> dedup -> reduce -> optimizer

with no size known to the host: it is issued without synchronising and can be replayed from a CUDA
graph (GraphedSlotStep).  Same arithmetic and the same merge order as ShardedSparseStep.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import torch

from . import _lib
from . import kernels as K
from .exchange import _RawDeviceMemory
from .parameters import ShardParameters


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


def slot_layout(world: int, stride: int, dim: int, wire_bytes: int, mixed: bool = False):
    """Byte layout of one rank's receive buffer -> (offsets {(gen, region): byte offset}, flags_off,
    counter_off, err_off, nbytes).  Pure host arithmetic (tested without a GPU).  mixed: the narrow exact
    gradient wire, a float64 region (grads) plus an fp32 region (grads32) with the same slot geometry."""
    row_bytes = {"keys": 8, "rows": dim * 4, "grads": dim * wire_bytes, "grads32": dim * 4 if mixed else 0, "counts": 4}
    offs, off = {}, 0
    for gen in (0, 1):
        for r in SlotBuffers.REGIONS:
            offs[(gen, r)] = off
            rows = world * stride if r != "counts" else world
            off += _align(rows * row_bytes[r])
    flags_off = off                           # [world + 1] int64: arrival values, then the timeout mark
    counter_off = off + 8 * (world + 1)       # int64 fence round counter
    err_off = counter_off + 8                 # int32 overflow flag
    off += _align(8 * (world + 1) + 16, 1024)
    return offs, flags_off, counter_off, err_off, off, row_bytes


class SlotBuffers:
    """One rank's receive buffers: per generation (two, alternated step by step) a keys, a rows
    and a grads region of `world` slots x `stride` rows each plus `world` int32 row counts; then
    the fence flags, the fence round counter and the overflow flag.  Allocated with edl_peer_alloc
    (zeroed) so that other ranks can map it with CUDA IPC."""

    REGIONS = ("keys", "rows", "grads", "grads32", "counts")

    def __init__(self, world: int, rank: int, device, stride: int, dim: int, wire_dtype=torch.float64):
        # wire_dtype: torch.float64, torch.float32, or "mixed" (fp32 rows for ids with one local gradient
        # row, float64 partial sums for the others: exact, ~60 % of the float64 bytes on Zipf 1.05)
        self.lib = _lib.load()
        self.world, self.rank, self.device = int(world), int(rank), torch.device(device)
        self.stride, self.dim, self.wire_dtype = int(stride), int(dim), wire_dtype
        self.mixed = isinstance(wire_dtype, str) and wire_dtype == "mixed"
        self.grad_dtype = torch.float64 if self.mixed else wire_dtype
        wire_b = 8 if self.grad_dtype == torch.float64 else 4
        if self.world * self.stride >= 1 << 30:
            raise ValueError("slot buffers: world x stride must stay below 2^30 rows")
        (self.off, self.flags_off, self.counter_off, self.err_off, self.nbytes,
         self.row_bytes) = slot_layout(self.world, self.stride, self.dim, wire_b, self.mixed)
        ptr = C.c_void_p()
        _lib.check(self.lib.edl_peer_alloc(C.byref(ptr), self.nbytes), "edl_peer_alloc")
        self.own_ptr = ptr.value
        self.own = torch.as_tensor(_RawDeviceMemory(self.own_ptr, self.nbytes), device=self.device)
        self.base: List[int] = []
        self._imported = []

    # ---- wiring
    def connect_ipc(self, group):
        import torch.distributed as dist
        handle = C.create_string_buffer(64)
        _lib.check(self.lib.edl_ipc_export(self.own_ptr, handle), "edl_ipc_export")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        for r, h in enumerate(handles):
            if r == self.rank:
                self.base.append(self.own_ptr)
            else:
                p = C.c_void_p()
                _lib.check(self.lib.edl_ipc_import(h, C.byref(p)), "edl_ipc_import")
                self.base.append(p.value)
                self._imported.append(p.value)
        self._tables()
        dist.barrier(group=group)

    def connect_local(self, all_buffers: Sequence["SlotBuffers"]):
        """Virtual ranks of one process on one GPU (tests): the peers' buffers are plain pointers."""
        self.base = [b.own_ptr for b in all_buffers]
        self._tables()

    def close(self):
        for p in self._imported:
            self.lib.edl_ipc_close(p)
        self._imported = []
        if self.own_ptr:
            self.lib.edl_peer_free(self.own_ptr)
        self.own_ptr = None

    # ---- views and address tables
    def _slot_addr(self, peer: int, gen: int, region: str, slot: int) -> int:
        rb = self.row_bytes[region]
        rows = self.stride if region != "counts" else 1
        return self.base[peer] + self.off[(gen, region)] + slot * rows * rb

    def _tables(self):
        me, W, dev = self.rank, self.world, self.device
        def table(fn):
            return torch.tensor([fn(d) for d in range(W)], dtype=torch.int64, device=dev)
        self.keys_dst = [table(lambda d: self._slot_addr(d, g, "keys", me)) for g in (0, 1)]
        self.count_dst = [table(lambda d: self._slot_addr(d, g, "counts", me)) for g in (0, 1)]
        self.grads_dst = [table(lambda d: self._slot_addr(d, g, "grads", me)) for g in (0, 1)]
        self.grads32_dst = [table(lambda d: self._slot_addr(d, g, "grads32", me)) if self.mixed else
                            torch.zeros(W, dtype=torch.int64, device=dev) for g in (0, 1)]
        self.flag_ptrs = torch.tensor([b + self.flags_off for b in self.base], dtype=torch.int64, device=dev)
        self.own_flags = self.own[self.flags_off:self.flags_off + 8 * (W + 1)].view(torch.int64)
        self.counter = self.own[self.counter_off:self.counter_off + 8].view(torch.int64)
        self.err = self.own[self.err_off:self.err_off + 4].view(torch.int32)

    def own_region(self, gen: int, region: str) -> torch.Tensor:
        """My receive region as [world * stride, ...] (slot-space rows)."""
        rb = self.row_bytes[region]
        rows = self.world * self.stride if region != "counts" else self.world
        o = self.off[(gen, region)]
        raw = self.own[o:o + rows * rb]
        if region == "keys":
            return raw.view(torch.int64)
        if region == "counts":
            return raw.view(torch.int32)
        return raw.view(torch.float32 if region in ("rows", "grads32") else self.grad_dtype).view(rows, self.dim)

    def remote_rows_slot(self, peer: int, gen: int, n_rows: int) -> torch.Tensor:
        """My slot of the rows region of `peer` (peer-mapped memory), as [n_rows, dim] fp32."""
        ptr = self._slot_addr(peer, gen, "rows", self.rank)
        nb = n_rows * self.row_bytes["rows"]
        t = torch.as_tensor(_RawDeviceMemory(ptr, nb), device=self.device)
        return t.view(torch.float32).view(n_rows, self.dim)

    def fence(self):
        rc = self.lib.edl_peer_signal_wait_counter(self.flag_ptrs.data_ptr(), self.own_flags.data_ptr(), self.rank,
                                                   self.world, self.counter.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "edl_peer_signal_wait_counter")


class SlotShardedStep:
    """One rank of the host-sync-free sharded step.  `slot_limit` bounds the rows one rank may send
    to one owner in a step: it sizes the slots and the owner-side launches.  A longer group is cut at the
    limit by every store path, flagged, and reported by check()."""

    def __init__(self, params: ShardParameters, batch_size: int, buffers: Optional[SlotBuffers] = None, *,
                 group=None, n_fields: Optional[int] = None, wire_dtype=torch.float64,
                 slot_slack: float = 1.5, field_range=None, emb: Optional[torch.Tensor] = None):
        # field_range = (f0, f1): this object handles only fields f0..f1-1 of a [B, F_total] batch (one
        # lane of PipelinedSlotStep): its ids are the contiguous [B, f1 - f0] slice, its embeddings go
        # to rows b * F_total + f of the shared `emb`, its gradient rows are read from there too.
        self.params = params
        self.S, self.rank = params.num_shards, params.shard
        self.F_total = params.layout.n_tables if n_fields is None else n_fields
        self.f0, f1 = (0, self.F_total) if field_range is None else field_range
        self.F = f1 - self.f0
        self.windowed = self.F != self.F_total
        self.B = int(batch_size)
        self.N = self.B * self.F
        self.dim, self.wire_dtype = params.dim, wire_dtype
        dev = params.device
        self.device = dev
        self.limit = min(self.N, int(slot_slack * self.N / self.S) + 1024)
        if buffers is None:    # a slot is `limit` rows long: every store path clamps there and flags the overflow
            buffers = SlotBuffers(self.S, self.rank, dev, self.limit, self.dim, wire_dtype)
            if group is not None:
                buffers.connect_ipc(group)
        self.buf = buffers
        if buffers.stride < self.limit:
            raise ValueError("slot buffers are shorter than the step's slot limit")
        self.mixed = buffers.mixed
        self.gen = 0
        self.field_base = torch.from_numpy(params.layout.key_base[self.f0:self.f0 + self.F].copy()).to(dev)
        self.key_bits, self.field_bits = params.layout.key_bits, params.layout.field_bits
        lib = _lib.load()
        N, R = self.N, self.S * self.limit
        i32 = dict(dtype=torch.int32, device=dev)
        self.uq = K.UniqueResult(
            uniq=torch.empty(N, dtype=torch.int64, device=dev), inverse=torch.empty(N, **i32), counts=None,
            order=torch.empty(N, **i32), seg_offsets=torch.empty(N + 1, **i32),
            n_unique=torch.zeros(1, **i32), n=N, seg_first=torch.empty(N, **i32))
        self.uq2 = K.UniqueResult(
            uniq=torch.empty(R, dtype=torch.int64, device=dev), inverse=None, counts=None,
            order=torch.empty(R, **i32), seg_offsets=torch.empty(R + 1, **i32),
            n_unique=torch.zeros(1, **i32), n=R, seg_first=torch.empty(R, **i32))
        self.keys_c = torch.empty(R, dtype=torch.int64, device=dev)
        self.map_c = torch.empty(R, **i32)
        self.n_recv = torch.zeros(1, **i32)
        self.order2 = torch.empty(R, **i32)
        self.first2 = torch.empty(R, **i32)
        self.slotpos = torch.empty(N, **i32)
        if emb is None:
            emb = torch.empty((self.B * self.F_total, self.dim), dtype=torch.float32, device=dev)
        self.emb = emb
        if self.windowed:
            j = torch.arange(N, dtype=torch.int64, device=dev)
            self.grad_map = ((j // self.F) * self.F_total + self.f0 + j % self.F).to(torch.int32)
            self.order_w = torch.empty(N, **i32)     # order / seg_first as rows of the full [B * F_total] batch
            self.first_w = torch.empty(N, **i32)
        self.ws_u1 = torch.empty(lib.edl_unique_workspace_bytes(N, self.key_bits, self.F, self.field_bits),
                                 dtype=torch.uint8, device=dev)
        self.ws_u2 = torch.empty(max(lib.edl_unique_workspace_bytes(R, self.key_bits, 0, 0),
                                     lib.edl_slot_merge_unique_workspace_bytes(self.S, self.limit)),
                                 dtype=torch.uint8, device=dev)
        # owner-side dedup: merge of the S sorted slots (default) or compact + radix sort + remaps (EDL_SLOT_OWNER_MERGE=0)
        self.owner_merge = os.environ.get("EDL_SLOT_OWNER_MERGE", "1") == "1"
        self.uq2_slot = K.UniqueResult(uniq=self.uq2.uniq, inverse=None, counts=None, order=self.order2,
                                       seg_offsets=self.uq2.seg_offsets, n_unique=self.uq2.n_unique, n=R,
                                       seg_first=self.first2)
        self.ws_part = torch.empty(lib.edl_partition_mod_workspace_bytes(N, self.S), dtype=torch.uint8, device=dev)
        n_grad = self.B * self.F_total
        self.ws_seg = torch.empty(max(lib.edl_segreduce_workspace_bytes(N, n_grad, self.dim),
                                      lib.edl_segreduce_workspace_bytes(R, R, self.dim)), dtype=torch.uint8, device=dev)
        self.S_const = torch.tensor([self.S, self.limit], dtype=torch.int64, device=dev)   # head of the dst table
        # helper streams (real ranks): the owner-side dedup of the received keys runs beside the rest
        # of the forward half, the NVLink-bound gathers into the peers beside the local HBM-bound one
        self.overlap = os.environ.get("EDL_SLOT_OVERLAP", "1") == "1"
        self.side = torch.cuda.Stream(device=dev, priority=-1) if self.overlap else None
        self.side2 = torch.cuda.Stream(device=dev) if self.overlap else None
        self.side3 = torch.cuda.Stream(device=dev) if self.overlap else None
        self._prepared = None
        self.part = K.partition_buffers(N, self.S, dev)
        # second front (dedup + partition results) for run_prefetch: the NEXT batch's phase A beside this step
        self.uq_next = None
        self.part_next = None
        self.side_pf = None
        self._ahead = None
        self.last_n_unique = self.uq.n_unique
        # A: unique, partition 4, send 1; B: S gathers; C: positions, compose, expand, 2 torch ops, reduce 4;
        # D: compact, unique, 2 remaps, update 4; 3 fences
        self.launches_per_step = (K.unique_launches(N, self.F, self.key_bits, self.field_bits) + 4 + 1 + self.S + 3 + 2 + 4
                                  + (4 if self.owner_merge else 1 + K.unique_launches(N, 0, self.key_bits, n_dev=True) + 2)
                                  + 4 + 3)

    # ------------------------------------------------------------------ phases
    def phase_a(self, ids: torch.Tensor, front=None, gen: Optional[int] = None):
        """front = (UniqueResult, PartitionResult) to fill and gen = the buffer generation to send into
        (defaults: this step's own).  run_prefetch passes the second front and the other generation."""
        b = self.buf
        g = self.gen if gen is None else gen
        uq, part = (self.uq, self.part) if front is None else front
        K.unique(ids.reshape(-1), self.field_base, self.key_bits, field_bits=self.field_bits, want_inverse=True,
                 want_counts=False, out=uq, workspace=self.ws_u1)
        K.partition_mod(uq.uniq, self.params.layout_dev, self.S, n_dev=uq.n_unique, workspace=self.ws_part, out=part)
        K.slot_send(part.keys, part.offsets, b.keys_dst[g], self.limit, b.count_dst[g], b.err, self.N,
                    flag_seg=uq.seg_offsets if self.mixed else None, flag_perm=part.perm if self.mixed else None)
        if self.windowed:
            K.remap_index(uq.order, self.grad_map, None, out=self.order_w)
            K.remap_index(uq.seg_first, self.grad_map, uq.n_unique, out=self.first_w)

    def _gather_for(self, s: int):
        b, g = self.buf, self.gen
        k = b.own_region(g, "keys")[s * b.stride:s * b.stride + self.limit]
        self.params.lookup(k, n_dev=b.own_region(g, "counts")[s:s + 1], out=b.remote_rows_slot(s, g, self.limit))

    def phase_b(self, overlap: bool = False):
        """Owner side of the lookup: gather every requester's rows straight into that requester's slot.
        The requesters are served in ring order starting after this rank (rank + 1, rank + 2, ...): with
        every owner starting at requester 0, all S GPUs wrote into the SAME GPU at any one time and shared
        its NVLink ingress (measured at S = 8: 183 GB/s per sender, 0.87 ms; a quarter of the step)."""
        ring = [(self.rank + k) % self.S for k in range(1, self.S)]
        if not overlap:
            for s in ring:
                self._gather_for(s)
            self._gather_for(self.rank)
            return
        cur = torch.cuda.current_stream()
        lanes = [self.side2, self.side3] if len(ring) > 1 else [self.side2]
        for j, st in enumerate(lanes):       # two requesters in flight: one gather's tail under the other's body
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                for s in ring[j::len(lanes)]:
                    self._gather_for(s)
        self._gather_for(self.rank)
        for st in lanes:
            cur.wait_stream(st)

    def phase_c_expand(self) -> torch.Tensor:
        b, g, uq, part = self.buf, self.gen, self.uq, self.part
        K.slot_positions(part.pos, uq.n_unique, part.offsets, b.stride, out=self.slotpos)
        idx = K.compose_index(uq.inverse, self.slotpos)
        if self.windowed:
            K.expand_window(b.own_region(g, "rows"), idx, self.F, self.F_total, self.f0, out=self.emb)
        else:
            K.expand(b.own_region(g, "rows"), idx, out=self.emb)
        return self.emb

    def phase_c_reduce(self, grad: torch.Tensor):
        b, g, uq, part = self.buf, self.gen, self.uq, self.part
        table = torch.cat([self.S_const, part.offsets.to(torch.int64), b.grads_dst[g], b.grads32_dst[g]])
        order, first = (self.order_w, self.first_w) if self.windowed else (uq.order, uq.seg_first)
        K.segreduce(grad, order, uq.seg_offsets, uq.n_unique, self.N, seg_first=first, out_pos=part.pos,
                    out=None, out_dtype=self.wire_dtype, dst_table=table, workspace=self.ws_seg)

    def phase_c(self, grad: torch.Tensor) -> torch.Tensor:
        emb = self.phase_c_expand()
        self.phase_c_reduce(grad)
        return emb

    def phase_d(self, bump_step: bool = True):
        self.phase_d_prepare()
        self.phase_d_apply(bump_step)

    rate_on_device = False     # set by GraphedSlotStep: the optimizer rate is read from params.lr_dev

    def phase_d_prepare(self):
        """Needs only the received keys (fence 1): compact, dedup, slot-space order."""
        b, g, uq2 = self.buf, self.gen, self.uq2
        if self.owner_merge:      # every slot holds a sorted run: merge them (4 launches instead of 18)
            K.slot_merge_unique(b.own_region(g, "keys"), b.own_region(g, "counts"), b.stride, self.limit,
                                self.uq2_slot, self.n_recv, workspace=self.ws_u2)
            return
        K.slot_compact(b.own_region(g, "keys"), b.own_region(g, "counts"), b.stride, self.limit, self.keys_c,
                       self.map_c, self.n_recv)
        K.unique(self.keys_c, None, self.key_bits, want_inverse=False, want_counts=False, n_dev=self.n_recv,
                 out=uq2, workspace=self.ws_u2)
        K.remap_index(uq2.order, self.map_c, self.n_recv, out=self.order2)
        K.remap_index(uq2.seg_first, self.map_c, uq2.n_unique, out=self.first2)

    def phase_d_apply(self, bump_step: bool = True):
        b, g, uq2 = self.buf, self.gen, self.uq2
        slotted = K.UniqueResult(uniq=uq2.uniq, inverse=None, counts=None, order=self.order2,
                                 seg_offsets=uq2.seg_offsets, n_unique=uq2.n_unique, n=uq2.n, seg_first=self.first2)
        self.params.apply_gradients(slotted, b.own_region(g, "grads"), uq2.n, self.ws_seg, n_rows=uq2.n,
                                    bump_step=bump_step, rate_on_device=self.rate_on_device,
                                    grad32=b.own_region(g, "grads32") if self.mixed else None)
        self.gen ^= 1

    # ------------------------------------------------------------------ drivers
    def lookup(self, ids: torch.Tensor) -> torch.Tensor:
        """Forward half on a real rank (one process per GPU): emb [B*F, dim].  No host synchronisation."""
        self.phase_a(ids)
        self.buf.fence()
        if self.overlap:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                self.phase_d_prepare()
                self._prepared = torch.cuda.Event()
                self._prepared.record()
        self.phase_b(self.overlap)
        self.buf.fence()
        return self.phase_c_expand()

    def update(self, ids: Optional[torch.Tensor], grad: torch.Tensor, uq=None):
        """Backward half for the ids of the preceding lookup()."""
        self.phase_c_reduce(grad)
        self.buf.fence()
        if self._prepared is not None:
            torch.cuda.current_stream().wait_event(self._prepared)
            self._prepared = None
            self.phase_d_apply()
        else:
            self.phase_d()

    def run(self, ids: torch.Tensor, grad: torch.Tensor) -> torch.Tensor:
        emb = self.lookup(ids)
        self.update(ids, grad)
        return emb

    def run_prefetch(self, ids: torch.Tensor, grad: torch.Tensor, next_ids: torch.Tensor) -> torch.Tensor:
        """One step with phase A of the NEXT batch (dedup, partition, key send into the other buffer
        generation) on a helper stream beside this step's rounds -- the ids of batch k + 1 are known while
        batch k trains.  Safe from fence 1 on: every peer has then finished step k - 1, the last reader of
        that generation's key slots.  Same work per step, same results.  Every rank must use it in lock step
        with the same batch order; not for the lanes of PipelinedSlotStep."""
        if self.windowed:
            raise ValueError("run_prefetch is for the unpipelined slot step")
        if self.uq_next is None:
            N, dev = self.N, self.device
            i32 = dict(dtype=torch.int32, device=dev)
            self.uq_next = K.UniqueResult(
                uniq=torch.empty(N, dtype=torch.int64, device=dev), inverse=torch.empty(N, **i32), counts=None,
                order=torch.empty(N, **i32), seg_offsets=torch.empty(N + 1, **i32),
                n_unique=torch.zeros(1, **i32), n=N, seg_first=torch.empty(N, **i32))
            self.part_next = K.partition_buffers(N, self.S, dev)
            self.side_pf = torch.cuda.Stream(device=dev, priority=-1)
        cur = torch.cuda.current_stream()
        if self._ahead != ids.data_ptr():
            self.phase_a(ids)                               # first step: this batch's own phase A
        self.buf.fence()
        if self.overlap:
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                self.phase_d_prepare()
                self._prepared = torch.cuda.Event()
                self._prepared.record()
        self.side_pf.wait_stream(cur)
        with torch.cuda.stream(self.side_pf):
            self.phase_a(next_ids, front=(self.uq_next, self.part_next), gen=self.gen ^ 1)
        self.phase_b(self.overlap)
        self.buf.fence()
        emb = self.phase_c_expand()
        self.update(ids, grad)
        cur.wait_stream(self.side_pf)                       # the next step's keys are sent before its fence 1
        self.uq, self.uq_next = self.uq_next, self.uq
        self.part, self.part_next = self.part_next, self.part
        self.last_n_unique = self.uq.n_unique
        self._ahead = next_ids.data_ptr()
        return emb

    PHASES = ("A_dedup_partition_send", "fence1", "B_owner_gather", "fence2", "C_expand", "C_reduce_send", "fence3",
              "D_owner_dedup", "D_owner_update")

    def run_timed(self, ids: torch.Tensor, grad: torch.Tensor):
        """One step with the phases in a row on the current stream (no helper streams) and a CUDA event
        after each: -> [(phase name, event before, event after)].  For bench.py's per-phase times; every
        rank must call it in lock step (the fences pair up)."""
        marks = [torch.cuda.Event(enable_timing=True)]
        marks[0].record()

        def mark():
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append(e)
        self.phase_a(ids); mark()
        self.buf.fence(); mark()
        self.phase_b(False); mark()
        self.buf.fence(); mark()
        self.phase_c_expand(); mark()
        self.phase_c_reduce(grad); mark()
        self.buf.fence(); mark()
        self.phase_d_prepare(); mark()
        self.phase_d_apply(); mark()
        return [(n, marks[i], marks[i + 1]) for i, n in enumerate(self.PHASES)]

    def check(self):
        """Synchronise and raise if a slot overflowed or a fence timed out since the last check."""
        torch.cuda.synchronize(self.device)
        err = int(self.buf.err.item())
        late = int(self.buf.own_flags[self.S].item())
        if err:
            raise RuntimeError(f"SlotShardedStep: more than slot_limit={self.limit} ids for one owner in a step; "
                               "raise slot_slack")
        if late:
            raise RuntimeError("SlotShardedStep: a peer did not reach the exchange fence (timeout)")
        if hasattr(self.params, "materialised"):
            self.params.materialised()      # lazily materialised shard: raises if it ran out of rows


def run_virtual(steps: Sequence[SlotShardedStep], ids: Sequence[torch.Tensor], grads: Sequence[torch.Tensor]):
    """All ranks of a step as virtual ranks on one GPU and one stream (tests): phase by phase over
    the ranks, which orders the rounds the way the fences do between real GPUs."""
    for st, i in zip(steps, ids):
        st.phase_a(i)
    for st in steps:
        st.phase_b()
    embs = [st.phase_c(g) for st, g in zip(steps, grads)]
    for st in steps:
        st.phase_d()
    return embs


class GraphedSlotStep:
    """SlotShardedStep (or PipelinedSlotStep) replayed from CUDA graphs for one fixed (ids, grad) buffer
    pair: one graph per buffer generation, replayed alternately.  The optimizer rate is read from the
    device scalar params.lr_dev, refreshed before every replay (Adam's rate changes every step).
    Every rank must construct and replay in lock step."""

    def __init__(self, step: SlotShardedStep, ids: torch.Tensor, grad: torch.Tensor):
        self.step, self.ids, self.grad = step, ids, grad
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):        # both generations once, outside capture
            step.run(ids, grad)
            step.run(ids, grad)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        host_step = step.params.step
        self.graphs = [None, None]
        lanes = getattr(step, "lanes", [step])
        for l in lanes:
            l.rate_on_device = step.params.rate_changes
        try:
            for _ in (0, 1):                 # in the order the generations come up next
                gen = step.gen
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    step.run(ids, grad)
                self.graphs[gen] = g
        finally:
            for l in lanes:
                l.rate_on_device = False
        step.params.step = host_step

    def run(self) -> torch.Tensor:
        st = self.step                        # the generation lives on the step: several graphed
        if st.params.rate_changes:
            st.params.stage_rate(st.params.step + 1)
        self.graphs[st.gen].replay()          # (ids, grad) pairs may share one step
        st.gen ^= 1
        st.params.step += 1
        return st.emb


class GraphedSlotRing:
    """SlotShardedStep.run_prefetch over a fixed ring of id buffers (even length) and one gradient buffer,
    one CUDA graph per ring position.  Two eager steps (ring positions 0 and 1) come first -- they are real
    steps -- and the graphs are captured from position 2 on, in the order they will be replayed, so buffer
    generations, fronts and the keys already sent for the next step line up.  Every rank constructs and
    replays in lock step."""

    def __init__(self, step: SlotShardedStep, ids_ring, grad: torch.Tensor):
        n = len(ids_ring)
        if n % 2 or n < 2:
            raise ValueError("the ring of id batches must have an even length")
        self.step, self.ids_ring, self.grad = step, list(ids_ring), grad
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for i in (0, 1):
                step.run_prefetch(self.ids_ring[i], grad, self.ids_ring[(i + 1) % n])
        cur.wait_stream(side)
        torch.cuda.synchronize()
        host_step = step.params.step
        step.rate_on_device = step.params.rate_changes
        self.graphs = [None] * n
        try:
            for k in range(n):
                i = (2 + k) % n
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    step.run_prefetch(self.ids_ring[i], grad, self.ids_ring[(i + 1) % n])
                self.graphs[i] = g
        finally:
            step.rate_on_device = False
        step.params.step = host_step
        self.pos = 2 % n

    def run(self) -> torch.Tensor:
        st = self.step
        if st.params.rate_changes:
            st.params.stage_rate(st.params.step + 1)
        self.graphs[self.pos].replay()
        self.pos = (self.pos + 1) % len(self.graphs)
        st.gen ^= 1
        st.params.step += 1
        return st.emb


class PipelinedSlotStep:
    """SlotShardedStep cut into `n_lanes` groups of tables that move through the exchange rounds one
    stage apart, so that a lane's NVLink-bound round (gather into the peers, reduce into the owners)
    runs beside another lane's HBM-bound work (dedup, expand, optimizer).  Tables never share rows, so
    the lanes are independent and the result is bit-identical to the unpipelined step.

        forward   stage t: lane t dedup + send keys | lane t-1 owner gather | lane t-2 expand     | fence
        backward  stage t: lane t reduce into the owners | lane t-1 owner reduce + optimizer       | fence

    One fence per stage closes every round of that stage; the fences are issued on the calling stream
    in a fixed order on every rank.  Parity-checked on virtual and on real ranks; measured SLOWER than the
    unpipelined step on 2 real GPUs, so it stays an option."""

    def __init__(self, params: ShardParameters, batch_size: int, *, group=None, buffers=None, n_lanes: int = 2,
                 wire_dtype=torch.float64, slot_slack: float = 1.5):
        self.params = params
        self.S, self.rank = params.num_shards, params.shard
        self.F = params.layout.n_tables
        self.B = int(batch_size)
        self.N = self.B * self.F
        self.dim = params.dim
        dev = params.device
        self.device = dev
        if not 1 <= n_lanes <= self.F:
            raise ValueError("n_lanes must be between 1 and the number of tables")
        cuts = [self.F * g // n_lanes for g in range(n_lanes + 1)]
        self.emb = torch.empty((self.N, self.dim), dtype=torch.float32, device=dev)
        self.lanes: List[SlotShardedStep] = []
        for g in range(n_lanes):
            self.lanes.append(SlotShardedStep(params, batch_size, None if buffers is None else buffers[g], group=group,
                                              wire_dtype=wire_dtype, slot_slack=slot_slack,
                                              field_range=(cuts[g], cuts[g + 1]), emb=self.emb))
        self.cuts = cuts
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
        self.prep = torch.cuda.Stream(device=dev, priority=-1)
        self._prepared = [None] * n_lanes
        self.last_n_unique = self.lanes[0].uq.n_unique
        self.launches_per_step = sum(l.launches_per_step for l in self.lanes) + n_lanes

    # the buffer generation is per lane but always equal across lanes (GraphedSlotStep reads / sets it)
    @property
    def gen(self) -> int:
        return self.lanes[0].gen

    @gen.setter
    def gen(self, v: int):
        for l in self.lanes:
            l.gen = v

    def _lane_ids(self, ids: torch.Tensor):
        ids = ids.view(self.B, self.F)
        return [ids[:, a:b].contiguous() for a, b in zip(self.cuts[:-1], self.cuts[1:])]

    # ---- stage contents (shared by the real and the virtual driver)
    def _forward_stage(self, t: int, lane_ids):
        """[(callable, kind)] of forward stage t; kind "link" = NVLink-bound, "hbm" otherwise."""
        G, ops = len(self.lanes), []
        if 0 <= t - 1 < G:
            ops.append((lambda l=self.lanes[t - 1]: l.phase_b(l.overlap and self._real), "link"))
        if 0 <= t < G:
            ops.append((lambda l=self.lanes[t], i=lane_ids[t]: l.phase_a(i), "hbm"))
        if 0 <= t - 2 < G:
            ops.append((lambda l=self.lanes[t - 2]: l.phase_c_expand(), "hbm"))
        return ops

    def _backward_stage(self, t: int, grad):
        G, ops = len(self.lanes), []
        if 0 <= t < G:
            ops.append((lambda l=self.lanes[t]: l.phase_c_reduce(grad), "link"))
        if 0 <= t - 1 < G:
            ops.append((lambda k=t - 1: self._apply(k), "hbm"))
        return ops

    def _apply(self, k: int):
        lane = self.lanes[k]
        if self._prepared[k] is not None:
            torch.cuda.current_stream().wait_event(self._prepared[k])
            self._prepared[k] = None
            lane.phase_d_apply(bump_step=(k == 0))
        else:
            lane.phase_d(bump_step=(k == 0))

    def _run_stage(self, ops):
        """First op on the calling stream, the others on helper streams, all joined before returning."""
        cur = torch.cuda.current_stream()
        for j, (fn, _) in enumerate(ops[1:]):
            st = self.streams[j % len(self.streams)]
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                fn()
        if ops:
            ops[0][0]()
        for j in range(min(len(ops) - 1, len(self.streams))):
            cur.wait_stream(self.streams[j])

    # ---- real ranks
    _real = True

    def lookup(self, ids: torch.Tensor) -> torch.Tensor:
        lane_ids = self._lane_ids(ids)
        G = len(self.lanes)
        for t in range(G + 2):
            if 1 <= t <= G:       # lane t-1's keys arrived with the previous fence: start its owner-side dedup
                self.prep.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self.prep):
                    self.lanes[t - 1].phase_d_prepare()
                    ev = torch.cuda.Event()
                    ev.record()
                self._prepared[t - 1] = ev
            self._run_stage(self._forward_stage(t, lane_ids))
            if t < G + 1:
                self.lanes[0].buf.fence()
        return self.emb

    def update(self, ids: Optional[torch.Tensor], grad: torch.Tensor, uq=None):
        G = len(self.lanes)
        for t in range(G + 1):
            self._run_stage(self._backward_stage(t, grad))
            if t < G:
                self.lanes[0].buf.fence()

    def run(self, ids: torch.Tensor, grad: torch.Tensor) -> torch.Tensor:
        emb = self.lookup(ids)
        self.update(ids, grad)
        return emb

    def check(self):
        for l in self.lanes:
            l.check()


def run_virtual_pipelined(steps: Sequence[PipelinedSlotStep], ids: Sequence[torch.Tensor],
                          grads: Sequence[torch.Tensor]):
    """PipelinedSlotStep for all ranks as virtual ranks on one GPU and one stream (tests): stage by
    stage over the ranks, no fences, no helper streams."""
    G = len(steps[0].lanes)
    lane_ids = [st._lane_ids(i) for st, i in zip(steps, ids)]
    for st in steps:
        st._real = False
    for t in range(G + 2):
        for st, li in zip(steps, lane_ids):
            for fn, _ in st._forward_stage(t, li):
                fn()
    for t in range(G + 1):
        for st, g in zip(steps, grads):
            for fn, _ in st._backward_stage(t, g):
                fn()
    return [st.emb for st in steps]
