"""The fused sparse step of one worker + PS shard pair on one GPU (S == 1 path).

lookup : emb[j] = table[ids[j]]                              (A3, direct gather, no dedup)
update : dedup(ids) -> segmented reduce of grad -> optimizer (A1 + A4 + A5)

All buffers are allocated once; a step enqueues kernels on the current stream and never
synchronises with the host.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _lib
from . import kernels as K
from .parameters import ShardParameters


class SparseStep:
    def __init__(self, params: ShardParameters, batch_size: int, n_fields: Optional[int] = None,
                 overlap_dedup: Optional[bool] = None):
        if params.num_shards != 1:
            raise ValueError("SparseStep is the single-shard path; use ShardedSparseStep for S > 1")
        self.params = params
        self.F = params.layout.n_tables if n_fields is None else n_fields
        self.B = int(batch_size)
        self.N = self.B * self.F
        dev = params.device
        lib = _lib.load()
        self.field_base = torch.from_numpy(params.layout.key_base[:self.F].copy()).to(dev)
        self.key_bits = params.layout.key_bits
        self.field_bits = params.layout.field_bits
        N, D = self.N, params.dim
        self.emb = torch.empty((N, D), dtype=torch.float32, device=dev)
        self.uq = K.UniqueResult(
            uniq=torch.empty(N, dtype=torch.int64, device=dev), inverse=None, counts=None,
            order=torch.empty(N, dtype=torch.int32, device=dev),
            seg_offsets=torch.empty(N + 1, dtype=torch.int32, device=dev),
            n_unique=torch.zeros(1, dtype=torch.int32, device=dev), n=N,
            seg_first=torch.empty(N, dtype=torch.int32, device=dev))
        self.ws_unique = torch.empty(lib.edl_unique_workspace_bytes(N, self.key_bits, self.F, self.field_bits),
                                     dtype=torch.uint8, device=dev)
        # second dedup result + workspace: run_prefetch() dedups the NEXT batch while this batch updates
        self.uq_next = None
        self.ws_unique_next = None
        self._ahead = None          # data_ptr of the id batch whose dedup sits in self.uq
        self.ws_seg = torch.empty(lib.edl_segreduce_workspace_bytes(N, N, D), dtype=torch.uint8, device=dev)
        # launches per step (for bench.py's gpu_launches): gather 1; unique (K.unique_launches); update 4
        # (k_seg_classify, main kernel, k_seg_chunk, k_seg_long_final) + memset nodes
        self.launches_per_step = 1 + K.unique_launches(N, self.F, self.key_bits, self.field_bits) + 4
        self.last_n_unique = self.uq.n_unique
        # The lookup does not need the dedup (it gathers by id), so the two can run side by side:
        # the dedup works in shared memory (or in L2 for the radix passes), the gather is one long HBM-bound kernel.
        if overlap_dedup is None:
            overlap_dedup = os.environ.get("EDL_STEP_OVERLAP", "1") == "1"
        self.side = torch.cuda.Stream(device=dev, priority=-1) if overlap_dedup else None

    def lookup(self, ids: torch.Tensor) -> torch.Tensor:
        """ids int64 [B, F] on the device -> emb [B*F, dim]."""
        return K.gather_ids(self.params.slab, ids.view(-1), self.field_base, out=self.emb)

    def dedup(self, ids: torch.Tensor, out: Optional[K.UniqueResult] = None, workspace=None) -> K.UniqueResult:
        return K.unique(ids.view(-1), self.field_base, self.key_bits, field_bits=self.field_bits,
                        want_inverse=False, want_counts=False, out=self.uq if out is None else out,
                        workspace=self.ws_unique if workspace is None else workspace)

    def _second_buffers(self):
        if self.uq_next is None:
            N, dev = self.N, self.params.device
            self.uq_next = K.UniqueResult(
                uniq=torch.empty(N, dtype=torch.int64, device=dev), inverse=None, counts=None,
                order=torch.empty(N, dtype=torch.int32, device=dev),
                seg_offsets=torch.empty(N + 1, dtype=torch.int32, device=dev),
                n_unique=torch.zeros(1, dtype=torch.int32, device=dev), n=N,
                seg_first=torch.empty(N, dtype=torch.int32, device=dev))
            self.ws_unique_next = torch.empty_like(self.ws_unique)
            if self.side is None:
                self.side = torch.cuda.Stream(device=dev, priority=-1)

    def run_prefetch(self, ids: torch.Tensor, grad: torch.Tensor, next_ids: torch.Tensor,
                     rate_on_device: bool = False) -> torch.Tensor:
        """One step with the dedup of the NEXT batch's ids (known ahead, as an input pipeline knows them)
        running on the helper stream beside this batch's lookup and update: the dedup is a chain of short,
        latency-bound, L2-resident kernels, the update a long HBM-latency-bound one, and neither fills the
        machine.  Same work per step, same results; the first call dedups its own batch first."""
        self._second_buffers()
        cur = torch.cuda.current_stream()
        if self._ahead != ids.data_ptr():
            self.dedup(ids)
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side):
            self.dedup(next_ids, out=self.uq_next, workspace=self.ws_unique_next)
        emb = self.lookup(ids)
        self.params.apply_gradients(self.uq, grad, self.N, self.ws_seg, rate_on_device=rate_on_device)
        cur.wait_stream(self.side)
        self.uq, self.uq_next = self.uq_next, self.uq
        self.ws_unique, self.ws_unique_next = self.ws_unique_next, self.ws_unique
        self.last_n_unique = self.uq.n_unique
        self._ahead = next_ids.data_ptr()
        return emb

    def update(self, ids: torch.Tensor, grad: torch.Tensor, uq: Optional[K.UniqueResult] = None,
               rate_on_device: bool = False):
        """grad f32 [B*F, dim]: upstream gradient of emb."""
        if uq is None:
            uq = self.dedup(ids)
        self.params.apply_gradients(uq, grad, self.N, self.ws_seg, rate_on_device=rate_on_device)

    def run(self, ids: torch.Tensor, grad: torch.Tensor, rate_on_device: bool = False) -> torch.Tensor:
        if self.side is None:
            emb = self.lookup(ids)
            self.update(ids, grad, rate_on_device=rate_on_device)
            return emb
        cur = torch.cuda.current_stream()
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side):
            uq = self.dedup(ids)
        emb = self.lookup(ids)
        cur.wait_stream(self.side)
        self.update(ids, grad, uq, rate_on_device=rate_on_device)
        return emb


class GraphedStep:
    """A SparseStep captured into CUDA graphs for one fixed (ids, grad) buffer pair: the
    launches of a step cost one graph launch instead of one trip through the driver each.  Three
    graphs (lookup / dedup / update) so that a caller can time the phases with events between
    them, plus one graph of the whole step.  The optimizer rate is read from a device scalar
    (ShardParameters.lr_dev) that run() refreshes before every replay, so Adam's bias-corrected
    rate, which changes every step, replays correctly."""

    def __init__(self, step: SparseStep, ids: torch.Tensor, grad: torch.Tensor):
        self.step, self.ids, self.grad = step, ids, grad
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):   # warm-up outside capture (sets function attributes)
            step.run(ids, grad)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.g_lookup, self.g_dedup, self.g_update, self.g_all = (torch.cuda.CUDAGraph() for _ in range(4))
        host_step = step.params.step   # capture records launches without running them
        with torch.cuda.graph(self.g_lookup):
            step.lookup(ids)
        with torch.cuda.graph(self.g_dedup):
            step.dedup(ids)
        dev_rate = step.params.rate_changes
        with torch.cuda.graph(self.g_update):
            step.update(ids, grad, step.uq, rate_on_device=dev_rate)
        with torch.cuda.graph(self.g_all):
            step.run(ids, grad, rate_on_device=dev_rate)
        step.params.step = host_step

    def run(self):
        p = self.step.params
        if p.rate_changes:
            p.stage_rate(p.step + 1)
        self.g_all.replay()
        p.step += 1
        return self.step.emb


class GraphedPrefetchRing:
    """SparseStep.run_prefetch over a fixed ring of (ids) buffers and one gradient buffer, one CUDA graph per
    ring position: position i looks up and updates batch i while batch (i + 1) mod n is deduplicated beside
    it.  The ring length must be even (the two dedup result buffers alternate).  run() advances the ring."""

    def __init__(self, step: SparseStep, ids_ring, grad: torch.Tensor):
        if len(ids_ring) % 2:
            raise ValueError("the ring of id batches must have an even length")
        self.step, self.ids_ring, self.grad, self.pos = step, list(ids_ring), grad, 0
        n = len(self.ids_ring)
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):        # warm-up outside capture, twice round the buffer pair
            for i in range(2):
                step.run_prefetch(self.ids_ring[i], grad, self.ids_ring[(i + 1) % n])
        cur.wait_stream(side)
        torch.cuda.synchronize()
        # after two warm-up steps the dedup of batch 2 % n sits in step.uq; restart the ring at 0
        step.dedup(self.ids_ring[0])
        step._ahead = self.ids_ring[0].data_ptr()
        torch.cuda.synchronize()
        host_step = step.params.step
        self.graphs = []
        for i in range(n):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                step.run_prefetch(self.ids_ring[i], grad, self.ids_ring[(i + 1) % n],
                                  rate_on_device=step.params.rate_changes)
            self.graphs.append(g)
        step.params.step = host_step
        # capture ran the host-side buffer swaps n times (n even): the pair is back where it started, and
        # the dedup of batch 0 (done eagerly above) is in step.uq

    def run(self):
        p = self.step.params
        if p.rate_changes:
            p.stage_rate(p.step + 1)
        self.graphs[self.pos].replay()
        self.pos = (self.pos + 1) % len(self.graphs)
        p.step += 1
        return self.step.emb
