"""The sparse step with tables row-sharded over S GPUs: every rank is a worker
for its slice of the batch and the PS shard that owns ids with id mod S == rank.

forward : A1 dedup(local ids) -> A2 partition by owner -> [ids all-to-all-v] -> A3 gather on the
          owner -> [rows all-to-all-v] -> expand to emb[N_local, D]
backward: A4 reduce grad by local unique id, written in partition order -> [gradient
          all-to-all-v] -> owner: A1 dedup(received keys; stable, so same-id gradients stay in
          peer-rank order) -> A4 reduce -> A5 optimizer in place.

One host synchronisation per step: the split sizes of the all-to-all-v (counts round).
"""
from __future__ import annotations

from typing import Generator, Optional

import torch

from . import kernels as K
from .parameters import ShardParameters


class ShardedSparseStep:
    def __init__(self, params: ShardParameters, batch_size: int, exchange=None,
                 n_fields: Optional[int] = None, wire_dtype=torch.float64):
        # wire_dtype: element type of the per-worker partial gradient sums on the wire.
        # float64 (default) keeps the owner's total equal to one float64 pass over the global
        # batch (SPEC A4: independent of S and of the worker split); float32 halves the
        # gradient exchange bytes at the cost of one extra fp32 rounding per worker.
        self.wire_dtype = wire_dtype
        self.params = params
        self.S = params.num_shards
        self.rank = params.shard
        self.exchange = exchange
        self.F = params.layout.n_tables if n_fields is None else n_fields
        self.B = int(batch_size)
        self.N = self.B * self.F
        dev = params.device
        self.field_base = torch.from_numpy(params.layout.key_base[:self.F].copy()).to(dev)
        self.key_bits = params.layout.key_bits
        self.layout_dev = params.layout_dev          # T and key_base are what partition reads
        self.last_n_unique = torch.zeros(1, dtype=torch.int32, device=dev)
        passes = (self.key_bits + 7) // 8
        # unique x2, partition 4, gather 1, compose 1, expand 1, segreduce 3, update 3
        self.launches_per_step = 2 * (3 * passes + 4) + 4 + 1 + 1 + 1 + 3 + 3
        self._fw = None

    # ------------------------------------------------------------------ generators
    def gen_lookup(self, ids: torch.Tensor) -> Generator:
        S = self.S
        flat = ids.reshape(-1)
        uq = K.unique(flat, self.field_base, self.key_bits, want_inverse=True, want_counts=False)
        part = K.partition_mod(uq.uniq, self.layout_dev, S, n_dev=uq.n_unique)
        counts = (part.offsets[1:] - part.offsets[:-1]).to(torch.int64).reshape(S, 1)
        got = yield ("counts", counts)                       # host [S, 1]: what each peer sends us
        send = [int(x) for x in counts.cpu().reshape(-1)] if S == 1 else None
        if send is None:
            send = [int(x) for x in part.offsets.cpu().diff()]
        recv = [int(x) for x in got.reshape(-1)]
        U = sum(send)
        recv_keys = yield ("a2av", part.keys[:U], send, recv)
        rows_out = self.params.lookup(recv_keys)
        rows_in = yield ("a2av", rows_out, recv, send)       # [U, D], partition order
        index = K.compose_index(uq.inverse, part.pos)
        emb = K.expand(rows_in, index)
        self._fw = (uq, part, send, recv, recv_keys, U)
        self.last_n_unique = uq.n_unique
        return emb

    def gen_update(self, grad: torch.Tensor) -> Generator:
        uq, part, send, recv, recv_keys, U = self._fw
        g_send = torch.empty((max(U, 1), self.params.dim), dtype=self.wire_dtype, device=grad.device)
        K.segreduce(grad, uq.order, uq.seg_offsets, uq.n_unique, uq.n, seg_first=uq.seg_first,
                    out_pos=part.pos, out=g_send)
        g_recv = yield ("a2av", g_send[:U], send, recv)      # aligned with recv_keys
        R = recv_keys.numel()
        if R == 0:
            self.params.step += 1
            return None
        uq2 = K.unique(recv_keys, None, self.key_bits, want_inverse=False, want_counts=False)
        self.params.apply_gradients(uq2, g_recv, R)
        return None

    def gen_step(self, ids: torch.Tensor, grad: torch.Tensor) -> Generator:
        emb = yield from self.gen_lookup(ids)
        yield from self.gen_update(grad)
        return emb

    # ------------------------------------------------------------------ real-rank drivers
    def lookup(self, ids):
        return self.exchange.run(self.gen_lookup(ids))

    def update(self, ids, grad, uq=None):
        return self.exchange.run(self.gen_update(grad))

    def run(self, ids, grad):
        return self.exchange.run(self.gen_step(ids, grad))
