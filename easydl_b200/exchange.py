"""A7: the worker <-> PS exchange.  In the lineage this is gRPC between Pods; here every GPU is
both a worker and a PS shard and the three per-step exchanges (ids out, rows back, gradients
out) are all-to-all-v over NCCL / NVLink.

Steps are written as generators that `yield` exchange requests, so the same step code runs
  - on S real ranks (TorchExchange: torch.distributed all_to_all_single, NCCL or gloo), and
  - on S virtual ranks inside one process on one GPU (run_lockstep), which is how the
    multi-shard path is parity-tested on a single-GPU box and on CPU tensors.

Requests:
  ("counts", t)                 t: int64 [W, k] on the step's device; row r goes to rank r.
                                -> int64 [W, k] HOST tensor, row r = what rank r sent us.
  ("a2av", t, send, recv)       t: [sum(send), ...]; send/recv: per-rank row counts (lists).
                                -> [sum(recv), ...]
"""
from __future__ import annotations

from typing import Generator, List, Sequence

import torch


def _drive(gen: Generator, handle):
    try:
        req = next(gen)
        while True:
            req = gen.send(handle(req))
    except StopIteration as stop:
        return stop.value


class TorchExchange:
    """One real rank of a torch.distributed process group (backend nccl on GPUs, gloo on CPU)."""

    def __init__(self, group=None, device=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = device

    def handle(self, req):
        kind = req[0]
        if kind == "counts":
            t = req[1].contiguous()
            out = torch.empty_like(t)
            self.dist.all_to_all_single(out, t, group=self.group)
            return out.cpu()  # the one host synchronisation of an exchange round
        if kind == "a2av":
            _, t, send, recv = req
            t = t.contiguous()
            out = torch.empty((int(sum(recv)),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            self.dist.all_to_all_single(out, t, output_split_sizes=[int(x) for x in recv],
                                        input_split_sizes=[int(x) for x in send], group=self.group)
            return out
        raise ValueError(f"unknown exchange request {kind!r}")

    def run(self, gen: Generator):
        return _drive(gen, self.handle)


class SingleExchange:
    """World of one (S == 1): every exchange is the identity."""
    world, rank = 1, 0

    def handle(self, req):
        if req[0] == "counts":
            return req[1].cpu()
        return req[1]

    def run(self, gen: Generator):
        return _drive(gen, self.handle)


def run_lockstep(gens: Sequence[Generator]) -> List:
    """Advance W generators (virtual ranks 0..W-1) together, serving their exchange requests
    from one another's send buffers.  All ranks must issue the same sequence of request kinds."""
    W = len(gens)
    results = [None] * W
    reqs = [None] * W
    live = [True] * W
    for r, g in enumerate(gens):
        try:
            reqs[r] = next(g)
        except StopIteration as stop:
            results[r], live[r] = stop.value, False
    while any(live):
        if not all(live):
            raise RuntimeError("virtual ranks fell out of step (some finished, some still exchanging)")
        kinds = {q[0] for q in reqs}
        if len(kinds) != 1:
            raise RuntimeError(f"virtual ranks issued different requests: {kinds}")
        kind = kinds.pop()
        if kind == "counts":
            mats = [q[1] for q in reqs]
            replies = [torch.stack([mats[src][dst] for src in range(W)]).cpu() for dst in range(W)]
        elif kind == "a2av":
            offs = []
            for q in reqs:
                o = [0]
                for c in q[2]:
                    o.append(o[-1] + int(c))
                offs.append(o)
            replies = []
            for dst in range(W):
                parts = [reqs[src][1][offs[src][dst]:offs[src][dst + 1]] for src in range(W)]
                for src in range(W):
                    assert parts[src].shape[0] == int(reqs[dst][3][src]), "send/recv split mismatch"
                replies.append(torch.cat(parts) if parts else reqs[dst][1][:0])
        else:
            raise ValueError(kind)
        for r, g in enumerate(gens):
            try:
                reqs[r] = g.send(replies[r])
            except StopIteration as stop:
                results[r], live[r] = stop.value, False
    return results
