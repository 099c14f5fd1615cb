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
  ("a2av", t, send, recv[, direction, tag])
                                t: [sum(send), ...]; send/recv: per-rank row counts (lists).
                                -> [sum(recv), ...]
                                direction "fwd" = the pattern of the last counts round, "rev" = its
                                transpose (answers travelling back); tag names the receive buffer.
                                Both are hints for PeerExchange; the other drivers ignore them.

PeerExchange replaces the NCCL all-to-all-v by direct writes into the peers' receive buffers over
NVLink (CUDA IPC mapped memory): measured on 2 B200s the three NCCL all-to-all-v rounds of a step
took 1.17 ms of a 2.28 ms step at ~300 GB/s; a peer copy of the same bytes runs at link speed.
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
            t, send, recv = req[1], req[2], req[3]
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


class _RawDeviceMemory:
    """Lets torch view a raw device pointer (zero copy) through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2}


class PeerExchange:
    """All-to-all-v by peer writes.  Every rank owns one receive buffer (two generations, used
    alternately step by step, so a fast rank that has already started the next step never
    overwrites what a slow rank is still reading); the buffers are mapped into every other rank
    with CUDA IPC at set-up.  A round = each rank copies its S slices straight into the S
    receivers' buffers at the offsets implied by the gathered count matrix, then one tiny NCCL
    all-reduce orders "all writes issued" before "any read" on every rank's stream."""

    REGIONS = ("keys", "rows", "grads")

    def __init__(self, group, device, capacity_rows: int, dim: int, wire_dtype=torch.float64):
        import ctypes as C

        import torch.distributed as dist

        from . import _lib
        self.lib = _lib.load()
        self.C, self.dist, self.group = C, dist, group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = torch.device(device)
        self.cap = int(capacity_rows)
        wire_b = 8 if wire_dtype == torch.float64 else 4
        self.row_bytes = {"keys": 8, "rows": dim * 4, "grads": dim * wire_b}
        self.dtypes = {"keys": torch.int64, "rows": torch.float32, "grads": wire_dtype}
        self.dim = dim
        self.region_off, off = {}, 0
        for gen in (0, 1):
            for r in self.REGIONS:
                self.region_off[(gen, r)] = off
                off += (self.cap * self.row_bytes[r] + 255) // 256 * 256
        self.flags_off = off                       # [world + 1] int64 arrival counters, then padding
        off += 1024
        self.nbytes = off
        ptr = C.c_void_p()
        _lib.check(self.lib.edl_peer_alloc(C.byref(ptr), self.nbytes), "edl_peer_alloc")
        self.own_ptr = ptr.value
        handle = C.create_string_buffer(64)
        _lib.check(self.lib.edl_ipc_export(self.own_ptr, handle), "edl_ipc_export")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        self.base = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.base.append(self.own_ptr)
            else:
                p = C.c_void_p()
                _lib.check(self.lib.edl_ipc_import(h, C.byref(p)), "edl_ipc_import")
                self.base.append(p.value)
        self.own = torch.as_tensor(_RawDeviceMemory(self.own_ptr, self.nbytes), device=self.device)
        self.flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.own_flags = self.own[self.flags_off:self.flags_off + 8 * (self.world + 1)].view(torch.int64)
        self.peer_flag_ptrs = torch.tensor([b + self.flags_off for b in self.base], dtype=torch.int64,
                                           device=self.device)
        self.round = 0
        self.gen = 1
        self.M = None          # [W, W] rows sent s -> d in the current step
        dist.barrier(group=group)

    def close(self):
        for r, p in enumerate(self.base):
            if r != self.rank and p:
                self.lib.edl_ipc_close(p)
        self.lib.edl_peer_free(self.own_ptr)
        self.base, self.own_ptr = [], None

    def _fence(self):
        """Order "every rank has issued its writes of this round" before "any rank reads"."""
        self.round += 1
        rc = self.lib.edl_peer_signal_wait(self.peer_flag_ptrs.data_ptr(), self.own_flags.data_ptr(),
                                           self.rank, self.world, self.round,
                                           torch.cuda.current_stream().cuda_stream)
        if rc:
            from . import _lib
            _lib.check(rc, "edl_peer_signal_wait")

    def handle(self, req):
        kind = req[0]
        if kind == "counts":
            t = req[1].contiguous()
            allc = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
            self.dist.all_gather_into_tensor(allc, t, group=self.group)
            allc = allc.cpu()                                   # [src, dst, k]: the step's one host sync
            self.M = allc[:, :, 0].clone()
            # Every rank holds the whole matrix here, so every rank sees the same overflow and raises
            # before ANY copy of the step is issued: no rank writes past another rank's region while
            # that rank has already given up.  (Both directions: rows received forward = column sums,
            # rows received on the way back = row sums.)
            worst = int(max(self.M.sum(0).max(), self.M.sum(1).max())) if self.M.numel() else 0
            if worst > self.cap:
                raise RuntimeError(f"PeerExchange: a rank would receive {worst} rows in one round, the "
                                   f"receive buffers hold {self.cap}; raise capacity_rows")
            self.late_check()
            self.gen ^= 1
            return allc[:, self.rank].contiguous()
        if kind != "a2av":
            raise ValueError(f"unknown exchange request {kind!r}")
        t, send, recv = req[1].contiguous(), req[2], req[3]
        direction = req[4] if len(req) > 4 else "fwd"
        tag = req[5] if len(req) > 5 else None
        if tag not in self.row_bytes or t.dtype != self.dtypes[tag]:
            raise ValueError(f"PeerExchange: no receive buffer for tag {tag!r} / dtype {t.dtype}")
        rb = self.row_bytes[tag]
        if t.numel() and t.element_size() * (t.numel() // t.shape[0]) != rb:
            raise ValueError("PeerExchange: row size does not match the receive buffer")
        n_recv = int(sum(recv))
        if n_recv > self.cap:
            raise RuntimeError(f"PeerExchange: {n_recv} rows to receive, buffer holds {self.cap}")
        M = self.M if direction == "fwd" else self.M.t()
        me, st = self.rank, torch.cuda.current_stream().cuda_stream
        roff = self.region_off[(self.gen, tag)]
        soff = 0
        for d in range(self.world):
            n = int(send[d])
            if n:
                dst_row = int(M[:me, d].sum())                  # rows that lower ranks put before mine
                rc = self.lib.edl_peer_copy(self.base[d] + roff + dst_row * rb, t.data_ptr() + soff * rb,
                                            n * rb, st)
                if rc:
                    from . import _lib
                    _lib.check(rc, "edl_peer_copy")
            soff += n
        self._fence()
        return self._recv_view(tag, n_recv)

    def late_check(self):
        """Raise if a fence of an EARLIER step timed out (the kernel leaves a mark in own_flags[world]);
        read at the counts round, where the host is synchronised with the device anyway."""
        if int(self.own_flags[self.world].item()) != 0:
            raise RuntimeError("PeerExchange: a peer did not reach an exchange fence in time; the data of that "
                               "round is incomplete")

    def _recv_view(self, tag, n_recv):
        rb, roff = self.row_bytes[tag], self.region_off[(self.gen, tag)]
        out = self.own[roff:roff + n_recv * rb].view(self.dtypes[tag])
        return out if tag == "keys" else out.view(n_recv, -1)

    # ---- fused pack -> send: kernels write straight into the receivers' buffers
    def remote_rows(self, tag: str, direction: str, dst: int, n_rows: int):
        """A tensor view of the slice of rank `dst`'s receive buffer that this rank fills in the
        current round (peer-mapped memory: stores to it travel over NVLink)."""
        rb = self.row_bytes[tag]
        M = self.M if direction == "fwd" else self.M.t()
        dst_row = int(M[:self.rank, dst].sum())
        ptr = self.base[dst] + self.region_off[(self.gen, tag)] + dst_row * rb
        t = torch.as_tensor(_RawDeviceMemory(ptr, max(n_rows * rb, 1)), device=self.device)
        return t[:n_rows * rb].view(self.dtypes[tag]).view(n_rows, -1)

    def dst_table(self, tag: str, direction: str, offsets):
        """Destination table for a fused reduce -> send (kernels.segreduce dst_table): group s of
        the partition order goes to rank s's receive buffer, after the rows of lower ranks."""
        rb = self.row_bytes[tag]
        M = self.M if direction == "fwd" else self.M.t()
        roff = self.region_off[(self.gen, tag)]
        bases = [self.base[d] + roff + int(M[:self.rank, d].sum()) * rb for d in range(self.world)]
        t = [self.world, 0] + [int(x) for x in offsets] + bases + [0] * self.world   # no slot limit, no fp32 region
        return torch.tensor(t, dtype=torch.int64, device=self.device)

    def finish_round(self, tag: str, recv):
        """Close a round whose data was written by kernels (remote_rows): fence, then the view."""
        n_recv = int(sum(recv))
        if n_recv > self.cap:
            raise RuntimeError(f"PeerExchange: {n_recv} rows to receive, buffer holds {self.cap}")
        self._fence()
        return self._recv_view(tag, n_recv)

    def run(self, gen: Generator):
        return _drive(gen, self.handle)
