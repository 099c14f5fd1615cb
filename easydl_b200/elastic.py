"""A6, device-to-device form: move every materialised row, its optimizer slots and the step
counter from an S-shard layout to an S'-shard layout with one all-to-all-v round (no files).

The world has W ranks (W >= max(S, S')); ranks >= S hold no old shard and send nothing, ranks
>= S' receive nothing and end with no shard.  4 -> 8 -> 6 on an 8-GPU box therefore keeps the
8 processes and their communicator alive and only changes the active shard set.
"""
from __future__ import annotations

from typing import Generator, Optional

import torch

from . import kernels as K
from .layout import ShardLayout
from .parameters import ShardParameters, ShardRecords


def gen_reshard(old, new_S: int, rank: int, world: int, *, rows, dim, opt, seed, device,
                factory=None) -> Generator:
    """Returns this rank's new shard (or None when rank >= new_S).  `old` is any object with
    export_records() / layout_dev (ShardParameters or LazyShardParameters); `factory(shard)` builds
    the new shard (default: a dense ShardParameters)."""
    dev = torch.device(device)
    n_slots = opt.n_slots
    counts = torch.zeros((world, 2), dtype=torch.int64, device=dev)   # [rows to send, step]
    if old is not None:
        rec = old.export_records()
        n = len(rec)
        part = K.partition_mod(rec.keys, old.layout_dev, new_S, want_pos=False)
        offs = part.offsets.to(torch.int64)
        counts[:new_S, 0] = offs[1:] - offs[:-1]
        counts[:, 1] = rec.step
        send_keys = part.keys
        # rows / slots in partition order: out[i] = src[perm[i]]
        send_rows = K.expand(rec.rows, part.perm) if n else rec.rows
        send_slots = [K.expand(s, part.perm) if n else s for s in rec.slots]
    else:
        send_keys = torch.empty(0, dtype=torch.int64, device=dev)
        send_rows = torch.empty((0, dim), dtype=torch.float32, device=dev)
        send_slots = [torch.empty((0, dim), dtype=torch.float32, device=dev) for _ in range(n_slots)]
    got = yield ("counts", counts)                       # host [world, 2]
    send = [int(x) for x in counts[:, 0].cpu()]
    recv = [int(x) for x in got[:, 0]]
    step = int(got[:, 1].max()) if world else 0
    keys_in = yield ("a2av", send_keys, send, recv)
    rows_in = yield ("a2av", send_rows, send, recv)
    slots_in = []
    for s in send_slots:
        slots_in.append((yield ("a2av", s, send, recv)))
    if rank >= new_S:
        return None
    new = factory(rank) if factory is not None else ShardParameters(rows, dim, opt, new_S, rank, seed=seed, device=dev)
    new.import_records(ShardRecords(keys_in, rows_in, slots_in, step))
    new.step = step
    return new
