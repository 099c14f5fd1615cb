"""Owner-side dedup of the slot step on ONE GPU with synthetic slots of the cfg2 shape: S sorted runs of about
U / S keys each (U = 708 K unique ids per sender, ids spread by id mod S), timed through
`edl_slot_merge_unique` and through the chain it replaces (`edl_slot_compact` -> `edl_unique` ->
`edl_remap_index` x 2).  Prints one JSON line per S.

    python tools/merge_probe.py [S ...]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from easydl_b200 import kernels as K  # noqa: E402

dev = torch.device("cuda:0")
i32 = dict(dtype=torch.int32, device=dev)
U, KEYSPACE = 708_000, 260_000_000


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


for S in [int(x) for x in sys.argv[1:]] or [2, 4, 8]:
    rng = np.random.default_rng(S)
    per = U // S
    limit = int(per * 1.5) + 1024
    stride = limit
    slots = np.zeros(S * stride, dtype=np.int64)
    counts = np.zeros(S, dtype=np.int32)
    hot = rng.integers(0, KEYSPACE, size=per // 2, dtype=np.int64)          # ids several senders ask for
    for s in range(S):
        own = rng.integers(0, KEYSPACE, size=per, dtype=np.int64)
        run = np.unique(np.concatenate([own[: per // 2], hot]))
        slots[s * stride:s * stride + len(run)] = run
        counts[s] = len(run)
    R = S * limit
    d_slots, d_counts = torch.from_numpy(slots).to(dev), torch.from_numpy(counts).to(dev)

    def result():
        return K.UniqueResult(uniq=torch.empty(R, dtype=torch.int64, device=dev), inverse=None, counts=None,
                              order=torch.empty(R, **i32), seg_offsets=torch.empty(R + 1, **i32),
                              n_unique=torch.zeros(1, **i32), n=R, seg_first=torch.empty(R, **i32))
    lib = K._lib.load()
    ws = torch.empty(max(lib.edl_unique_workspace_bytes(R, 28, 0, 0), lib.edl_slot_merge_unique_workspace_bytes(S, limit)),
                     dtype=torch.uint8, device=dev)
    keys_c, map_c, n_recv = torch.empty(R, dtype=torch.int64, device=dev), torch.empty(R, **i32), torch.zeros(1, **i32)
    ref, got, n_out = result(), result(), torch.zeros(1, **i32)
    o2, f2 = torch.empty(R, **i32), torch.empty(R, **i32)

    def chain():
        K.slot_compact(d_slots, d_counts, stride, limit, keys_c, map_c, n_recv)
        K.unique(keys_c, None, 28, want_inverse=False, want_counts=False, n_dev=n_recv, out=ref, workspace=ws)
        K.remap_index(ref.order, map_c, n_recv, out=o2)
        K.remap_index(ref.seg_first, map_c, ref.n_unique, out=f2)

    def merge():
        K.slot_merge_unique(d_slots, d_counts, stride, limit, got, n_out, workspace=ws)

    t_chain, t_merge = timed(chain), timed(merge)
    chain()
    merge()
    torch.cuda.synchronize()
    n, Uq = int(n_recv.item()), ref.count()
    same = (got.count() == Uq and torch.equal(got.uniq[:Uq], ref.uniq[:Uq]) and torch.equal(got.order[:n], o2[:n])
            and torch.equal(got.seg_first[:Uq], f2[:Uq]) and torch.equal(got.seg_offsets[:Uq + 1], ref.seg_offsets[:Uq + 1]))
    print(json.dumps({"S": S, "keys": n, "unique": Uq, "sort_chain_us": round(t_chain, 1), "merge_us": round(t_merge, 1),
                      "same_result": bool(same)}), flush=True)
