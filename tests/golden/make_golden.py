#!/usr/bin/env python
"""Generates tests/golden/*.json from the NumPy oracle (run from the repo root):

    python tests/golden/make_golden.py

The reference snapshot has no fixtures for this path, so these are
SELF-GENERATED known answers: they freeze the oracle's behaviour at small shapes so that a
later change to the oracle, the SPEC or the kernels shows up as a diff.  Floats are stored as
their IEEE-754 bit patterns (uint32) so the comparison is bit-exact.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import edl_oracle as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).reshape(a.shape).tolist()


def dump(name, obj):
    with open(os.path.join(OUT, name), "w") as f:
        json.dump(obj, f, separators=(",", ":"), sort_keys=True)
        f.write("\n")


def main():
    # ---- row initialiser
    cases = []
    for (seed, t, ids, dim) in [(O.DEFAULT_SEED, 0, [0, 1, 2, 9_999_999], 8), (O.DEFAULT_SEED, 25, [7, 123456], 16),
                                (1, 3, [0], 4), (2 ** 63 + 5, 63, [2 ** 31, 39_999_999], 4)]:
        cases.append(dict(seed=seed, t=t, ids=ids, dim=dim, bits=bits(O.init_rows(seed, t, np.array(ids), dim))))
    dump("init_rows.json", cases)

    # ---- synthetic ids
    rows = [1000, 37, 50001, 8]
    dump("criteo_ids.json", dict(rows=rows, batch=12, alpha=1.05, seed=O.DEFAULT_SEED, step=3,
                                 ids=O.criteo_batch(rows, 12, 1.05, O.DEFAULT_SEED, 3)[1].tolist(),
                                 ids_alpha1=O.criteo_batch(rows, 12, 1.0, O.DEFAULT_SEED, 0)[1].tolist()))

    # ---- A1 / A2
    keys = np.array([5, 3, 5, 5, 1038, 0, 3, 1037, 51037, 51037, 1000, 999], dtype=np.int64)
    u, inv, cnt, order, seg = O.dedup(keys)
    bases = O.key_bases(rows)
    part = {}
    for S in (1, 2, 3, 6, 8):
        pk, off, perm = O.partition_mod(u, bases, S)
        part[str(S)] = dict(keys=pk.tolist(), offsets=off.tolist(), perm=perm.tolist())
    dump("dedup_partition.json", dict(rows=rows, keys=keys.tolist(), uniq=u.tolist(), inverse=inv.tolist(),
                                      counts=cnt.tolist(), order=order.tolist(), seg_offsets=seg.tolist(),
                                      partition=part))

    # ---- whole steps (A3 + A4 + A5), Adagrad and Adam, then re-shard 1 -> 4 -> 8 -> 6
    dim, batch = 8, 64
    for name, opt in (("adagrad", O.OptimizerSpec.adagrad()), ("adam", O.OptimizerSpec.adam())):
        sh = O.OracleShard(rows, dim, opt)
        embs = []
        for step in range(3):
            _, ids, _ = O.criteo_batch(rows, batch, 1.2, step=step)
            k = O.fuse_keys(ids, bases)
            g = O.synthetic_grad(len(k), dim, step=step)
            emb, _, gu = O.sparse_step(sh, k, g)
            embs.append(hashlib.sha256(emb.tobytes()).hexdigest())
        s4 = O.reshard([sh], 4)
        s8 = O.reshard(s4, 8)
        s6 = O.reshard(s8, 6)
        dump(f"steps_{name}.json", dict(
            rows=rows, dim=dim, batch=batch, alpha=1.2, n_steps=3, emb_sha256=embs,
            keys=sh.keys.tolist(), w_bits=bits(sh.w), slot_bits=[bits(s) for s in sh.slots], step=sh.step,
            shard_sizes={"4": [len(s.keys) for s in s4], "8": [len(s.keys) for s in s8],
                         "6": [len(s.keys) for s in s6]},
            shard6_keys=[s.keys.tolist() for s in s6]))
        # checkpoint bytes of shard 1 of 6
        p = os.path.join(OUT, "_tmp.ckpt")
        O.write_checkpoint(p, s6[1])
        blob = open(p, "rb").read()
        os.remove(p)
        dump(f"checkpoint_{name}.json", dict(shard=1, S=6, n_bytes=len(blob), sha256=hashlib.sha256(blob).hexdigest(),
                                             head_hex=blob[:32].hex()))


if __name__ == "__main__":
    main()
