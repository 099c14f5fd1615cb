"""Checkpoint files of a PS shard and checkpoint-based re-sharding (SPEC "Checkpoint file",
A6).  Lineage names: CheckpointSaver, restore_params_from_checkpoint.

File `version-<step>/variables-<s>-of-<S>.ckpt`, little-endian:
    magic "EDLS", version u32, S u32, shard u32, T u32, step u64, optimizer u32,
    then per table: t u32, D u32, n u64, ids[n] i64, rows[n*D] f32, slot0[n*D] f32,
    (slot1[n*D] f32 for Adam).
Only materialised (updated) rows are stored.  Restore on shard s' of S' reads ALL files of the
version and keeps the records with id mod S' == s' -- the filter runs on the GPU inside
edl_import_shard.
"""
from __future__ import annotations

import glob
import os
import re
import struct
from typing import List, Optional, Tuple

import numpy as np
import torch

from .layout import ShardLayout
from .parameters import ShardParameters, ShardRecords

MAGIC = b"EDLS"
VERSION = 1
_HEAD = "<IIIIQI"
_THEAD = "<IIQ"


def shard_file(directory: str, step: int, shard: int, num_shards: int) -> str:
    return os.path.join(directory, f"version-{step}", f"variables-{shard}-of-{num_shards}.ckpt")


def write_shard_file(path: str, layout: ShardLayout, dim: int, opt_kind: int, rec: ShardRecords):
    """rec: host tensors, ascending by fused key."""
    keys = rec.keys.numpy()
    rows = rec.rows.numpy()
    slots = [s.numpy() for s in rec.slots]
    kb = layout.key_base
    bounds = np.searchsorted(keys, kb)  # records of table t: [bounds[t], bounds[t+1])
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack(_HEAD, VERSION, layout.num_shards, layout.shard, layout.n_tables,
                            rec.step, opt_kind))
        for t in range(layout.n_tables):
            b, e = int(bounds[t]), int(bounds[t + 1])
            f.write(struct.pack(_THEAD, t, dim, e - b))
            f.write((keys[b:e] - kb[t]).astype("<i8").tobytes())
            f.write(np.ascontiguousarray(rows[b:e], dtype="<f4").tobytes())
            for s in slots:
                f.write(np.ascontiguousarray(s[b:e], dtype="<f4").tobytes())
    os.replace(tmp, path)  # a crashed writer never leaves a half-written shard file behind


def read_shard_file(path: str, key_base: np.ndarray) -> Tuple[dict, ShardRecords]:
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:4] != MAGIC:
        raise ValueError(f"{path}: bad magic {buf[:4]!r}")
    ver, S, shard, T, step, kind = struct.unpack_from(_HEAD, buf, 4)
    if ver != VERSION:
        raise ValueError(f"{path}: unsupported version {ver}")
    if T != len(key_base) - 1:
        raise ValueError(f"{path}: {T} tables in file, {len(key_base) - 1} expected")
    off = 4 + struct.calcsize(_HEAD)
    n_slots = 1 if kind == 0 else 2
    ks, rs, ss = [], [], [[] for _ in range(n_slots)]
    dim = None
    for _ in range(T):
        t, D, n = struct.unpack_from(_THEAD, buf, off)
        off += struct.calcsize(_THEAD)
        dim = D
        ks.append(np.frombuffer(buf, "<i8", n, off) + key_base[t])
        off += 8 * n
        rs.append(np.frombuffer(buf, "<f4", n * D, off).reshape(n, D))
        off += 4 * n * D
        for k in range(n_slots):
            ss[k].append(np.frombuffer(buf, "<f4", n * D, off).reshape(n, D))
            off += 4 * n * D
    if off != len(buf):
        raise ValueError(f"{path}: {len(buf) - off} trailing bytes")
    rec = ShardRecords(torch.from_numpy(np.concatenate(ks).astype(np.int64)),
                       torch.from_numpy(np.concatenate(rs).astype(np.float32)),
                       [torch.from_numpy(np.concatenate(s).astype(np.float32)) for s in ss], int(step))
    return dict(S=S, shard=shard, T=T, step=int(step), opt_kind=kind, dim=dim), rec


class CheckpointSaver:
    """Writes one shard's file per version; keeps at most `keep_max` versions."""

    def __init__(self, directory: str, keep_max: int = 0):
        self.directory = directory
        self.keep_max = keep_max

    def save(self, params: ShardParameters) -> str:
        rec = params.export_records().cpu()
        path = shard_file(self.directory, rec.step, params.shard, params.num_shards)
        write_shard_file(path, params.layout, params.dim, params.opt.kind, rec)
        if self.keep_max > 0 and params.shard == 0:
            self.prune()
        return path

    def prune(self):
        """Keep the newest `keep_max` COMPLETE versions (every shard's file present) and whatever is newer
        than the oldest of them (versions other shards are still writing).  Shard 0 calls this after its
        own write without waiting for the others, so an older complete version must survive until a newer
        one is whole: otherwise a crash of any other shard would leave nothing to restore from."""
        vs = list_versions(self.directory)
        complete = [v for v in vs if version_complete(self.directory, v)]
        if len(complete) <= self.keep_max:
            return
        oldest_kept = complete[-self.keep_max]
        for v in vs:
            if v < oldest_kept:
                for p in glob.glob(os.path.join(self.directory, f"version-{v}", "*.ckpt")):
                    os.remove(p)
                try:
                    os.rmdir(os.path.join(self.directory, f"version-{v}"))
                except OSError:
                    pass


def list_versions(directory: str) -> List[int]:
    out = []
    for d in glob.glob(os.path.join(directory, "version-*")):
        m = re.fullmatch(r"version-(\d+)", os.path.basename(d))
        if m:
            out.append(int(m.group(1)))
    return sorted(out)


def version_files(directory: str, version: int) -> List[str]:
    files = glob.glob(os.path.join(directory, f"version-{version}", "variables-*-of-*.ckpt"))
    parsed = []
    for p in files:
        m = re.fullmatch(r"variables-(\d+)-of-(\d+)\.ckpt", os.path.basename(p))
        if m:
            parsed.append((int(m.group(1)), int(m.group(2)), p))
    if not parsed:
        raise FileNotFoundError(f"no shard files under {directory}/version-{version}")
    S = parsed[0][1]
    have = sorted(s for s, _, _ in parsed)
    if any(n != S for _, n, _ in parsed) or have != list(range(S)):
        raise ValueError(f"incomplete checkpoint version {version}: shards {have} of {S}")
    return [p for _, _, p in sorted(parsed)]


def version_complete(directory: str, version: int) -> bool:
    try:
        version_files(directory, version)
        return True
    except (FileNotFoundError, ValueError):
        return False


def latest_complete_version(directory: str) -> int:
    """Newest version whose shard files are all present (a version that some shard was still writing
    when the job died is skipped)."""
    vs = list_versions(directory)
    for v in reversed(vs):
        if version_complete(directory, v):
            return v
    raise FileNotFoundError(f"no complete checkpoint version in {directory} (versions seen: {vs})")


def restore_params_from_checkpoint(directory: str, params: ShardParameters,
                                   version: Optional[int] = None) -> int:
    """Re-initialise `params` (any shard count) and load its rows from ALL shard files of the
    version (default: the newest complete one); returns the restored step.  This is scale-up /
    scale-down through checkpoints, and the recovery of a failed PS shard."""
    if version is None:
        version = latest_complete_version(directory)
    params.reset()
    step = 0
    for p in version_files(directory, version):
        meta, rec = read_shard_file(p, params.layout.key_base)
        if meta["opt_kind"] != params.opt.kind or meta["dim"] not in (None, params.dim):
            raise ValueError(f"{p}: optimizer / dim mismatch")
        params.import_records(rec.to(params.device))
        step = max(step, rec.step)
    params.step = step
    return step
