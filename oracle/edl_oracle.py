"""CPU oracle for the EasyDL ElasticTrainer sparse-embedding step.  TEST INFRASTRUCTURE.

PARITY UNPINNED.  The reference snapshot (/root/reference) holds no implementation,
tests, golden vectors or fixtures for this path: ElasticTrainer
exists only as one README bullet (ref:README.md:11) and as PS / worker role names in
the operator design (ref:docs/design/elastic-training-operator.md:39-44,65-85).  This
module is therefore a NumPy restatement of the written contract in SURVEY.md
section 8-SPEC (copied to SPEC.md), which BASELINE.json's north_star explicitly allows
("a documented NumPy/TF-CPU re-implementation of the same semantics").  It follows no
reference file because none exists; each function cites the SPEC clause it restates.
It is pinned only by (i) SPEC.md, (ii) published known-answer values for SplitMix64,
(iii) independent cross-checks against torch.unique / torch.optim.Adagrad on CPU,
(iv) self-generated golden fixtures under tests/golden/ (script committed there).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product package (easydl_b200/) never does.

Vocabulary follows the parameter-server domain: tables, rows, shards, slots, workers.
"""
from __future__ import annotations

import io
import os
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

DEFAULT_SEED = 20240607
_U64 = np.uint64
_MASK64 = (1 << 64) - 1

ADAGRAD = 0
ADAM = 1


# --------------------------------------------------------------------------- hashing
def splitmix64(x: np.ndarray) -> np.ndarray:
    """SplitMix64 finaliser over uint64 (wrapping).  SPEC 'Row initialisation'.

    Published algorithm (Steele, Lea, Flood 2014; the java.util.SplittableRandom mixer):
    state += 0x9E3779B97F4A7C15, then two xor-shift-multiply rounds.
    """
    with np.errstate(over="ignore"):
        z = np.asarray(x, dtype=_U64) + _U64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> _U64(30))) * _U64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> _U64(27))) * _U64(0x94D049BB133111EB)
        return z ^ (z >> _U64(31))


def init_rows(seed: int, t: int, ids: np.ndarray, dim: int) -> np.ndarray:
    """Initial value of table t's rows `ids` -> [n, dim] fp32.  SPEC 'Row initialisation'.

    h = splitmix64(seed ^ (t << 56) ^ (id*D + c)); k = h >> 40; x = fp32(k) * 2^-24;
    w = x * 0.1f - 0.05f with two separately rounded fp32 operations (no FMA).
    """
    ids = np.asarray(ids, dtype=np.int64).astype(_U64)
    with np.errstate(over="ignore"):
        idx = ids[:, None] * _U64(dim) + np.arange(dim, dtype=_U64)[None, :]
        salt = _U64((int(seed) ^ (int(t) << 56)) & _MASK64)
        h = splitmix64(salt ^ idx)
    k = (h >> _U64(40)).astype(np.float32)  # 24 bits: exact in fp32
    x = k * np.float32(2.0 ** -24)
    w = x * np.float32(0.1)
    w = w - np.float32(0.05)
    return w.astype(np.float32)


# --------------------------------------------------------------------------- shard math
def int_to_id(ids, num_shards: int):
    """shard(t, id) = id mod S on the per-table id.  SPEC 'Shard function'."""
    return np.asarray(ids, dtype=np.int64) % np.int64(num_shards)


def local_row(ids, num_shards: int):
    """local_row(t, id) = id div S.  SPEC 'Shard function'."""
    return np.asarray(ids, dtype=np.int64) // np.int64(num_shards)


def shard_rows(rows_t: int, shard: int, num_shards: int) -> int:
    """Shard s of table t holds ceil((rows_t - s) / S) local rows."""
    if shard >= rows_t:
        return 0
    return (rows_t - shard + num_shards - 1) // num_shards


def key_bases(rows: Sequence[int]) -> np.ndarray:
    """Fused-key bases: key(t, id) = base[t] + id; base = exclusive prefix sum of rows_t.

    Keys of different tables occupy disjoint ascending ranges, so a dedup on the fused
    key equals the per-table dedups concatenated (SPEC A1 'results must be identical
    per table')."""
    b = np.zeros(len(rows) + 1, dtype=np.int64)
    b[1:] = np.cumsum(np.asarray(rows, dtype=np.int64))
    return b


def fuse_keys(ids: np.ndarray, bases: np.ndarray) -> np.ndarray:
    """ids[B, F] (field f looks up table f) -> flat keys[B*F], sample-major."""
    ids = np.asarray(ids, dtype=np.int64)
    return (ids + bases[None, : ids.shape[1]]).reshape(-1)


def split_key(keys: np.ndarray, bases: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """fused key -> (table, per-table id)."""
    keys = np.asarray(keys, dtype=np.int64)
    t = np.searchsorted(bases, keys, side="right") - 1
    return t.astype(np.int64), keys - bases[t]


# --------------------------------------------------------------------------- A1
def dedup(keys: np.ndarray):
    """A1: uniq ascending int64, inverse int32, counts int32 (== np.unique).

    Also returns the two derived arrays the segmented reduce consumes:
    order = stable argsort(keys) (int32) and seg_offsets = exclusive cumsum(counts)
    with seg_offsets[U] = N (int32)."""
    keys = np.asarray(keys, dtype=np.int64).reshape(-1)
    uniq, inverse, counts = np.unique(keys, return_inverse=True, return_counts=True)
    order = np.argsort(keys, kind="stable").astype(np.int32)
    seg = np.zeros(len(uniq) + 1, dtype=np.int32)
    seg[1:] = np.cumsum(counts)
    return (uniq.astype(np.int64), inverse.reshape(-1).astype(np.int32),
            counts.astype(np.int32), order, seg)


# --------------------------------------------------------------------------- A2
def partition_mod(keys: np.ndarray, bases: np.ndarray, num_shards: int):
    """A2: stable partition of `keys` by (per-table id mod S).

    Returns keys grouped by shard (ascending within each group when the input is
    ascending), offsets[S+1] int32, perm[U] int32 = original position of each output."""
    keys = np.asarray(keys, dtype=np.int64)
    _, ids = split_key(keys, bases)
    shard = ids % np.int64(num_shards)
    perm = np.argsort(shard, kind="stable").astype(np.int32)
    offsets = np.zeros(num_shards + 1, dtype=np.int32)
    offsets[1:] = np.cumsum(np.bincount(shard, minlength=num_shards))
    return keys[perm], offsets, perm


# --------------------------------------------------------------------------- A4
def segment_sum_f64_reduceat(grad: np.ndarray, order: np.ndarray, seg: np.ndarray) -> np.ndarray:
    """A4, definitional form: np.add.reduceat over the float64 copy of the sorted rows."""
    n_u = len(seg) - 1
    if n_u == 0:
        return np.zeros((0, grad.shape[1]), dtype=np.float32)
    g64 = grad[order].astype(np.float64)
    out = np.add.reduceat(g64, seg[:-1].astype(np.int64), axis=0)
    return out.astype(np.float32)


def segment_sum_f64(grad: np.ndarray, order: np.ndarray, seg: np.ndarray) -> np.ndarray:
    """A4: g_uniq[u] = sum of grad rows of segment u, accumulated in float64 and
    rounded once to fp32.  SPEC 'A4 gradient reduce'.

    Same sums as segment_sum_f64_reduceat, evaluated as a CSR (U x N, all ones) times dense
    product in float64 when SciPy is present: its C loop adds the rows of a segment left to
    right, 20x faster than reduceat over axis 0, which matters for the timed CPU baseline."""
    n_u = len(seg) - 1
    if n_u == 0:
        return np.zeros((0, grad.shape[1]), dtype=np.float32)
    try:
        from scipy.sparse import csr_matrix
    except ImportError:  # pragma: no cover
        return segment_sum_f64_reduceat(grad, order, seg)
    n = len(order)
    m = csr_matrix((np.ones(n, dtype=np.float64), order.astype(np.int32), seg.astype(np.int32)),
                   shape=(n_u, grad.shape[0]))
    return np.asarray(m @ grad.astype(np.float64)).astype(np.float32)


# --------------------------------------------------------------------------- A5
@dataclass
class OptimizerSpec:
    """SPEC 'A5 optimizer'.  kind: ADAGRAD or ADAM."""
    kind: int = ADAGRAD
    lr: float = 0.01
    eps: float = 1e-7
    init_accum: float = 0.1
    beta1: float = 0.9
    beta2: float = 0.999

    @staticmethod
    def adagrad(lr=0.01, eps=1e-7, init_accum=0.1):
        return OptimizerSpec(ADAGRAD, lr, eps, init_accum)

    @staticmethod
    def adam(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8):
        return OptimizerSpec(ADAM, lr, eps, 0.0, beta1, beta2)

    @property
    def n_slots(self):
        return 1 if self.kind == ADAGRAD else 2


def adam_lr_t(lr: float, beta1: float, beta2: float, step: int) -> np.float32:
    """lr_t = lr*sqrt(1-b2^t)/(1-b1^t) in float64 on the host, passed as fp32."""
    return np.float32(float(lr) * np.sqrt(1.0 - float(beta2) ** step) / (1.0 - float(beta1) ** step))


def adagrad_rows(p, acc, g, lr, eps):
    """acc += g*g; p -= lr*g/(sqrt(acc)+eps), every op rounded to fp32 separately."""
    lr = np.float32(lr)
    eps = np.float32(eps)
    acc = acc + g * g
    p = p - (lr * g) / (np.sqrt(acc) + eps)
    return p.astype(np.float32), acc.astype(np.float32)


def adam_rows(p, m, v, g, lr_t, beta1, beta2, eps):
    """m = b1*m + (1-b1)*g; v = b2*v + ((1-b2)*g)*g; p -= lr_t*m/(sqrt(v)+eps), fp32."""
    b1 = np.float32(beta1)
    b2 = np.float32(beta2)
    omb1 = np.float32(1.0) - b1
    omb2 = np.float32(1.0) - b2
    eps = np.float32(eps)
    lr_t = np.float32(lr_t)
    m = b1 * m + omb1 * g
    v = b2 * v + (omb2 * g) * g
    p = p - (lr_t * m) / (np.sqrt(v) + eps)
    return p.astype(np.float32), m.astype(np.float32), v.astype(np.float32)


# --------------------------------------------------------------------------- tables
class OracleShard:
    """One PS shard's lazily materialised state for all tables, keyed by fused key.

    A row is 'materialised' once it has received an update (SPEC A6).  Untouched rows
    are a pure function of (seed, t, id) so they are never stored.  Mirrors the
    lineage's Parameters / EmbeddingTable.get lazy initialisation [RECALL, SURVEY 2d].
    """

    def __init__(self, rows: Sequence[int], dim: int, opt: OptimizerSpec,
                 num_shards: int = 1, shard: int = 0, seed: int = DEFAULT_SEED):
        self.rows = [int(r) for r in rows]
        self.dim = int(dim)
        self.opt = opt
        self.S = int(num_shards)
        self.shard = int(shard)
        self.seed = int(seed)
        self.bases = key_bases(self.rows)
        self.step = 0  # Adam's 1-based global step counter: number of applied batches
        self.keys = np.zeros(0, dtype=np.int64)  # ascending fused keys of materialised rows
        self.w = np.zeros((0, dim), dtype=np.float32)
        self.slots = [np.zeros((0, dim), dtype=np.float32) for _ in range(opt.n_slots)]

    # -- helpers
    def _init_w(self, keys: np.ndarray) -> np.ndarray:
        t, ids = split_key(keys, self.bases)
        out = np.empty((len(keys), self.dim), dtype=np.float32)
        for tt in np.unique(t):
            sel = t == tt
            out[sel] = init_rows(self.seed, int(tt), ids[sel], self.dim)
        return out

    def _init_slot(self, n: int, k: int) -> np.ndarray:
        val = self.opt.init_accum if self.opt.kind == ADAGRAD else 0.0
        return np.full((n, self.dim), np.float32(val), dtype=np.float32)

    def owns(self, keys: np.ndarray) -> np.ndarray:
        _, ids = split_key(keys, self.bases)
        return (ids % self.S) == self.shard

    def materialise(self, keys: np.ndarray):
        """Insert rows for ascending unique `keys` that are not yet stored."""
        pos = np.searchsorted(self.keys, keys)
        have = np.zeros(len(keys), dtype=bool)
        inb = pos < len(self.keys)
        have[inb] = self.keys[pos[inb]] == keys[inb]
        new = keys[~have]
        if len(new) == 0:
            return
        merged = np.concatenate([self.keys, new])
        o = np.argsort(merged, kind="stable")
        self.keys = merged[o]
        self.w = np.concatenate([self.w, self._init_w(new)])[o]
        self.slots = [np.concatenate([s, self._init_slot(len(new), k)])[o]
                      for k, s in enumerate(self.slots)]

    # -- A3
    def lookup(self, keys: np.ndarray) -> np.ndarray:
        """rows[i] = table[key_i]; materialised value if present else the init value."""
        keys = np.asarray(keys, dtype=np.int64)
        out = self._init_w(keys) if len(self.keys) == 0 else None
        if out is not None:
            return out
        pos = np.searchsorted(self.keys, keys)
        inb = pos < len(self.keys)
        have = np.zeros(len(keys), dtype=bool)
        have[inb] = self.keys[pos[inb]] == keys[inb]
        out = np.empty((len(keys), self.dim), dtype=np.float32)
        out[have] = self.w[pos[have]]
        if (~have).any():
            out[~have] = self._init_w(keys[~have])
        return out

    def slot_lookup(self, k: int, keys: np.ndarray) -> np.ndarray:
        keys = np.asarray(keys, dtype=np.int64)
        out = self._init_slot(len(keys), k)
        if len(self.keys):
            pos = np.searchsorted(self.keys, keys)
            inb = pos < len(self.keys)
            have = np.zeros(len(keys), dtype=bool)
            have[inb] = self.keys[pos[inb]] == keys[inb]
            out[have] = self.slots[k][pos[have]]
        return out

    # -- A5
    def apply(self, uniq: np.ndarray, g_uniq: np.ndarray, bump_step: bool = True,
              materialise: bool = True):
        """Sparse optimizer step on rows `uniq` (ascending, unique, owned by this shard).
        bump_step=False / materialise=False let a per-table driver call this once per table
        inside one global step, on rows that already exist (disjoint in-place updates)."""
        if bump_step:
            self.step += 1
        if len(uniq) == 0:
            return
        if materialise:
            self.materialise(uniq)
        pos = np.searchsorted(self.keys, uniq)
        if self.opt.kind == ADAGRAD:
            p, acc = adagrad_rows(self.w[pos], self.slots[0][pos], g_uniq,
                                  self.opt.lr, self.opt.eps)
            self.w[pos] = p
            self.slots[0][pos] = acc
        else:
            lr_t = adam_lr_t(self.opt.lr, self.opt.beta1, self.opt.beta2, self.step)
            p, m, v = adam_rows(self.w[pos], self.slots[0][pos], self.slots[1][pos], g_uniq,
                                lr_t, self.opt.beta1, self.opt.beta2, self.opt.eps)
            self.w[pos] = p
            self.slots[0][pos] = m
            self.slots[1][pos] = v

    # -- A6
    def export(self):
        """(keys ascending, rows, slots..., step) of materialised rows."""
        return (self.keys.copy(), self.w.copy(), [s.copy() for s in self.slots], self.step)


def sparse_step(shard: OracleShard, keys: np.ndarray, grad: np.ndarray):
    """One whole single-shard step on fused keys[N] with upstream grad[N, D]:
    A1 dedup -> A3 lookup + expand -> A4 reduce -> A5 apply.  Returns (emb[N, D], dedup
    outputs, g_uniq)."""
    uniq, inverse, counts, order, seg = dedup(keys)
    rows_u = shard.lookup(uniq)
    emb = rows_u[inverse]
    g_uniq = segment_sum_f64(grad, order, seg)
    shard.apply(uniq, g_uniq)
    return emb, (uniq, inverse, counts, order, seg), g_uniq


def sparse_step_per_table(shard: OracleShard, ids: np.ndarray, grad: np.ndarray, pool=None):
    """The same step as sparse_step, one table at a time (tables are independent), optionally
    on a thread pool: NumPy releases the GIL inside sort / take / ufunc loops, so this is the
    multi-core form of the CPU parameter-server path used by bench.py --impl reference.
    ids[B, F] per-table ids, grad[B*F, D].  Rows must already be materialised when pool is
    given (threads then only update disjoint slices in place)."""
    B, F = ids.shape
    D = shard.dim
    g3 = grad.reshape(B, F, D)
    emb = np.empty((B, F, D), dtype=np.float32)
    shard.step += 1

    def one(t):
        keys = ids[:, t] + shard.bases[t]
        uniq, inverse, counts, order, seg = dedup(keys)
        emb[:, t, :] = shard.lookup(uniq)[inverse]
        g_uniq = segment_sum_f64(np.ascontiguousarray(g3[:, t, :]), order, seg)
        shard.apply(uniq, g_uniq, bump_step=False, materialise=pool is None)
        return len(uniq)

    if pool is None:
        n_u = [one(t) for t in range(F)]
    else:
        n_u = list(pool.map(one, range(F)))
    return emb.reshape(B * F, D), int(sum(n_u))


def sharded_step(shards: List[OracleShard], keys_per_worker: List[np.ndarray],
                 grads_per_worker: List[np.ndarray]):
    """Multi-worker / multi-shard step: every worker looks its keys up on the
    owning shards; gradients of the WHOLE global batch are reduced in one float64 pass
    per id (SPEC A4: the answer does not depend on S) and applied by the owner.
    Returns the per-worker emb arrays."""
    S = len(shards)
    bases = shards[0].bases
    embs = []
    for keys in keys_per_worker:
        _, ids = split_key(keys, bases)
        own = ids % S
        emb = np.empty((len(keys), shards[0].dim), dtype=np.float32)
        for s in range(S):
            sel = own == s
            emb[sel] = shards[s].lookup(keys[sel])
        embs.append(emb)
    all_keys = np.concatenate(keys_per_worker)  # worker-rank order == peer-rank order
    all_grad = np.concatenate(grads_per_worker)
    uniq, _, _, order, seg = dedup(all_keys)
    g_uniq = segment_sum_f64(all_grad, order, seg)
    _, uid = split_key(uniq, bases)
    own = uid % S
    for s in range(S):
        sel = own == s
        shards[s].apply(uniq[sel], g_uniq[sel])
    return embs


def reshard(shards: List[OracleShard], new_S: int) -> List[OracleShard]:
    """A6: S -> S' re-shard.  New shard s' receives every materialised record with
    id mod S' == s', ascending by key, bit-exact, plus the step counter."""
    old = shards[0]
    keys = np.concatenate([s.keys for s in shards])
    w = np.concatenate([s.w for s in shards])
    slots = [np.concatenate([s.slots[k] for s in shards]) for k in range(old.opt.n_slots)]
    steps = {s.step for s in shards}
    assert len(steps) == 1, "shards disagree on the step counter"
    _, ids = split_key(keys, old.bases)
    out = []
    for sp in range(new_S):
        ns = OracleShard(old.rows, old.dim, old.opt, new_S, sp, old.seed)
        sel = np.nonzero(ids % new_S == sp)[0]
        o = sel[np.argsort(keys[sel], kind="stable")]
        ns.keys = keys[o]
        ns.w = w[o]
        ns.slots = [sl[o] for sl in slots]
        ns.step = old.step
        out.append(ns)
    return out


# --------------------------------------------------------------------------- checkpoint
CKPT_MAGIC = b"EDLS"
CKPT_VERSION = 1


def write_checkpoint(path: str, shard: OracleShard):
    """SPEC 'Checkpoint file (per shard, little-endian)'."""
    keys, w, slots, step = shard.export()
    t_of, ids = split_key(keys, shard.bases)
    T = len(shard.rows)
    with open(path, "wb") as f:
        f.write(CKPT_MAGIC)
        f.write(struct.pack("<IIIIQI", CKPT_VERSION, shard.S, shard.shard, T, step, shard.opt.kind))
        for t in range(T):
            sel = t_of == t
            n = int(sel.sum())
            f.write(struct.pack("<IIQ", t, shard.dim, n))
            f.write(ids[sel].astype("<i8").tobytes())
            f.write(w[sel].astype("<f4").tobytes())
            for s in slots:
                f.write(s[sel].astype("<f4").tobytes())


def read_checkpoint(path: str):
    """-> dict(S, shard, T, step, opt_kind, tables=[(t, D, ids, rows, [slots])])."""
    with open(path, "rb") as f:
        buf = f.read()
    assert buf[:4] == CKPT_MAGIC, "bad magic"
    ver, S, shard, T, step, kind = struct.unpack_from("<IIIIQI", buf, 4)
    assert ver == CKPT_VERSION
    off = 4 + struct.calcsize("<IIIIQI")
    n_slots = 1 if kind == ADAGRAD else 2
    tables = []
    for _ in range(T):
        t, D, n = struct.unpack_from("<IIQ", buf, off)
        off += struct.calcsize("<IIQ")
        ids = np.frombuffer(buf, dtype="<i8", count=n, offset=off).copy()
        off += 8 * n
        rows = np.frombuffer(buf, dtype="<f4", count=n * D, offset=off).reshape(n, D).copy()
        off += 4 * n * D
        slots = []
        for _k in range(n_slots):
            slots.append(np.frombuffer(buf, dtype="<f4", count=n * D, offset=off).reshape(n, D).copy())
            off += 4 * n * D
        tables.append((t, D, ids, rows, slots))
    assert off == len(buf)
    return dict(S=S, shard=shard, T=T, step=step, opt_kind=kind, tables=tables)


def restore_from_checkpoints(paths: Sequence[str], rows: Sequence[int], dim: int,
                             opt: OptimizerSpec, new_S: int, new_shard: int,
                             seed: int = DEFAULT_SEED) -> OracleShard:
    """Restore = read ALL old shard files, keep records with id mod S' == s'.  SPEC A6."""
    ns = OracleShard(rows, dim, opt, new_S, new_shard, seed)
    ks, ws, sls = [], [], [[] for _ in range(opt.n_slots)]
    step = None
    for p in paths:
        ck = read_checkpoint(p)
        step = ck["step"] if step is None else step
        assert step == ck["step"]
        for (t, D, ids, r, slots) in ck["tables"]:
            sel = ids % new_S == new_shard
            ks.append(ids[sel] + ns.bases[t])
            ws.append(r[sel])
            for k in range(opt.n_slots):
                sls[k].append(slots[k][sel])
    keys = np.concatenate(ks) if ks else np.zeros(0, np.int64)
    o = np.argsort(keys, kind="stable")
    ns.keys = keys[o]
    ns.w = (np.concatenate(ws) if ws else np.zeros((0, dim), np.float32))[o]
    ns.slots = [(np.concatenate(s) if s else np.zeros((0, dim), np.float32))[o] for s in sls]
    ns.step = int(step or 0)
    return ns


# --------------------------------------------------------------------------- inputs
def zipf_ids(rows: int, alpha: float, t: int, batch: int, seed: int) -> np.ndarray:
    """SPEC 'Synthetic Criteo-shape input': bounded-Zipf rank, scattered by a
    multiplicative hash.  Restated here so the oracle is self-contained; the product's
    generator (easydl_b200/criteo.py) is tested against it."""
    u = np.random.Generator(np.random.PCG64(seed + t)).random(batch)
    if abs(alpha - 1.0) < 1e-12:
        x = np.power(float(rows + 1), u)
    else:
        a = 1.0 - alpha
        x = np.power((np.power(float(rows + 1), a) - 1.0) * u + 1.0, 1.0 / a)
    r = np.clip(np.floor(x).astype(np.int64) - 1, 0, rows - 1)
    return ((r * np.int64(2654435761) + np.int64(t)) % np.int64(rows)).astype(np.int64)


def criteo_batch(rows: Sequence[int], batch: int, alpha: float = 1.05,
                 seed: int = DEFAULT_SEED, step: int = 0):
    """-> dense[B,13] f32, ids[B,F] i64, labels[B] f32.  `step` offsets the seed by
    7919*step so successive batches differ (extension of the SPEC for multi-step runs)."""
    s = seed + 7919 * step
    F = len(rows)
    ids = np.stack([zipf_ids(int(rows[t]), alpha, t, batch, s) for t in range(F)], axis=1)
    g = np.random.Generator(np.random.PCG64(s + 500))
    dense = np.log1p(g.lognormal(0.0, 1.0, size=(batch, 13))).astype(np.float32)
    labels = (g.random(batch) < 0.25).astype(np.float32)
    return dense, ids, labels


def synthetic_grad(n: int, dim: int, seed: int = DEFAULT_SEED, step: int = 0) -> np.ndarray:
    """Upstream gradient for benchmarks: N(0, sigma=1e-2) fp32, seed + 1000."""
    g = np.random.Generator(np.random.PCG64(seed + 1000 + 7919 * step))
    return (g.standard_normal((n, dim), dtype=np.float32) * np.float32(1e-2)).astype(np.float32)
