"""ctypes binding of oracle/libedloracle.so (edl_oracle_c.c).  TEST INFRASTRUCTURE.

The C file is a second, independent restatement of SPEC.md; this module gives it the NumPy oracle's call
shapes so that tests/test_oracle_c.py can compare the two bit for bit, and so that bench.py can time the
compiled multi-threaded CPU parameter-server path.  Only tests/ and bench.py's CPU legs import it.
Build the library with `make -C oracle` (done by __graft_entry__.build()).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libedloracle.so")
_lib = None

_p, _i64, _int, _u64, _f, _d = C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_float, C.c_double
_PROTOS = {
    "edlo_splitmix64": (_u64, [_u64]),
    "edlo_init_rows": (None, [_u64, _int, _p, _i64, _int, _p]),
    "edlo_dedup": (_i64, [_p, _i64, _p, _p, _p, _p, _p]),
    "edlo_partition_mod": (None, [_p, _i64, _p, _int, _int, _p, _p, _p]),
    "edlo_segment_sum_f64": (None, [_p, _int, _p, _p, _i64, _p]),
    "edlo_adagrad_rows": (None, [_p, _p, _p, _i64, _f, _f]),
    "edlo_adam_rows": (None, [_p, _p, _p, _p, _i64, _f, _f, _f, _f]),
    "edlo_adam_lr_t": (_f, [_d, _d, _d, _i64]),
    "edlo_shard_create": (_p, [_p, _int, _int, _int, _d, _d, _d, _d, _d, _u64]),
    "edlo_shard_destroy": (None, [_p]),
    "edlo_shard_step_count": (_i64, [_p]),
    "edlo_shard_rows": (_i64, [_p]),
    "edlo_shard_step": (_i64, [_p, _p, _i64, _int, _p, _p, _int]),
    "edlo_shard_export": (None, [_p, _p, _p, _p, _p]),
    "edlo_shard_write_checkpoint": (_int, [_p, C.c_char_p, _int, _int]),
}


def available() -> bool:
    return os.path.exists(LIB_PATH)


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{LIB_PATH} is missing: run `make -C oracle`")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def _a(x, dtype):
    return np.ascontiguousarray(x, dtype=dtype)


def _ptr(a):
    return None if a is None else a.ctypes.data


def splitmix64(x: int) -> int:
    return int(load().edlo_splitmix64(int(x) & (2 ** 64 - 1)))


def init_rows(seed: int, t: int, ids, dim: int) -> np.ndarray:
    ids = _a(ids, np.int64)
    out = np.empty((len(ids), dim), dtype=np.float32)
    load().edlo_init_rows(int(seed) & (2 ** 64 - 1), int(t), _ptr(ids), len(ids), dim, _ptr(out))
    return out


def dedup(keys):
    """-> (uniq, inverse, counts, order, seg_offsets), the shapes and dtypes of edl_oracle.dedup."""
    keys = _a(keys, np.int64).reshape(-1)
    n = len(keys)
    uniq = np.empty(max(n, 1), dtype=np.int64)
    inverse, counts, order = (np.empty(max(n, 1), dtype=np.int32) for _ in range(3))
    seg = np.empty(n + 1, dtype=np.int32)
    U = int(load().edlo_dedup(_ptr(keys), n, _ptr(uniq), _ptr(inverse), _ptr(counts), _ptr(order), _ptr(seg)))
    return uniq[:U].copy(), inverse[:n].copy(), counts[:U].copy(), order[:n].copy(), seg[:U + 1].copy()


def partition_mod(keys, bases, num_shards: int):
    keys, bases = _a(keys, np.int64), _a(bases, np.int64)
    n = len(keys)
    out = np.empty(max(n, 1), dtype=np.int64)
    perm = np.empty(max(n, 1), dtype=np.int32)
    offsets = np.empty(num_shards + 1, dtype=np.int32)
    load().edlo_partition_mod(_ptr(keys), n, _ptr(bases), len(bases) - 1, int(num_shards), _ptr(out), _ptr(offsets),
                              _ptr(perm))
    return out[:n].copy(), offsets, perm[:n].copy()


def segment_sum_f64(grad, order, seg) -> np.ndarray:
    grad, order, seg = _a(grad, np.float32), _a(order, np.int32), _a(seg, np.int32)
    U, dim = len(seg) - 1, grad.shape[1]
    out = np.empty((max(U, 0), dim), dtype=np.float32)
    if U > 0:
        load().edlo_segment_sum_f64(_ptr(grad), dim, _ptr(order), _ptr(seg), U, _ptr(out))
    return out


def adagrad_rows(p, acc, g, lr, eps):
    p, acc, g = _a(p, np.float32).copy(), _a(acc, np.float32).copy(), _a(g, np.float32)
    load().edlo_adagrad_rows(_ptr(p), _ptr(acc), _ptr(g), p.size, float(np.float32(lr)), float(np.float32(eps)))
    return p, acc


def adam_rows(p, m, v, g, lr_t, beta1, beta2, eps):
    p, m, v, g = _a(p, np.float32).copy(), _a(m, np.float32).copy(), _a(v, np.float32).copy(), _a(g, np.float32)
    f = lambda x: float(np.float32(x))
    load().edlo_adam_rows(_ptr(p), _ptr(m), _ptr(v), _ptr(g), p.size, f(lr_t), f(beta1), f(beta2), f(eps))
    return p, m, v


def adam_lr_t(lr, beta1, beta2, step) -> np.float32:
    return np.float32(load().edlo_adam_lr_t(float(lr), float(beta1), float(beta2), int(step)))


class CShard:
    """One single-shard (S == 1) parameter server held by the C library: hash map id -> row per table,
    rows materialised on their first update.  `opt` is an edl_oracle.OptimizerSpec."""

    def __init__(self, rows, dim: int, opt, seed: int = 20240607):
        self.lib = load()
        self.rows = _a(rows, np.int64)
        self.dim, self.opt = int(dim), opt
        self.n_slots = 1 if opt.kind == 0 else 2
        self.h = self.lib.edlo_shard_create(_ptr(self.rows), len(self.rows), self.dim, int(opt.kind), float(opt.lr),
                                            float(opt.eps), float(opt.init_accum), float(opt.beta1), float(opt.beta2),
                                            int(seed) & (2 ** 64 - 1))

    def close(self):
        if self.h:
            self.lib.edlo_shard_destroy(self.h)
            self.h = None

    __del__ = close

    @property
    def step(self) -> int:
        return int(self.lib.edlo_shard_step_count(self.h))

    def sparse_step(self, ids, grad, threads: int = 1, want_emb: bool = True):
        """ids[B, F] per-table ids, grad[B * F, dim] -> (emb[B * F, dim] or None, unique ids summed over tables)."""
        ids, grad = _a(ids, np.int64), _a(grad, np.float32)
        B, F = ids.shape
        emb = np.empty((B * F, self.dim), dtype=np.float32) if want_emb else None
        n_u = self.lib.edlo_shard_step(self.h, _ptr(ids), B, F, _ptr(grad), _ptr(emb), int(threads))
        return emb, int(n_u)

    def write_checkpoint(self, path: str, new_S: int = 1, new_shard: int = 0):
        """The EDLS file of shard `new_shard` of `new_S` after a re-shard of this shard (SPEC A6 + file format)."""
        if self.lib.edlo_shard_write_checkpoint(self.h, os.fsencode(path), int(new_S), int(new_shard)) != 0:
            raise OSError(f"could not write {path}")

    def export(self):
        """(keys ascending, rows, [slots...], step) like OracleShard.export()."""
        n = int(self.lib.edlo_shard_rows(self.h))
        keys = np.empty(n, dtype=np.int64)
        w = np.empty((n, self.dim), dtype=np.float32)
        s0 = np.empty((n, self.dim), dtype=np.float32)
        s1 = np.empty((n, self.dim), dtype=np.float32) if self.n_slots == 2 else None
        self.lib.edlo_shard_export(self.h, _ptr(keys), _ptr(w), _ptr(s0), _ptr(s1))
        return keys, w, [s0] + ([s1] if s1 is not None else []), self.step
