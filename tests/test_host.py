"""CPU tests of the host layer: C-ABI library loads and exports what the header declares,
layout math, the product's input generator and checkpoint writer against the oracle's."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from easydl_b200 import _lib
    lib = _lib.load()  # raises if not built
    hdr = open(os.path.join(ROOT, "include", "edlsparse.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(edl_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in edlsparse.h but not exported"
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    assert lib.edl_version() >= 100
    # argument errors come back as codes + message, without touching the GPU
    assert lib.edl_unique_workspace_bytes(1000, 28, 0, 0) > 0
    rc = lib.edl_gather_keys(None, 6, None, 4, None, None, None, None)
    assert rc == -1 and b"keys is null" in lib.edl_last_error_string()


def test_missing_library_fails_loudly(monkeypatch):
    from easydl_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libedlsparse.so")
    with pytest.raises(_lib.EdlError, match="no CPU"):
        _lib.load()


def test_cpu_tensor_is_rejected_not_computed_on_cpu():
    from easydl_b200 import _lib, kernels as K
    with pytest.raises(_lib.EdlError, match="CUDA tensor"):
        K.gather_keys(torch.zeros(4, 8), torch.zeros(2, dtype=torch.int64), None)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "easydl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "edl_oracle" not in src and "oracle/" not in src, f"{f} reaches into oracle/"


def test_layout_matches_oracle_shard_math():
    from easydl_b200.layout import ShardLayout, shard_rows
    rows = [10, 7, 40_000_000, 1]
    for S in (1, 2, 4, 6, 8):
        tot = 0
        for s in range(S):
            lay = ShardLayout(rows, S, s)
            for t, r in enumerate(rows):
                n = shard_rows(r, s, S)
                assert n == O.shard_rows(r, s, S) == len(range(s, r, S))
                assert lay.local_base[t + 1] - lay.local_base[t] == n
            tot += lay.local_rows
            a = lay.to_array()
            assert a[0] == len(rows) and a[1] == S and a[2] == s and len(a) == 4 + 2 * (len(rows) + 1)
            assert np.array_equal(lay.key_base, O.key_bases(rows))
        assert tot == sum(rows)
    assert ShardLayout([10_000_000] * 26).key_bits == 28
    assert ShardLayout([40_000_000] * 26).key_bits == 30


def test_generator_matches_oracle_generator():
    from easydl_b200.criteo import CriteoSynthetic
    rows = [1000, 37, 50001, 8]
    for alpha in (0.5, 1.0, 1.05, 1.2):
        src = CriteoSynthetic(rows, 257, alpha)
        for step in (0, 5):
            d, ids, lab = src.batch(step)
            od, oids, olab = O.criteo_batch(rows, 257, alpha, step=step)
            assert np.array_equal(ids, oids) and np.array_equal(d, od) and np.array_equal(lab, olab)
            assert np.array_equal(src.grad(8, step), O.synthetic_grad(257 * 4, 8, step=step))


def test_unique_fraction_matches_survey_table():
    # batch 64K over 10M rows, alpha 1.05 -> U/N ~ 0.418 per table
    from easydl_b200.criteo import zipf_field
    ids = zipf_field(10_000_000, 1.05, 0, 65536, O.DEFAULT_SEED)
    frac = len(np.unique(ids)) / 65536
    assert 0.39 < frac < 0.45, frac


@pytest.mark.parametrize("name", ["adagrad", "adam"])
def test_checkpoint_writer_is_byte_identical_to_oracle(name, tmp_path):
    from easydl_b200 import checkpoint as ck
    from easydl_b200.layout import ShardLayout
    from easydl_b200.parameters import ShardRecords
    opt = O.OptimizerSpec.adagrad() if name == "adagrad" else O.OptimizerSpec.adam()
    rows, dim = [1000, 37, 50001, 8], 8
    bases = O.key_bases(rows)
    sh = O.OracleShard(rows, dim, opt)
    for step in range(2):
        _, ids, _ = O.criteo_batch(rows, 64, 1.2, step=step)
        k = O.fuse_keys(ids, bases)
        O.sparse_step(sh, k, O.synthetic_grad(len(k), dim, step=step))
    for s in O.reshard([sh], 3):
        a, b = str(tmp_path / f"o{s.shard}.ckpt"), str(tmp_path / f"p{s.shard}.ckpt")
        O.write_checkpoint(a, s)
        rec = ShardRecords(torch.from_numpy(s.keys), torch.from_numpy(s.w),
                           [torch.from_numpy(x) for x in s.slots], s.step)
        ck.write_shard_file(b, ShardLayout(rows, 3, s.shard), dim, opt.kind, rec)
        assert open(a, "rb").read() == open(b, "rb").read()
        meta, back = ck.read_shard_file(b, bases)
        assert meta["S"] == 3 and meta["shard"] == s.shard and meta["step"] == s.step
        assert np.array_equal(back.keys.numpy(), s.keys) and np.array_equal(back.rows.numpy(), s.w)
    assert ck.shard_file("d", 12, 1, 3).endswith("version-12/variables-1-of-3.ckpt")


def test_adam_lr_t_matches_oracle():
    from easydl_b200.optimizer import OptimizerSpec
    spec = OptimizerSpec.adam()
    for t in (1, 2, 10, 1000):
        assert np.float32(spec.params(t).lr) == O.adam_lr_t(1e-3, 0.9, 0.999, t)
