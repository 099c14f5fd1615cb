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
    rc = lib.edl_gather_keys(None, 6, 0, None, 4, None, None, None, None)
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


def test_slot_buffer_layout_regions_do_not_overlap():
    """easydl_b200.slot_step.slot_layout: every region of both generations, the flags, the round
    counter and the overflow flag are disjoint, aligned, and sized for world x stride rows."""
    from easydl_b200.slot_step import slot_layout
    for world, stride, dim, wb in [(1, 7, 4, 8), (2, 1000, 64, 8), (8, 1703936, 64, 8), (6, 333, 128, 4)]:
        mixed = wb == 8 and world % 2 == 0
        offs, flags_off, counter_off, err_off, nbytes, rb = slot_layout(world, stride, dim, wb, mixed)
        spans = []
        for (gen, region), o in offs.items():
            rows = world * stride if region != "counts" else world
            assert o % 256 == 0
            spans.append((o, o + rows * rb[region]))
        spans.append((flags_off, flags_off + 8 * (world + 1)))
        spans.append((counter_off, counter_off + 8))
        spans.append((err_off, err_off + 4))
        spans.sort()
        for (a0, a1), (b0, b1) in zip(spans[:-1], spans[1:]):
            assert a1 <= b0
        assert spans[-1][1] <= nbytes
        assert rb == {"keys": 8, "rows": dim * 4, "grads": dim * wb, "grads32": dim * 4 if mixed else 0, "counts": 4}
        assert world * stride < 2 ** 31           # slot-space rows are int32


def test_slot_round_model_matches_packed_all_to_all():
    """NumPy model of the slot rounds (what csrc/slots.cu computes): sending partition groups into
    per-sender slots and compacting them in slot order gives the owner exactly the concatenation, in
    peer-rank order, that the packed all-to-all-v delivers -- the merge order SPEC A4 fixes."""
    rng = np.random.default_rng(5)
    S, stride = 4, 50
    sent = []                                    # sent[src][dst] = keys src sends to dst
    for src in range(S):
        sizes = rng.integers(0, 12, size=S)
        sent.append([rng.integers(0, 1 << 30, size=int(n), dtype=np.int64) for n in sizes])
    for dst in range(S):
        slots = np.full(S * stride, -1, dtype=np.int64)
        counts = np.zeros(S, dtype=np.int32)
        for src in range(S):                     # edl_slot_send on every sender
            k = sent[src][dst]
            slots[src * stride:src * stride + k.size] = k
            counts[src] = k.size
        start = np.concatenate([[0], np.cumsum(counts)])   # edl_slot_compact on the owner
        keys_c = np.concatenate([slots[s * stride:s * stride + counts[s]] for s in range(S)])
        map_c = np.concatenate([s * stride + np.arange(counts[s]) for s in range(S)]).astype(np.int64)
        packed = np.concatenate([sent[src][dst] for src in range(S)])
        assert np.array_equal(keys_c, packed) and start[-1] == packed.size
        assert np.array_equal(slots[map_c], packed)        # edl_remap_index: compact -> slot-space rows


def _c_host_cmd(out):
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    return root, ["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(root, "include"), "-I",
                  os.path.join(cuda, "include"), os.path.join(root, "examples", "c_host.c"), "-L",
                  os.path.join(root, "easydl_b200"), "-ledlsparse", "-L", os.path.join(cuda, "lib64"), "-lcudart",
                  "-lm", "-o", out]


def test_header_is_plain_c_and_c_host_links(tmp_path):
    """include/edlsparse.h compiles as C99 (no C++ in the boundary) and a C program using it links
    against libedlsparse.so without anything from Python or torch."""
    import os, subprocess
    root, cmd = _c_host_cmd(str(tmp_path / "c_host"))
    r = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c",
                        os.path.join(root, "include", "edlsparse.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert os.path.exists(tmp_path / "c_host")


def test_ctypes_prototypes_match_the_header_types():
    """Every parameter and return type of _lib.PROTOTYPES has the kind (pointer / integer / float) and the
    width of the C declaration in include/edlsparse.h: a drifted binding would pass garbage to a kernel."""
    import ctypes as C
    from easydl_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "edlsparse.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    decls = re.findall(r"([A-Za-z_][\w\s\*]*?)\b(edl_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", hdr)
    scalars = {"int": ("int", C.sizeof(C.c_int)), "int32_t": ("int", 4), "int64_t": ("int", 8), "uint64_t": ("int", 8),
               "size_t": ("int", C.sizeof(C.c_size_t)), "float": ("float", 4), "edl_stream_t": ("ptr", 8)}

    def c_kind(t):
        t = re.sub(r"\b(const|unsigned)\b", "", t).strip()
        return ("ptr", 8) if "*" in t else scalars[t]

    def py_kind(a):
        if a is None:
            return None
        if a in (C.c_void_p, C.c_char_p) or issubclass(a, C._Pointer):
            return ("ptr", 8)
        return ("float" if a in (C.c_float, C.c_double) else "int", C.sizeof(a))

    assert len(decls) == len(_lib.PROTOTYPES)
    for ret, name, args in decls:
        res, argtypes = _lib.PROTOTYPES[name]
        want = [c_kind(re.match(r"(.*?)\w+$", x.strip()).group(1)) for x in args.split(",") if x.strip() not in ("", "void")]
        assert [py_kind(x) for x in argtypes] == want, name
        assert py_kind(res) == c_kind(ret), name
    assert C.sizeof(_lib.OptParams) == 32       # edl_opt_params: int + 4 floats, padded, then a pointer


def _touch_version(d, v, shards, S):
    import os
    os.makedirs(os.path.join(d, f"version-{v}"), exist_ok=True)
    for s in shards:
        open(os.path.join(d, f"version-{v}", f"variables-{s}-of-{S}.ckpt"), "wb").close()


def test_checkpoint_prune_keeps_a_complete_version(tmp_path):
    """Shard 0 prunes right after its own write: an incomplete newer version (another shard has not
    written yet, or died) must not cost the last complete one (checkpoint.py)."""
    from easydl_b200 import checkpoint as ck
    d = str(tmp_path)
    _touch_version(d, 10, range(4), 4)          # complete
    _touch_version(d, 20, range(4), 4)          # complete
    _touch_version(d, 30, [0], 4)               # shard 0 only: in progress / crashed
    saver = ck.CheckpointSaver(d, keep_max=1)
    saver.prune()
    assert ck.list_versions(d) == [20, 30]      # 10 is behind the newest complete one; 20 survives
    assert ck.latest_complete_version(d) == 20
    _touch_version(d, 30, [1, 2, 3], 4)         # the other shards finish
    saver.prune()
    assert ck.list_versions(d) == [30]
    assert ck.latest_complete_version(d) == 30


def test_restore_skips_incomplete_versions(tmp_path):
    from easydl_b200 import checkpoint as ck
    import pytest as _pt
    d = str(tmp_path)
    _touch_version(d, 5, [0, 2], 3)
    with _pt.raises(FileNotFoundError, match="no complete checkpoint"):
        ck.latest_complete_version(d)
    _touch_version(d, 3, range(3), 3)
    assert ck.latest_complete_version(d) == 3


def test_unique_launch_count_follows_the_dedup_path():
    """K.unique_launches mirrors sort.cu's choice of path (bench.py's gpu_launches is built from it)."""
    from easydl_b200 import _lib
    from easydl_b200 import kernels as K
    saved = _lib.tuning_get("dedup_impl")
    try:
        _lib.tuning_set("dedup_impl", 2)
        assert K.unique_launches(65536 * 26, 26, 28, 24) == 2          # transpose + cluster sort (writes everything)
        assert K.unique_launches(65536 * 26, 26, 28, 24, want_counts=True) == 3
        assert K.unique_launches(65537 * 26, 26, 28, 24) == 3 * 3 + 3  # field too long: global passes, one saved
        assert K.unique_launches(1000, 0, 28, n_dev=True) == 4 * 3 + 3
        _lib.tuning_set("dedup_impl", 1)
        assert K.unique_launches(65536 * 26, 26, 28, 24) == 4          # + scan + unique write, head count fused
        assert K.unique_launches(1000 * 26, 26, 28, 24) == 5
        _lib.tuning_set("dedup_impl", 0)
        assert K.unique_launches(65536 * 26, 26, 28, 24) == 12
        assert K.unique_launches(32768 * 26, 26, 30, 26) == 15         # no pass saved: flat sort
    finally:
        _lib.tuning_set("dedup_impl", saved)


def test_reference_arm_prints_the_contract_line():
    """bench.py --impl reference on a small workload (no GPU): one JSON line with the contract's keys, the compiled
    CPU path as the primary leg and the NumPy port beside it, both counting the same unique ids per step."""
    import json, subprocess, sys
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--rows", "20000",
                        "--batch", "512", "--dim", "16", "--steps", "3", "--warmup", "1", "--n-batches", "2"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "impl", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
              "scaling", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in line, k
    assert line["impl"] == "reference" and line["steps"] == 3 and line["warmup"] == 1 and line["gpu_launches"] == 0
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "compiled C" in line["cpu_baseline"]["sample"] and line["cpu_baseline"]["value"] == line["value"]
    assert "NumPy" in line["cpu_baseline_numpy"]["sample"] and line["cpu_baseline_numpy"]["value"] > 0
    # the same batches, so the same unique ids per step as the NumPy oracle counts
    rows = [20000] * 26
    u = [len(np.unique(O.fuse_keys(O.criteo_batch(rows, 512, 1.05, step=i)[1], O.key_bases(rows)))) for i in range(2)]
    assert line["config"]["unique_per_step"] == round(float(np.mean([u[(1 + i) % 2] for i in range(3)])), 1)
