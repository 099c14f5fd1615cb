"""ctypes binding of libedlsparse.so (include/edlsparse.h).

The product path has no CPU fallback: if the shared library is missing, or a call returns a
negative code, this module raises.  Build with `python -c "import __graft_entry__ as g;
g.build()"` or `make -C easydl_b200/csrc`.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libedlsparse.so")

EDL_OPT_ADAGRAD = 0
EDL_OPT_ADAM = 1


class EdlError(RuntimeError):
    """A libedlsparse call returned a negative EDL_E* code."""


class OptParams(C.Structure):
    _fields_ = [("kind", C.c_int), ("lr", C.c_float), ("eps", C.c_float),
                ("beta1", C.c_float), ("beta2", C.c_float), ("lr_dev", C.c_void_p)]


_p = C.c_void_p
_i64 = C.c_int64
_sz = C.c_size_t
_int = C.c_int

# name -> (restype, argtypes); mirrors include/edlsparse.h one to one
PROTOTYPES = {
    "edl_version": (_int, []),
    "edl_last_error_string": (C.c_char_p, []),
    "edl_tuning_set": (_int, [C.c_char_p, _int]),
    "edl_tuning_get": (_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "edl_init_rows": (_int, [_p, _int, _i64, _p, _i64, _i64, C.c_uint64, _p]),
    "edl_fill_f32": (_int, [_p, _i64, C.c_float, _p]),
    "edl_fill_rows_f32": (_int, [_p, _int, _i64, _i64, C.c_float, _p]),
    "edl_unique_workspace_bytes": (_sz, [_i64, _int, _int, _int]),
    "edl_unique": (_int, [_p, _i64, _p, _p, _int, _int, _int, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "edl_unique_static": (_int, [_p, _i64, _p, _p, _int, _int, _int, _p, _p, _p, _p, _p, _p, _p, _p]),
    "edl_partition_mod_workspace_bytes": (_sz, [_i64, _int]),
    "edl_partition_mod": (_int, [_p, _i64, _p, _p, _int, _p, _p, _p, _p, _p, _sz, _p]),
    "edl_gather_keys": (_int, [_p, _int, _i64, _p, _i64, _p, _p, _p, _p]),
    "edl_gather_ids": (_int, [_p, _int, _i64, _p, _i64, _p, _int, _p, _p]),
    "edl_expand": (_int, [_p, _int, _p, _i64, _p, _p]),
    "edl_compose_index": (_int, [_p, _p, _i64, _p, _p]),
    "edl_segreduce_workspace_bytes": (_sz, [_i64, _i64, _int]),
    "edl_segreduce": (_int, [_p, _int, _p, _p, _p, _p, _i64, _i64, _p, _p, _int, _p, _p, _sz, _p]),
    "edl_segreduce_update": (_int, [_p, _p, _p, _int, _i64, _p, _p, _p, _p, _p, _i64, _i64, _p, _int, _p, _p,
                                    C.POINTER(OptParams), _p, _p, _sz, _p]),
    "edl_selftest_rounding": (_int, [_i64, C.c_uint64, _p, _p]),
    "edl_export_workspace_bytes": (_sz, [_i64]),
    "edl_export_count": (_int, [_p, _i64, _p, _p, _sz, _p]),
    "edl_export_shard": (_int, [_p, _p, _p, _int, _i64, _p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "edl_import_shard": (_int, [_p, _p, _p, _int, _i64, _p, _p, _p, _p, _p, _p, _i64, _p]),
    "edl_scatter_rows": (_int, [_p, _int, _p, _i64, _p, _p, _i64, _p]),
    "edl_lazy_index_bytes": (_sz, [_i64]),
    "edl_lazy_clear": (_int, [_p, _i64, _p]),
    "edl_lazy_gather": (_int, [_p, _i64, _p, _int, _i64, _p, _i64, _p, _p, C.c_uint64, _p, _p]),
    "edl_lazy_insert": (_int, [_p, _i64, _p, _p, _p, _int, _i64, _i64, _p, _i64, _p, _p, C.c_uint64,
                               C.c_float, C.c_float, _p, _p]),
    "edl_lazy_stamp": (_int, [_p, _p, _i64, _p, C.c_int32, _p]),
    "edl_lazy_status_ptr": (_p, [_p, _i64]),
    "edl_lazy_dump": (_int, [_p, _i64, _p, _p, _p]),
    "edl_peer_alloc": (_int, [C.POINTER(C.c_void_p), _sz]),
    "edl_peer_free": (_int, [_p]),
    "edl_ipc_export": (_int, [_p, C.c_char_p]),
    "edl_ipc_import": (_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "edl_ipc_close": (_int, [_p]),
    "edl_peer_copy": (_int, [_p, _p, _sz, _p]),
    "edl_peer_signal_wait": (_int, [_p, _p, _int, _int, _i64, _p]),
    "edl_peer_signal_wait_counter": (_int, [_p, _p, _int, _int, _p, _p]),
    "edl_slot_send": (_int, [_p, _int, _p, _int, _p, _i64, _p, _p, _i64, _p, _p, _p]),
    "edl_slot_positions": (_int, [_p, _i64, _p, _p, _int, _i64, _p, _p]),
    "edl_slot_compact": (_int, [_p, _p, _int, _i64, _i64, _p, _p, _p, _p]),
    "edl_remap_index": (_int, [_p, _p, _i64, _p, _p, _p]),
    "edl_slot_merge_unique_workspace_bytes": (_sz, [_int, _i64]),
    "edl_slot_merge_unique": (_int, [_p, _p, _int, _i64, _i64, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "edl_expand_window": (_int, [_p, _int, _p, _i64, _int, _int, _int, _p, _p]),
}

_lib = None


def load() -> C.CDLL:
    """Load the library once; raise loudly when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EdlError(
            f"{LIB_PATH} is missing: the CUDA extension is not built and there is no CPU "
            "fallback. Run `make -C easydl_b200/csrc` (nvcc, sm_100a).")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def tuning_set(name: str, value: int):
    """Set a launch-shape knob of the library (include/edlsparse.h lists them)."""
    check(load().edl_tuning_set(name.encode(), int(value)), "edl_tuning_set")


def tuning_get(name: str) -> int:
    v = C.c_int(0)
    check(load().edl_tuning_get(name.encode(), C.byref(v)), "edl_tuning_get")
    return int(v.value)


def check(rc: int, what: str):
    if rc != 0:
        msg = load().edl_last_error_string().decode("utf-8", "replace")
        raise EdlError(f"{what} failed with code {rc}: {msg}")
