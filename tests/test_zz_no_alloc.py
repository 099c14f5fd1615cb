"""edl_unique_static: the dedup for callers that cannot allocate device memory.

Some users of libedlsparse cannot call cudaMalloc: a parameter-server process that has
already filled the GPU with its tables, a C host embedded in a framework whose allocator owns
all device memory, or a serving thread that cannot afford the device-wide synchronisation an
allocation implies.  For them the library keeps a fixed scratch array inside its own image.

This file is named test_zz_* so pytest collects it last: if it fails under `-x`, every other
test has already run.  It calls the entry point through the C ABI (ctypes, raw device
pointers) and compares every output with edl_unique's, bit for bit (integer work), on two
calls in a row -- the second call finds the first call's leftovers in the shared scratch and
must not be affected by them.  It passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_static(lib, ids, field_base, key_bits, field_bits=0):
    """One edl_unique_static call through ctypes.  Outputs are allocated by the caller (this
    test) up front; the entry point itself takes no workspace and allocates nothing."""
    dev = ids.device
    n = ids.numel()
    out = dict(
        uniq=torch.full((max(n, 1),), -1, dtype=torch.int64, device=dev),
        inverse=torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev),
        counts=torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev),
        order=torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev),
        seg=torch.full((n + 1,), -1, dtype=torch.int32, device=dev),
        first=torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev),
        nu=torch.full((1,), -1, dtype=torch.int32, device=dev),
    )
    fb_ptr = None if field_base is None else field_base.data_ptr()
    nf = 0 if field_base is None else field_base.numel()
    rc = lib.edl_unique_static(
        ids.data_ptr() if n else None, n, None, fb_ptr, nf, key_bits, field_bits,
        out["uniq"].data_ptr(), out["inverse"].data_ptr(), out["counts"].data_ptr(),
        out["order"].data_ptr(), out["seg"].data_ptr(), out["first"].data_ptr(), out["nu"].data_ptr(),
        torch.cuda.current_stream().cuda_stream)
    return rc, out


def _compare(out, ref, n):
    U = int(out["nu"].item())
    assert U == ref.count()
    assert torch.equal(out["uniq"][:U], ref.uniq[:U])
    assert torch.equal(out["inverse"][:n], ref.inverse[:n])
    assert torch.equal(out["counts"][:U], ref.counts[:U])
    assert torch.equal(out["order"][:n], ref.order[:n])
    assert torch.equal(out["seg"][:U + 1], ref.seg_offsets[:U + 1])
    assert torch.equal(out["first"][:U], ref.seg_first[:U])


def test_static_scratch_unique_matches_edl_unique_twice(cuda):
    from easydl_b200 import _lib, kernels as K
    lib = _lib.load()
    rng = np.random.default_rng(123)
    # call 1: a full-size batch for the built-in scratch (4096 samples x 26 fields, 28-bit keys)
    rows = 10_000_000
    ids1 = torch.from_numpy(rng.integers(0, rows, size=(4096, 26), dtype=np.int64)).to(cuda)
    fb = torch.arange(26, dtype=torch.int64, device=cuda) * rows
    # call 2: fewer, heavily repeated, wider keys -- different scratch carving than call 1
    ids2 = torch.from_numpy(rng.integers(0, 1 << 40, size=300, dtype=np.int64)
                            [rng.integers(0, 300, size=50_001)]).to(cuda)
    ref1 = K.unique(ids1.view(-1), fb, 28)
    ref2 = K.unique(ids2, None, 41)
    torch.cuda.synchronize()

    rc, out1 = _call_static(lib, ids1.view(-1), fb, 28)
    assert rc == 0, lib.edl_last_error_string()
    rc, out2 = _call_static(lib, ids2, None, 41)      # straight after, same stream, same scratch
    assert rc == 0, lib.edl_last_error_string()
    torch.cuda.synchronize()
    _compare(out1, ref1, ids1.numel())
    _compare(out2, ref2, ids2.numel())

    # and once more with the first input: leftovers of call 2 must not matter either
    rc, out3 = _call_static(lib, ids1.view(-1), fb, 28)
    assert rc == 0, lib.edl_last_error_string()
    _compare(out3, ref1, ids1.numel())


def test_static_scratch_rejects_oversize_and_handles_empty(cuda):
    from easydl_b200 import _lib
    lib = _lib.load()
    too_many = torch.zeros(131072 + 1, dtype=torch.int64, device=cuda)
    rc, _ = _call_static(lib, too_many, None, 16)
    assert rc == -4  # EDL_ELIMIT
    msg = lib.edl_last_error_string().decode()
    assert "131072" in msg and "edl_unique" in msg
    at_limit = torch.arange(131072, dtype=torch.int64, device=cuda) % 1000
    rc, out = _call_static(lib, at_limit, None, 64)   # widest keys at the limit still fit
    assert rc == 0, lib.edl_last_error_string()
    assert int(out["nu"].item()) == 1000
    rc, out = _call_static(lib, torch.empty(0, dtype=torch.int64, device=cuda), None, 16)
    assert rc == 0 and int(out["nu"].item()) == 0 and int(out["seg"][0].item()) == 0


def test_static_scratch_field_major_against_the_oracle_and_allocates_nothing(cuda):
    """The field-major sort (field_bits > 0: every field sorted on its own 24 bits, padded to whole
    tiles) carves the fixed array differently and needs more of it.  Checked against the NumPy oracle
    directly, twice in a row, with the free device memory read before and after: the call itself must
    not take a byte (a caller whose tables fill the GPU has none to give)."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import edl_oracle as O
    from easydl_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(77)
    rows, B, F = 10_000_000, 4096, 26
    fb = torch.arange(F, dtype=torch.int64, device=cuda) * rows
    batches = [rng.integers(0, rows, size=(B, F), dtype=np.int64) // d for d in (1, 1000)]   # second: many repeats
    rc, _ = _call_static(lib, torch.from_numpy(batches[0]).to(cuda).view(-1), fb, 28, 24)      # first call: set-up
    assert rc == 0, lib.edl_last_error_string()
    torch.cuda.synchronize()
    for ids_np in batches:
        ids = torch.from_numpy(ids_np).to(cuda).view(-1)
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        # outputs are the caller's: allocate them before the measurement window
        n = ids.numel()
        bufs = [torch.empty(n, dtype=torch.int64, device=cuda)] + [torch.empty(n + 1, dtype=torch.int32, device=cuda) for _ in range(5)]
        nu = torch.zeros(1, dtype=torch.int32, device=cuda)
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        rc = lib.edl_unique_static(ids.data_ptr(), n, None, fb.data_ptr(), F, 28, 24, bufs[0].data_ptr(),
                                   bufs[1].data_ptr(), bufs[2].data_ptr(), bufs[3].data_ptr(), bufs[4].data_ptr(),
                                   bufs[5].data_ptr(), nu.data_ptr(), torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.edl_last_error_string()
        torch.cuda.synchronize()
        free2, _ = torch.cuda.mem_get_info()
        assert free2 == free1, f"edl_unique_static changed free device memory by {free1 - free2} bytes"
        keys = O.fuse_keys(ids_np, np.arange(F + 1, dtype=np.int64) * rows)
        uniq, inverse, counts, order, seg = O.dedup(keys.reshape(-1))
        U = int(nu.item())
        assert U == len(uniq)
        assert np.array_equal(bufs[0][:U].cpu().numpy(), uniq)
        assert np.array_equal(bufs[1][:n].cpu().numpy(), inverse)
        assert np.array_equal(bufs[2][:U].cpu().numpy(), counts)
        assert np.array_equal(bufs[3][:n].cpu().numpy(), order)
        assert np.array_equal(bufs[4][:U + 1].cpu().numpy(), seg)
        assert np.array_equal(bufs[5][:U].cpu().numpy(), order[seg[:-1]])
        del free0


def test_static_scratch_field_major_carve_over_the_limit_is_refused(cuda):
    """n is within the id limit, but the padded field-major carve at 64-bit keys needs 3.8 MB of the
    3 MiB array: the call must say so (EDL_ELIMIT) before any launch."""
    from easydl_b200 import _lib
    lib = _lib.load()
    B, F = 5000, 26
    ids = torch.zeros(B * F, dtype=torch.int64, device=cuda)
    fb = torch.arange(F, dtype=torch.int64, device=cuda) * (1 << 40)
    rc, out = _call_static(lib, ids, fb, 64, 24)
    assert rc == -4, lib.edl_last_error_string()
    assert "scratch bytes" in lib.edl_last_error_string().decode()
    assert int(out["nu"].item()) == -1      # nothing was launched: the output still holds the fill value
