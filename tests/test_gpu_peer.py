"""The NVLink peer-write exchange pieces that can be exercised on ONE GPU (the destination
table of the fused reduce -> send, the gather into a foreign buffer, the fence kernel with a world
of one), plus the real multi-rank run when the box has at least two GPUs."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import edl_oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def dev(a, cuda):
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


@pytest.mark.parametrize("wire", [torch.float32, torch.float64])
def test_segreduce_dst_table_equals_plain_scatter(cuda, wire):
    """dst_table: output row pos of group s lands at base[s] + (pos - offsets[s]) rows -- here the
    bases point into three separate local buffers standing in for three peers."""
    from easydl_b200 import kernels as K
    rng = np.random.default_rng(3)
    n, dim = 20000, 32
    keys = rng.integers(0, 3000, size=n, dtype=np.int64)
    grad = (rng.standard_normal((n, dim)) * 1e-2).astype(np.float32)
    r = K.unique(dev(keys, cuda), None, 12)
    U = r.count()
    part = K.partition_mod(r.uniq, None, 3, n_dev=r.n_unique)
    offs = [int(x) for x in part.offsets.cpu()]
    plain = K.segreduce(dev(grad, cuda), r.order, r.seg_offsets, r.n_unique, n, seg_first=r.seg_first,
                        out_pos=part.pos, out_dtype=wire)[:U]
    bufs = [torch.zeros((offs[s + 1] - offs[s] + 5, dim), dtype=wire, device=cuda) for s in range(3)]
    shift = [2, 0, 5]                                   # rows of "lower ranks" already in each buffer
    esz = 8 if wire == torch.float64 else 4
    table = [3, 0] + offs + [bufs[s].data_ptr() + shift[s] * dim * esz for s in range(3)] + [0, 0, 0]
    K.segreduce(dev(grad, cuda), r.order, r.seg_offsets, r.n_unique, n, seg_first=r.seg_first,
                out_pos=part.pos, out=None, out_dtype=wire,
                dst_table=torch.tensor(table, dtype=torch.int64, device=cuda))
    for s in range(3):
        cnt = offs[s + 1] - offs[s]
        got = bufs[s][shift[s]:shift[s] + cnt]
        assert torch.equal(got, plain[offs[s]:offs[s + 1]])
    _, _, _, order, seg = O.dedup(keys)
    want = O.segment_sum_f64(grad, order, seg)[part.perm[:U].cpu().numpy()]
    np.testing.assert_allclose(plain.float().cpu().numpy(), want, rtol=1e-6, atol=1e-9)


def test_peer_alloc_copy_and_fence_world_of_one(cuda):
    from easydl_b200 import _lib
    from easydl_b200.exchange import _RawDeviceMemory
    lib = _lib.load()
    p = C.c_void_p()
    assert lib.edl_peer_alloc(C.byref(p), 4096) == 0
    handle = C.create_string_buffer(64)
    assert lib.edl_ipc_export(p.value, handle) == 0 and any(handle.raw)
    buf = torch.as_tensor(_RawDeviceMemory(p.value, 4096), device=cuda)
    assert int(buf.sum()) == 0                                   # edl_peer_alloc clears
    src = torch.arange(256, dtype=torch.uint8, device=cuda)
    st = torch.cuda.current_stream().cuda_stream
    assert lib.edl_peer_copy(p.value + 1024, src.data_ptr(), 256, st) == 0
    flags = buf[2048:2048 + 16].view(torch.int64)               # [world + 1] = 2 slots
    ptrs = torch.tensor([p.value + 2048], dtype=torch.int64, device=cuda)
    for round_no in (1, 2, 3):
        assert lib.edl_peer_signal_wait(ptrs.data_ptr(), flags.data_ptr(), 0, 1, round_no, st) == 0
    torch.cuda.synchronize()
    assert torch.equal(buf[1024:1280], src)
    assert flags.cpu().tolist() == [3, 0]                        # no timeout marker
    assert lib.edl_peer_signal_wait(ptrs.data_ptr(), flags.data_ptr(), 3, 1, 1, st) == -1
    assert lib.edl_peer_free(p.value) == 0


@pytest.mark.parametrize("mode", ["peer", "nccl", "slots", "slots-pipe"])
def test_real_ranks_match_oracle(cuda, mode):
    """tools/parity.py on 2 (or 4) real GPUs: Adagrad and Adam against the oracle, graph replay == eager."""
    import json
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs (run with gpurun --gpus 2)")
    world = 2 if n < 4 else 4
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29577",
                        os.path.join(ROOT, "tools", "mgpu_check.py"), mode],
                       capture_output=True, text=True, timeout=400, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert res["ok"] and res["max_rel_err"] <= 1e-6


@pytest.mark.parametrize("opt", ["adagrad", "adam"])
def test_elastic_reshard_on_real_ranks(cuda, opt):
    """configs[3] on real ranks: S0 -> S1 -> S2 over NCCL with idle ranks (tools/mgpu_elastic.py)."""
    import json
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs (run with gpurun --gpus 2)")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29578",
                        os.path.join(ROOT, "tools", "mgpu_elastic.py"), opt],
                       capture_output=True, text=True, timeout=400, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert res["ok"]
