#!/usr/bin/env python
"""bench.py -- Criteo-shape sparse step samples/s on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N ...            # the CPU parameter-server path

A "step" is one pass of the hot path over one synthetic Criteo-shape batch:
lookup (A3) -> [upstream gradient, synthetic, resident] -> dedup (A1) -> segmented reduce +
Adagrad row update (A4 + A5).  Default workload = BASELINE.json configs[1]: DeepFM-shape,
26 tables x 10M rows x dim 64, batch 64K, Adagrad, one B200.  For N > 1 the tables are
row-sharded over the N GPUs (id mod N) and every GPU brings its own 64K batch (weak scaling),
exchanging ids / rows / gradients through fixed per-pair slots in peer-mapped buffers over NVLink
(`--exchange slots`, the default; `peer` and `nccl` are the packed all-to-all-v forms).

One JSON line on stdout (rank 0).  `value` is device-resident throughput; `e2e` is the same
step through the host-buffer PS API (ids H2D, embeddings D2H, gradients H2D each step).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "criteo_sparse_step_samples_per_s"
UNIT = "samples/s"
N_FIELDS = 26


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000, help="rows per table (global)")
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--batch", type=int, default=65536, help="samples per GPU per step")
    ap.add_argument("--alpha", type=float, default=1.05)
    ap.add_argument("--optimizer", default="adagrad", choices=["adagrad", "adam"])
    ap.add_argument("--n-batches", type=int, default=8, help="distinct id batches cycled through")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--cpu-steps", type=int, default=3, help="timed oracle steps for cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-compiled-child", type=int, default=0, metavar="THREADS",
                    help="internal: time the compiled CPU path (oracle/edl_oracle_c.c) on THREADS threads and print one "
                         "JSON object; bench.py runs this in a child process so that a fault there cannot cost the line")
    ap.add_argument("--cpu-budget-s", type=float, default=60.0, help="time budget of one CPU baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-prefetch", action="store_true",
                    help="N=1: time only the strictly sequential step (no dedup of the next batch beside the update)")
    ap.add_argument("--no-parity", action="store_true", help="N>1: skip the oracle parity check before the warm-up")
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly instead of replaying CUDA graphs")
    ap.add_argument("--lazy-rows", type=int, default=0,
                    help="N>1, slot exchanges: shards that store only updated rows (LazyShardParameters) with room for "
                         "this many rows each -- for tables whose dense shards do not fit (BASELINE.json configs[2]: "
                         "26 x 40M x dim 128 with Adam state is 200 GB per GPU on eight GPUs)")
    ap.add_argument("--force-sharded", action="store_true",
                    help="N=1 only: run the multi-shard code path (dedup/partition/exchange/expand) with S=1")
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl", "slots", "slots-pipe"],
                    help="N>1: slots = fixed per-pair slots in peer-mapped buffers, no host round trip, step "
                         "replayed from CUDA graphs; peer = packed all-to-all-v by NVLink peer writes (CUDA "
                         "IPC, fused with the producing kernels); nccl = all-to-all-v by NCCL; auto = slots "
                         "(measured fastest at 2 and 8 GPUs)")
    ap.add_argument("--lanes", type=int, default=2, help="--exchange slots-pipe: number of table groups")
    ap.add_argument("--wire", default="auto", choices=["auto", "mixed", "f64", "f32"],
                    help="exchanged partial gradient sums (N>1).  mixed (default with the slot exchanges): an id with "
                         "one local gradient row travels as that fp32 row, the others as float64 partial sums -- bit "
                         "for bit the result of the all-float64 wire at ~60 %% of its bytes; f64: all float64 "
                         "(default of the peer / nccl exchanges); f32 halves the bytes but breaks the 1e-6 tolerance "
                         "for very hot ids and for Adam")
    return ap.parse_args()


def workload_name(a):
    return (f"deepfm-shape {N_FIELDS} tables x {a.rows} rows x dim {a.dim}, batch {a.batch}/GPU, "
            f"{a.optimizer}, zipf {a.alpha}")


def algorithmic_bytes(N, U, D, optimizer):
    """compulsory HBM traffic of one step (ids 8 B, inverse 4 B, fp32)."""
    lookup = 8 * N + 4 * D * U + 4 * D * N
    if optimizer == "adagrad":
        update = 4 * D * N + 4 * N + 16 * D * U
    else:
        update = 4 * D * N + 4 * N + 24 * D * U
    return lookup, update


def _single_share(step):
    """Share of this rank's unique ids (last step) that have exactly one gradient row (device scalar)."""
    import torch
    u = int(step.uq.n_unique.item())
    seg = step.uq.seg_offsets[:u + 1]
    return ((seg[1:] - seg[:-1]) == 1).sum().to(torch.float64) / max(u, 1)


NVLINK_GBS = 770.0   # measured peer copy per direction per GPU on this pool (B200_PROFILING.md); nominal 900


def sharded_roofline(a, world, N, U, owner_counts, ms_step, peak, peak_kind):
    """Per-rank byte counts of one sharded step (per GPU) against the two ceilings it can
    hit: bytes that must leave the GPU over NVLink, and bytes that must cross its own HBM.  frac = the
    larger of the two ideal times / the measured step time."""
    D, S = a.dim, world
    R, Uo, single_frac = owner_counts            # keys received per owner, unique among them, share of one-row ids
    wire = {"f64": 8.0, "f32": 4.0, "mixed": 4.0 * single_frac + 8.0 * (1.0 - single_frac)}[a.wire]
    out_frac = (S - 1) / S
    nv = out_frac * (8 * U + 4 * D * R + wire * D * U)
    hbm = (8 * N                                   # A: ids (the sort's working set stays in L2)
           + 4 * D * R + 4 * D * R / S             # B: owner gather reads, local share of the writes
           + 4 * D * U + 4 * D * N + 4 * N         # C: expand
           + 4 * D * N + 4 * N + wire * D * U / S  # C: reduce, local share of the partial sums
           + wire * D * R + (16 if a.optimizer == "adagrad" else 24) * D * Uo)   # D: owner reduce + update
    t_nv, t_hbm = nv / (NVLINK_GBS * 1e9), hbm / (peak * 1e9)
    bound = "nvlink" if t_nv >= t_hbm else "hbm"
    t_step = ms_step * 1e-3
    return {"kernel": "whole sharded step of one rank (slot exchange; NVLink-bound rounds and HBM-bound phases overlap "
                      "only partly, so the ideal is the larger of the two)",
            "bound": bound, "achieved": (nv if bound == "nvlink" else hbm) / t_step / 1e9,
            "peak": NVLINK_GBS if bound == "nvlink" else peak, "unit": "GB/s",
            "frac": max(t_nv, t_hbm) / t_step, "traffic": None, "peak_kind": peak_kind,
            "nvlink_out_bytes_per_rank": nv, "hbm_bytes_per_rank": hbm,
            "ideal_nvlink_ms": t_nv * 1e3, "ideal_hbm_ms": t_hbm * 1e3,
            "nvlink_peak_note": "770 GB/s = measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
            "keys_received_per_owner": R, "unique_keys_per_owner": Uo,
            "gradient_wire_bytes_per_element": wire, "single_row_id_share": single_frac}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while `active` is set."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.active, self._stop = [], set(), False, False
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)
        self.t = threading.Thread(target=self._run, daemon=True)
        if self.ok:
            self.t.start()

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop:
            if self.active:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, nm in names.items():
                        if r & bit:
                            self.reasons.add(nm)
                except Exception:
                    pass
            time.sleep(0.01)

    def result(self):
        self._stop = True
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------ CPU arm
def cpu_reference(a, threads: int, steps: int, warmup: int, nb_max: int = 2, state=None):
    """The CPU parameter-server path (oracle port) on the host cores.  Tables are lazily
    materialised; one untimed pass per distinct batch creates the rows so the timed steps
    measure lookup + reduce + update, not row initialisation."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import edl_oracle as O
    rows = [a.rows] * N_FIELDS
    opt = O.OptimizerSpec.adagrad() if a.optimizer == "adagrad" else O.OptimizerSpec.adam()
    shard = O.OracleShard(rows, a.dim, opt)
    nb = max(1, min(a.n_batches, nb_max))
    batches = [O.criteo_batch(rows, a.batch, a.alpha, step=i)[1] for i in range(nb)]
    if state is not None:      # unique ids per step over the steps that will be timed (as the GPU arm counts them)
        u = [len(np.unique(O.fuse_keys(b, shard.bases))) for b in batches]
        state["nb"] = nb
        state["U_of"] = lambda w, k: float(np.mean([u[(w + i) % nb] for i in range(k)]))
    grad = O.synthetic_grad(a.batch * N_FIELDS, a.dim)
    for ids in batches:  # materialise
        shard.materialise(np.unique(O.fuse_keys(ids, shard.bases)))
    pool = ThreadPoolExecutor(threads) if threads > 1 else None
    for i in range(warmup):
        O.sparse_step_per_table(shard, batches[i % nb], grad, pool)
    t0 = time.perf_counter()
    for i in range(steps):
        O.sparse_step_per_table(shard, batches[(warmup + i) % nb], grad, pool)
    dt = time.perf_counter() - t0
    return a.batch * steps / dt, dt / steps


def cpu_compiled_child(a):
    """Child process: the compiled CPU parameter-server path (oracle/edl_oracle_c.c: one table per thread, a hash
    map id -> row per table) on the same workload.  Honours --steps / --warmup up to --cpu-budget-s."""
    from oracle import edl_oracle as O
    from oracle import edl_oracle_c as OC
    threads = max(1, a.cpu_compiled_child)
    rows = [a.rows] * N_FIELDS
    opt = O.OptimizerSpec.adagrad() if a.optimizer == "adagrad" else O.OptimizerSpec.adam()
    shard = OC.CShard(rows, a.dim, opt)
    nb = max(1, a.n_batches)
    batches = [O.criteo_batch(rows, a.batch, a.alpha, step=i)[1] for i in range(nb)]
    grad = O.synthetic_grad(a.batch * N_FIELDS, a.dim)
    emb = None
    t0 = time.perf_counter()
    for ids in batches:                     # untimed: materialises the rows of every batch
        shard.sparse_step(ids, grad, threads=threads, want_emb=False)
    probe = (time.perf_counter() - t0) / nb
    steps = max(1, min(a.steps, int(a.cpu_budget_s / max(probe, 1e-3))))
    warm = max(0, min(a.warmup, int(0.2 * a.cpu_budget_s / max(probe, 1e-3))))
    for i in range(warm):
        shard.sparse_step(batches[i % nb], grad, threads=threads)
    n_u = []
    t0 = time.perf_counter()
    for i in range(steps):
        emb, u = shard.sparse_step(batches[(warm + i) % nb], grad, threads=threads)
        n_u.append(u)
    dt = time.perf_counter() - t0
    del emb
    print(json.dumps({"value": a.batch * steps / dt, "ms_per_step": dt / steps * 1e3, "steps": steps, "warmup": warm,
                      "threads": threads, "nb": nb, "unique_per_step": float(np.mean(n_u))}), flush=True)


def cpu_compiled(a, threads, steps, warmup, budget_s, timeout_s=420):
    """Run cpu_compiled_child in a child process -> its JSON object, or None (library not built, fault, time-out)."""
    import subprocess
    if not os.path.exists(os.path.join(ROOT, "oracle", "libedloracle.so")):
        return None
    cmd = [sys.executable, os.path.abspath(__file__), "--cpu-compiled-child", str(threads), "--steps", str(steps),
           "--warmup", str(warmup), "--rows", str(a.rows), "--dim", str(a.dim), "--batch", str(a.batch),
           "--alpha", str(a.alpha), "--optimizer", a.optimizer, "--n-batches", str(a.n_batches),
           "--cpu-budget-s", str(budget_s)]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout_s)
        if r.returncode != 0:
            return None
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        return None


def config_for(a, world, U_mean=None, nb=None):
    """The `config` object of the JSON line: the WORKLOAD only, so that both arms print the same object
    (how an arm runs it goes into `implementation`)."""
    F, D, B = N_FIELDS, a.dim, a.batch
    N = B * F
    table_gb = F * a.rows * D * 4 / 1e9
    slots = 1 if a.optimizer == "adagrad" else 2
    return {"workload": workload_name(a), "global_batch": B * world, "tables": F, "rows_per_table": a.rows,
            "dim": D, "zipf_alpha": a.alpha, "optimizer": a.optimizer,
            "unique_per_step": None if U_mean is None else round(float(U_mean), 1),
            "sharding": f"row-wise id mod {world}",
            "cache": f"inputs larger than L2: {table_gb * (1 + slots):.0f} GB of table + optimizer rows touched at "
                     f"random, {nb if nb else a.n_batches} id batches cycled, {N * D * 4 / 1e6:.0f} MB gradient + "
                     f"{N * D * 4 / 1e6:.0f} MB output per GPU per step"}


def run_reference_arm(a):
    """The CPU parameter-server path on the host cores, all threads.  Primary leg: the compiled restatement
    (oracle/edl_oracle_c.c, in a child process); the NumPy oracle port is timed beside it on a short sample and is
    the primary leg when the C library is not built or its child fails."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    threads = min(cores, N_FIELDS)
    c = cpu_compiled(a, threads, a.steps, a.warmup, budget_s=90.0)
    # NumPy port: one probe step sizes the run (30 s beside the compiled leg, 150 s when it is the only one)
    probe_v, probe_spp = cpu_reference(a, threads, 1, 1)
    budget_s = 30.0 if c else 150.0
    steps = max(1, min(a.steps, int(budget_s / max(probe_spp, 1e-3))))
    warm = max(0, min(a.warmup, int(0.2 * budget_s / max(probe_spp, 1e-3))))
    st = {}
    v, spp = cpu_reference(a, threads, steps, warm, nb_max=a.n_batches, state=st)
    numpy_leg = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "ms_per_step": spp * 1e3,
                 "sample": f"{steps} full steps (batch {a.batch}, {N_FIELDS} tables, dim {a.dim}) of the NumPy oracle, "
                           f"one table per thread on {threads} threads, rows pre-materialised"}
    if c:
        value, ms, k_steps, k_warm = c["value"], c["ms_per_step"], c["steps"], c["warmup"]
        cfg = config_for(a, a.gpus, c["unique_per_step"], c["nb"])
        primary = {"value": value, "unit": UNIT, "cores": c["threads"], "kind": "port", "ms_per_step": ms,
                   "sample": f"{k_steps} full steps (batch {a.batch}, {N_FIELDS} tables, dim {a.dim}) of the compiled C "
                             f"restatement (oracle/edl_oracle_c.c: hash map id -> row per table, gcc -O3, SSE2 / AVX2), "
                             f"one table per thread on {c['threads']} threads, rows pre-materialised"}
        path = "CPU parameter-server path (compiled C port of the oracle), one host, all cores; the GPUs are unused"
    else:
        value, ms, k_steps, k_warm = v, spp * 1e3, steps, warm
        cfg = config_for(a, a.gpus, st["U_of"](warm, steps), st["nb"])
        primary = numpy_leg
        path = "CPU parameter-server path (NumPy oracle port), one host, all cores; the GPUs are unused"
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": a.gpus,
        "steps": k_steps, "warmup": k_warm, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "implementation": {"code_path": path},
        "cpu_baseline": primary,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if c:
        line["cpu_baseline_numpy"] = numpy_leg
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ GPU arm
def run_b200_arm(a):
    import torch
    import torch.distributed as dist

    from easydl_b200 import _lib
    from easydl_b200.criteo import CriteoSynthetic
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters

    _lib.load()  # no fallback: raises if the CUDA extension is missing
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    rows = [a.rows] * N_FIELDS
    opt = OptimizerSpec.adagrad() if a.optimizer == "adagrad" else OptimizerSpec.adam()
    D, B, F = a.dim, a.batch, N_FIELDS
    N = B * F
    if a.exchange == "auto":
        a.exchange = "slots" if world > 1 else "peer"
    if a.wire == "auto":
        a.wire = "mixed" if (world > 1 and a.exchange in ("slots", "slots-pipe")) else "f64"
    if a.wire == "mixed" and not (world > 1 and a.exchange in ("slots", "slots-pipe")):
        raise SystemExit("--wire mixed is implemented by the slot exchanges (N > 1)")
    wire_dtype = {"f64": torch.float64, "f32": torch.float32, "mixed": "mixed"}[a.wire]

    # ---------------- real-rank parity against the oracle, before anything is timed (N > 1)
    parity = None
    if world > 1 and not a.no_parity:
        from tools.parity import check_exchange
        factory = None
        if a.lazy_rows:
            from easydl_b200.lazy_parameters import LazyShardParameters
            from tools.parity import ROWS as P_ROWS, DIM as P_DIM
            factory = lambda po: LazyShardParameters(P_ROWS, P_DIM, po, world, rank, max_rows=1 << 17, device=dev)
        parity = check_exchange(a.exchange, dev, dist.group.WORLD, wire_dtype, a.lanes, params_factory=factory)
        parity["wire"] = a.wire
        parity["shards"] = "lazy" if a.lazy_rows else "dense"
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "error": "parity check against the oracle failed",
                                  "parity": parity}), flush=True)
            dist.destroy_process_group()
            raise SystemExit(3)

    if a.lazy_rows:
        if not (world > 1 and a.exchange in ("slots", "slots-pipe")):
            raise SystemExit("--lazy-rows runs on the slot exchanges (N > 1)")
        from easydl_b200.lazy_parameters import LazyShardParameters
        params = LazyShardParameters(rows, D, opt, world, rank, max_rows=a.lazy_rows, device=dev)
    else:
        params = ShardParameters(rows, D, opt, world, rank, device=dev)

    sharded = world > 1 or a.force_sharded
    if not sharded:
        from easydl_b200.step import SparseStep
        step = SparseStep(params, B)
    else:
        from easydl_b200.sharded_step import ShardedSparseStep
        from easydl_b200.exchange import PeerExchange, SingleExchange, TorchExchange
        slots = world > 1 and a.exchange in ("slots", "slots-pipe")
        if slots:
            ex = None
        elif world == 1:
            ex = SingleExchange()
        elif a.exchange == "peer":
            # receive buffers sized for 1.5x the balanced share of unique ids (hash-mod spreads evenly)
            ex = PeerExchange(dist.group.WORLD, dev, int(1.5 * N), D, wire_dtype)
        else:
            ex = TorchExchange(dist.group.WORLD, dev)
        if slots:
            # fixed per-pair slots, no host round trip: the step is replayed from CUDA graphs below
            from easydl_b200.slot_step import PipelinedSlotStep, SlotShardedStep
            if a.exchange == "slots-pipe":     # table groups one stage apart (measured slower than `slots`)
                step = PipelinedSlotStep(params, B, group=dist.group.WORLD, wire_dtype=wire_dtype, n_lanes=a.lanes)
            else:
                step = SlotShardedStep(params, B, group=dist.group.WORLD, wire_dtype=wire_dtype)
        else:
            step = ShardedSparseStep(params, B, ex, wire_dtype=wire_dtype)

    # distinct batches per rank: step index = batch_no * world + rank
    src = CriteoSynthetic(rows, B, a.alpha)
    nb = max(1, a.n_batches)
    ids_host = [torch.from_numpy(src.ids(i * world + rank)).pin_memory() for i in range(nb)]
    grad_host = torch.from_numpy(src.grad(D, rank)).pin_memory()
    ids_dev = [t.to(dev) for t in ids_host]
    grad_dev = grad_host.to(dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank) if rank == 0 else None

    # ---------------- device-resident timed region
    # S == 1, Adagrad: the step is replayed from CUDA graphs (one per id batch buffer), which
    # removes the host launch gaps between its kernels.  Otherwise: eager launches.
    # unique ids per batch (deterministic per batch buffer): read once, outside the timed region
    U_batch = []
    for i in range(nb):
        if not sharded:
            U_batch.append(int(step.dedup(ids_dev[i]).n_unique.item()))
        else:
            step.run(ids_dev[i], grad_dev)
            U_batch.append(int(step.last_n_unique.item()))
    graphed = None
    if not sharded and not a.no_graph:
        from easydl_b200.step import GraphedStep
        graphed = [GraphedStep(step, ids_dev[i], grad_dev) for i in range(nb)]
    elif sharded and world > 1 and a.exchange in ("slots", "slots-pipe") and not a.no_graph:
        from easydl_b200.slot_step import GraphedSlotStep
        graphed = [GraphedSlotStep(step, ids_dev[i], grad_dev) for i in range(nb)]

    def one_step(i):
        if graphed is not None:
            graphed[i % nb].run()
        else:
            step.run(ids_dev[i % nb], grad_dev)

    for i in range(a.warmup):
        one_step(i)
    barrier()
    t_begin = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    if sampler:
        sampler.active = True
    t_begin.record()
    for i in range(a.steps):
        one_step(a.warmup + i)
    t_end.record()
    barrier()
    if sampler:
        sampler.active = False
    ms = t_begin.elapsed_time(t_end)
    if hasattr(step, "check"):
        step.check()            # slot exchange: raises if a slot overflowed or a fence timed out
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_max = float(t_ms.item())
    value = B * world * a.steps / (ms_max * 1e-3)

    # ---------------- the same K steps with the NEXT batch's dedup running beside this batch's lookup and
    # update (SparseStep.run_prefetch: the ids of batch i + 1 are known while batch i trains, as an input
    # pipeline knows them).  Same work per step, bit-identical results; this is the headline `value`, the
    # strictly sequential step above is reported beside it.
    sequential = None
    if sharded and world > 1 and a.exchange == "slots" and graphed is not None and nb % 2 == 0 and not a.no_prefetch:
        # sharded: phase A (dedup, partition, key send) of batch i + 1 beside the rounds of batch i
        from easydl_b200.slot_step import GraphedSlotRing
        sequential = {"value": value, "ms_per_step": ms_max / a.steps}
        ring = GraphedSlotRing(step, ids_dev, grad_dev)
        for i in range(max(a.warmup, nb)):
            ring.run()
        barrier()
        r0 = torch.cuda.Event(enable_timing=True)
        r1 = torch.cuda.Event(enable_timing=True)
        if sampler:
            sampler.active = True
        r0.record()
        for i in range(a.steps):
            ring.run()
        r1.record()
        barrier()
        if sampler:
            sampler.active = False
        step.check()
        t_ms = torch.tensor([r0.elapsed_time(r1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        ms_max = float(t_ms.item())
        value = B * world * a.steps / (ms_max * 1e-3)
    if not sharded and graphed is not None and nb % 2 == 0 and not a.no_prefetch:
        from easydl_b200.step import GraphedPrefetchRing
        sequential = {"value": value, "ms_per_step": ms_max / a.steps}
        ring = GraphedPrefetchRing(step, ids_dev, grad_dev)
        for i in range(max(a.warmup, nb)):
            ring.run()
        while ring.pos != (a.warmup % nb):      # start the timed steps at the same batch as the sequential run
            ring.run()
        barrier()
        r0 = torch.cuda.Event(enable_timing=True)
        r1 = torch.cuda.Event(enable_timing=True)
        if sampler:
            sampler.active = True
        r0.record()
        for i in range(a.steps):
            ring.run()
        r1.record()
        barrier()
        if sampler:
            sampler.active = False
        ms_max = r0.elapsed_time(r1)
        value = B * world * a.steps / (ms_max * 1e-3)
    U_mean = float(np.mean([U_batch[(a.warmup + i) % nb] for i in range(a.steps)]))

    # ---------------- per-phase timing (same K steps again, events between the phases)
    phases = None
    if not sharded:
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(a.steps)]
        if sampler:
            sampler.active = True
        p_begin = torch.cuda.Event(enable_timing=True)
        p_end = torch.cuda.Event(enable_timing=True)
        p_begin.record()
        for i in range(a.steps):
            j = (a.warmup + i) % nb
            if graphed is not None and params.rate_changes:
                params.stage_rate(params.step + 1)       # Adam: the step's rate, before the timed phases
            ev[i][0].record()
            if graphed is not None:
                graphed[j].g_lookup.replay()
                ev[i][1].record()
                graphed[j].g_dedup.replay()
                ev[i][2].record()
                graphed[j].g_update.replay()
                params.step += 1
            else:
                step.lookup(ids_dev[j])
                ev[i][1].record()
                uq = step.dedup(ids_dev[j])
                ev[i][2].record()
                step.update(ids_dev[j], grad_dev, uq)
            ev[i][3].record()
        p_end.record()
        barrier()
        if sampler:
            sampler.active = False
        lk = [e[0].elapsed_time(e[1]) for e in ev]
        dd = [e[1].elapsed_time(e[2]) for e in ev]
        up = [e[2].elapsed_time(e[3]) for e in ev]
        phases = {"lookup_ms": statistics.mean(lk), "dedup_ms": statistics.mean(dd),
                  "update_ms": statistics.mean(up),
                  "phase_loop_ms_per_step": p_begin.elapsed_time(p_end) / a.steps}

    sharded_phases, owner_counts = None, None
    if sharded and world > 1 and a.exchange == "slots":
        # the same step issued eagerly, phases in a row on one stream (no helper-stream overlap), a CUDA
        # event after each: says where the time goes; the headline above is the overlapped graph replay
        acc = {n: 0.0 for n in step.PHASES}
        for i in range(3):
            step.run_timed(ids_dev[i % nb], grad_dev)
        barrier()
        k_ph = min(a.steps, 20)
        for i in range(k_ph):
            marks = step.run_timed(ids_dev[(a.warmup + i) % nb], grad_dev)
            torch.cuda.synchronize()
            for n, e0, e1 in marks:
                acc[n] += e0.elapsed_time(e1)
        step.check()
        t_ph = torch.tensor([acc[n] / k_ph for n in step.PHASES], dtype=torch.float64, device=dev)
        dist.all_reduce(t_ph, op=dist.ReduceOp.MAX)
        sharded_phases = {n + "_ms": float(t_ph[j].item()) for j, n in enumerate(step.PHASES)}
        oc = torch.stack([step.n_recv[0].to(torch.float64), step.uq2.n_unique[0].to(torch.float64),
                          _single_share(step).to(torch.float64)])
        dist.all_reduce(oc, op=dist.ReduceOp.SUM)
        owner_counts = tuple(float(x) / world for x in oc.tolist())   # received keys, owner-unique keys, one-row share
    elif sharded and world > 1 and a.exchange == "slots-pipe":
        torch.cuda.synchronize()
        oc = torch.stack([sum(l.n_recv[0].to(torch.float64) for l in step.lanes),
                          sum(l.uq2.n_unique[0].to(torch.float64) for l in step.lanes),
                          sum(_single_share(l) for l in step.lanes).to(torch.float64) / len(step.lanes)])
        dist.all_reduce(oc, op=dist.ReduceOp.SUM)
        owner_counts = tuple(float(x) / world for x in oc.tolist())

    # ---------------- end-to-end through the host-buffer PS API
    e2e = None
    if not a.no_e2e:
        from easydl_b200.ps_client import HostStepRunner
        runner = HostStepRunner(step, dev)
        for i in range(2):
            runner.run(ids_host[i % nb], grad_host)
        barrier()
        if sampler:
            sampler.active = True
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.e2e_steps):
            runner.run(ids_host[i % nb], grad_host)
        e1.record()
        barrier()
        if sampler:
            sampler.active = False
        t2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2e = {"value": B * world * a.e2e_steps / (float(t2.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": runner.h2d_bytes, "d2h_bytes_per_step": runner.d2h_bytes,
               "ms_per_step": float(t2.item()) / a.e2e_steps,
               "path": "HostStepRunner: ids H2D -> lookup -> emb D2H (stream 1) || grad H2D (stream 2) -> dedup + update -> n_unique D2H"}

    # ---------------- the same through the lineage API itself (N = 1): PSClient.pull_embedding_batch (host ids
    # -> host embeddings) + push_gradient_batch (host gradients), one in-process PS shard over the same tables
    lineage = None
    if not a.no_e2e and not sharded:
        from easydl_b200.ps import ParameterServer
        from easydl_b200.ps_client import PSClient
        ps = ParameterServer([f"C{i + 1}" for i in range(F)], rows, D, opt, params=params)
        client = PSClient([ps])
        g3 = grad_host.view(B, F, D)
        for i in range(2):
            client.pull_embedding_batch(ids_host[i % nb])
            client.push_gradient_batch(g3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(a.e2e_steps):
            client.pull_embedding_batch(ids_host[i % nb])
            client.push_gradient_batch(g3)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / a.e2e_steps
        lineage = {"value": B / dt, "unit": UNIT, "ms_per_step": dt * 1e3,
                   "path": "PSClient.pull_embedding_batch -> PSClient.push_gradient_batch (pinned host arrays in and "
                           "out, one wait per pull, none per push); host wall clock"}
        del client, ps

    clocks = sampler.result() if sampler else None

    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        lookup_b, update_b = algorithmic_bytes(N, U_mean, D, a.optimizer)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_max / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_for(a, world, U_mean, nb),
            "implementation": {
                "code_path": ("sharded, lazily materialised rows" if a.lazy_rows else "sharded") if sharded else "single-shard fused",
                "gradient_wire": a.wire if sharded else None, "exchange": (a.exchange if world > 1 else None),
                "owner_dedup": (("merge of the S sorted key slots" if os.environ.get("EDL_SLOT_OWNER_MERGE", "1") == "1"
                                 else "radix sort of the received keys") if sharded and a.exchange.startswith("slots") else None)},
            "gpu_launches": step.launches_per_step * a.steps,
            "launch_mode": "cuda_graph" if graphed is not None else "eager",
            "step_pipelining": (("phase A (dedup, partition, key send) of batch i+1 beside the rounds of batch i" if sharded
                                 else "dedup of batch i+1 (helper stream) beside lookup + update of batch i") +
                                f"; ring of {nb} captured graphs" if sequential is not None
                                else "none: each step strictly after the last"),
            "clocks": clocks,
        }
        if phases:
            traffic, traffic_src = None, None
            tp = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tp) and a.dim == 64 and a.batch == 65536 and a.optimizer == "adagrad":
                tj = json.load(open(tp))
                traffic, traffic_src = tj["update_dram_bytes_per_step"], tj["source"]
            t_up = phases["update_ms"] * 1e-3
            t_lk = phases["lookup_ms"] * 1e-3
            ach = update_b / t_up / 1e9
            line["roofline"] = {
                "kernel": ("k_seg_stream" if (a.dim >= 256 and N >= (256 << 10)) or (a.dim >= 128 and N >= (1 << 20))
                           else "k_seg_direct") +   # seg_direct.cuh, launch_direct: the library's own choice
                          " || (k_seg_classify -> k_seg_chunk -> k_seg_long_final) (segmented reduce + "
                          "row update: four launches per step, hot-id kernels on a forked branch, timed together)",
                "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_kind": peak_kind,
                "algorithmic_bytes_per_launch": update_b,
            }
            line["phases"] = dict(phases)
            line["phases"]["lookup_gbs"] = lookup_b / t_lk / 1e9
            line["phases"]["lookup_frac"] = lookup_b / t_lk / 1e9 / peak
            line["phases"]["lookup_update_gbs_whole_step"] = (lookup_b + update_b) / (ms_max / a.steps * 1e-3) / 1e9
        if sequential is not None:
            line["sequential_step"] = sequential
        if parity is not None:
            line["parity"] = parity
        if sharded_phases is not None:
            line["phases"] = sharded_phases
        if owner_counts is not None:
            line["roofline"] = sharded_roofline(a, world, N, U_mean, owner_counts, ms_max / a.steps, peak, peak_kind)
        if e2e:
            line["e2e"] = e2e
        if lineage:
            line["e2e_lineage_api"] = lineage
        if not a.no_cpu_baseline and world == 1:     # the CPU port beside the N = 1 line only
            v, spp = cpu_reference(a, 1, a.cpu_steps, 1)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"{a.cpu_steps} full steps of the same workload through the NumPy oracle "
                          f"(1 thread; {os.cpu_count()} cores present), rows pre-materialised",
                "ms_per_step": spp * 1e3}
            cores = os.cpu_count() or 1
            c = cpu_compiled(a, min(cores, N_FIELDS), 20, 2, budget_s=20.0, timeout_s=240)
            if c:
                line["cpu_baseline_compiled"] = {
                    "value": c["value"], "unit": UNIT, "cores": c["threads"], "kind": "port", "ms_per_step": c["ms_per_step"],
                    "sample": f"{c['steps']} full steps of the same workload through the compiled C restatement "
                              f"(oracle/edl_oracle_c.c), one table per thread on {c['threads']} threads, rows pre-materialised"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse_args()
    if a.cpu_compiled_child:
        cpu_compiled_child(a)
    elif a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)


if __name__ == "__main__":
    main()
