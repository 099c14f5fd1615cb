"""Times the single-shard step's phases (CUDA events around graph replays) for a list of launch-shape
knob settings of libedlsparse in ONE process, and checks on a small table that every setting gives the
same bits as the first.  One gpurun call ranks the kernel variants:

    python tools/knob_sweep.py [--rows 10000000] [--dim 64] [--batch 65536] [--optimizer adagrad]
                               [--steps 60] [--out gpurun_out/knob_sweep.json] [--set name=value,...;...]
"""
import argparse, hashlib, json, os, statistics, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from easydl_b200 import _lib
from easydl_b200.criteo import CriteoSynthetic
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.step import SparseStep

DEFAULT_SETS = (
    "seg_impl=0,classify_side=0,gather_impl=0;"        # the round-1 shapes
    "seg_impl=0,classify_side=1,gather_impl=1;"
    "seg_impl=1,stream_ctas=3,long_thr=16;"
    "seg_impl=1,stream_ctas=2,long_thr=16;"
    "seg_impl=1,stream_ctas=3,long_thr=32;"
    "seg_impl=1,stream_ctas=3,long_thr=64;"
    "seg_impl=1,stream_ctas=3,long_thr=8;"
    "seg_impl=1,stream_ctas=3,long_thr=16,classify_side=0;"
    "seg_impl=1,stream_ctas=3,long_thr=16,seg_fork=0"
)
KNOBS = ("long_thr", "seg_fork", "classify_side", "seg_impl", "stream_ctas", "stream_stages", "direct_batch", "gather_impl", "dedup_impl", "dedup_shape")


def apply(setting, defaults):
    for k in KNOBS:
        _lib.tuning_set(k, setting.get(k, defaults[k]))


def digest(opt_name, dev):
    rows = [50_000] * 26
    opt = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    P = ShardParameters(rows, 64, opt, device=dev)
    src = CriteoSynthetic(rows, 8192, alpha=1.2)
    st = SparseStep(P, 8192)
    h = hashlib.sha256()
    for k in range(3):
        ids = torch.from_numpy(src.ids(k)).to(dev)
        grad = torch.from_numpy(src.grad(64, k)).to(dev)
        emb = st.run(ids, grad)
        torch.cuda.synchronize()
        h.update(emb.cpu().numpy().tobytes())
    h.update(P.slab.cpu().numpy().tobytes())
    for s in P.slots:
        h.update(s.cpu().numpy().tobytes())
    return h.hexdigest()[:16]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--alpha", type=float, default=1.05)
    ap.add_argument("--optimizer", default="adagrad")
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--n-batches", type=int, default=4)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "knob_sweep.json"))
    ap.add_argument("--set", default=DEFAULT_SETS)
    ap.add_argument("--interleaved", action="store_true", help="[p | slots] records instead of separate dense arrays")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    _lib.load()
    defaults = {k: _lib.tuning_get(k) for k in KNOBS}
    settings = []
    for part in a.set.split(";"):
        part = part.strip()
        if part:
            settings.append({kv.split("=")[0]: int(kv.split("=")[1]) for kv in part.split(",")})
    peak = 6550.0
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = float(json.load(open(pk))["hbm_gbs"])
    F, D, B = 26, a.dim, a.batch
    N = B * F
    rows = [a.rows] * F
    opt = OptimizerSpec.adagrad() if a.optimizer == "adagrad" else OptimizerSpec.adam()
    P = ShardParameters(rows, D, opt, device=dev, interleaved=a.interleaved)
    step = SparseStep(P, B)
    src = CriteoSynthetic(rows, B, a.alpha)
    ids = [torch.from_numpy(src.ids(i)).to(dev) for i in range(a.n_batches)]
    grad = torch.from_numpy(src.grad(D, 0)).to(dev)
    U = statistics.mean(int(step.dedup(i).n_unique.item()) for i in ids)
    lookup_b = 8 * N + 4 * D * U + 4 * D * N
    update_b = 4 * D * N + 4 * N + (16 if a.optimizer == "adagrad" else 24) * D * U
    results = []
    ref_digest = None
    print(f"# {F} x {a.rows} x dim {D}, batch {B}, {a.optimizer}, zipf {a.alpha}; U = {U:.0f}; peak {peak} GB/s")
    print("| setting | digest ok | lookup us (frac) | dedup us | update us (frac) | step us |")
    print("|---|---|---|---|---|---|")
    for sset in settings:
        apply(sset, defaults)
        try:
            dg = digest(a.optimizer, dev)
            if ref_digest is None:
                ref_digest = dg
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for i in ids[:1]:
                    step.run(i, grad)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            host_step = P.step
            graphs = []
            for i in ids:
                gl, gd, gu, ga = (torch.cuda.CUDAGraph() for _ in range(4))
                with torch.cuda.graph(gl):
                    step.lookup(i)
                with torch.cuda.graph(gd):
                    step.dedup(i)
                with torch.cuda.graph(gu):
                    step.update(i, grad, step.uq, rate_on_device=P.rate_changes)
                with torch.cuda.graph(ga):
                    step.run(i, grad, rate_on_device=P.rate_changes)
                graphs.append((gl, gd, gu, ga))
            P.step = host_step

            def timed(which):
                ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
                for k in range(a.steps + 5):
                    g = graphs[k % len(graphs)]
                    if which == 2:
                        g[1].replay()         # the update reads the dedup of ITS batch
                    if which >= 2 and P.rate_changes:
                        P.stage_rate(P.step + 1)
                    if k >= 5:
                        ev[k - 5][0].record()
                    g[which].replay()
                    if k >= 5:
                        ev[k - 5][1].record()
                    if which >= 2:
                        P.step += 1
                torch.cuda.synchronize()
                return statistics.mean(x.elapsed_time(y) for x, y in ev) * 1e-3
            t_l, t_d, t_u, t_a = timed(0), timed(1), timed(2), timed(3)
            r = {"setting": sset, "digest_ok": dg == ref_digest, "lookup_us": t_l * 1e6, "dedup_us": t_d * 1e6,
                 "update_us": t_u * 1e6, "step_us": t_a * 1e6, "lookup_frac": lookup_b / t_l / 1e9 / peak,
                 "update_frac": update_b / t_u / 1e9 / peak}
            del graphs
        except Exception as e:  # keep going: one broken variant must not hide the others
            r = {"setting": sset, "error": repr(e)[:400]}
            print(f"| {sset} | ERROR {r['error']} |", flush=True)
            results.append(r)
            continue
        results.append(r)
        print(f"| {sset} | {r['digest_ok']} | {r['lookup_us']:.1f} ({r['lookup_frac']:.3f}) | {r['dedup_us']:.1f} | "
              f"{r['update_us']:.1f} ({r['update_frac']:.3f}) | {r['step_us']:.1f} |", flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump({"workload": vars(a), "unique": U, "peak_gbs": peak, "results": results}, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
