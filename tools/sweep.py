"""BASELINE.json configs[4]: sparse lookup+update sweep on one B200 -- batch x dim x Zipf skew,
phase times by CUDA events (median of 7 eager runs), achieved GB/s on the algorithmic bytes of
SURVEY.md 8(d), fraction of the measured HBM peak.  Table rows are scaled so each slab stays
near 33 GB.  Prints a markdown table."""
import json, os, statistics, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from easydl_b200.criteo import CriteoSynthetic
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.step import SparseStep

dev = torch.device('cuda:0')
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
batches = [int(x) for x in os.environ.get("SWEEP_B", "4096,65536,262144,1048576").split(",")]
dims = [int(x) for x in os.environ.get("SWEEP_D", "16,64,128,256").split(",")]
alphas = [float(x) for x in os.environ.get("SWEEP_A", "0.5,1.05,1.2").split(",")]
opt_name = os.environ.get("SWEEP_OPT", "adagrad")
F = 26

def med(fn, reps=7):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts) * 1e-3

print(f"| batch | dim | zipf | U/N | lookup us | GB/s (frac) | dedup us | update us | GB/s (frac) | step samples/s |")
print("|---|---|---|---|---|---|---|---|---|---|")
for D in dims:
    rows_t = min(10_000_000, int(33e9 / (F * 4 * D)))
    rows = [rows_t] * F
    opt = OptimizerSpec.adagrad() if opt_name == "adagrad" else OptimizerSpec.adam()
    P = ShardParameters(rows, D, opt, device=dev)
    for B in batches:
        if B * F * D * 4 * 2 > 60e9:
            continue
        step = SparseStep(P, B)
        N = B * F
        grad = torch.randn((N, D), device=dev) * 1e-2
        for al in alphas:
            ids = torch.from_numpy(CriteoSynthetic(rows, B, al).ids(0)).to(dev)
            t_l = med(lambda: step.lookup(ids))
            t_d = med(lambda: step.dedup(ids))
            uq = step.dedup(ids)
            U = int(uq.n_unique.item())
            t_u = med(lambda: step.update(ids, grad, uq))
            lb = 8 * N + 4 * D * U + 4 * D * N
            ub = 4 * D * N + 4 * N + (16 if opt_name == "adagrad" else 24) * D * U
            gl, gu = lb / t_l / 1e9, ub / t_u / 1e9
            print(f"| {B} | {D} | {al} | {U/N:.3f} | {t_l*1e6:.0f} | {gl:.0f} ({gl/peak:.2f}) | {t_d*1e6:.0f} | "
                  f"{t_u*1e6:.0f} | {gu:.0f} ({gu/peak:.2f}) | {B/(t_l+t_d+t_u)/1e6:.1f} M |", flush=True)
        del step, grad
    del P
    torch.cuda.empty_cache()
