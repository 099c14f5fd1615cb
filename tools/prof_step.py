"""A few eager single-shard steps at the cfg2 shape for ncu captures (no graphs, no timing):
    python tools/prof_step.py [--rows N] [--steps K] [--set knob=value,...]"""
import argparse, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from easydl_b200 import _lib
from easydl_b200.criteo import CriteoSynthetic
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters
from easydl_b200.step import SparseStep

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_000_000)
ap.add_argument("--dim", type=int, default=64)
ap.add_argument("--batch", type=int, default=65536)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--optimizer", default="adagrad")
ap.add_argument("--set", default="")
a = ap.parse_args()
dev = torch.device("cuda:0")
_lib.load()
for kv in filter(None, a.set.split(",")):
    k, v = kv.split("=")
    _lib.tuning_set(k, int(v))
rows = [a.rows] * 26
opt = OptimizerSpec.adagrad() if a.optimizer == "adagrad" else OptimizerSpec.adam()
P = ShardParameters(rows, a.dim, opt, device=dev)
step = SparseStep(P, a.batch, overlap_dedup=False)
src = CriteoSynthetic(rows, a.batch, 1.05)
grad = torch.from_numpy(src.grad(a.dim, 0)).to(dev)
for k in range(a.steps):
    ids = torch.from_numpy(src.ids(k)).to(dev)
    step.run(ids, grad)
torch.cuda.synchronize()
print("ok", int(step.uq.n_unique.item()))
