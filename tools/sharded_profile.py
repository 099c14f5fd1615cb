"""GPU time of each operation of the sharded step at the cfg2 shape (one rank's work, S virtual
shards' worth of keys handled locally; no NCCL).  CUDA events around each op, median of 10."""
import statistics, sys, torch
sys.path.insert(0, '.')
from easydl_b200 import kernels as K
from easydl_b200.criteo import CriteoSynthetic
from easydl_b200.optimizer import OptimizerSpec
from easydl_b200.parameters import ShardParameters

dev = torch.device('cuda:0')
rows = [int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000] * 26
D, B = 64, 65536
P = ShardParameters(rows, D, OptimizerSpec.adagrad(), device=dev)
src = CriteoSynthetic(rows, B)
ids = torch.from_numpy(src.ids(0)).to(dev)
grad = torch.from_numpy(src.grad(D, 0)).to(dev)
fb = torch.from_numpy(P.layout.key_base[:26].copy()).to(dev)
kb = P.layout.key_bits
S = 8

def timed(name, fn, reps=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    print(f"{name:34s} {statistics.median(ts)*1e3:8.1f} us")
    return out

uq = timed("unique(ids, inverse)", lambda: K.unique(ids.view(-1), fb, kb, want_inverse=True, want_counts=False))
U = uq.count()
part = timed("partition_mod(uniq, S=8)", lambda: K.partition_mod(uq.uniq, P.layout_dev, S, n_dev=uq.n_unique))
keys = part.keys[:U].contiguous()
rows_out = timed("gather_keys (owner lookup, U rows)", lambda: K.gather_keys(P.slab, keys, None))
idx = timed("compose_index", lambda: K.compose_index(uq.inverse, part.pos))
emb = timed("expand (U rows -> N rows)", lambda: K.expand(rows_out, idx))
g64 = timed("segreduce -> f64 send rows", lambda: K.segreduce(grad, uq.order, uq.seg_offsets, uq.n_unique, uq.n, seg_first=uq.seg_first, out_pos=part.pos, out_dtype=torch.float64))
g32 = timed("segreduce -> f32 send rows", lambda: K.segreduce(grad, uq.order, uq.seg_offsets, uq.n_unique, uq.n, seg_first=uq.seg_first, out_pos=part.pos, out_dtype=torch.float32))
uq2 = timed("unique(recv keys)", lambda: K.unique(keys, None, kb, want_inverse=False, want_counts=False))
timed("segreduce_update (f64 rows in)", lambda: P.apply_gradients(uq2, g64[:U].contiguous(), U))
timed("segreduce_update (f32 rows in)", lambda: P.apply_gradients(uq2, g32[:U].contiguous(), U))
print("U =", U, "N =", uq.n)
