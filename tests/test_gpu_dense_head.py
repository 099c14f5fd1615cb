"""The seam between the sparse step and a dense head: real training steps with a falling loss for
the three model families of BASELINE.json (Wide&Deep, DeepFM, DLRM)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("family,dim,opt", [("wide_deep", 16, "adagrad"), ("deepfm", 64, "adagrad"), ("dlrm", 32, "adam")])
def test_training_loss_falls(cuda, family, dim, opt):
    from easydl_b200.criteo import CriteoSynthetic
    from easydl_b200.dense_head import HEADS, SparseDenseTrainer
    from easydl_b200.optimizer import OptimizerSpec
    from easydl_b200.parameters import ShardParameters
    from easydl_b200.step import SparseStep
    torch.manual_seed(0)
    rows, B = [2000] * 26, 1024
    # the embedding gradient carries 1/B, so let Adagrad normalise it (tiny initial accumulator)
    spec = OptimizerSpec.adagrad(lr=0.05, init_accum=1e-8) if opt == "adagrad" else OptimizerSpec.adam(lr=5e-3)
    P = ShardParameters(rows, dim, spec, device=cuda)
    head = HEADS[family](26, dim).to(cuda)
    tr = SparseDenseTrainer(SparseStep(P, B), head, torch.optim.Adam(head.parameters(), lr=2e-3))
    src = CriteoSynthetic(rows, B, 1.05)
    losses = []
    for k in range(120):
        dense, ids, _ = src.batch(k)
        # a target only the embeddings can learn: a function of the id of one field
        labels = ((ids[:, 0] % 5) < 2).astype(np.float32)
        loss = tr.train_step(torch.from_numpy(dense).to(cuda), torch.from_numpy(ids).to(cuda),
                             torch.from_numpy(labels).to(cuda))
        losses.append(float(loss))
    first, last = np.mean(losses[:5]), np.mean(losses[-5:])
    assert np.isfinite(losses).all()
    assert last < 0.8 * first, (first, last)
    assert P.step == 120


def test_sparse_gradient_seam_matches_autograd(cuda):
    """The gradient handed to the sparse update is d(mean loss)/d(emb): applying it with plain SGD
    semantics through torch (index_add on a dense copy) gives the same direction the kernels see."""
    from easydl_b200.dense_head import DeepFMHead
    torch.manual_seed(1)
    B, Fn, D = 64, 26, 8
    head = DeepFMHead(Fn, D).to(cuda)
    emb = torch.randn(B, Fn, D, device=cuda, requires_grad=True)
    dense = torch.rand(B, 13, device=cuda)
    labels = (torch.rand(B, device=cuda) < 0.3).float()
    loss = torch.nn.functional.binary_cross_entropy_with_logits(head(dense, emb), labels)
    loss.backward()
    g = emb.grad
    # mean reduction: per-sample gradients carry the 1/B factor (SPEC A4: no extra division later)
    loss_sum = torch.nn.functional.binary_cross_entropy_with_logits(head(dense, emb.detach()), labels, reduction="sum")
    emb2 = emb.detach().clone().requires_grad_(True)
    torch.nn.functional.binary_cross_entropy_with_logits(head(dense, emb2), labels, reduction="sum").backward()
    torch.testing.assert_close(g, emb2.grad / B, rtol=1e-4, atol=1e-7)
