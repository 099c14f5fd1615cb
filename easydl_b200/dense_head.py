"""Minimal dense heads over the sparse step: Wide&Deep, DeepFM and DLRM-style
interaction + MLP in plain PyTorch, so that the BASELINE.json configs are real training steps
with a falling loss.  The dense math is library code (cuBLAS through torch) and is NOT part of the
hot path this repository implements; what matters here is the seam: the head consumes the
looked-up embeddings and hands their gradient (already scaled by 1/B of the mean loss, SPEC A4)
to the sparse update.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

N_DENSE = 13


def _mlp(sizes, last_act=False):
    layers = []
    for i in range(len(sizes) - 1):
        layers.append(nn.Linear(sizes[i], sizes[i + 1]))
        if i < len(sizes) - 2 or last_act:
            layers.append(nn.ReLU())
    return nn.Sequential(*layers)


class WideDeepHead(nn.Module):
    def __init__(self, n_fields: int, dim: int, hidden=(256, 128)):
        super().__init__()
        self.wide = nn.Linear(N_DENSE, 1)
        self.deep = _mlp([N_DENSE + n_fields * dim, *hidden, 1])

    def forward(self, dense, emb):          # dense [B, 13], emb [B, F, D] -> logits [B]
        x = torch.cat([dense, emb.flatten(1)], dim=1)
        return (self.wide(dense) + self.deep(x)).squeeze(1)


class DeepFMHead(nn.Module):
    def __init__(self, n_fields: int, dim: int, hidden=(400, 400)):
        super().__init__()
        self.first = nn.Linear(N_DENSE, 1)
        self.deep = _mlp([N_DENSE + n_fields * dim, *hidden, 1])

    def forward(self, dense, emb):
        s = emb.sum(dim=1)                                   # FM second order: 0.5*((sum v)^2 - sum v^2)
        fm = 0.5 * (s * s - (emb * emb).sum(dim=1)).sum(dim=1, keepdim=True)
        x = torch.cat([dense, emb.flatten(1)], dim=1)
        return (self.first(dense) + fm + self.deep(x)).squeeze(1)


class DLRMHead(nn.Module):
    def __init__(self, n_fields: int, dim: int, bottom=(512, 256), top=(512, 256)):
        super().__init__()
        self.bottom = _mlp([N_DENSE, *bottom, dim], last_act=True)
        n = n_fields + 1
        self.register_buffer("tri", torch.triu_indices(n, n, offset=1), persistent=False)
        self.top = _mlp([dim + n * (n - 1) // 2, *top, 1])

    def forward(self, dense, emb):
        z = self.bottom(dense)                                # [B, D]
        v = torch.cat([z.unsqueeze(1), emb], dim=1)           # [B, F+1, D]
        dots = torch.bmm(v, v.transpose(1, 2))[:, self.tri[0], self.tri[1]]
        return self.top(torch.cat([z, dots], dim=1)).squeeze(1)


HEADS = {"wide_deep": WideDeepHead, "deepfm": DeepFMHead, "dlrm": DLRMHead}


class SparseDenseTrainer:
    """One training step = sparse lookup (A3) -> dense head forward/backward (torch) -> sparse
    update (A1 + A4 + A5) -> dense optimizer step.  `sparse_step` is a SparseStep (one GPU) or a
    ShardedSparseStep (row-sharded tables; the head then runs data-parallel on each rank's slice
    and its gradients are all-reduced by the caller / DDP)."""

    def __init__(self, sparse_step, head: nn.Module, dense_opt: Optional[torch.optim.Optimizer] = None):
        self.sparse = sparse_step
        self.head = head
        self.dense_opt = dense_opt or torch.optim.Adagrad(head.parameters(), lr=0.01)
        self.F = sparse_step.F
        self.D = sparse_step.params.dim

    def train_step(self, dense: torch.Tensor, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """dense [B, 13] f32, ids [B, F] int64, labels [B] f32, all on the device -> loss (device)."""
        B = ids.shape[0]
        emb = self.sparse.lookup(ids).view(B, self.F, self.D).detach().requires_grad_(True)
        logits = self.head(dense, emb)
        loss = F.binary_cross_entropy_with_logits(logits, labels)
        self.dense_opt.zero_grad(set_to_none=True)
        loss.backward()
        self.sparse.update(ids, emb.grad.reshape(B * self.F, self.D).contiguous())
        self.dense_opt.step()
        return loss.detach()
