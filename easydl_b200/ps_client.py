"""Worker-side entry points with HOST buffers (the boundary a worker <-> PS RPC would cross).

HostStepRunner is what bench.py's `e2e` number times: every step the ids and the upstream
gradient come from pinned host memory and the looked-up embeddings go back to pinned host
memory, all on one CUDA stream.
"""
from __future__ import annotations

import torch


class HostStepRunner:
    def __init__(self, step, device):
        p = step.params
        self.step = step
        N, D = step.N, p.dim
        self.ids_dev = torch.empty((step.B, step.F), dtype=torch.int64, device=device)
        self.grad_dev = torch.empty((N, D), dtype=torch.float32, device=device)
        self.emb_host = torch.empty((N, D), dtype=torch.float32).pin_memory()
        self.nu_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.h2d_bytes = N * 8 + N * D * 4
        self.d2h_bytes = N * D * 4 + 4

    def run(self, ids_host: torch.Tensor, grad_host: torch.Tensor):
        """pull_embedding_vectors(ids) -> host, then push_gradients(ids, grad) from host."""
        st = self.step
        self.ids_dev.copy_(ids_host, non_blocking=True)
        emb = st.lookup(self.ids_dev)
        self.emb_host.copy_(emb, non_blocking=True)
        self.grad_dev.copy_(grad_host, non_blocking=True)
        st.update(self.ids_dev, self.grad_dev)
        self.nu_host.copy_(st.last_n_unique, non_blocking=True)
        return self.emb_host
