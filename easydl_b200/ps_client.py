"""Worker-side entry points with HOST buffers (the boundary a worker <-> PS RPC would cross).

HostStepRunner is what bench.py's `e2e` number times: every step the ids and the upstream
gradient come from pinned host memory and the looked-up embeddings go back to pinned host
memory.
"""
from __future__ import annotations

import torch


class HostStepRunner:
    """Both PCIe directions are kept busy: the gradient upload runs on a second stream while the
    first carries ids up, the lookup, and the embeddings down; the update waits for both."""

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
        self.up = torch.cuda.Stream(device=device)
        self.ev_up = torch.cuda.Event()
        self.ev_free = torch.cuda.Event()
        self.ev_free.record()

    def run(self, ids_host: torch.Tensor, grad_host: torch.Tensor):
        """pull_embedding_vectors(ids) -> host, then push_gradients(ids, grad) from host."""
        st = self.step
        main = torch.cuda.current_stream()
        # Do not submit this step's copies before the previous update has finished: copies queued
        # behind an unfinished dependency block the copy engines' FIFOs and the two directions stop
        # overlapping (measured: 13.9 ms per step when the host runs ahead, 10.6 ms with this wait).
        self.ev_free.synchronize()
        # ids first: both uploads share one copy engine queue, and the lookup waits for the ids
        self.ids_dev.copy_(ids_host, non_blocking=True)
        self.up.wait_event(self.ev_free)             # previous update has finished reading grad_dev
        with torch.cuda.stream(self.up):
            self.grad_dev.copy_(grad_host, non_blocking=True)
            self.ev_up.record()
        emb = st.lookup(self.ids_dev)
        self.emb_host.copy_(emb, non_blocking=True)
        main.wait_event(self.ev_up)
        st.update(self.ids_dev, self.grad_dev)
        self.ev_free.record()
        self.nu_host.copy_(st.last_n_unique, non_blocking=True)
        return self.emb_host


class PSClient:
    """Worker-side view of S parameter-server shards living in this process (one per GPU, or
    several virtual shards on one GPU).  Lineage surface: `pull_embedding_vectors(name, ids)`
    and `push_gradients(...)` take and return HOST arrays, exactly what a worker <-> PS RPC
    would carry; the fused forms take a whole [B, F] id batch.

    Per call: ids H2D -> A1 dedup -> A2 partition by owner -> per-shard A3 gather ->
    reassemble -> expand -> D2H."""

    def __init__(self, servers):
        from .hash_utils import scatter_embedding_vector
        self._scatter = scatter_embedding_vector
        self.servers = list(servers)
        self.S = len(self.servers)
        p0 = self.servers[0].embedding_params
        self.layout = p0.layout
        self.device = p0.device
        self.key_bits = p0.layout.key_bits
        self.layout_dev = p0.layout_dev
        self.field_base = torch.from_numpy(p0.layout.key_base[:-1].copy()).to(self.device)
        self._last = None
        self._fast = None          # S == 1: buffers of the synchronisation-free batch path

    # ---- S == 1: one PS shard in this process.  No dedup is needed to look rows up, nothing has to be
    # partitioned, and no size has to come back to the host: ids go up, rows come down into a pinned buffer,
    # gradients go up and the dedup + reduce + optimizer are enqueued behind them.  The only host wait is the
    # one the caller asked for (the embeddings of pull_embedding_batch).
    def _fast_buffers(self, B: int, F: int):
        from .step import SparseStep
        key = (B, F)
        if self._fast is None or self._fast["key"] != key:
            p = self.servers[0].embedding_params
            step = SparseStep(p, B, F, overlap_dedup=False)
            self._fast = {"key": key, "step": step,
                          "ids": torch.empty((B, F), dtype=torch.int64, device=self.device),
                          "grad": torch.empty((B * F, p.dim), dtype=torch.float32, device=self.device),
                          "emb_host": torch.empty((B * F, p.dim), dtype=torch.float32).pin_memory(),
                          "done": torch.cuda.Event()}
        return self._fast

    def _pull_batch_fast(self, ids_t: torch.Tensor):
        B, F = ids_t.shape
        f = self._fast_buffers(B, F)
        f["ids"].copy_(ids_t, non_blocking=True)
        emb = f["step"].lookup(f["ids"])
        f["emb_host"].copy_(emb, non_blocking=True)
        f["done"].record()
        f["done"].synchronize()                      # the caller gets host data: the one wait of the call
        return f["emb_host"].numpy().reshape(B, F, -1)   # a view of the pinned buffer, valid until the next pull

    def _push_batch_fast(self, g_t: torch.Tensor):
        f = self._fast
        f["grad"].copy_(g_t.reshape(f["grad"].shape), non_blocking=True)
        f["step"].update(f["ids"], f["grad"])        # dedup + reduce + optimizer, enqueued; no host wait

    # ---- helpers
    def _pull_keys(self, flat_ids_dev, field_base):
        from . import kernels as K
        uq = K.unique(flat_ids_dev, field_base, self.key_bits, want_inverse=True, want_counts=False)
        U = uq.count()
        keys_by, offs, _, perm = self._scatter(None, uq.uniq[:U], self.S, self.layout_dev)
        rows = torch.empty((max(U, 1), self.servers[0].embedding_params.dim), dtype=torch.float32,
                           device=self.device)
        for s, ps in enumerate(self.servers):
            a, b = offs[s], offs[s + 1]
            if b > a:
                dev_s = ps.embedding_params.device
                r = ps.pull_embedding_keys(keys_by[a:b].to(dev_s))
                rows[a:b] = r.to(self.device)
        pos = torch.empty(max(U, 1), dtype=torch.int32, device=self.device)
        pos[perm[:U].long()] = torch.arange(U, dtype=torch.int32, device=self.device)
        emb = K.expand(rows, K.compose_index(uq.inverse, pos))
        return emb, (uq, U, keys_by, offs, pos)

    # ---- lineage calls (host arrays)
    def pull_embedding_vectors(self, layer_name: str, ids):
        """ids: 1-D host array of per-table ids (duplicates allowed) -> [n, dim] host array."""
        t = self.servers[0].names.index(layer_name)
        d = torch.as_tensor(ids, dtype=torch.int64).reshape(-1).to(self.device)
        fb = self.field_base[t:t + 1].contiguous()
        emb, _ = self._pull_keys(d, fb)
        return emb.cpu().numpy()

    def pull_embedding_batch(self, ids):
        """ids: host [B, F] (field f reads table f) -> host [B, F, dim]; remembers the dedup for
        the matching push_gradient_batch."""
        ids_t = torch.as_tensor(ids, dtype=torch.int64)
        B, F = ids_t.shape
        if self.S == 1 and self.servers[0].embedding_params.num_shards == 1:
            self._last = "fast"
            return self._pull_batch_fast(ids_t)
        emb, st = self._pull_keys(ids_t.to(self.device).reshape(-1), self.field_base[:F].contiguous())
        self._last = st
        return emb.cpu().numpy().reshape(B, F, -1)

    def push_gradient_batch(self, grads):
        """grads: host [B, F, dim], gradient of the last pull_embedding_batch's output.
        One in-process shard (S == 1): the call returns once the upload and the update are ENQUEUED.  A pageable
        array has been read by then; a PINNED array is read by the copy engine later, so leave it unchanged until
        the next pull_embedding_batch returns (its wait covers this upload) or the device is synchronised."""
        from . import kernels as K
        if self._last == "fast":
            return self._push_batch_fast(torch.as_tensor(grads, dtype=torch.float32))
        uq, U, keys_by, offs, pos = self._last
        g = torch.as_tensor(grads, dtype=torch.float32).reshape(uq.n, -1).to(self.device)
        g_send = torch.empty((max(U, 1), g.shape[1]), dtype=torch.float64, device=self.device)
        K.segreduce(g, uq.order, uq.seg_offsets, uq.n_unique, uq.n, seg_first=uq.seg_first,
                    out_pos=pos, out=g_send)
        for s, ps in enumerate(self.servers):
            a, b = offs[s], offs[s + 1]
            dev_s = ps.embedding_params.device
            ps.push_gradients(keys_by[a:b].to(dev_s), g_send[a:b].to(dev_s))

    def push_gradients(self, layer_name: str, ids, grads):
        """(ids, grads) of one table as host arrays, duplicates allowed (an IndexedSlices)."""
        from . import kernels as K
        t = self.servers[0].names.index(layer_name)
        keys = torch.as_tensor(ids, dtype=torch.int64).reshape(-1).to(self.device) + int(self.layout.key_base[t])
        g = torch.as_tensor(grads, dtype=torch.float32).to(self.device).contiguous()
        keys_by, offs, rows_by, _ = self._scatter(g, keys, self.S, self.layout_dev)
        for s, ps in enumerate(self.servers):
            a, b = offs[s], offs[s + 1]
            dev_s = ps.embedding_params.device
            ps.push_gradients(keys_by[a:b].to(dev_s).contiguous(),
                              (rows_by[a:b] if rows_by is not None else g[:0]).to(dev_s).contiguous())

    def partition_dense_parameters(self, names):
        from .hash_utils import string_to_id
        out = {s: [] for s in range(self.S)}
        for n in names:
            out[string_to_id(n, self.S)].append(n)
        return out
