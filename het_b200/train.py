"""Full-graph training step — the call a user makes.

Host-side mirror of ``HET_RGNN_train`` (hrt/python/RGNNUtils/RGNNUtils.py:199-345):
one "epoch" is one full-graph step = ``model(g, node_embed)`` → ``nll_loss(log_softmax)``
→ ``loss.backward()`` → ``optimizer.step()`` (Adam over the model and the learnable node
embedding, RGNNUtils.py:78-121). Two differences, both B200-first:

* the whole step (forward, loss, backward, Adam) is captured once in a CUDA graph and
  replayed, because at BGS size every kernel is a few microseconds and the step is
  otherwise launch-bound;
* with ``world > 1`` the graph is sharded by destination range (het_b200/parallel.py).

``step()`` keeps everything resident in HBM (the reference's embedding path).
``step_from_host(x, labels)`` is the end-to-end path for graphs that come with node
features: it copies this step's features and labels from pinned host memory, runs the same
forward/backward/Adam (the feature gradient is still produced, as the embedding path's
is) and returns the loss as a Python float (a device→host read).
"""
from __future__ import annotations

import contextlib
import os

import torch
import torch.distributed as dist
import torch.nn.functional as Fn

from . import backend as B
from . import parallel
from .graph import HetGraph
from .graphs import HeteroCOO
from .models import HET_HGT_DGLHetero, HET_RelGraphEmbed, HET_RGATModel, HET_RGCNModel


_NVTX = os.environ.get("HETB_NVTX", "0") == "1"


@contextlib.contextmanager
def _nvtx_range(name):
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


@contextlib.contextmanager
def _no_range(name):
    yield


class HetbAdam:
    """torch.optim.Adam's update (no weight decay, no amsgrad) as one hand-written kernel per tensor
    (csrc/optim.cu), step count on the device so that the optimizer step is part of the captured step
    graph. Same interface subset the trainer uses (step / zero_grad)."""

    def __init__(self, params, lr=1e-2, betas=(0.9, 0.999), eps=1e-8):
        from .kernels import K
        self.K = K
        self.params = [p for p in params]
        self.lr, self.betas, self.eps = lr, betas, eps
        self.state = {}
        self.step_count = None

    def _state(self, p):
        if p not in self.state:
            self.state[p] = (torch.zeros_like(p, memory_format=torch.contiguous_format),
                             torch.zeros_like(p, memory_format=torch.contiguous_format))
        return self.state[p]

    @torch.no_grad()
    def step(self):
        live = [p for p in self.params if p.grad is not None]
        if not live:
            return
        if self.step_count is None:
            self.step_count = torch.zeros(1, dtype=torch.float32, device=live[0].device)
        self.K.hetb200_adam_bump_step(self.step_count)
        for p in live:
            m, v = self._state(p)
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            self.K.hetb200_adam_step(p, g, m, v, self.step_count, self.lr, self.betas[0], self.betas[1], self.eps)

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()


class FullGraphTrainer:
    def __init__(self, kind: str, coo: HeteroCOO, device, *, hidden=64, out_dim=64, heads=1,
                 layers=1, lr=1e-2, rank=0, world=1, group=None, cuda_graph=True, seed=0,
                 compact_as_of_node=False, fused=False, bf16_tables=False):
        assert kind in ("rgat", "hgt", "rgcn")
        self.kind, self.device, self.rank, self.world, self.group = kind, torch.device(device), rank, world, group
        self.num_nodes_total = coo.num_nodes
        self.num_edges_total = coo.num_edges
        if world > 1:
            self.plan = parallel.make_shard(coo, rank, world)
            self.g = parallel.shard_graph(self.plan, coo, self.device)
            self.n_local = self.plan.num_local_nodes
            self.owned = torch.from_numpy(self.plan.owned)   # global ids of the owned rows, local order
        else:
            self.plan = None
            self.g = HetGraph(coo, self.device)
            self.n_local = coo.num_nodes
            self.owned = torch.arange(coo.num_nodes)
        self.num_edges_local = self.g.num_edges
        if compact_as_of_node:  # build the unique (relation, node) lists now, outside any graph capture
            self.g.get_separate_unique_node_indices_single_sided()

        torch.manual_seed(seed)  # identical replicated weights on every rank
        if kind == "rgat":
            self.model = HET_RGATModel(coo.num_rels, hidden, hidden, out_dim, heads, layers,
                                       compact_as_of_node_flag=compact_as_of_node,
                                       linear_reorder_flag=bool(fused and compact_as_of_node)).to(self.device)
        elif kind == "rgcn":
            self.model = HET_RGCNModel(coo.num_rels, hidden, hidden, out_dim, layers,
                                       compact_as_of_node_flag=compact_as_of_node, compact_direct_indexing_flag=True,
                                       csr_aggregation_flag=bool(fused and compact_as_of_node)).to(self.device)
            # edge norm 1 / in-degree(dst), indexed by canonical edge id (= separate-COO position)
            col = self.g.get_separate_coo_original()["col_indices"]
            deg = torch.bincount(col, minlength=self.g.get_num_nodes()).clamp(min=1).to(torch.float32)
            self.edge_norm = (1.0 / deg[col]).view(-1, 1).contiguous()
        else:
            self.model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, hidden, out_dim, heads, layers,
                                           compact_as_of_node_flag=compact_as_of_node,
                                           fused_attention_aggregation_flag=bool(fused and compact_as_of_node),
                                           linear_reorder_flag=bool(fused and compact_as_of_node),
                                           bf16_tables_flag=bool(bf16_tables)).to(self.device)
        gen = torch.Generator().manual_seed(seed + 1)
        embed_full = torch.rand(coo.num_nodes, hidden, generator=gen) * 2 - 1
        labels_full = torch.randint(0, out_dim, (coo.num_nodes,), generator=gen)
        self.embed = HET_RelGraphEmbed(self.n_local, hidden).to(self.device)
        with torch.no_grad():
            self.embed.embeds.copy_(embed_full[self.owned])
        self.labels = labels_full[self.owned].to(self.device)
        self.x_in = torch.zeros(self.n_local, hidden, device=self.device, requires_grad=True)
        self.labels_in = torch.zeros(self.n_local, dtype=torch.int64, device=self.device)

        cap = bool(cuda_graph)
        self.opt_embed = HetbAdam(list(self.model.parameters()) + list(self.embed.parameters()), lr=lr)
        self.opt_feat = HetbAdam(list(self.model.parameters()), lr=lr)
        self.cuda_graph = cap
        self._graphs = {}
        self._loss = {}

    # ------------------------------------------------------------------
    @property
    def _gather(self):
        if getattr(self, "_row_exchange", None) is None:
            self._row_exchange = parallel.RowExchange(self.plan, self.group)
        return self._row_exchange

    def forward(self, h0):
        if self.kind == "rgcn":
            return self.model(self.g, h0, self.edge_norm, None if self.world == 1 else self._gather)
        if self.world == 1:
            return self.model(self.g, h0)
        return self.model(self.g, h0, self._gather)

    def _step_body(self, mode):
        if mode == "embed":
            h0, labels, opt = self.embed(), self.labels, self.opt_embed
        else:
            h0, labels, opt = self.x_in, self.labels_in, self.opt_feat
        # NVTX ranges (the reference brackets its phases with nvtx.range_push/pop,
        # hrt/python/RGNNUtils/RGNNUtils.py:291-345) when HETB_NVTX=1: host-side markers, legal under graph capture
        rng = _nvtx_range if _NVTX else _no_range
        with rng("hetb200/forward"):
            logits = self.forward(h0)
        with rng("hetb200/loss"):
            # = Fn.nll_loss(Fn.log_softmax(logits, dim=-1), labels, reduction="sum") / num_nodes_total, fused
            loss = B.softmax_cross_entropy(logits, labels, self.num_nodes_total)
        with rng("hetb200/backward"):
            loss.backward()
            if self.world > 1:
                parallel.all_reduce_grads(list(self.model.parameters()), self.group)
        with rng("hetb200/adam"):
            opt.step()
        return loss.detach()

    def _zero(self, mode):
        opt = self.opt_embed if mode == "embed" else self.opt_feat
        opt.zero_grad(set_to_none=True)
        self.x_in.grad = None

    def _run(self, mode):
        if not self.cuda_graph:
            self._zero(mode)
            self._loss[mode] = self._step_body(mode)
            return self._loss[mode]
        if mode not in self._graphs:
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                for _ in range(3):  # warm-up outside capture (lazy inits, Adam state)
                    self._zero(mode)
                    self._step_body(mode)
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            self._zero(mode)
            with torch.cuda.graph(graph):
                self._loss[mode] = self._step_body(mode)
            self._graphs[mode] = graph
        self._graphs[mode].replay()
        return self._loss[mode]

    # ------------------------------------------------------------------
    def step(self) -> torch.Tensor:
        """One epoch with the learnable node embedding (everything resident)."""
        return self._run("embed")

    def step_from_host(self, x_host: torch.Tensor, labels_host: torch.Tensor) -> float:
        """One epoch with this step's node features / labels coming from pinned host memory."""
        with torch.no_grad():
            self.x_in.copy_(x_host, non_blocking=True)
            self.labels_in.copy_(labels_host, non_blocking=True)
        loss = self._run("feat")
        if self.world > 1:
            loss = loss.clone()
            dist.all_reduce(loss, group=self.group)
        return float(loss.item())

    def run_from_host(self, batches) -> list:
        """Pipelined epochs over an iterable of (x_host, labels_host) pinned-memory pairs: while step
        i computes, a copy stream already moves step i+1's inputs into the other of two device
        staging buffers, and step i's loss comes back through a pinned slot that is read one step
        later. Every step still pays its own H2D copy and its own D2H loss read; they just overlap
        the previous / next step's kernels instead of serialising with them. Returns the losses."""
        dev = self.device
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_pipe", None) is None:
            self._pipe = {
                "copy": torch.cuda.Stream(dev),
                "x": [torch.empty_like(self.x_in, requires_grad=False) for _ in range(2)],
                "y": [torch.empty_like(self.labels_in) for _ in range(2)],
                "loss": [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)],
                "ready": [torch.cuda.Event() for _ in range(2)],   # H2D into slot done
                "free": [torch.cuda.Event() for _ in range(2)],    # slot consumed by the step
                "done": [torch.cuda.Event() for _ in range(2)],    # loss of the step in the pinned slot
            }
        P = self._pipe
        copy = P["copy"]
        copy.wait_stream(main)

        def prefetch(i, batch):
            j = i & 1
            with torch.cuda.stream(copy), torch.no_grad():
                if i >= 2:
                    copy.wait_event(P["free"][j])
                P["x"][j].copy_(batch[0], non_blocking=True)
                P["y"][j].copy_(batch[1], non_blocking=True)
                P["ready"][j].record(copy)

        losses = []
        it = iter(batches)
        nxt = next(it, None)
        if nxt is not None:
            prefetch(0, nxt)
        i = 0
        while nxt is not None:
            cur, j = nxt, i & 1
            nxt = next(it, None)
            if nxt is not None:
                prefetch(i + 1, nxt)
            main.wait_event(P["ready"][j])
            with torch.no_grad():
                self.x_in.copy_(P["x"][j])
                self.labels_in.copy_(P["y"][j])
            P["free"][j].record(main)
            if i >= 2:  # the pinned loss slot j is about to be reused: collect step i-2 first
                P["done"][j].synchronize()
                losses.append(float(P["loss"][j].item()))
            loss = self._run("feat")
            if self.world > 1:
                loss = loss.clone()
                dist.all_reduce(loss, group=self.group)
            P["loss"][j].copy_(loss.reshape(1), non_blocking=True)
            P["done"][j].record(main)
            i += 1
        for k in range(max(0, i - 2), i):
            P["done"][k & 1].synchronize()
            losses.append(float(P["loss"][k & 1].item()))
        return losses

    def h2d_bytes_per_step(self) -> int:
        return self.n_local * self.x_in.shape[1] * 4 + self.n_local * 8

    @staticmethod
    def d2h_bytes_per_step() -> int:
        return 4
