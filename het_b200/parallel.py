"""Destination sharding of full-graph training over the GPUs of one box.

The reference has no distributed code at all; this is the one parallelism
BASELINE.json's north star adds. Every rank owns a contiguous destination range OF EVERY NODE TYPE
(rank p holds slice p of each type's id range), all in-edges of those destinations, their output
rows and the per-edge tensors of those edges. Weights are replicated. Per layer there is ONE
exchange: the layer input ``h`` is all-gathered (``ncclAllGather`` over NVLink 5 / NVSwitch); its
backward is the matching ``ncclReduceScatter``. Weight gradients are all-reduced once per step.

Why one slice per type rather than one range of the whole id space (round 1): node types differ
10x in in-degree, so ranges balanced by in-edges differed 2.0-2.4x in node count — the typed linears
were unbalanced and the exchange had to be a variable-size group of broadcasts / rooted reduces.
With a slice of every type each rank holds N/P rows (+-T) AND, because hub ids are not clustered
inside a type, ~E/P in-edges; the exchange is a plain equal-size all-gather / reduce-scatter, the
collectives NCCL (and NVSwitch in-network reduction) is fastest at.

Layout of the gathered matrix: rank q's rows at [q*max_rows, q*max_rows + n_q), type by type
(``max_rows`` = the largest n_q; the <= T padding rows per rank are zeros). Source ids of a shard's
edges are remapped into that layout once, at plan time. HGT gathers only ``h`` and computes K for
the gathered rows locally (round 1 gathered K and h: twice the bytes).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist

from .graphs import HeteroCOO


def type_slices(node_type_offsets: np.ndarray, world: int) -> np.ndarray:
    """cuts[t, p] .. cuts[t, p+1]: the id range of type t owned by rank p ([T, world+1], int64)."""
    offs = np.asarray(node_type_offsets, dtype=np.int64)
    T = offs.size - 1
    cuts = np.zeros((T, world + 1), dtype=np.int64)
    for t in range(T):
        n = offs[t + 1] - offs[t]
        cuts[t] = offs[t] + (n * np.arange(world + 1, dtype=np.int64)) // world
    return cuts


@dataclass
class ShardPlan:
    rank: int
    world: int
    cuts: np.ndarray               # [T, world+1] per-type ownership ranges (global ids)
    max_rows: int                  # rows of the largest shard = the per-rank slot of the gathered layout
    owned: np.ndarray              # [n_local] global ids of the owned nodes, in local order (type by type)
    src: np.ndarray                # local edges: source ids in the gathered layout
    dst: np.ndarray                # local edges: destination ids in local order
    etype: np.ndarray
    node_type_offsets: np.ndarray  # [T+1] type offsets of the owned rows (local ids)
    gathered_type_ptrs: np.ndarray  # [world*(T+1)+1] segments of the gathered layout: per rank T types + padding
    edges_per_rank: np.ndarray     # [world] in-edges of every shard (balance report)
    exchange: str = "all_gather"   # "all_gather": every row of every rank; "halo": only the rows a shard references
    halo_ids: np.ndarray = None    # [n_halo] rows of the all-gather layout this rank reads, ascending (owner-major)
    recv_counts: list = None       # [world] rows received from every rank (halo_ids grouped by owner)
    send_idx: np.ndarray = None    # local rows to send, grouped by receiving rank
    send_counts: list = None       # [world] rows sent to every rank
    segments_per_rank: int = 0     # typed-linear segments per owner block of the gathered matrix (T, or T+1 with padding)

    @property
    def num_local_nodes(self) -> int:
        return int(self.owned.size)

    @property
    def num_gathered_rows(self) -> int:
        return int(self.halo_ids.size) if self.exchange == "halo" else self.world * self.max_rows

    @property
    def gathered_offset(self) -> int:
        """Row of this rank's first node inside the gathered matrix (all-gather layout)."""
        return self.rank * self.max_rows

    @property
    def num_ntypes(self) -> int:
        return int(self.node_type_offsets.size - 1)


def gathered_ids(cuts: np.ndarray, max_rows: int, ids: np.ndarray) -> np.ndarray:
    """global node id -> row of the gathered matrix"""
    T, world = cuts.shape[0], cuts.shape[1] - 1
    ids = np.asarray(ids, dtype=np.int64)
    t = np.searchsorted(cuts[:, 0], ids, side="right") - 1                     # node type
    sizes = np.diff(cuts, axis=1)                                              # [T, world]
    before = np.concatenate([np.zeros((1, world), dtype=np.int64), np.cumsum(sizes, axis=0)[:-1]], axis=0)
    p = np.empty_like(ids)
    for tt in range(T):                                                        # owner rank inside the type
        m = t == tt
        if m.any():
            p[m] = np.searchsorted(cuts[tt], ids[m], side="right") - 1
    return p * max_rows + before[t, p] + (ids - cuts[t, p])


def make_shard(coo: HeteroCOO, rank: int, world: int, exchange: str = None) -> ShardPlan:
    """``exchange``: "all_gather" gathers every rank's rows with one equal-size all-gather; "halo"
    exchanges only the rows each shard's in-edges reference, with one all-to-all (at 8 ranks a shard of
    the mag-shaped graph reads 60 % of the nodes, at 4 ranks 76 %, at 2 ranks 90 %). Default (env
    HETB_EXCHANGE overrides): halo from 8 ranks up, where it is measured faster (8.1 vs 9.5 ms/step on
    HGT-mag; equal at 4 ranks), all-gather below."""
    import os
    exchange = exchange or os.environ.get("HETB_EXCHANGE") or ("halo" if world >= 8 else "all_gather")
    assert exchange in ("all_gather", "halo")
    cuts = type_slices(coo.node_type_offsets, world)
    T = cuts.shape[0]
    sizes = np.diff(cuts, axis=1)                       # [T, world]
    n_per_rank = sizes.sum(0)
    max_rows = int(n_per_rank.max())
    owned = np.concatenate([np.arange(cuts[t, rank], cuts[t, rank + 1], dtype=np.int64) for t in range(T)])
    g_dst = gathered_ids(cuts, max_rows, coo.dst)
    owner = g_dst // max_rows
    mine = owner == rank
    local_offs = np.zeros(T + 1, dtype=np.int64)
    np.cumsum(sizes[:, rank], out=local_offs[1:])
    ptrs = [0]
    for q in range(world):
        base = q * max_rows
        ptrs += [base + int(v) for v in np.cumsum(sizes[:, q])]
        ptrs.append(base + max_rows)                    # the padding rows of rank q (zero length when n_q == max_rows)
    plan = ShardPlan(rank, world, cuts, max_rows, owned,
                     gathered_ids(cuts, max_rows, coo.src[mine]), (g_dst[mine] - rank * max_rows).astype(np.int64),
                     coo.etype[mine].astype(np.int64), local_offs, np.asarray(ptrs, dtype=np.int64),
                     np.bincount(owner, minlength=world).astype(np.int64), segments_per_rank=T + 1)
    if exchange == "halo":
        # every (reading rank q, row of the all-gather layout) pair, once; sorted q-major, then owner-major
        ng = world * max_rows
        uniq = np.unique(owner * ng + gathered_ids(cuts, max_rows, coo.src))
        uq, ug = uniq // ng, uniq % ng
        uo = ug // max_rows                                  # owner of the row
        halo = ug[uq == rank]
        plan.exchange = "halo"
        plan.halo_ids = halo
        plan.recv_counts = np.bincount(uo[uq == rank], minlength=world).tolist()
        sends = uo == rank                                   # rows this rank owns, grouped by the rank that reads them
        plan.send_idx = (ug[sends] - rank * max_rows).astype(np.int64)
        plan.send_counts = np.bincount(uq[sends], minlength=world).tolist()
        plan.src = np.searchsorted(halo, plan.src).astype(np.int64)
        # (owner, type) segments of the halo table: owner blocks keep their rows type-sorted
        bounds = []
        for q in range(world):
            bounds += [q * max_rows + int(v) for v in np.concatenate([[0], np.cumsum(sizes[:, q])])[:-1]]
        bounds.append(ng)
        plan.gathered_type_ptrs = np.searchsorted(halo, np.asarray(bounds, dtype=np.int64)).astype(np.int64)
        plan.segments_per_rank = T
    return plan


def shard_graph(plan: ShardPlan, coo: HeteroCOO, device):
    from .graph import HetGraph
    g = HetGraph.from_arrays(
        f"{coo.name}[{plan.rank}/{plan.world}]", torch.from_numpy(plan.src), torch.from_numpy(plan.dst),
        torch.from_numpy(plan.etype), plan.num_gathered_rows, plan.num_local_nodes, coo.num_rels,
        torch.from_numpy(plan.node_type_offsets), device)
    g.rel_src_type = torch.from_numpy(coo.rel_src_type).to(device)
    g.rel_dst_type = torch.from_numpy(coo.rel_dst_type).to(device)
    # segments of the gathered matrix for typed linears on gathered rows: (T + 1) per rank, the last = padding
    g.gathered_type_ptrs = torch.from_numpy(plan.gathered_type_ptrs).to(device)
    g.num_ntypes_global = plan.num_ntypes
    g.world = plan.world
    g.gathered_segments_per_rank = plan.segments_per_rank
    return g


class _AllGatherRows(torch.autograd.Function):
    """[n_local, ...] -> [world*max_rows, ...] (rank p at row p*max_rows; padding rows zero);
    backward = reduce-scatter of the gradient of the gathered matrix."""

    @staticmethod
    def forward(ctx, x, max_rows, group):
        world = dist.get_world_size(group)
        ctx.n, ctx.max_rows, ctx.group = x.shape[0], max_rows, group
        if x.shape[0] == max_rows:
            part = x.contiguous()
        else:
            part = x.new_zeros((max_rows,) + tuple(x.shape[1:]))
            part[: x.shape[0]] = x
        out = x.new_empty((world * max_rows,) + tuple(x.shape[1:]))
        dist.all_gather_into_tensor(out, part, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous()
        mine = g.new_empty((ctx.max_rows,) + tuple(g.shape[1:]))
        if dist.get_backend(ctx.group) == "nccl":
            dist.reduce_scatter_tensor(mine, g, op=dist.ReduceOp.SUM, group=ctx.group)
        else:  # gloo (CPU tests) has no reduce_scatter
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=ctx.group)
            r = dist.get_rank(ctx.group)
            mine = g[r * ctx.max_rows:(r + 1) * ctx.max_rows]
        return mine[: ctx.n], None, None


class _HaloExchange(torch.autograd.Function):
    """[n_local, D] -> [n_halo, D]: the rows this rank's in-edges reference, owner by owner, through ONE
    all-to-all (``ncclSend/Recv`` group over NVLink); backward = the reverse all-to-all and a
    scatter-add of the returned gradient rows into the local gradient (a row is sent to every rank
    that reads it). Pack / unpack are this library's row-gather and row-scatter-add kernels."""

    @staticmethod
    def forward(ctx, x, send_idx, send_counts, recv_counts, group):
        ctx.group, ctx.send_counts, ctx.recv_counts, ctx.n = group, send_counts, recv_counts, x.shape[0]
        ctx.save_for_backward(send_idx)
        x = x.contiguous()
        send = _gather_rows(x, send_idx)
        out = x.new_empty((sum(recv_counts),) + tuple(x.shape[1:]))
        dist.all_to_all_single(out, send, recv_counts, send_counts, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        (send_idx,) = ctx.saved_tensors
        g = g.contiguous()
        back = g.new_empty((sum(ctx.send_counts),) + tuple(g.shape[1:]))
        dist.all_to_all_single(back, g, ctx.send_counts, ctx.recv_counts, group=ctx.group)
        grad = g.new_zeros((ctx.n,) + tuple(g.shape[1:]))
        _scatter_add_rows(grad, send_idx, back)
        return grad, None, None, None, None


def _gather_rows(x, idx):
    if x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.shape[1] % 4 == 0:
        from .kernels import K
        out = x.new_empty((idx.numel(), x.shape[1]))
        K.hetb200_gather_rows(idx, x, out)
        return out
    return x.index_select(0, idx)


def _scatter_add_rows(dst, idx, src):
    if dst.is_cuda and dst.dtype == torch.float32 and dst.dim() == 2 and dst.shape[1] % 4 == 0:
        from .kernels import K
        K.hetb200_scatter_add_rows(idx, src, dst)
    else:
        dst.index_add_(0, idx, src)


def all_gather_rows(x: torch.Tensor, max_rows: int, group=None) -> torch.Tensor:
    return _AllGatherRows.apply(x, max_rows, group)


def gather_rows(x: torch.Tensor, plan: "ShardPlan", group=None) -> torch.Tensor:
    """The layer-input exchange."""
    if plan.exchange == "halo":
        if getattr(plan, "_send_idx_t", None) is None or plan._send_idx_t.device != x.device:
            plan._send_idx_t = torch.from_numpy(plan.send_idx).to(x.device)
        return _HaloExchange.apply(x, plan._send_idx_t, plan.send_counts, plan.recv_counts, group)
    return all_gather_rows(x, plan.max_rows, group)


class RowExchange:
    """The layer-input exchange of one shard as an object: calling it is the autograd form
    (``gather_rows``); the ``*_start`` / ``*_finish`` pairs are the same collectives issued asynchronously
    (on NCCL's stream, ordered after the work already queued on the current stream) so that a caller can
    put independent kernels between the two — ``backend.HGTReorderedProjectionsSharded`` runs the
    destination-side typed linear under the forward exchange and the weight-gradient kernels under the
    backward one. No autograd inside: the caller is an autograd node."""

    def __init__(self, plan: "ShardPlan", group=None):
        self.plan, self.group = plan, group

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        return gather_rows(x, self.plan, self.group)

    def _send_idx(self, device):
        plan = self.plan
        if getattr(plan, "_send_idx_t", None) is None or plan._send_idx_t.device != device:
            plan._send_idx_t = torch.from_numpy(plan.send_idx).to(device)
        return plan._send_idx_t

    def forward_start(self, x: torch.Tensor):
        plan = self.plan
        x = x.contiguous()
        if plan.exchange == "halo":
            send = _gather_rows(x, self._send_idx(x.device))
            out = x.new_empty((sum(plan.recv_counts),) + tuple(x.shape[1:]))
            work = dist.all_to_all_single(out, send, plan.recv_counts, plan.send_counts, group=self.group, async_op=True)
            return work, out, send
        if x.shape[0] == plan.max_rows:
            part = x
        else:
            part = x.new_zeros((plan.max_rows,) + tuple(x.shape[1:]))
            part[: x.shape[0]] = x
        out = x.new_empty((plan.world * plan.max_rows,) + tuple(x.shape[1:]))
        work = dist.all_gather_into_tensor(out, part, group=self.group, async_op=True)
        return work, out, part

    @staticmethod
    def forward_finish(handle) -> torch.Tensor:
        work, out, _keep = handle
        work.wait()
        return out

    def backward_start(self, g: torch.Tensor):
        plan = self.plan
        g = g.contiguous()
        if plan.exchange == "halo":
            back = g.new_empty((sum(plan.send_counts),) + tuple(g.shape[1:]))
            work = dist.all_to_all_single(back, g, plan.send_counts, plan.recv_counts, group=self.group, async_op=True)
            return work, back, g
        if dist.get_backend(self.group) == "nccl":
            mine = g.new_empty((plan.max_rows,) + tuple(g.shape[1:]))
            work = dist.reduce_scatter_tensor(mine, g, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            return work, mine, g
        work = dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self.group, async_op=True)  # gloo has no reduce_scatter
        r = dist.get_rank(self.group)
        return work, g[r * plan.max_rows:(r + 1) * plan.max_rows], g

    def backward_finish(self, handle, grad_x: torch.Tensor) -> None:
        """Adds the exchanged gradient rows into ``grad_x`` ([n_local, D])."""
        work, buf, _keep = handle
        work.wait()
        if self.plan.exchange == "halo":
            _scatter_add_rows(grad_x, self._send_idx(grad_x.device), buf)
        else:
            grad_x += buf[: grad_x.shape[0]]


class LocalRowExchange(RowExchange):
    """Single-process stand-in with the all-gather layout (tests, tools/profile_shard.py): the own rows at
    their place, the other ranks' rows zero; backward = the own rows of the gradient."""

    def __init__(self, plan: "ShardPlan"):
        assert plan.exchange == "all_gather"
        super().__init__(plan, None)

    def _place(self, x):
        out = x.new_zeros((self.plan.num_gathered_rows,) + tuple(x.shape[1:]))
        lo = self.plan.rank * self.plan.max_rows
        out[lo: lo + x.shape[0]] = x
        return out

    def __call__(self, x):
        return self._place(x)  # slice assignment into zeros: differentiable

    def forward_start(self, x):
        return None, self._place(x.detach()), None

    @staticmethod
    def forward_finish(handle):
        return handle[1]

    def backward_start(self, g):
        return None, g, None

    def backward_finish(self, handle, grad_x):
        lo = self.plan.rank * self.plan.max_rows
        grad_x += handle[1][lo: lo + grad_x.shape[0]]


def all_reduce_grads(params, group=None) -> None:
    """One flat all-reduce (sum) of every parameter gradient (replicated weights)."""
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for g in grads:
        g.copy_(flat[off:off + g.numel()].view_as(g))
        off += g.numel()
