"""Destination-range sharding of full-graph training over the GPUs of one box.

The reference has no distributed code at all; this is the one
parallelism BASELINE.json's north star adds: every rank owns a contiguous range
of destination nodes (balanced by in-edge count), all in-edges of those
destinations, the output rows and the per-edge tensors of those edges. Weights
are replicated. Per layer there is ONE exchange step: the rows a layer gathers by
*source* id (RGAT: the layer input; HGT: K and V) are all-gathered over NCCL
(NVLink 5 / NVSwitch); its backward is the matching reduce-scatter. Weight
gradients are all-reduced once per step.

Layout of the gathered matrix (``layout`` of make_shard):
* "global" (default): the gathered matrix is the full [N, D] matrix in global node-id order; rank
  p contributes rows [bounds[p], bounds[p+1]) through a variable-size all-gather (one coalesced
  group of NCCL broadcasts straight into the row slices, no staging copy). Destination ranges
  balanced by in-edges have very different node counts on a power-law graph (2.0-2.4x between the
  largest range and the mean at 4-8 ranks), so nothing is padded and the collective moves N rows.
* "padded": rank p's rows sit at [p*max_rows, p*max_rows+n_p), one ``all_gather_into_tensor``;
  the shard's source indices are remapped into that layout once, at plan time. Moves
  world*max_rows rows; kept for equal-sized ranges and as the fallback.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist

from .graphs import HeteroCOO


def partition_destinations(dst: np.ndarray, num_nodes: int, world: int) -> np.ndarray:
    """bounds[world+1]: contiguous destination ranges with ~equal in-edge counts
    (every range non-empty when num_nodes >= world)."""
    indeg = np.bincount(dst, minlength=num_nodes).astype(np.int64)
    csum = np.concatenate([[0], np.cumsum(indeg)])
    total = int(csum[-1])
    bounds = np.zeros(world + 1, dtype=np.int64)
    bounds[world] = num_nodes
    for p in range(1, world):
        b = int(np.searchsorted(csum, total * p / world, side="left"))
        bounds[p] = min(max(b, bounds[p - 1] + 1), num_nodes - (world - p))
    return bounds


@dataclass
class ShardPlan:
    rank: int
    world: int
    bounds: np.ndarray      # [world+1] destination ranges
    max_rows: int           # rows of the largest range (the per-rank slot of the padded layout)
    lo: int
    hi: int
    src: np.ndarray         # local edges: source ids in the gathered layout
    dst: np.ndarray         # local edges: destination ids relative to lo
    etype: np.ndarray
    node_type_offsets: np.ndarray  # [T+1] offsets of the owned rows, relative to lo
    layout: str = "global"

    @property
    def num_local_nodes(self) -> int:
        return self.hi - self.lo

    @property
    def sizes(self) -> list:
        return [int(v) for v in np.diff(self.bounds)]

    @property
    def num_gathered_rows(self) -> int:
        return int(self.bounds[-1]) if self.layout == "global" else self.world * self.max_rows

    @property
    def gathered_offset(self) -> int:
        """Row of this rank's first node inside the gathered matrix."""
        return self.lo if self.layout == "global" else self.rank * self.max_rows


def make_shard(coo: HeteroCOO, rank: int, world: int, layout: str = None) -> ShardPlan:
    import os
    layout = layout or os.environ.get("HETB_GATHER_LAYOUT", "global")
    assert layout in ("global", "padded")
    bounds = partition_destinations(coo.dst, coo.num_nodes, world)
    sizes = np.diff(bounds)
    max_rows = int(sizes.max())
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    mine = (coo.dst >= lo) & (coo.dst < hi)
    src_g = coo.src[mine]
    if layout == "padded":
        owner = np.searchsorted(bounds, src_g, side="right") - 1
        src_p = owner * max_rows + (src_g - bounds[owner])
    else:
        src_p = src_g
    offs = np.clip(coo.node_type_offsets, lo, hi) - lo
    return ShardPlan(rank, world, bounds, max_rows, lo, hi, src_p.astype(np.int64),
                     (coo.dst[mine] - lo).astype(np.int64), coo.etype[mine].astype(np.int64),
                     offs.astype(np.int64), layout)


def shard_graph(plan: ShardPlan, coo: HeteroCOO, device):
    from .graph import HetGraph
    g = HetGraph.from_arrays(
        f"{coo.name}[{plan.rank}/{plan.world}]", torch.from_numpy(plan.src), torch.from_numpy(plan.dst),
        torch.from_numpy(plan.etype), plan.num_gathered_rows, plan.num_local_nodes, coo.num_rels,
        torch.from_numpy(plan.node_type_offsets), device)
    g.rel_src_type = torch.from_numpy(coo.rel_src_type).to(device)
    g.rel_dst_type = torch.from_numpy(coo.rel_dst_type).to(device)
    return g


class _AllGatherRows(torch.autograd.Function):
    """[n_local, ...] -> [world*max_rows, ...] (rank p at row p*max_rows); backward is the
    reduce-scatter of the gradient of the gathered matrix."""

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


class _AllGatherRowsV(torch.autograd.Function):
    """[n_p, ...] on rank p -> [sum n_p, ...] in rank order (variable sizes); backward is the
    matching variable-size reduce-scatter. NCCL runs both as one coalesced group of
    broadcasts / reduces on the row slices of the big tensor (no staging copy)."""

    @staticmethod
    def forward(ctx, x, sizes, group):
        ctx.sizes, ctx.group = list(sizes), group
        rank = dist.get_rank(group)
        assert x.shape[0] == ctx.sizes[rank]
        out = x.new_empty((sum(ctx.sizes),) + tuple(x.shape[1:]))
        if dist.get_backend(group) == "nccl":
            dist.all_gather(list(out.split(ctx.sizes)), x.contiguous(), group=group)
        else:  # gloo (CPU tests): equal-size all_gather of padded parts, then slice
            m = max(ctx.sizes)
            part = x.new_zeros((m,) + tuple(x.shape[1:]))
            part[: x.shape[0]] = x
            parts = [torch.empty_like(part) for _ in ctx.sizes]
            dist.all_gather(parts, part, group=group)
            for dst, src, n in zip(out.split(ctx.sizes), parts, ctx.sizes):
                dst.copy_(src[:n])
        return out

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous()
        rank = dist.get_rank(ctx.group)
        if dist.get_backend(ctx.group) == "nccl":
            mine = g.new_empty((ctx.sizes[rank],) + tuple(g.shape[1:]))
            dist.reduce_scatter(mine, list(g.split(ctx.sizes)), op=dist.ReduceOp.SUM, group=ctx.group)
        else:  # gloo has no reduce_scatter
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=ctx.group)
            mine = g.split(ctx.sizes)[rank]
        return mine, None, None


def all_gather_rows(x: torch.Tensor, max_rows: int, group=None) -> torch.Tensor:
    return _AllGatherRows.apply(x, max_rows, group)


def all_gather_rows_v(x: torch.Tensor, sizes, group=None) -> torch.Tensor:
    return _AllGatherRowsV.apply(x, sizes, group)


def gather_rows(x: torch.Tensor, plan: "ShardPlan", group=None) -> torch.Tensor:
    """The layer-input exchange for ``plan``'s layout."""
    if plan.layout == "global":
        return all_gather_rows_v(x, plan.sizes, group)
    return all_gather_rows(x, plan.max_rows, group)


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
