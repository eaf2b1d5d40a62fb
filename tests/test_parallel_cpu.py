"""CPU (gloo, world_size 2): host logic of the destination-range sharding — the partition,
the source-id remap into the gathered layout, and the all-gather / reduce-scatter autograd
pair — checked against the unsharded computation."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_partition_balances_in_edges():
    from het_b200 import graphs, parallel
    coo = graphs.bgs_shaped(scale=0.05)
    for world in (2, 4, 8):
        b = parallel.partition_destinations(coo.dst, coo.num_nodes, world)
        assert b[0] == 0 and b[-1] == coo.num_nodes and np.all(np.diff(b) > 0)
        per = np.array([((coo.dst >= b[p]) & (coo.dst < b[p + 1])).sum() for p in range(world)])
        assert per.sum() == coo.num_edges
        hub = np.bincount(coo.dst).max()
        assert per.max() <= coo.num_edges / world + hub + 1  # balanced up to one hub row


def test_shards_cover_every_edge_once_and_remap_sources():
    from het_b200 import graphs, parallel
    coo = graphs.mag_shaped(scale=0.002)
    world = 3
    seen = 0
    for r in range(world):
        g = parallel.make_shard(coo, r, world)  # default layout: global ids, nothing padded
        mine = (coo.dst >= g.lo) & (coo.dst < g.hi)
        assert g.layout == "global" and np.array_equal(g.src, coo.src[mine])
        assert g.num_gathered_rows == coo.num_nodes and sum(g.sizes) == coo.num_nodes
        assert g.gathered_offset == g.lo
        p = parallel.make_shard(coo, r, world, layout="padded")
        seen += p.src.size
        assert p.dst.min() >= 0 and p.dst.max() < p.num_local_nodes
        # padded source id -> (owner, local row) -> global id round-trips
        owner, local = p.src // p.max_rows, p.src % p.max_rows
        glob = p.bounds[owner] + local
        mine = (coo.dst >= p.lo) & (coo.dst < p.hi)
        assert np.array_equal(glob, coo.src[mine])
        assert np.all(local < np.diff(p.bounds)[owner])
        # per-shard node type offsets are the global ones clipped to the range
        assert p.node_type_offsets[0] == 0 and p.node_type_offsets[-1] == p.num_local_nodes
    assert seen == coo.num_edges


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q, layout="global"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from het_b200 import graphs, parallel
        coo = graphs.aifb_shaped(scale=0.2)
        plan = parallel.make_shard(coo, rank, world, layout=layout)
        torch.manual_seed(0)
        x_full = torch.randn(coo.num_nodes, 8, dtype=torch.float64)
        w = torch.randn(8, 4, dtype=torch.float64, requires_grad=True)
        x_loc = x_full[plan.lo:plan.hi].clone().requires_grad_(True)
        xg = parallel.gather_rows(x_loc, plan)
        assert xg.shape[0] == plan.num_gathered_rows
        # a toy "layer": out[dst] += (x[src] @ w), loss = sum(out * c)
        src, dst = torch.from_numpy(plan.src), torch.from_numpy(plan.dst)
        out = torch.zeros(plan.num_local_nodes, 4, dtype=torch.float64).index_add(0, dst, xg[src] @ w)
        c = torch.arange(coo.num_nodes * 4, dtype=torch.float64).view(-1, 4)[plan.lo:plan.hi] / 100
        (out * c).sum().backward()
        parallel.all_reduce_grads([w])
        # unsharded reference
        xf = x_full.clone().requires_grad_(True)
        wf = w.detach().clone().requires_grad_(True)
        of = torch.zeros(coo.num_nodes, 4, dtype=torch.float64).index_add(
            0, torch.from_numpy(coo.dst), xf[torch.from_numpy(coo.src)] @ wf)
        (of * (torch.arange(coo.num_nodes * 4, dtype=torch.float64).view(-1, 4) / 100)).sum().backward()
        ok = (torch.allclose(out, of[plan.lo:plan.hi].detach()) and torch.allclose(w.grad, wf.grad)
              and torch.allclose(x_loc.grad, xf.grad[plan.lo:plan.hi]))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
@pytest.mark.parametrize("layout", ["global", "padded"])
def test_all_gather_rows_autograd_gloo_world2(layout):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, layout)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
        assert p.exitcode == 0
    res = sorted(q.get(timeout=5) for _ in range(2))
    assert res == [(0, True), (1, True)]
