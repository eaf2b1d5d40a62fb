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


def test_type_slices_balance_nodes_and_edges():
    """Every rank owns a slice of every node type: N/P rows (+-T) and, with hub ids spread inside a
    type, ~E/P in-edges (up to the largest hubs)."""
    from het_b200 import graphs, parallel
    coo = graphs.mag_shaped(scale=0.01)
    for world in (2, 4, 8):
        cuts = parallel.type_slices(coo.node_type_offsets, world)
        assert np.array_equal(cuts[:, 0], coo.node_type_offsets[:-1]) and np.array_equal(cuts[:, -1], coo.node_type_offsets[1:])
        assert np.all(np.diff(cuts, axis=1) >= 0)
        n_per = np.diff(cuts, axis=1).sum(0)
        assert n_per.sum() == coo.num_nodes and n_per.max() - n_per.min() <= coo.num_ntypes
        plan = parallel.make_shard(coo, 0, world)
        per = plan.edges_per_rank
        assert per.sum() == coo.num_edges
        hub = np.bincount(coo.dst).max()
        assert per.max() <= 1.25 * coo.num_edges / world + 4 * hub


def test_shards_cover_every_edge_once_and_remap_sources():
    from het_b200 import graphs, parallel
    coo = graphs.mag_shaped(scale=0.002)
    world = 3
    seen = 0
    owned_all = []
    for r in range(world):
        p = parallel.make_shard(coo, r, world)
        seen += p.src.size
        owned_all.append(p.owned)
        assert p.dst.min() >= 0 and p.dst.max() < p.num_local_nodes
        # every local edge's destination is owned, and maps back to the global edge list
        mine = np.isin(coo.dst, p.owned)
        assert mine.sum() == p.src.size
        assert np.array_equal(p.owned[p.dst], coo.dst[mine])
        # gathered source id -> (owner, local row) -> global id round-trips
        owner, local = p.src // p.max_rows, p.src % p.max_rows
        glob = np.empty_like(p.src)
        for q in range(world):
            pq = parallel.make_shard(coo, q, world)
            m = owner == q
            assert np.all(local[m] < pq.num_local_nodes)
            glob[m] = pq.owned[local[m]]
        assert np.array_equal(glob, coo.src[mine])
        # local rows are sorted by type: per-shard type offsets partition them
        offs = p.node_type_offsets
        assert offs[0] == 0 and offs[-1] == p.num_local_nodes
        for t in range(coo.num_ntypes):
            ids = p.owned[offs[t]:offs[t + 1]]
            assert np.all((ids >= coo.node_type_offsets[t]) & (ids < coo.node_type_offsets[t + 1]))
        # segments of the gathered layout: per rank T types + its padding, covering world * max_rows rows
        gp = p.gathered_type_ptrs
        assert gp.size == world * (coo.num_ntypes + 1) + 1 and gp[0] == 0 and gp[-1] == p.num_gathered_rows
        assert np.all(np.diff(gp) >= 0)
    assert seen == coo.num_edges
    assert np.array_equal(np.sort(np.concatenate(owned_all)), np.arange(coo.num_nodes))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q, exchange="all_gather"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from het_b200 import graphs, parallel
        coo = graphs.aifb_shaped(scale=0.2)
        plan = parallel.make_shard(coo, rank, world, exchange=exchange)
        torch.manual_seed(0)
        x_full = torch.randn(coo.num_nodes, 8, dtype=torch.float64)
        w = torch.randn(8, 4, dtype=torch.float64, requires_grad=True)
        own = torch.from_numpy(plan.owned)
        x_loc = x_full[own].clone().requires_grad_(True)
        xg = parallel.gather_rows(x_loc, plan)
        assert xg.shape[0] == plan.num_gathered_rows
        # a toy "layer": out[dst] += (x[src] @ w), loss = sum(out * c)
        src, dst = torch.from_numpy(plan.src), torch.from_numpy(plan.dst)
        out = torch.zeros(plan.num_local_nodes, 4, dtype=torch.float64).index_add(0, dst, xg[src] @ w)
        c = torch.arange(coo.num_nodes * 4, dtype=torch.float64).view(-1, 4)[own] / 100
        (out * c).sum().backward()
        parallel.all_reduce_grads([w])
        # unsharded reference
        xf = x_full.clone().requires_grad_(True)
        wf = w.detach().clone().requires_grad_(True)
        of = torch.zeros(coo.num_nodes, 4, dtype=torch.float64).index_add(
            0, torch.from_numpy(coo.dst), xf[torch.from_numpy(coo.src)] @ wf)
        (of * (torch.arange(coo.num_nodes * 4, dtype=torch.float64).view(-1, 4) / 100)).sum().backward()
        ok = (torch.allclose(out, of[own].detach()) and torch.allclose(w.grad, wf.grad)
              and torch.allclose(x_loc.grad, xf.grad[own]))
        # the split (asynchronous) form of the same exchange gives the same rows and the same gradient
        xch = parallel.RowExchange(plan)
        xg2 = xch.forward_finish(xch.forward_start(x_loc.detach()))
        ok = ok and torch.equal(xg2, xg.detach())
        gg = torch.randn(xg.shape, dtype=torch.float64, generator=torch.Generator().manual_seed(5 + rank))
        (want,) = torch.autograd.grad(parallel.gather_rows(x_loc, plan), x_loc, gg.clone())  # gloo reduces in place
        got = torch.zeros_like(want)
        xch.backward_finish(xch.backward_start(gg.clone()), got)
        ok = ok and torch.allclose(got, want)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_halo_plan_is_consistent_across_ranks():
    """exchange="halo": what rank p sends to rank q is what q expects from p, the halo table holds
    exactly the rows q's edges reference (ascending in the all-gather layout), and the remapped
    source ids point at the right global nodes."""
    from het_b200 import graphs, parallel
    coo = graphs.mag_shaped(scale=0.002)
    world = 4
    plans = [parallel.make_shard(coo, r, world, exchange="halo") for r in range(world)]
    full = [parallel.make_shard(coo, r, world, exchange="all_gather") for r in range(world)]
    for q in range(world):
        pq = plans[q]
        assert pq.exchange == "halo" and pq.num_gathered_rows == pq.halo_ids.size == sum(pq.recv_counts)
        assert np.all(np.diff(pq.halo_ids) > 0)
        assert np.array_equal(np.unique(full[q].src), pq.halo_ids)           # exactly the referenced rows
        assert np.array_equal(pq.halo_ids[pq.src], full[q].src)              # remap round-trips
        off = 0
        for p in range(world):
            n = pq.recv_counts[p]
            assert plans[p].send_counts[q] == n
            so = int(np.sum(plans[p].send_counts[:q]))
            sent_local = plans[p].send_idx[so:so + n]
            assert np.array_equal(sent_local + p * pq.max_rows, pq.halo_ids[off:off + n])
            off += n
        gp = pq.gathered_type_ptrs
        assert gp.size == world * coo.num_ntypes + 1 and gp[0] == 0 and gp[-1] == pq.halo_ids.size
        assert np.all(np.diff(gp) >= 0) and pq.segments_per_rank == coo.num_ntypes
        # every halo row of segment (owner p, type t) is a node of type t
        owned = [plans[p].owned for p in range(world)]
        for p in range(world):
            for t in range(coo.num_ntypes):
                ids = pq.halo_ids[gp[p * coo.num_ntypes + t]:gp[p * coo.num_ntypes + t + 1]]
                glob = owned[p][ids - p * pq.max_rows]
                assert np.all((glob >= coo.node_type_offsets[t]) & (glob < coo.node_type_offsets[t + 1]))


@pytest.mark.timeout(180)
@pytest.mark.parametrize("exchange", ["all_gather", "halo"])
def test_all_gather_rows_autograd_gloo_world2(exchange):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, exchange)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
        assert p.exitcode == 0
    res = sorted(q.get(timeout=5) for _ in range(2))
    assert res == [(0, True), (1, True)]
