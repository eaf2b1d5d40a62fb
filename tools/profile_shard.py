"""Per-shard cost of the destination-range sharded HGT / RGAT step, measured on ONE GPU: builds the shard
of every rank of a `world`-way partition in turn, runs the layer stack forward + backward on it with a local
stand-in for the all-gather (zero-padded rows), and prints the time per rank and the top kernels of the
slowest one. Usage (GPU box): python tools/profile_shard.py hgt 4 [scale]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from het_b200 import graphs, parallel
from het_b200.models import HET_HGT_DGLHetero, HET_RGATModel

kind = sys.argv[1] if len(sys.argv) > 1 else "hgt"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 4
scale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
dev = torch.device("cuda:0")
coo = graphs.mag_shaped(scale=scale) if kind == "hgt" else graphs.bgs_shaped(scale=scale)
torch.manual_seed(0)
if kind == "hgt":
    model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, 64, 64, 4, 2, compact_as_of_node_flag=True,
                              fused_attention_aggregation_flag=True, linear_reorder_flag=True).to(dev)
else:
    model = HET_RGATModel(coo.num_rels, 64, 64, 64, 1, 1, compact_as_of_node_flag=True, linear_reorder_flag=True).to(dev)


def run(rank, profile=False):
    plan = parallel.make_shard(coo, rank, world, exchange="all_gather")
    g = parallel.shard_graph(plan, coo, dev)
    u = g.get_separate_unique_node_indices_single_sided()
    n = plan.num_local_nodes
    x = (torch.rand(n, 64, device=dev) * 2 - 1).requires_grad_(True)

    gather = parallel.LocalRowExchange(plan)  # stand-in for the exchange: own rows at their place, zeros elsewhere

    def step():
        x.grad = None
        out = model(g, x, gather)
        out.sum().backward()

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        step()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    print(f"rank {rank}/{world}: {ms:8.2f} ms  local nodes {n}  edges {g.num_edges}  max_rows {plan.max_rows}  "
          f"U_row {u['node_indices_row'].numel()}  U_col {u['node_indices_col'].numel()}", flush=True)
    if profile:
        from torch.profiler import ProfilerActivity, profile as tprofile
        with tprofile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=60), flush=True)
    return ms


if len(sys.argv) > 4 and sys.argv[4] == "one":   # rank 0 only, with the kernel table
    run(0, profile=True)
else:
    times = [run(r) for r in range(world)]
    worst = max(range(world), key=lambda r: times[r])
    run(worst, profile=True)
