"""Row-dot (N = 1) and expand (K = 1) typed linears at wikikg2 size: ms and GB/s per launch.
usage (GPU box): python tools/bench_rowdot.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200.kernels import K

dev = torch.device("cuda")
M, N_nodes, R = 13_839_842, 2_500_000, 535
flush = torch.zeros(64 << 20, device=dev)


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


w = torch.rand(R, device=dev).pow(3)
sizes = (w / w.sum() * M).long()
sizes[0] += M - sizes.sum()
ptrs = torch.cat([torch.zeros(1, dtype=torch.long, device=dev), sizes.cumsum(0)])
x = torch.randn(M, 64, device=dev)
wl = torch.randn(R, 1, 64, 1, device=dev)
y = torch.empty(M, 1, device=dev)
ms = timeit(lambda: K.rgnn_relational_matmul_no_scatter_gather_list(ptrs, wl, x, y))
print(f"rowdot contiguous  {M} rows: {ms:.3f} ms  {M * 260 / ms / 1e6:.0f} GB/s")
xn = torch.randn(N_nodes, 1, 64, device=dev)
idx = torch.randint(0, N_nodes, (M,), device=dev)
args = {"unique_srcs_and_dests_rel_ptrs": ptrs, "unique_srcs_and_dests_node_indices": idx}
y3 = torch.empty(M, 1, 1, device=dev)
ms = timeit(lambda: K.rgnn_relational_matmul(args, 1, wl, xn, y3, True))
print(f"rowdot gathered    {M} rows: {ms:.3f} ms  {M * 268 / ms / 1e6:.0f} GB/s")
