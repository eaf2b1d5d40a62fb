"""How the tcgen05 typed linear behaves at BGS size (2e5 rows): one segment vs 207 equal segments vs
207 Zipf-sized segments, cold (L2 flushed) and warm. Usage (GPU box): python tools/bench_segmm_small.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200.kernels import K

dev = torch.device("cuda:0")
flush = torch.zeros(64 * 1024 * 1024, device=dev)


def timeit(fn, cold, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(reps):
        if cold:
            flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2] * 1e3


def case(name, N, M, R, zipf):
    w = 1.0 / torch.arange(1, R + 1, dtype=torch.float64) ** zipf
    sizes = torch.clamp((w / w.sum() * M).floor().long(), min=1)
    sizes[0] += M - sizes.sum()
    rel_ptrs = torch.cat([torch.zeros(1, dtype=torch.int64), sizes.cumsum(0)]).to(dev)
    g = torch.Generator().manual_seed(0)
    nodes = torch.cat([torch.sort(torch.randint(0, N, (int(s),), generator=g)).values for s in sizes]).to(dev)
    X = torch.randn(N, 64, device=dev)
    W = torch.randn(R, 1, 64, 64, device=dev) / 8
    ret = torch.empty(M, 1, 64, device=dev)
    args = {"unique_srcs_and_dests_rel_ptrs": rel_ptrs, "unique_srcs_and_dests_node_indices": nodes}
    gout = torch.randn(M, 1, 64, device=dev)
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    WT = W.transpose(2, 3).contiguous()
    for cold in (True, False):
        f = timeit(lambda: K.rgnn_relational_matmul(args, 1, W, X, ret, True), cold)
        b = timeit(lambda: K.backward_rgnn_relational_matmul(args, 1, WT, X, gout, gX, gW, True), cold)
        print(f"{name:34s} {'cold' if cold else 'warm'}  fwd {f:7.1f} us   bwd(dX+dW) {b:7.1f} us", flush=True)


for M in (199_000, 1_000_000):
    case(f"M={M} 1 segment", 94_000, M, 1, 0.0)
    case(f"M={M} 207 equal segments", 94_000, M, 207, 0.0)
    case(f"M={M} 207 Zipf(1.0) segments", 94_000, M, 207, 1.0)
