"""A/B of the 64-wide tcgen05 typed linear with the A tile delivered by TMA (csrc/segment_mm_tc_tma.cu) against
the register-path kernel (csrc/segment_mm_tc.cu), through the C ABI: first that both give the same bits on a set
of shapes (contiguous / gathered rows, scatter, accumulation, block-diagonal and stacked weights, bf16 output,
ragged tails), then the time of the mag-sized launches of one HGT layer.
Usage (GPU box): python tools/bench_segmm_tma.py [check|time|all]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from het_b200.kernels import KERNEL_LIB

dev = torch.device("cuda:0")
lib = ctypes.CDLL(str(KERNEL_LIB))
vp, i64, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
lib.hetb_segment_mm.argtypes = [vp, vp, ci, vp, vp, vp, ci, vp, i64, i64, ci, ci, ci, ci, vp]
lib.hetb_segment_mm_bf16out.argtypes = [vp, vp, ci, vp, vp, vp, vp, i64, i64, ci, ci, ci, vp]


def ptr(t):
    return None if t is None else t.data_ptr()


def run(A, a_idx, a_heads, W, Y, c_idx, c_heads, seg, rows, H, K, N, acc, bf16=False):
    st = torch.cuda.current_stream().cuda_stream
    if bf16:
        rc = lib.hetb_segment_mm_bf16out(ptr(A), ptr(a_idx), a_heads, ptr(W), ptr(Y), ptr(c_idx), ptr(seg),
                                         seg.numel() - 1, rows, H, K, N, st)
    else:
        rc = lib.hetb_segment_mm(ptr(A), ptr(a_idx), a_heads, ptr(W), ptr(Y), ptr(c_idx), c_heads, ptr(seg),
                                 seg.numel() - 1, rows, H, K, N, acc, st)
    assert rc == 0, f"rc {rc}"


def seg_ptrs(rows, nseg, g, zipf=0.7):
    w = 1.0 / torch.arange(1, nseg + 1, dtype=torch.float64) ** zipf
    sizes = (w / w.sum() * rows).floor().long()
    sizes[0] += rows - sizes.sum()
    sizes = sizes[torch.randperm(nseg, generator=g)]
    return torch.cat([torch.zeros(1, dtype=torch.int64), sizes.cumsum(0)]).to(dev)


def check():
    g = torch.Generator().manual_seed(1)
    lib.hetb_set_segment_mm_impl(2)  # tensor-core kernels at every size
    # (table rows, rows, segments, gather, scatter, accumulate, mode, bf16)
    cases = [
        (5000, 5000, 4, False, False, 0, "dense", False),
        (5000, 4999, 1, False, False, 0, "dense", False),
        (300, 100, 3, False, False, 0, "dense", False),
        (70000, 70001, 37, False, False, 0, "blockdiag", False),
        (9000, 9000, 5, False, False, 1, "dense", False),
        (9000, 9000, 5, False, True, 0, "dense", False),
        (3000, 20000, 11, True, False, 0, "dense", False),
        (3000, 20003, 11, True, False, 0, "blockdiag", False),
        (3000, 20003, 207, True, True, 1, "stacked", False),
        (3000, 20003, 11, True, True, 1, "blockdiag", False),
        (3000, 33, 2, True, False, 0, "dense", False),
        (3000, 20003, 11, True, False, 0, "blockdiag", True),
        (6000, 6000, 3, False, False, 0, "dense", True),
        (400000, 1000000, 8, True, False, 0, "heads_out", False),
    ]
    for table, rows, nseg, gather, scatter, acc, mode, bf16 in cases:
        seg = seg_ptrs(rows, nseg, g)
        A = torch.randn(table, 64, device=dev)
        a_idx = torch.randint(0, table, (rows,), generator=g).to(dev) if gather else None
        if mode == "dense":
            H, K, N, ah, ch = 1, 64, 64, 1, 1
        elif mode == "blockdiag":
            H, K, N, ah, ch = 4, 16, 16, 4, 4
        elif mode == "heads_out":
            H, K, N, ah, ch = 4, 64, 16, 1, 4
        else:  # stacked: per-head input rows summed into one 64-wide output row
            H, K, N, ah, ch = 4, 16, 64, 4, 1
        W = torch.randn(nseg, H, K, N, device=dev) / K ** 0.5
        out_rows = table if scatter else rows
        c_idx = torch.randperm(out_rows, generator=g)[:rows].to(dev) if scatter and not acc else (
            torch.randint(0, out_rows, (rows,), generator=g).to(dev) if scatter else None)
        outs = []
        for tma in (0, 1):
            lib.hetb_set_segment_mm_tma(tma)
            dt = torch.bfloat16 if bf16 else torch.float32
            Y = torch.zeros(out_rows, 64, device=dev, dtype=dt) if acc else torch.full((out_rows, 64), float("nan"), device=dev, dtype=dt)
            run(A, a_idx, ah, W, Y, c_idx, ch, seg, rows, H, K, N, acc, bf16)
            torch.cuda.synchronize()
            outs.append(Y.float())
        a, b = outs
        same = torch.equal(torch.nan_to_num(a, nan=-7.0), torch.nan_to_num(b, nan=-7.0))
        err = (torch.nan_to_num(a) - torch.nan_to_num(b)).abs().max().item()
        ok = same if not (acc and scatter) else err <= 1e-5 * a.abs().max().item()  # red.add order differs
        print(f"table {table:7d} rows {rows:8d} segs {nseg:4d} gather {int(gather)} scatter {int(scatter)} acc {acc} "
              f"{mode:9s} bf16 {int(bf16)}: {'bit-identical' if same else f'max diff {err:.3e}'} {'OK' if ok else 'FAIL'}",
              flush=True)
        assert ok
    lib.hetb_set_segment_mm_impl(0)
    lib.hetb_set_segment_mm_tma(1)
    print("TMA A/B CHECK OK")


flush = None


def timeit(fn, reps=10):
    global flush
    if flush is None:
        flush = torch.zeros(64 * 1024 * 1024, device=dev)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(reps):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def time_cases():
    g = torch.Generator().manual_seed(2)
    N = 1939700
    types = torch.tensor([0, 736389, 1870635, 1879375, N], dtype=torch.int64, device=dev)  # mag node types
    A = torch.randn(N, 64, device=dev)
    Wt = torch.randn(4, 1, 64, 64, device=dev) / 8
    Y = torch.empty(N, 64, device=dev)
    U = 4460000
    seg8 = seg_ptrs(U, 8, g, zipf=0.3)
    # compact (relation, source) rows: sorted by source inside a relation, as the plan lists them
    a_idx = torch.cat([torch.sort(torch.randint(0, N, (int(seg8[i + 1] - seg8[i]),), generator=g))[0] for i in range(8)]).to(dev)
    Wr = torch.randn(8, 4, 16, 16, device=dev) / 4
    Yu = torch.empty(U, 64, device=dev)
    Yb = torch.empty(U, 64, device=dev, dtype=torch.bfloat16)
    G = torch.randn(U, 64, device=dev)
    dA = torch.zeros(N, 64, device=dev)
    rows = [
        ("contiguous N rows, dense, plain store", lambda: run(A, None, 1, Wt, Y, None, 1, types, N, 1, 64, 64, 0), 2 * N * 256),
        ("contiguous N rows, dense, accumulate", lambda: run(A, None, 1, Wt, Y, None, 1, types, N, 1, 64, 64, 1), 3 * N * 256),
        ("gathered 4.46 M rows, block-diag, plain store", lambda: run(A, a_idx, 4, Wr, Yu, None, 4, seg8, U, 4, 16, 16, 0), N * 256 + U * 264),
        ("gathered 4.46 M rows, block-diag, bf16 store", lambda: run(A, a_idx, 4, Wr, Yb, None, 4, seg8, U, 4, 16, 16, 0, True), N * 256 + U * 136),
        ("delta-X: contiguous 4.46 M rows, scatter red.add into N", lambda: run(G, None, 4, Wr, dA, a_idx, 4, seg8, U, 4, 16, 16, 1), U * 264 + 2 * N * 256),
    ]
    for name, fn, nbytes in rows:
        res = []
        for tma in (0, 1):
            lib.hetb_set_segment_mm_tma(tma)
            res.append(timeit(fn))
        print(f"{name:58s} registers {res[0]:7.4f} ms ({nbytes / res[0] / 1e6:6.0f} GB/s)   TMA {res[1]:7.4f} ms "
              f"({nbytes / res[1] / 1e6:6.0f} GB/s)   x{res[0] / res[1]:.3f}", flush=True)
    lib.hetb_set_segment_mm_tma(1)


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("check", "all"):
    check()
if which in ("time", "all"):
    time_cases()
