"""Micro-benchmark of hetb_segment_mm (through the torch_hrt ops): SIMT vs tcgen05 path.
Usage (GPU box): python tools/bench_segmm.py [bgs|mag|all]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200.kernels import K

dev = torch.device("cuda:0")
flush = torch.zeros(64 * 1024 * 1024, device=dev)


def timeit(fn, reps=10):
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


def case(name, N, E, R, H, Kd, Nd, head_one, zipf=1.0):
    g = torch.Generator().manual_seed(0)
    w = 1.0 / torch.arange(1, R + 1, dtype=torch.float64) ** zipf
    sizes = torch.clamp((w / w.sum() * E).floor().long(), min=1)
    sizes[0] += E - sizes.sum()
    rel_ptrs = torch.cat([torch.zeros(1, dtype=torch.int64), sizes.cumsum(0)]).to(dev)
    gather = torch.randint(0, N, (E,), generator=g).to(dev)
    eids = torch.arange(E, device=dev)
    a_heads = 1 if head_one else H
    X = torch.randn(N, a_heads, Kd, device=dev)
    W = torch.randn(R, H, Kd, Nd, device=dev) / Kd ** 0.5
    ret = torch.empty(E, H, Nd, device=dev)
    args = {"separate_coo_rel_ptrs": rel_ptrs, "separate_coo_node_indices": gather, "separate_coo_eids": eids}
    uniq = int(torch.unique(gather).numel())
    nbytes = uniq * a_heads * Kd * 4 + E * H * Nd * 4 + R * H * Kd * Nd * 4 + E * 8
    res = {}
    PROBES = [(1, "simt"), (2, "tc")] + [(16 * ((m << 4) | ps), f"tc p{ps} mask{m}") for ps, m in
              ((3, 2), (3, 4), (3, 6), (3, 8), (3, 14), (0, 14))]
    for impl, nm in PROBES:
        K.hetb200_set_segment_mm_impl(impl)
        ms = timeit(lambda: K.rgnn_relational_matmul(args, 0, W, X, ret, head_one))
        res[nm] = ms
        print(f"{name:28s} fwd {nm:12s} {ms:8.4f} ms  {nbytes / ms / 1e6:8.1f} GB/s (algorithmic)")
    # backward (dX accumulate + dW)
    gout = torch.randn(E, H, Nd, device=dev)
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    WT = W.transpose(2, 3).contiguous()
    for impl, nm in ((1, "simt"), (2, "tc")):
        K.hetb200_set_segment_mm_impl(impl)
        ms = timeit(lambda: K.backward_rgnn_relational_matmul(args, 0, WT, X, gout, gX, gW, head_one))
        print(f"{name:28s} bwd {nm:5s} {ms:8.4f} ms")
    # the two halves of the backward on their own, through the C ABI (ctypes on libhetb200.so)
    import ctypes
    from het_b200.kernels import KERNEL_LIB
    lib = ctypes.CDLL(str(KERNEL_LIB))
    vp, i64, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    lib.hetb_segment_mm_grad_w.argtypes = [vp, vp, ci, vp, vp, ci, vp, vp, i64, i64, ci, ci, ci, vp]
    lib.hetb_segment_mm.argtypes = [vp, vp, ci, vp, vp, vp, ci, vp, i64, i64, ci, ci, ci, ci, vp]
    st = torch.cuda.current_stream().cuda_stream
    rows_bytes = E * (a_heads * Kd + H * Nd) * 4
    for impl, nm in ((1, "simt"), (2, "tc"), (304, "tc raw-hi"), (48 + 512, "tc noMMA"), (48 + 1024, "tc noLoad"),
                     (48 + 2048, "tc noSplit"), (48 + 256 * 14, "tc skeleton"), (48 + 256 * 30, "tc skel-nocp"), (48 + 256 * 16, "tc nocp")):
        K.hetb200_set_segment_mm_impl(impl)
        ms = timeit(lambda: lib.hetb_segment_mm_grad_w(X.data_ptr(), gather.data_ptr(), a_heads, gout.data_ptr(),
                                                       eids.data_ptr(), H, gW.data_ptr(), rel_ptrs.data_ptr(), R, E,
                                                       H, Kd, Nd, st))
        print(f"{name:28s} dW  {nm:10s} {ms:8.4f} ms  {rows_bytes / ms / 1e6:8.1f} GB/s (rows read: gathered X + dY)")
    for impl, nm in ((1, "simt"), (2, "tc")):
        K.hetb200_set_segment_mm_impl(impl)
        ms = timeit(lambda: lib.hetb_segment_mm(gout.data_ptr(), eids.data_ptr(), H, WT.data_ptr(), gX.data_ptr(),
                                                gather.data_ptr(), a_heads, rel_ptrs.data_ptr(), R, E, H, Nd, Kd, 1, st))
        print(f"{name:28s} dX  {nm:10s} {ms:8.4f} ms")
    K.hetb200_set_segment_mm_impl(0)
    # reference point: a plain copy of the output-sized tensor (read + write E*D*4 each)
    src = torch.empty(E, H * Nd, device=dev)
    ms = timeit(lambda: ret.view(E, -1).copy_(src))
    print(f"{name:28s} copy_ of the output tensor {ms:8.4f} ms  {2 * E * H * Nd * 4 / ms / 1e6:8.1f} GB/s")


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("bgs", "all"):
    case("bgs rgat X.W (64x64)", 94000, 670000, 207, 1, 64, 64, True)
if which in ("mag", "all"):
    case("mag hgt q.W_att (4x16x16)", 1939700, 21100000, 8, 4, 16, 16, False, zipf=0.3)
    case("mag dense 64x64", 1939700, 21100000, 8, 1, 64, 64, True, zipf=0.3)
