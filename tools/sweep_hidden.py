"""Typed linear (rgnn_relational_matmul_no_scatter_gather_list) at hidden 64 / 128 / 256 / 512 on the
mag-shaped node set (1.94 M rows, 4 types): ms, TFLOP/s, % of the measured tensor peak, GB/s, % of the
measured HBM peak. hidden 64 runs the resident-B tcgen05 kernel, wider shapes the streamed-operand one
(both 3xTF32: 3 MMAs per useful product) — the line says which.
usage (GPU box): python tools/sweep_hidden.py [rows] [hidden,hidden,...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200.kernels import K

dev = torch.device("cuda")
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_939_700
peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
except OSError:
    pass
hbm, tc = peaks.get("hbm_gbs", 6650.0), peaks.get("bf16_tflops_sustained", 1400.0)
flush = torch.zeros(64 * 1024 * 1024, device=dev)
offs = torch.tensor([0, rows * 38 // 100, rows * 96 // 100, rows * 97 // 100, rows], device=dev)


def timeit(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


hids = [int(v) for v in sys.argv[2].split(',')] if len(sys.argv) > 2 else [64, 128, 256, 512]
for hid in hids:
    n = rows if hid <= 128 else rows // (hid // 128)   # keep the tensors under ~4 GB
    o = (offs.double() * n / rows).long()
    o[-1] = n
    x = torch.randn(n, hid, device=dev)
    w = torch.randn(4, 1, hid, hid, device=dev) / hid ** 0.5
    y = torch.empty(n, hid, device=dev)
    ms = timeit(lambda: K.rgnn_relational_matmul_no_scatter_gather_list(o, w, x, y))
    flops = 2.0 * n * hid * hid
    nbytes = 2.0 * n * hid * 4
    path = "tcgen05 3xTF32, B resident (segment_mm_tc.cu)" if hid == 64 else "tcgen05 3xTF32, streamed operands (segment_mm_tc_wide.cu)"
    # the tensor pipe executes 3 tf32 MMAs per useful product; tf32 dense peak = half the bf16 figure
    exec_tf = 3 * flops / ms / 1e9
    print(f"hidden {hid:4d} rows {n:8d}: {ms:7.3f} ms  {flops / ms / 1e9:7.1f} TFLOP/s useful, {exec_tf:7.1f} executed (3xTF32) = "
          f"{100 * exec_tf / (tc / 2):5.1f}% of the tf32 peak ({tc / 2:.0f} TF/s = half the measured bf16 sustained)  "
          f"{nbytes / ms / 1e6:7.0f} GB/s = {100 * nbytes / ms / 1e6 / hbm:5.1f}% of HBM  [{path}]", flush=True)
