"""Wait-time accounting of the tcgen05 weight-gradient kernel: which role waits on which
barrier for how long (cycles, per warp, averaged over CTAs). The role that hardly waits is the
bottleneck. Usage (GPU box): python tools/prof_gradw.py [impl codes...]"""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200.kernels import K, KERNEL_LIB

dev = torch.device("cuda:0")
lib = ctypes.CDLL(str(KERNEL_LIB))
vp, i64, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
lib.hetb_segment_mm_grad_w.argtypes = [vp, vp, ci, vp, vp, ci, vp, vp, i64, i64, ci, ci, ci, vp]
lib.hetb_debug_set_gradw_buffer.argtypes = [vp]
N, E, R = 1939700, 21100000, 8
g = torch.Generator().manual_seed(0)
w = 1.0 / torch.arange(1, R + 1, dtype=torch.float64) ** 0.3
sizes = torch.clamp((w / w.sum() * E).floor().long(), min=1); sizes[0] += E - sizes.sum()
rel_ptrs = torch.cat([torch.zeros(1, dtype=torch.int64), sizes.cumsum(0)]).to(dev)
gather = torch.randint(0, N, (E,), generator=g).to(dev)
eids = torch.arange(E, device=dev)
X = torch.randn(N, 64, device=dev); G = torch.randn(E, 64, device=dev)
gW = torch.zeros(R, 1, 64, 64, device=dev)
st = torch.cuda.current_stream().cuda_stream
NW = 13
dbg = torch.zeros(148 * NW * 5, dtype=torch.int64, device=dev)
names = ["flush0", "flush1", "flush2", "flush3", "mma"] + [f"prod{i}" for i in range(8)]
for impl in [int(a) for a in sys.argv[1:]] or [2]:
    K.hetb200_set_segment_mm_impl(impl)
    call = lambda: lib.hetb_segment_mm_grad_w(X.data_ptr(), gather.data_ptr(), 1, G.data_ptr(), eids.data_ptr(), 1,
                                              gW.data_ptr(), rel_ptrs.data_ptr(), R, E, 1, 64, 64, st)
    for _ in range(2):
        call()
    lib.hetb_debug_set_gradw_buffer(dbg.data_ptr())
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); call(); b.record(); torch.cuda.synchronize()
    lib.hetb_debug_set_gradw_buffer(None)
    d = dbg.view(148, NW, 5).double().mean(0)
    stages = E / 148 / 64
    print(f"impl {impl}: {a.elapsed_time(b):.3f} ms; cycles/stage {d[4,0] / stages:.0f}")
    for i, nm in enumerate(names):
        if nm not in ("flush0", "mma", "prod0", "prod7"):
            continue
        print(f"   {nm:7s} total {d[i,0]:10.0f}  wait0 {d[i,1] / d[i,0] * 100:5.1f}%  store {d[i,2] / d[i,0] * 100:5.1f}%  issue {d[i,3] / d[i,0] * 100:5.1f}%  fence {d[i,4] / d[i,0] * 100:5.1f}%")
K.hetb200_set_segment_mm_impl(0)
