"""Debug helper for the tcgen05 weight-gradient kernel: structured inputs, prints how the
output relates to the expected X^T G."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200.kernels import K
torch.set_printoptions(linewidth=200, precision=1, sci_mode=False)
dev = torch.device("cuda:0")
K.hetb200_set_segment_mm_impl(int(sys.argv[1]) if len(sys.argv) > 1 else 2)


def run(X, G):
    M = X.shape[0]
    offs = torch.tensor([0, M], dtype=torch.int64, device=dev)
    W = torch.zeros(1, 1, 64, 64, device=dev)
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    K.backward_rgnn_relational_matmul_no_scatter_gather_list(offs, W, X, G, gX, gW)
    torch.cuda.synchronize()
    return gW[0, 0]


def report(name, got, want):
    print(f"== {name}: got nonzeros {int((got != 0).sum())}, want nonzeros {int((want != 0).sum())}, "
          f"max|got| {got.abs().max():.3f} max|want| {want.abs().max():.3f} "
          f"err {(got - want).abs().max():.3e} errT {(got - want.t()).abs().max():.3e}")
    print("got[:4,:10]\n", got[:4, :10].cpu())
    print("got[32:36,:10]\n", got[32:36, :10].cpu())
    nz = (got != 0).nonzero()
    if len(nz):
        print("first nonzero idx", nz[:6].tolist(), "last", nz[-3:].tolist())


f = torch.arange(1, 65, device=dev, dtype=torch.float32)
for r in (0, 1, 9, 31):
    X = torch.zeros(32, 64, device=dev); G = torch.zeros(32, 64, device=dev)
    X[r] = f; G[r] = f * 64
    report(f"row {r} only", run(X, G), X.t() @ G)
for rows, r in ((64, 5), (64, 40), (64, 63), (128, 70), (256, 200), (512, 300)):
    X = torch.zeros(rows, 64, device=dev); G = torch.zeros(rows, 64, device=dev)
    X[r] = f; G[r] = f * 64
    report(f"{rows} rows, row {r} only", run(X, G), X.t() @ G)
X = torch.zeros(32, 64, device=dev); G = torch.zeros(32, 64, device=dev)
X[:, 5] = 1.0; G[:, 7] = 1.0
report("col 5 x col 7 ones", run(X, G), X.t() @ G)
torch.manual_seed(0)
for rows in (64, 128, 192, 256, 512, 1024):
    X = torch.randn(rows, 64, device=dev); G = torch.randn(rows, 64, device=dev)
    report(f"random {rows}", run(X, G), (X.double().t() @ G.double()).float())
X = torch.randn(4000, 64, device=dev); G = torch.randn(4000, 64, device=dev)
report("random 4000", run(X, G), (X.double().t() @ G.double()).float())
