"""One-off probe (GPU): which rows does the reference's outer-product (dW) SGEMM kernel actually sum?
Prints the relative error of the reference dW against candidate models, for even / odd input-dim rows."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import ref_kernels as RK
from oracle import het_ops_ref as R

dev = torch.device("cuda")
torch.manual_seed(0)
n, e, r, K_, N_ = 300, 5000, 7, 64, 64
sizes = torch.tensor([700, 900, 33, 1200, 5, 1500, 662])
rel_ptrs = torch.cat([torch.zeros(1, dtype=torch.long), sizes.cumsum(0)]).to(dev)
gather = torch.randint(0, n, (e,), device=dev)
eids = torch.randperm(e, device=dev)
W = torch.randn(r, 1, K_, N_, device=dev) * 0.2
X = torch.randn(n, K_, device=dev)
gy = torch.randn(e, 1, N_, device=dev)
gx_r, gw_r = RK.relational_matmul_bwd(rel_ptrs, gather, eids, W, X, gy, True)
gw_r = gw_r.double().cpu()[:, 0]
a = X.cpu().double()[gather.cpu()]
g = gy.cpu().double()[eids.cpu()][:, 0]
ptr = rel_ptrs.cpu()


def model(frac_rule):
    out = torch.zeros(r, K_, N_, dtype=torch.float64)
    for s in range(r):
        lo, hi = int(ptr[s]), int(ptr[s + 1])
        m = frac_rule(hi - lo)
        out[s] = a[lo:lo + m].T @ g[lo:lo + m]
    return out


cands = {
    "all rows": model(lambda L: L),
    "first ceil(L/32)*8 rows": model(lambda L: min(L, -(-L // 32) * 8)),
    "first ceil(L/8) * 8 /4": model(lambda L: min(L, (-(-L // 8) * 8) // 4)),
}
for name, c in cands.items():
    for par, sl in (("even k", slice(0, None, 2)), ("odd k", slice(1, None, 2))):
        err = (gw_r[:, sl] - c[:, sl]).abs().max() / c[:, sl].abs().max()
        print(f"{name:28s} {par}: rel err {float(err):.3e}")
    per_rel = [(float((gw_r[s, 0::2] - c[s, 0::2]).abs().max() / (c[s, 0::2].abs().max() + 1e-30))) for s in range(r)]
    print("   per relation (even k):", " ".join(f"{v:.1e}" for v in per_rel))
