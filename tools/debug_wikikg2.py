"""One-off probe (GPU): localise the wikikg2-shaped RGAT grad-x mismatch between the tcgen05 and the SIMT segment MM."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200 import graphs, backend as B
from het_b200.graph import HetGraph
from het_b200.kernels import K
from het_b200.models import HET_RGATModel
from oracle.testing import unit_gain_init

dev = torch.device("cuda")
coo = graphs.SHAPES["wikikg2"](scale=1.0 / 16)
g = HetGraph(coo, dev)
u = g.get_separate_unique_node_indices_single_sided()
torch.manual_seed(3)
model = HET_RGATModel(coo.num_rels, 64, 64, 64, 1, 1, compact_as_of_node_flag=True, linear_reorder_flag=True).to(dev)
unit_gain_init(model, seed=5)
x0 = (torch.rand(coo.num_nodes, 64, device=dev) * 2 - 1)
cap = {}
orig = B.relational_matmul_compact_pair


def patched(gr, x, w_row, w_col):
    a, b = orig(gr, x, w_row, w_col)
    cap["feat"], cap["er"] = a.detach().clone(), b.detach().clone()
    a.register_hook(lambda t: cap.__setitem__("g_feat", t.detach().clone()))
    b.register_hook(lambda t: cap.__setitem__("g_er", t.detach().clone()))
    return a, b


B.relational_matmul_compact_pair = patched
runs = {}
gout = None
for impl in (1, 0):
    K.hetb200_set_segment_mm_impl(impl)
    x = x0.clone().requires_grad_(True)
    for p in model.parameters():
        p.grad = None
    out = model(g, x)
    if gout is None:
        torch.manual_seed(9)
        gout = torch.randn_like(out)
    out.backward(gout)
    runs[impl] = dict(cap, out=out.detach().clone(), gx=x.grad.clone(), gw=model.layers[0].conv_weights.grad.clone(),
                      gal=model.layers[0].attn_l.grad.clone(), gar=model.layers[0].attn_r.grad.clone())
    cap = {}
    B.relational_matmul_compact_pair = patched
for k in runs[1]:
    a, b = runs[1][k].double(), runs[0][k].double()
    e = (a - b).abs()
    rows = torch.nonzero(e.reshape(e.shape[0], -1).max(1).values > 1e-4 * a.abs().max()).view(-1)
    print(f"{k:7s}: max |SIMT - TC| {float(e.max()):.3e} vs max {float(a.abs().max()):.3e}; rows off: {rows.tolist()[:8]} ({rows.numel()})")
    if k in ("feat", "g_feat") and rows.numel():
        p = u["rel_ptrs_row"].cpu()
        for r in rows.tolist()[:4]:
            s = int(torch.searchsorted(p, torch.tensor(r), right=True)) - 1
            print(f"     compact row {r}: node {int(u['node_indices_row'][r])} rel {s} seg [{int(p[s])},{int(p[s+1])}) tile {r//128} off {r%128}")
K.hetb200_set_segment_mm_impl(0)
