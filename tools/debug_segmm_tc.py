"""One-off probe (GPU): tcgen05 segment MM vs SIMT on the wikikg2-shaped unique (relation, source) list:
which rows differ, and where they sit relative to 128-row tiles and segment boundaries."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from het_b200 import graphs
from het_b200.graph import HetGraph
from het_b200.kernels import K

dev = torch.device("cuda")
coo = graphs.SHAPES["wikikg2"](scale=1.0 / 16)
g = HetGraph(coo, dev)
u = g.get_separate_unique_node_indices_single_sided()
ptr, nodes = u["rel_ptrs_row"], u["node_indices_row"]
U = nodes.numel()
torch.manual_seed(0)
W = torch.randn(coo.num_rels, 1, 64, 64, device=dev) * 0.2
gy = torch.randn(U, 1, 64, device=dev)
ident = torch.arange(U, device=dev)
d = {"unique_srcs_and_dests_rel_ptrs": ptr, "unique_srcs_and_dests_node_indices": ident}
res = {}
for mode, accumulate_target in (("store-like (identity scatter)", ident), ("scatter to nodes", nodes)):
    outs = []
    for impl in (1, 2):
        K.hetb200_set_segment_mm_impl(impl)
        for rep in range(3 if impl == 2 else 1):
            gx = torch.zeros(U if accumulate_target is ident else coo.num_nodes, 64, device=dev)
            gw = torch.zeros_like(W)
            dd = {"unique_srcs_and_dests_rel_ptrs": ptr, "unique_srcs_and_dests_node_indices": accumulate_target}
            torch.manual_seed(5)
            X = torch.randn(gx.shape[0], 64, device=dev)
            K.backward_rgnn_relational_matmul(dd, 1, W.transpose(2, 3).contiguous(), X, gy, gx, gw, True)
            torch.cuda.synchronize()
            outs.append(gx.double().cpu())
            gws = globals().setdefault('gws', {}); gws.setdefault(mode, []).append(gw.double().cpu())
    ref = outs[0]
    for i, o in enumerate(outs[1:]):
        err = (o - ref).abs().max(1).values
        bad = torch.nonzero(err > 1e-4 * ref.abs().max()).view(-1)
        print(f"{mode}: TC run {i}: max err {float(err.max()):.3e}, bad rows {bad.tolist()[:12]}")
        if accumulate_target is ident:
            p = ptr.cpu()
            for b in bad.tolist()[:6]:
                s = int(torch.searchsorted(p, torch.tensor(b), right=True)) - 1
                print(f"    row {b}: tile {b // 128} offset {b % 128}; segment {s} = [{int(p[s])}, {int(p[s+1])}) "
                      f"len {int(p[s+1]-p[s])}; tiles of CTA: num_tiles {-(-U // 128)}")
for mode, l in gws.items():
    for i, o in enumerate(l[1:]):
        e = (o - l[0]).abs()
        print(f"{mode}: dW TC run {i} vs SIMT: max err {float(e.max()):.3e} vs max {float(l[0].abs().max()):.3e}; worst relation {int(e.view(e.shape[0], -1).max(1).values.argmax())}")
K.hetb200_set_segment_mm_impl(0)
print("U", U, "segments", ptr.numel() - 1, "min/median seg", int((ptr[1:] - ptr[:-1]).min()), int((ptr[1:] - ptr[:-1]).median()))
