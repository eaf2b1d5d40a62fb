"""Layer-level parity: HET_RGATLayer / HET_HGTLayerHetero forward + backward on the
GPU (torch_hrt ops) against the differentiable fp64 CPU restatement of the same
layer on the integrated COO (oracle/models_ref.py). Checks the output, the input
gradient and EVERY parameter gradient."""
import pytest
import torch

from oracle import models_ref as M
from oracle.testing import close, unit_gain_init

pytestmark = pytest.mark.gpu

# Tolerances, all relative to the tensor's own max|ref| (tests.test_gpu_ops.close has no floor):
# outputs at the north star's rtol 1e-4 (+1e-5 of the scale for elements that cancel to ~0);
# gradients w.r.t. activations 1e-4 + 2e-5 of the scale; gradients of weights are fp32 sums over
# 1e3..1e5 rows whose summation order differs from the fp64 oracle's: 2e-4 + 3e-5 of the scale.
GRAD_TOL = dict(rtol=1e-4, atol=2e-5)
WGRAD_TOL = dict(rtol=2e-4, atol=3e-5)


def _graph(cuda, shape, scale):
    from het_b200 import graphs
    from het_b200.graph import HetGraph
    coo = graphs.SHAPES[shape](scale=scale)
    return coo, HetGraph(coo, cuda)


def _coo_t(coo):
    return torch.from_numpy(coo.src), torch.from_numpy(coo.dst), torch.from_numpy(coo.etype)


def _rgat_oracle(g, src, dst, et, x, layers, gout):
    """fp64 oracle of a stack of RGAT layers (ReLU between them). For the few edges whose logit is
    within 2e-5 of the leaky-ReLU kink the oracle takes the branch the fp32 layer took
    (HET_RGATLayer.logit_signs; see oracle/models_ref.py: rgat_layer). Returns the output, the input
    gradient and the per-layer parameter gradients."""
    x64 = x.detach().double().cpu().requires_grad_(True)
    h, ps = x64, []
    h32 = x.detach()
    for i, layer in enumerate(layers):
        branch = layer.logit_signs(g, h32)[g.orig_to_canonical_eid].cpu()   # [E, H] in the order of src / dst
        p = {k: v.detach().double().cpu().requires_grad_(True) for k, v in layer.named_parameters()}
        ps.append(p)
        h = M.rgat_layer(src, dst, et, h, p["conv_weights"], p["attn_l"], p["attn_r"], 0.2, p["h_bias"], branch=branch)
        if i + 1 < len(layers):
            h = torch.relu(h)
            with torch.no_grad():
                h32 = torch.relu(layer(g, h32))
    h.backward(gout.double().cpu())
    return h.detach(), x64.grad, [{k: v.grad for k, v in p.items()} for p in ps]


@pytest.mark.parametrize("shape,scale,heads,edge_parallel,compact", [
    ("aifb", 1.0, 1, False, False), ("bgs", 0.1, 1, False, False), ("bgs", 0.05, 4, False, False),
    ("aifb", 0.5, 1, True, False),
    # compact_as_of_node_flag: X.W_r once per unique (relation, node) pair, direct indexing
    ("aifb", 1.0, 1, False, True), ("bgs", 0.1, 1, False, True), ("bgs", 0.05, 4, False, True),
    # + linear operator reordering of the destination logits: er = X.(W_r.a_r)
    ("aifb", 1.0, 1, False, "reorder"), ("bgs", 0.1, 1, False, "reorder"), ("bgs", 0.05, 4, False, "reorder"),
])
def test_rgat_layer_fwd_bwd(cuda, shape, scale, heads, edge_parallel, compact):
    from het_b200.models import HET_RGATLayer
    coo, g = _graph(cuda, shape, scale)
    torch.manual_seed(0)
    layer = HET_RGATLayer(64, 64, coo.num_rels, heads, gat_edge_parallel_flag=edge_parallel,
                          compact_as_of_node_flag=bool(compact),
                          linear_reorder_flag=(compact == "reorder")).to(cuda)
    unit_gain_init(layer)
    x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
    out = layer(g, x)
    gout = torch.randn_like(out)
    out.backward(gout)

    src, dst, et = _coo_t(coo)
    ref, gx, gp = _rgat_oracle(g, src, dst, et, x, [layer], gout)
    close(out, ref, "rgat out")
    close(x.grad, gx, "rgat grad x", **GRAD_TOL)
    for k, v in layer.named_parameters():
        close(v.grad, gp[0][k], f"rgat grad {k}", **WGRAD_TOL)


@pytest.mark.parametrize("scale,heads,layers,compact", [
    (0.002, 4, 1, False), (0.005, 1, 1, False), (0.002, 4, 2, False),
    (0.002, 4, 1, True), (0.005, 1, 1, True), (0.002, 4, 2, True),   # compact_as_of_node_flag
    (0.002, 4, 1, 2), (0.005, 1, 1, 2), (0.002, 4, 2, 2),            # + fused score/softmax/aggregation pass
    (0.002, 4, 1, 3), (0.005, 1, 1, 3), (0.002, 4, 2, 3),            # + linear operator reordering (W_q.W_att, W_v.W_msg)
])
def test_hgt_layer_fwd_bwd(cuda, scale, heads, layers, compact):
    from het_b200.models import HET_HGT_DGLHetero
    coo, g = _graph(cuda, "mag", scale)
    torch.manual_seed(1)
    model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, 64, 64, heads, num_layers=layers,
                              compact_as_of_node_flag=bool(compact),
                              fused_attention_aggregation_flag=(compact in (2, 3)),
                              linear_reorder_flag=(compact == 3)).to(cuda)
    unit_gain_init(model)
    x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
    out = model(g, x)
    gout = torch.randn_like(out)
    out.backward(gout)

    src, dst, et = _coo_t(coo)
    offs = torch.from_numpy(coo.node_type_offsets)
    x64 = x.detach().double().cpu().requires_grad_(True)
    h = x64
    ps = []
    for layer in model.layers:
        p = {k: v.detach().double().cpu().requires_grad_(True) for k, v in layer.named_parameters()}
        ps.append(p)
        h = M.hgt_layer(src, dst, et, offs, h, p["k_linears"], p["q_linears"], p["v_linears"], p["a_linears"],
                        p["relation_pri"], p["relation_att"], p["relation_msg"], p["skip"])
    h.backward(gout.double().cpu())
    assert float(h.abs().max()) > 0.05, "test signal too small to detect a wrong kernel"
    close(out, h, "hgt out")
    close(x.grad, x64.grad, "hgt grad x", **GRAD_TOL)
    for layer, p in zip(model.layers, ps):
        for k, v in layer.named_parameters():
            assert float(p[k].grad.abs().max()) > 1e-3, f"gradient of {k} too small to test"
            close(v.grad, p[k].grad, f"hgt grad {k}", **WGRAD_TOL)


def test_unique_node_indices_single_sided(cuda):
    """The compact-as-of-node lists (per relation: sorted unique sources / destinations, and the
    edge -> compact row map) are bit-exact against the numpy restatement of the reference's
    per-relation torch.unique (hrt/python/utils_lite/mydgl_graph_methods.py:11-157)."""
    from oracle import converters_ref as C
    coo, g = _graph(cuda, "bgs", 0.05)
    u = g.get_separate_unique_node_indices_single_sided()
    sc = g.get_separate_coo_original()
    ref = C.unique_node_indices_single_sided_ref(sc["rel_ptrs"].cpu().numpy(), sc["row_indices"].cpu().numpy(),
                                                 sc["col_indices"].cpu().numpy())
    for k, v in ref.items():
        assert torch.equal(u[k].cpu(), torch.from_numpy(v)), k


def test_rgat_compact_backward_row_walker_vs_edge_parallel(cuda):
    """The compact RGAT backward has two implementations (walker over compact source rows, default;
    edge-parallel over the out-CSR): same gradients on the same inputs."""
    from het_b200 import backend as B
    from het_b200.models import HET_RGATLayer
    coo, g = _graph(cuda, "bgs", 0.2)
    grads = []
    for by_rows in (True, False):
        B.RGAT_COMPACT_BACKWARD_BY_ROWS = by_rows
        try:
            torch.manual_seed(0)
            layer = HET_RGATLayer(64, 64, coo.num_rels, 4, compact_as_of_node_flag=True, linear_reorder_flag=True).to(cuda)
            unit_gain_init(layer)
            x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
            out = layer(g, x)
            torch.manual_seed(1)
            out.backward(torch.randn_like(out))
            grads.append([x.grad] + [p.grad for p in layer.parameters()])
        finally:
            B.RGAT_COMPACT_BACKWARD_BY_ROWS = True
    for a, b in zip(*grads):
        close(a, b, "row walker vs edge parallel", **WGRAD_TOL)


@pytest.mark.parametrize("shape,scale,mode,num_bases", [
    ("aifb", 1.0, "per_edge", None), ("bgs", 0.1, "per_edge", None),
    ("aifb", 1.0, "compact_coo", None), ("bgs", 0.1, "compact_coo", 8),
    ("aifb", 1.0, "compact_csr", None), ("bgs", 0.1, "compact_csr", None), ("mag", 0.002, "compact_csr", 4),
])
def test_rgcn_layer_fwd_bwd(cuda, shape, scale, mode, num_bases):
    """HET_EglRelGraphConv_EdgeParallel (RGCN) against the DGL RelGraphConv restatement
    in fp64 (oracle/models_ref.py: dgl_rgcn_layer): output, input gradient, weight / basis gradients
    and (per-edge path, the only one whose reference op returns it) the edge-norm gradient."""
    from het_b200.models import HET_EglRelGraphConv_EdgeParallel
    coo, g = _graph(cuda, shape, scale)
    torch.manual_seed(0)
    layer = HET_EglRelGraphConv_EdgeParallel(
        64, 64, coo.num_rels, coo.num_edges, num_bases=num_bases, compact_as_of_node_flag=(mode != "per_edge"),
        compact_direct_indexing_flag=True, csr_aggregation_flag=(mode == "compact_csr")).to(cuda)
    unit_gain_init(layer)
    x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
    # norm is indexed by (canonical) edge id = separate-COO position
    norm = (torch.rand(coo.num_edges, 1, device=cuda) * 0.9 + 0.1).requires_grad_(mode == "per_edge")
    out = layer(g, x, norm)
    gout = torch.randn_like(out)
    out.backward(gout)

    sc = g.get_separate_coo_original()
    rel_ptrs = sc["rel_ptrs"].cpu()
    et = torch.repeat_interleave(torch.arange(coo.num_rels), rel_ptrs[1:] - rel_ptrs[:-1])
    src, dst = sc["row_indices"].cpu(), sc["col_indices"].cpu()
    p = {k: v.detach().double().cpu().requires_grad_(True) for k, v in layer.named_parameters()}
    w = p["weight"]
    if "w_comp" in p:
        w = (p["w_comp"] @ w.view(w.shape[0], -1)).view(coo.num_rels, 64, 64)
    x64 = x.detach().double().cpu().requires_grad_(True)
    n64 = norm.detach().double().cpu().view(-1).requires_grad_(True)
    ref = M.dgl_rgcn_layer(src, dst, et, x64, w, n64) + p["h_bias"]
    ref.backward(gout.double().cpu())
    close(out, ref, "rgcn out")
    close(x.grad, x64.grad, "rgcn grad x", **GRAD_TOL)
    for k, v in layer.named_parameters():
        close(v.grad, p[k].grad, f"rgcn grad {k}", **WGRAD_TOL)
    if mode == "per_edge":
        close(norm.grad.view(-1), n64.grad, "rgcn grad norm", **GRAD_TOL)


def test_hgt_fused_on_padded_shard_sources(cuda):
    """A destination shard indexes its sources in the all-gather layout: every other rank's rows are
    there, and this rank's edges reference only some of them. The plan-based fused path (one
    exchange, k computed on the gathered rows) must give what the op-by-op path gives on the same shard."""
    from het_b200 import graphs, parallel
    from het_b200.models import HET_HGT_DGLHetero
    coo = graphs.SHAPES["mag"](scale=0.005)
    plan = parallel.make_shard(coo, 3, 4)
    g = parallel.shard_graph(plan, coo, cuda)
    n = plan.num_local_nodes
    assert plan.num_gathered_rows > 3 * n  # sources index the gathered matrix: long runs of rows without out-edges

    gather = parallel.LocalRowExchange(plan)  # own rows at their place, the other ranks' rows zero

    outs = []
    for fused in (True, False):
        torch.manual_seed(2)
        model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, 64, 64, 4, num_layers=1, compact_as_of_node_flag=True,
                                  fused_attention_aggregation_flag=fused, linear_reorder_flag=fused).to(cuda)
        unit_gain_init(model)
        torch.manual_seed(3)
        x = (torch.rand(n, 64, device=cuda) * 2 - 1).requires_grad_(True)
        out = model(g, x, gather)
        torch.manual_seed(4)
        out.backward(torch.randn_like(out))
        outs.append([out.detach(), x.grad] + [p.grad for p in model.parameters()])
    for a, b in zip(*outs):
        close(a, b.double().cpu(), "fused vs op-by-op on a padded shard", **WGRAD_TOL)


@pytest.mark.parametrize("kind,shape,scale,heads,layers", [
    ("rgat", "bgs", 1.0, 1, 1),          # BASELINE configs[1] at FULL size
    ("hgt", "mag", 1.0 / 16, 4, 2),      # configs[2] (and [4]'s graph) at 1/16: E = 1.3 M
    ("rgat", "wikikg2", 1.0 / 16, 1, 1),  # configs[3] at 1/16: 535 Zipf(1.2) relations, most of them tiny
])
def test_baseline_configs_default_path(cuda, kind, shape, scale, heads, layers):
    """The bench's default path (compact-as-of-node + linear operator reordering + plan-based fused
    edge passes) on the BASELINE.json graphs themselves, against the fp64 layer oracle on the host:
    output, input gradient and every parameter gradient."""
    from het_b200.models import HET_HGT_DGLHetero, HET_RGATModel
    coo, g = _graph(cuda, shape, scale)
    src, dst, et = _coo_t(coo)
    torch.manual_seed(3)
    if kind == "rgat":
        model = HET_RGATModel(coo.num_rels, 64, 64, 64, heads, layers, compact_as_of_node_flag=True,
                              linear_reorder_flag=True).to(cuda)
    else:
        model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, 64, 64, heads, num_layers=layers,
                                  compact_as_of_node_flag=True, fused_attention_aggregation_flag=True,
                                  linear_reorder_flag=True).to(cuda)
    unit_gain_init(model, seed=5)
    x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
    out = model(g, x)
    gout = torch.randn_like(out)
    out.backward(gout)
    if kind == "rgat":
        h, gx, gp = _rgat_oracle(g, src, dst, et, x, list(model.layers), gout)
        assert float(h.abs().max()) > 0.05
        close(out, h, f"{kind} {shape} out")
        close(x.grad, gx, f"{kind} {shape} grad x", **GRAD_TOL)
        for i, layer in enumerate(model.layers):
            for k, v in layer.named_parameters():
                close(v.grad, gp[i][k], f"{kind} {shape} grad {k}", **WGRAD_TOL)
        return
    x64 = x.detach().double().cpu().requires_grad_(True)
    h, ps = x64, []
    offs = torch.from_numpy(coo.node_type_offsets)
    for i, layer in enumerate(model.layers):
        p = {k: v.detach().double().cpu().requires_grad_(True) for k, v in layer.named_parameters()}
        ps.append(p)
        h = M.hgt_layer(src, dst, et, offs, h, p["k_linears"], p["q_linears"], p["v_linears"], p["a_linears"],
                        p["relation_pri"], p["relation_att"], p["relation_msg"], p["skip"])
    h.backward(gout.double().cpu())
    assert float(h.abs().max()) > 0.05
    close(out, h, f"{kind} {shape} out")
    close(x.grad, x64.grad, f"{kind} {shape} grad x", **GRAD_TOL)
    for layer, p in zip(model.layers, ps):
        for k, v in layer.named_parameters():
            close(v.grad, p[k].grad, f"{kind} {shape} grad {k}", **WGRAD_TOL)


@pytest.mark.parametrize("scale,layers", [(0.002, 1), (0.005, 2), (1.0 / 16, 2)])
def test_hgt_bf16_table_storage(cuda, scale, layers):
    """x1: bf16 storage of the gathered tables (k, W_att.q, W_msg.v) with fp32 arithmetic, outputs and
    gradients. Tolerance is the north star's bf16 bar: rtol 2e-2 (+ 2e-2 of the tensor's max for
    elements that cancel), against the same fp64 oracle as the fp32 path."""
    from het_b200.models import HET_HGT_DGLHetero
    coo, g = _graph(cuda, "mag", scale)
    torch.manual_seed(1)
    model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, 64, 64, 4, num_layers=layers, compact_as_of_node_flag=True,
                              fused_attention_aggregation_flag=True, linear_reorder_flag=True,
                              bf16_tables_flag=True).to(cuda)
    unit_gain_init(model)
    x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
    out = model(g, x)
    assert out.dtype == torch.float32
    gout = torch.randn_like(out)
    out.backward(gout)
    src, dst, et = _coo_t(coo)
    offs = torch.from_numpy(coo.node_type_offsets)
    x64 = x.detach().double().cpu().requires_grad_(True)
    h, ps = x64, []
    for layer in model.layers:
        p = {k: v.detach().double().cpu().requires_grad_(True) for k, v in layer.named_parameters()}
        ps.append(p)
        h = M.hgt_layer(src, dst, et, offs, h, p["k_linears"], p["q_linears"], p["v_linears"], p["a_linears"],
                        p["relation_pri"], p["relation_att"], p["relation_msg"], p["skip"])
    h.backward(gout.double().cpu())
    tol = dict(rtol=2e-2, atol=2e-2)
    close(out, h, "hgt bf16-table out", **tol)
    err = (out.double().cpu() - h).abs().max() / h.abs().max()
    assert float(err) > 1e-5, "bf16 storage should be visible in the output: is the bf16 path running?"
    close(x.grad, x64.grad, "hgt bf16-table grad x", **tol)
    assert x.grad.dtype == torch.float32
    for layer, p in zip(model.layers, ps):
        for k, v in layer.named_parameters():
            assert v.grad.dtype == torch.float32
            close(v.grad, p[k].grad, f"hgt bf16-table grad {k}", **tol)
