"""P1 / N2 / N4 rows of SURVEY.md §8: the separate-COO HGT ops, the binary-search and dual-list
compact kinds, and the model flags that route to them — each against the fp64 oracle; the P2 names
are registered and raise."""
import numpy as np
import pytest
import torch

from oracle import het_ops_ref as R
from oracle import models_ref as M
from oracle.testing import close, unit_gain_init
from tests.test_gpu_ops import build

pytestmark = pytest.mark.gpu

GRAD_TOL = dict(rtol=1e-4, atol=2e-5)
WGRAD_TOL = dict(rtol=2e-4, atol=3e-5)


@pytest.mark.parametrize("flags,transpose_att", [
    (dict(hgt_fused_attn_score_flag=True), True),
    (dict(fused_message_mean_aggregation_flag=True), False),
    (dict(hgt_fused_attn_score_flag=True, fused_message_mean_aggregation_flag=True), True),
    (dict(compact_as_of_node_flag=True, compact_direct_indexing_flag=False), False),     # binary-search kind 1
    (dict(multiply_among_weights_first_flag=True, fused_attention_aggregation_flag=True), False),
])
@pytest.mark.parametrize("heads", [1, 4])
def test_hgt_layer_reference_flags(cuda, flags, transpose_att, heads):
    from het_b200 import graphs
    from het_b200.graph import HetGraph
    from het_b200.models import HET_HGT_DGLHetero
    coo = graphs.SHAPES["mag"](scale=0.003)
    g = HetGraph(coo, cuda)
    torch.manual_seed(1)
    model = HET_HGT_DGLHetero(coo.num_ntypes, coo.num_rels, 64, 64, heads, num_layers=1, **flags).to(cuda)
    unit_gain_init(model)
    x = (torch.rand(coo.num_nodes, 64, device=cuda) * 2 - 1).requires_grad_(True)
    out = model(g, x)
    gout = torch.randn_like(out)
    out.backward(gout)
    src, dst, et = (torch.from_numpy(a) for a in (coo.src, coo.dst, coo.etype))
    offs = torch.from_numpy(coo.node_type_offsets)
    x64 = x.detach().double().cpu().requires_grad_(True)
    p = {k: v.detach().double().cpu().requires_grad_(True) for k, v in model.layers[0].named_parameters()}
    # the one-op attention score contracts W_att as k.W.q^T, the two-op path as q.W.k^T (same op family,
    # transposed weight): the oracle is given the transpose so that it states the function the flag computes
    att = p["relation_att"].transpose(2, 3) if transpose_att else p["relation_att"]
    h = M.hgt_layer(src, dst, et, offs, x64, p["k_linears"], p["q_linears"], p["v_linears"], p["a_linears"],
                    p["relation_pri"], att, p["relation_msg"], p["skip"])
    h.backward(gout.double().cpu())
    close(out, h, "hgt out")
    close(x.grad, x64.grad, "hgt grad x", **GRAD_TOL)
    for k, v in model.layers[0].named_parameters():
        close(v.grad, p[k].grad, f"hgt grad {k}", **WGRAD_TOL)


@pytest.mark.parametrize("n,e,r,H", [(50, 300, 3, 1), (200, 3000, 7, 4), (64, 2000, 5, 2)])
@pytest.mark.parametrize("fmt", ["csr", "coo"])
def test_enorm_to_unnormalized_backward(cuda, n, e, r, H, fmt):
    """backward_hgt_full_graph_enorm_to_unnormalized_attn_score_{csr,separate_coo}: the softmax Jacobian from
    the gradient of the normalised score, against torch autograd of softmax(mu * a) in fp64."""
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=17 * e + H, hub=True)
    ic, sc = g.get_in_csr(), g.get_separate_coo_original()
    torch.manual_seed(e)
    attn = torch.randn(e, H, device=cuda)
    mu = torch.rand(r, H, device=cuda) + 0.5
    gnorm = torch.randn(e, H, device=cuda)
    rel_by_eid = torch.repeat_interleave(torch.arange(r), (sc["rel_ptrs"][1:] - sc["rel_ptrs"][:-1]).cpu())
    dst_by_eid = sc["col_indices"].cpu()  # eids are separate-COO positions
    a64 = attn.double().cpu().requires_grad_(True)
    mu64 = mu.double().cpu().requires_grad_(True)
    t = torch.exp(a64 * mu64[rel_by_eid])
    ssum = torch.zeros(n, H, dtype=torch.float64).index_add(0, dst_by_eid, t)
    norm64 = t / ssum[dst_by_eid]
    norm64.backward(gnorm.double().cpu())
    norm = norm64.detach().float().to(cuda)
    ga, gmu = torch.zeros_like(attn), torch.zeros_like(mu)
    if fmt == "csr":
        K.backward_hgt_full_graph_enorm_to_unnormalized_attn_score_csr(
            ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], attn, norm, gnorm, mu, ga, gmu)
    else:
        scratch = torch.full((n, H), float("nan"), device=cuda)
        K.backward_hgt_full_graph_enorm_to_unnormalized_attn_score_separate_coo(
            sc["row_indices"], sc["col_indices"], sc["eids"], sc["rel_ptrs"], attn, norm, gnorm, mu, ga, gmu, scratch)
    close(ga, a64.grad, "grad of the raw score", atol=2e-5)
    close(gmu, mu64.grad, "grad_mu", rtol=2e-4, atol=5e-5)


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_gather_compact_to_non_compact(cuda, kind):
    from het_b200.kernels import K
    g = build(cuda, 80, 1500, 4, seed=3)
    sc = g.get_separate_coo_original()
    u = g.get_separate_unique_node_indices_single_sided()
    e = g.num_edges
    torch.manual_seed(0)
    rows = e if kind == 0 else u["node_indices_col"].numel()
    data = torch.randn(rows, 2, 32, device=cuda)
    out = torch.full((e, 2, 32), float("nan"), device=cuda)
    args = {} if kind == 0 else ({"edata_idx_to_inverse_idx": u["inverse_idx_col"]} if kind == 2 else
                                 {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_col"],
                                  "unique_srcs_and_dests_node_indices": u["node_indices_col"]})
    K.rgnn_gather_compact_to_non_compact(args, kind, sc["rel_ptrs"], sc["eids"], sc["row_indices"], sc["col_indices"],
                                         data, out)
    want = data if kind == 0 else data[u["inverse_idx_col"]]
    assert torch.equal(out, want)


def _rgat_expanded_oracle(g, feat_e, el_e, er_e, gout):
    ic, oc = g.get_in_csr(), g.get_out_csr()
    n = g.num_nodes
    dst_in = R._rows_of(ic["row_ptrs"].cpu())
    s_o, ex_o, ret_o = R.rgat_fwd(dst_in, ic["eids"].cpu(), n, feat_e, el_e, er_e, 0.2)
    grads = R.rgat_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), feat_e, el_e, er_e, s_o, ex_o, ret_o,
                       gout.cpu().double(), 0.2)
    return s_o, ex_o, ret_o, grads


def _scatter(per_edge, idx, rows):
    out = torch.zeros((rows,) + tuple(per_edge.shape[1:]), dtype=per_edge.dtype)
    return out.index_add_(0, idx, per_edge)


@pytest.mark.parametrize("n,e,r,H,F", [(60, 900, 3, 1, 64), (120, 2500, 5, 4, 16)])
def test_rgat_csr_compact_as_of_node_flag(cuda, n, e, r, H, F):
    """relational_fused_gat_csr / backward_… with CompactAsOfNodeFlag = true (CompactAsOfNodeKind::Enabled,
    OpExport/RGATOps.inc.h:251-277, 430-461): one row per (relation, unique node) of the DOUBLE-SIDED list,
    located by binary search."""
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=5 * e + H, hub=True)
    ic, oc, sc = g.get_in_csr(), g.get_out_csr(), g.get_separate_coo_original()
    ud = g.get_separate_unique_node_indices()
    U = ud["node_indices"].numel()
    torch.manual_seed(e)
    feat, el, er = torch.randn(U, H, F, device=cuda), torch.randn(U, H, 1, device=cuda), torch.randn(U, H, 1, device=cuda)
    gout = torch.randn(n, H, F, device=cuda)
    s = torch.full((n, H), float("nan"), device=cuda)
    ex = torch.full((e, H, 1), float("nan"), device=cuda)
    ret = torch.full((n, H, F), float("nan"), device=cuda)
    K.relational_fused_gat_csr(ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], ud["rel_ptrs"],
                               ud["node_indices"], feat, el, er, s, ex, ret, 0.2, True)
    # compact row of (relation, node) per edge id, by numpy search in the lists
    up, un = ud["rel_ptrs"].cpu().numpy(), ud["node_indices"].cpu().numpy()
    rel = np.repeat(np.arange(r), np.diff(sc["rel_ptrs"].cpu().numpy()))
    find = lambda nodes: torch.from_numpy(np.array(
        [up[q] + np.searchsorted(un[up[q]:up[q + 1]], v) for q, v in zip(rel, nodes)], dtype=np.int64))
    isrc, idst = find(sc["row_indices"].cpu().numpy()), find(sc["col_indices"].cpu().numpy())
    f64, el64, er64 = feat.cpu().double(), el.cpu().double(), er.cpu().double()
    s_o, ex_o, ret_o, (gf_e, gel_e, ger_e) = _rgat_expanded_oracle(g, f64[isrc], el64[isrc], er64[idst], gout)
    close(s, s_o, "sum")
    close(ex, ex_o, "exp")
    close(ret, ret_o, "ret")
    gf, gel, ger = torch.zeros_like(feat), torch.zeros_like(el), torch.zeros_like(er)
    K.backward_relational_fused_gat_csr(oc["row_ptrs"], oc["col_indices"], oc["eids"], oc["rel_types"], ud["rel_ptrs"],
                                        ud["node_indices"], feat, el, er, s, ex, ret, gout, gf, gel, ger, 0.2, True)
    close(gf, _scatter(gf_e, isrc, U), "grad_feat (compact rows)", atol=2e-5)
    close(gel.view(U, H), _scatter(gel_e, isrc, U), "grad_el (compact rows)", atol=2e-5)
    close(ger.view(U, H), _scatter(ger_e, idst, U), "grad_er (compact rows)", atol=2e-5)


@pytest.mark.parametrize("kind", [1, 3, 4])
def test_rgat_separate_coo_compact_kinds(cuda, kind):
    """relational_fused_gat_separate_coo / backward_… with IntKind 1 (one double-sided list, binary
    search), 3 (dual lists, binary search), 4 (dual direct maps) — OpExport/RGATOps.inc.h:170-245, 465-551."""
    from het_b200.kernels import K
    n, e, r, H, F = 90, 2000, 4, 2, 32
    g = build(cuda, n, e, r, seed=23 + kind, hub=True)
    sc = g.get_separate_coo_original()
    u = g.get_separate_unique_node_indices_single_sided()
    if kind == 1:
        ud = g.get_separate_unique_node_indices()
        args = {"unique_srcs_and_dests_rel_ptrs": ud["rel_ptrs"], "unique_srcs_and_dests_node_indices": ud["node_indices"]}
        up, un = ud["rel_ptrs"].cpu().numpy(), ud["node_indices"].cpu().numpy()
        rel = np.repeat(np.arange(r), np.diff(sc["rel_ptrs"].cpu().numpy()))
        find = lambda nodes: torch.from_numpy(np.array(
            [up[q] + np.searchsorted(un[up[q]:up[q + 1]], v) for q, v in zip(rel, nodes)], dtype=np.int64))
        isrc, idst = find(sc["row_indices"].cpu().numpy()), find(sc["col_indices"].cpu().numpy())
        Us = Ud = ud["node_indices"].numel()
    else:
        isrc, idst = u["inverse_idx_row"].cpu(), u["inverse_idx_col"].cpu()
        Us, Ud = u["node_indices_row"].numel(), u["node_indices_col"].numel()
        if kind == 3:
            args = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"], "unique_srcs_and_dests_rel_ptrs_col": u["rel_ptrs_col"],
                    "unique_srcs_and_dests_node_indices_row": u["node_indices_row"],
                    "unique_srcs_and_dests_node_indices_col": u["node_indices_col"]}
        else:
            args = {"edata_idx_to_inverse_idx_row": u["inverse_idx_row"], "edata_idx_to_inverse_idx_col": u["inverse_idx_col"]}
    torch.manual_seed(kind)
    feat, el, er = torch.randn(Us, H, F, device=cuda), torch.randn(Us, H, 1, device=cuda), torch.randn(Ud, H, 1, device=cuda)
    gout = torch.randn(n, H, F, device=cuda)
    s = torch.full((n, H), float("nan"), device=cuda)
    ex = torch.full((e, H, 1), float("nan"), device=cuda)
    ret = torch.full((n, H, F), float("nan"), device=cuda)
    K.relational_fused_gat_separate_coo(sc["eids"], sc["rel_ptrs"], sc["row_indices"], sc["col_indices"], kind, args,
                                        feat, el, er, s, ex, ret, 0.2)
    s_o, ex_o, ret_o, (gf_e, gel_e, ger_e) = _rgat_expanded_oracle(
        g, feat.cpu().double()[isrc], el.cpu().double()[isrc], er.cpu().double()[idst], gout)
    close(s, s_o, "sum")
    close(ret, ret_o, "ret")
    gf, gel, ger = torch.zeros_like(feat), torch.zeros_like(el), torch.zeros_like(er)
    K.backward_relational_fused_gat_separate_coo(sc["eids"], sc["rel_ptrs"], sc["row_indices"], sc["col_indices"], kind,
                                                 args, feat, el, er, s, ex, ret, gout, gf, gel, ger, 0.2)
    close(gf, _scatter(gf_e, isrc, Us), "grad_feat (compact rows)", atol=2e-5)
    close(gel.view(Us, H), _scatter(gel_e, isrc, Us), "grad_el (compact rows)", atol=2e-5)
    close(ger.view(Ud, H), _scatter(ger_e, idst, Ud), "grad_er (compact rows)", atol=2e-5)


def test_p2_names_are_registered_and_raise(cuda):
    from het_b200.kernels import K
    t = torch.zeros(1, device=cuda)
    for name, nt, has_slope in [("seastar_rgcn_layer1_csr", 8, False), ("seastar_backward_rgcn_layer1_coo", 11, False),
                                ("fused_gat_csr", 9, True), ("backward_fused_gat_csr", 13, True)]:
        args = [t] * nt + ([0.2] if has_slope else [])
        with pytest.raises(RuntimeError, match="outside the hot path"):
            getattr(K, name)(*args)
