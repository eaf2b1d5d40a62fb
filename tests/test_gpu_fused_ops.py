"""Op-level parity of the ops the benchmark actually runs (the compact + fused default path):

* ``hgt_fused_attention_and_aggregation_csr_compact``  (hetb200_hgt_fused_attention_aggregation_csr + backward)
* ``relational_fused_gat_csr_compact`` / ``…_with_logits``  (hetb200_relational_fused_gat_csr_compact,
  hetb200_backward_relational_fused_gat_compact_rows / …_csr_compact)

with O(1) random inputs, against COMPOSITIONS of the op-level fp64 oracle (oracle/het_ops_ref.py:
inner_product → hgt_edge_softmax → hgt_aggregate and their backward functions; rgat_fwd / rgat_bwd),
which tests/test_gpu_reference_kernels.py pins to the reference's own CUDA kernels. Every output
and every gradient is compared relative to its own magnitude (tests.test_gpu_ops.close: no
absolute floor). Graph cases: random, a hub destination (half of all edges) with empty rows, and a
source id space with multi-million-row gaps (a destination-range shard's view).
"""
import numpy as np
import pytest
import torch

from oracle import het_ops_ref as R
from tests.test_gpu_ops import close

pytestmark = pytest.mark.gpu


def _graph(cuda, n_dst, n_src, e, r, seed, hub=False, gaps=False):
    from het_b200.graph import HetGraph
    g = torch.Generator().manual_seed(seed)
    if gaps:  # few distinct sources spread over a huge id space
        pool = torch.randint(0, n_src, (max(8, e // 6),), generator=g)
        src = pool[torch.randint(0, pool.numel(), (e,), generator=g)]
    else:
        src = torch.randint(0, n_src, (e,), generator=g)
    dst = torch.randint(0, n_dst, (e,), generator=g)
    if hub:
        dst[: e // 2] = 3
    rel = torch.randint(0, r, (e,), generator=g)
    offs = torch.tensor([0, n_dst // 2, n_dst])
    return HetGraph.from_arrays("t", src, dst, rel, n_src, n_dst, r, offs, cuda)


def _scatter_rows(per_edge, idx, rows):
    out = torch.zeros((rows,) + tuple(per_edge.shape[1:]), dtype=per_edge.dtype)
    return out.index_add_(0, idx, per_edge)


HGT_CASES = [
    # n_dst, n_src, e, r, H, F, hub, gaps
    (200, 200, 3000, 7, 4, 16, False, False),
    (200, 200, 3000, 7, 4, 16, True, False),
    (64, 64, 2000, 3, 1, 64, True, False),
    (100, 100, 1500, 5, 2, 32, False, False),
    (300, 2_000_000, 4000, 4, 4, 16, False, True),
    (300, 2_000_000, 4000, 4, 4, 16, True, True),
    (40, 40, 37, 2, 8, 8, False, False),
]


@pytest.mark.parametrize("n_dst,n_src,e,r,H,F,hub,gaps", HGT_CASES)
@pytest.mark.parametrize("impl", ["plan", "plan_saved_scores", "plan_two_pass", "csr"])
def test_hgt_fused_attention_aggregation_op(cuda, n_dst, n_src, e, r, H, F, hub, gaps, impl, monkeypatch):
    """plan = the default (one backward walk that recomputes the raw score; the forward saves nothing per
    edge); plan_saved_scores = the forward stores the score, the backward reads it; plan_two_pass =
    hetb_hgt_fused_bwd_plan2 (grad_left owned by a second, destination-major walk: no atomics)."""
    from het_b200 import backend as B
    monkeypatch.setattr(B, "SAVE_HGT_SCORES", impl == "plan_saved_scores")
    monkeypatch.setattr(B, "TWO_PASS_HGT_BACKWARD", impl == "plan_two_pass")
    fused = (B.hgt_fused_attention_and_aggregation_plan if impl.startswith("plan")
             else B.hgt_fused_attention_and_aggregation_csr_compact)
    g = _graph(cuda, n_dst, n_src, e, r, seed=e + H + int(hub), hub=hub, gaps=gaps)
    u = g.get_separate_unique_node_indices_single_sided()
    sc, ic, oc = g.get_separate_coo_original(), g.get_in_csr(), g.get_out_csr()
    Ur, Uc = u["node_indices_row"].numel(), u["node_indices_col"].numel()
    torch.manual_seed(e + 3)
    left = torch.randn(Uc, H, F, device=cuda, requires_grad=True)
    k = torch.randn(n_src, H, F, device=cuda, requires_grad=True)
    msg = torch.randn(Ur, H, F, device=cuda, requires_grad=True)
    mu = (torch.rand(r, H, device=cuda) + 0.5).requires_grad_(True)
    out = fused(g, left, k, msg, mu)
    gout = torch.randn_like(out)
    out.backward(gout)

    # ---- oracle composition on per-edge expansions of the compact tensors ----
    inv_col, inv_row = u["inverse_idx_col"].cpu(), u["inverse_idx_row"].cpu()
    eids, row = sc["eids"].cpu(), sc["row_indices"].cpu()
    left_e = left.detach().cpu().double()[inv_col]   # indexed by edge id
    msg_e = msg.detach().cpu().double()[inv_row]
    k64, mu64 = k.detach().cpu().double(), mu.detach().cpu().double()
    score = R.inner_product(eids, row, left_e, k64)                                # a5
    dst_in = R._rows_of(ic["row_ptrs"].cpu())
    s_o, t_o, norm_o = R.hgt_edge_softmax(dst_in, ic["eids"].cpu(), ic["rel_types"].cpu(), n_dst, score, mu64)  # a10
    out_o = R.hgt_aggregate(dst_in, ic["eids"].cpu(), n_dst, msg_e, t_o, s_o)      # a11
    close(out, out_o, "fused out")
    g64 = gout.cpu().double()
    gm_e = R.hgt_aggregate_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), norm_o, g64)          # a12
    ga, gmu_o = R.hgt_edge_softmax_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), oc["rel_types"].cpu(), msg_e, score,
                                       norm_o, out_o, g64, mu64)
    gl_e, gk_o = R.inner_product_bwd(eids, row, left_e, k64, ga)                  # a6
    close(msg.grad, _scatter_rows(gm_e, inv_row, Ur), "fused grad_msg (compact rows)", atol=2e-5)
    close(left.grad, _scatter_rows(gl_e, inv_col, Uc), "fused grad_left (compact rows)", atol=2e-5)
    close(k.grad, gk_o, "fused grad_k", atol=2e-5)
    close(mu.grad, gmu_o, "fused grad_mu", rtol=2e-4, atol=5e-5)
    # empty destination rows are exactly zero, not garbage
    empty = (ic["row_ptrs"][1:] == ic["row_ptrs"][:-1])
    assert float(out[empty].abs().sum()) == 0.0


def test_hgt_two_pass_backward_is_deterministic_and_needs_no_zero_fill(cuda):
    """grad_left of the two-pass backward: NaN-filled on entry (every row is written), bit-identical
    across runs (no atomics outside the hub runs the plan lists), equal to the single walk to fp32 rounding."""
    from het_b200.kernels import K
    n, e, r, H, F = 500, 60000, 5, 4, 16
    g = _graph(cuda, n, n, e, r, seed=9, hub=True)
    u = g.get_separate_unique_node_indices_single_sided()
    pl = g.get_edge_plan()
    assert pl["fwd"].meta[6] > 0, "the hub row should cut (relation, destination) runs"
    Ur, Uc = u["node_indices_row"].numel(), u["node_indices_col"].numel()
    torch.manual_seed(4)
    left, k, msg = (torch.randn(m, H, F, device=cuda) for m in (Uc, n, Ur))
    mu = torch.rand(r, H, device=cuda) + 0.5
    sc = g.plan_scratch(H, H * F)
    out, lse = torch.empty(n, H, F, device=cuda), torch.empty(n, H, device=cuda)
    attn = torch.empty(e, H, device=cuda)
    K.hetb200_hgt_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu, attn, lse, out, sc["part_m"],
                                 sc["part_s"], sc["part_acc"], sc["ssum"])
    gout = torch.randn_like(out)
    res = []
    for _ in range(2):
        a2 = attn.clone()
        gl = torch.full((Uc, H, F), float("nan"), device=cuda)
        gk, gm = torch.full_like(k, float("nan")), torch.full_like(msg, float("nan"))
        gmu = torch.zeros_like(mu)
        K.hetb200_hgt_fused_bwd_plan2(pl["bwd"].blob, pl["bwd"].meta, pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu,
                                      a2, lse, out, gout, sc["gdot"], gl, gk, gm, gmu)
        res.append((gl, gk, gm))
    unsplit = torch.ones(Uc, dtype=torch.bool, device=cuda)
    fix = pl["fwd"].blob[pl["fwd"].meta[7 + 8]: pl["fwd"].meta[7 + 8] + pl["fwd"].meta[6]].long()
    unsplit[fix] = False
    assert torch.equal(res[0][0][unsplit], res[1][0][unsplit])
    # the single walk, with the saved scores and with the scores recomputed (the default): the recomputed
    # score is the forward's bit for bit, so grad_k / grad_msg (no atomics on unsplit rows) are identical
    outs = []
    for a_in in (attn, torch.empty(0, H, device=cuda)):
        gl1 = torch.zeros(Uc, H, F, device=cuda)
        gk1, gm1, gmu1 = torch.empty_like(k), torch.empty_like(msg), torch.zeros_like(mu)
        K.hetb200_hgt_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, left, k, msg, mu, a_in, lse, out, gout, sc["gdot"],
                                     gl1, gk1, gm1, gmu1)
        outs.append((gl1, gk1, gm1))
        close(res[0][0], gl1.cpu(), "two-pass vs single-walk grad_left", rtol=2e-5, atol=2e-6)
    bw = pl["bwd"]
    rows_ok = torch.ones(n, dtype=torch.bool, device=cuda)
    rows_ok[bw.blob[bw.meta[7 + 4]: bw.meta[7 + 4] + bw.meta[3]].long()] = False          # fix_rows: reduced
    runs_ok = torch.ones(Ur, dtype=torch.bool, device=cuda)
    runs_ok[bw.blob[bw.meta[7 + 8]: bw.meta[7 + 8] + bw.meta[6]].long()] = False          # fix_runs: reduced
    assert int(runs_ok.sum()) > 0
    for name, got in (("recomputed scores", outs[1]), ("two-pass", res[0])):
        close(got[1], outs[0][1].cpu(), f"grad_k ({name})", rtol=2e-5, atol=2e-6)
        close(got[2], outs[0][2].cpu(), f"grad_msg ({name})", rtol=2e-5, atol=2e-6)
        assert torch.equal(got[1][rows_ok], outs[0][1][rows_ok]), f"grad_k of unsplit rows ({name})"
        assert torch.equal(got[2][runs_ok], outs[0][2][runs_ok]), f"grad_msg of unsplit runs ({name})"


RGAT_CASES = [
    (200, 200, 3000, 7, 1, 64, False, False),
    (200, 200, 3000, 7, 4, 16, True, False),
    (64, 64, 2000, 3, 2, 32, True, False),
    (300, 2_000_000, 4000, 4, 1, 64, False, True),
    (300, 2_000_000, 4000, 4, 4, 16, True, True),
]


@pytest.mark.parametrize("n_dst,n_src,e,r,H,F,hub,gaps", RGAT_CASES)
@pytest.mark.parametrize("by_rows", ["plan", True, False])
@pytest.mark.parametrize("with_logits", [False, True])
def test_rgat_compact_edge_op(cuda, n_dst, n_src, e, r, H, F, hub, gaps, by_rows, with_logits):
    """by_rows: "plan" = the plan-based walkers (bench default); True / False = the two backward
    implementations behind the compact in-CSR op of csrc/edge_ops.cu."""
    from het_b200 import backend as B
    g = _graph(cuda, n_dst, n_src, e, r, seed=2 * e + H + int(hub), hub=hub, gaps=gaps)
    u = g.get_separate_unique_node_indices_single_sided()
    ic, oc = g.get_in_csr(), g.get_out_csr()
    Ur, Uc = u["node_indices_row"].numel(), u["node_indices_col"].numel()
    torch.manual_seed(e + 5)
    feat = torch.randn(Ur, H, F, device=cuda, requires_grad=True)
    er = torch.randn(Uc, H, 1, device=cuda, requires_grad=True)
    attn_l = torch.randn(r, H, F, device=cuda, requires_grad=True)
    el = torch.randn(Ur, H, 1, device=cuda, requires_grad=True)
    B.RGAT_COMPACT_BACKWARD_BY_ROWS = by_rows is not False
    try:
        if with_logits:
            fn = B.relational_fused_gat_plan_with_logits if by_rows == "plan" else B.relational_fused_gat_csr_compact_with_logits
            out = fn(g, feat, er, attn_l, 0.2)
        else:
            fn = B.relational_fused_gat_plan if by_rows == "plan" else B.relational_fused_gat_csr_compact
            out = fn(g, feat, el, er, 0.2)
        gout = torch.randn_like(out)
        out.backward(gout)
    finally:
        B.RGAT_COMPACT_BACKWARD_BY_ROWS = True

    inv_col, inv_row = u["inverse_idx_col"].cpu(), u["inverse_idx_row"].cpu()
    feat64 = feat.detach().cpu().double()
    if with_logits:
        rel_of_urow = torch.repeat_interleave(torch.arange(r), (u["rel_ptrs_row"][1:] - u["rel_ptrs_row"][:-1]).cpu())
        al64 = attn_l.detach().cpu().double()
        el64 = (feat64 * al64[rel_of_urow]).sum(-1, keepdim=True)
    else:
        el64 = el.detach().cpu().double()
    feat_e, el_e, er_e = feat64[inv_row], el64[inv_row], er.detach().cpu().double()[inv_col]
    dst_in = R._rows_of(ic["row_ptrs"].cpu())
    s_o, ex_o, ret_o = R.rgat_fwd(dst_in, ic["eids"].cpu(), n_dst, feat_e, el_e, er_e, 0.2)      # a7
    close(out, ret_o, "compact rgat out")
    gf_e, gel_e, ger_e = R.rgat_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), feat_e, el_e, er_e, s_o, ex_o, ret_o,
                                    gout.cpu().double(), 0.2)                                   # a9
    gfeat = _scatter_rows(gf_e, inv_row, Ur)
    gel = _scatter_rows(gel_e, inv_row, Ur)
    close(er.grad, _scatter_rows(ger_e, inv_col, Uc).view_as(er), "compact rgat grad_er", atol=2e-5)
    if with_logits:
        gfeat = gfeat + gel.unsqueeze(-1) * al64[rel_of_urow]
        gal = torch.zeros_like(al64).index_add_(0, rel_of_urow, gel.unsqueeze(-1) * feat64)
        close(attn_l.grad, gal, "compact rgat grad_attn_l", rtol=2e-4, atol=2e-5)
    else:
        close(el.grad, gel.view_as(el), "compact rgat grad_el", atol=2e-5)
    close(feat.grad, gfeat, "compact rgat grad_feat", atol=2e-5)
    empty = (ic["row_ptrs"][1:] == ic["row_ptrs"][:-1])
    assert float(out[empty].abs().sum()) == 0.0


def test_softmax_is_stable_at_large_logits(cuda):
    """north star: max-subtracted online softmax. Logits around +-100 overflow exp() in fp32
    (the reference's kernels return inf/NaN there: GAT/FusedGAT.cu.h:193-203,
    HGT/HGTForwardKernels.cu.h:499-508); the fused / compact ops must stay finite and match the fp64
    softmax."""
    from het_b200 import backend as B
    g = _graph(cuda, 64, 64, 2000, 3, seed=9, hub=True)
    u = g.get_separate_unique_node_indices_single_sided()
    sc, ic = g.get_separate_coo_original(), g.get_in_csr()
    Ur, Uc = u["node_indices_row"].numel(), u["node_indices_col"].numel()
    H, F, r = 4, 16, 3
    torch.manual_seed(3)
    left = torch.randn(Uc, H, F, device=cuda) * 6
    k = torch.randn(64, H, F, device=cuda) * 6     # scores ~ N(0, 36*4) -> |a| up to ~450
    msg = torch.randn(Ur, H, F, device=cuda)
    mu = torch.ones(r, H, device=cuda)
    left.requires_grad_(True); k.requires_grad_(True); msg.requires_grad_(True)
    out = B.hgt_fused_attention_and_aggregation_plan(g, left, k, msg, mu)
    out.backward(torch.randn_like(out))
    assert bool(torch.isfinite(out).all()) and bool(torch.isfinite(left.grad).all())
    assert bool(torch.isfinite(k.grad).all()) and bool(torch.isfinite(msg.grad).all())
    inv_col, inv_row = u["inverse_idx_col"].cpu(), u["inverse_idx_row"].cpu()
    score = R.inner_product(sc["eids"].cpu(), sc["row_indices"].cpu(), left.detach().cpu().double()[inv_col],
                            k.detach().cpu().double())
    assert float(score.abs().max()) > 100.0
    dst_pos = R._rows_of(ic["row_ptrs"].cpu())
    a = torch.zeros_like(score)
    eids = ic["eids"].cpu()
    # fp64 softmax with max subtraction per (destination, head)
    mx = torch.full((64, H), -float("inf"), dtype=torch.float64).scatter_reduce(
        0, dst_pos.unsqueeze(-1).expand(-1, H), score[eids], "amax")
    w = torch.exp(score[eids] - mx[dst_pos])
    ssum = torch.zeros(64, H, dtype=torch.float64).index_add_(0, dst_pos, w)
    a[eids] = w / ssum[dst_pos]
    ref = torch.zeros(64, H, F, dtype=torch.float64).index_add_(
        0, dst_pos, a[eids].unsqueeze(-1) * msg.detach().cpu().double()[inv_row][eids])
    close(out, ref, "fused out at |logit| > 100", rtol=2e-4, atol=2e-5)

    # RGAT compact op at logits ~ +-120
    feat = torch.randn(Ur, 1, 64, device=cuda)
    el = (torch.randn(Ur, 1, 1, device=cuda) * 60)
    er = (torch.randn(Uc, 1, 1, device=cuda) * 60)
    feat.requires_grad_(True); el.requires_grad_(True); er.requires_grad_(True)
    o2 = B.relational_fused_gat_plan(g, feat, el, er, 0.2)
    o2.backward(torch.randn_like(o2))
    assert bool(torch.isfinite(o2).all()) and bool(torch.isfinite(feat.grad).all())
    assert bool(torch.isfinite(el.grad).all()) and bool(torch.isfinite(er.grad).all())
    x = (el.detach().cpu().double()[inv_row] + er.detach().cpu().double()[inv_col]).view(-1, 1)
    x = torch.where(x > 0, x, 0.2 * x)
    assert float(x.max()) > 100.0
    mx = torch.full((64, 1), -float("inf"), dtype=torch.float64).scatter_reduce(0, dst_pos.unsqueeze(-1), x[eids], "amax")
    w = torch.exp(x[eids] - mx[dst_pos])
    ssum = torch.zeros(64, 1, dtype=torch.float64).index_add_(0, dst_pos, w)
    ref2 = torch.zeros(64, 1, 64, dtype=torch.float64).index_add_(
        0, dst_pos, (w / ssum[dst_pos]).unsqueeze(-1) * feat.detach().cpu().double()[inv_row][eids])
    close(o2, ref2, "compact rgat out at |logit| > 100", rtol=2e-4, atol=2e-5)
