"""Op-level parity: every torch_hrt op (CUDA, through the C ABI) against the fp64
CPU oracle on the same seeded inputs. fp32 tolerance rtol 1e-4 (north star)."""
import numpy as np
import pytest
import torch

from oracle import het_ops_ref as R

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-4, 1e-5


from oracle.testing import close  # noqa: E402,F401  (re-exported: other test modules import it from here)


def make_graph(n, e, r, seed, hub=False):
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, n, (e,), generator=g)
    if hub and e > 10:  # a heavy destination row (> 256 in-edges) and empty rows
        dst[: e // 2] = 3
    rel = torch.randint(0, r, (e,), generator=g)
    return src, dst, rel


def build(cuda, n, e, r, seed, hub=False):
    from het_b200.graph import HetGraph
    from het_b200.graphs import HeteroCOO
    src, dst, rel = make_graph(n, e, r, seed, hub)
    coo = HeteroCOO("t", n, r, src.numpy(), dst.numpy(), rel.numpy(), np.array([0, n // 2, n], dtype=np.int64),
                    np.zeros(r, dtype=np.int64), np.zeros(r, dtype=np.int64))
    return HetGraph(coo, cuda)


def cpu(d):
    return {k: v.cpu() for k, v in d.items()}


CASES = [(50, 300, 3, 1, 64), (200, 3000, 7, 4, 16), (64, 2000, 5, 2, 32), (33, 700, 2, 1, 128), (40, 500, 3, 2, 8)]


@pytest.mark.parametrize("n,e,r,H,F", CASES)
@pytest.mark.parametrize("hub", [False, True])
def test_rgat_edge_ops(cuda, n, e, r, H, F, hub):
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=e + H, hub=hub)
    ic, oc, sc = g.get_in_csr(), g.get_out_csr(), g.get_separate_coo_original()
    torch.manual_seed(e)
    feat = torch.randn(e, H, F, device=cuda)
    el = torch.randn(e, H, 1, device=cuda)
    er = torch.randn(e, H, 1, device=cuda)
    empty = torch.empty(0, dtype=torch.int64, device=cuda)
    dst_in = R._rows_of(ic["row_ptrs"].cpu())
    s_ref, ex_ref, ret_ref = R.rgat_fwd(dst_in, ic["eids"].cpu(), n, feat.cpu(), el.cpu(), er.cpu(), 0.2)
    for variant in ("csr", "coo"):
        s = torch.full((n, H), float("nan"), device=cuda)  # outputs arrive uninitialised
        ex = torch.full((e, H, 1), float("nan"), device=cuda)
        ret = torch.full((n, H, F), float("nan"), device=cuda)
        if variant == "csr":
            K.relational_fused_gat_csr(ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], empty, empty,
                                       feat, el, er, s, ex, ret, 0.2)
        else:
            K.relational_fused_gat_separate_coo(sc["eids"], sc["rel_ptrs"], sc["row_indices"], sc["col_indices"], 0,
                                                {}, feat, el, er, s, ex, ret, 0.2)
        close(s, s_ref, f"sum/{variant}")
        close(ex, ex_ref, f"exp/{variant}")
        close(ret, ret_ref, f"ret/{variant}")
    gout = torch.randn(n, H, F, device=cuda)
    gf_ref, gel_ref, ger_ref = R.rgat_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), feat.cpu(), el.cpu(), er.cpu(),
                                          s_ref, ex_ref, ret_ref, gout.cpu(), 0.2)
    for variant in ("csr", "coo"):
        gf, gel, ger = torch.zeros_like(feat), torch.zeros_like(el), torch.zeros_like(er)
        if variant == "csr":
            K.backward_relational_fused_gat_csr(oc["row_ptrs"], oc["col_indices"], oc["eids"], oc["rel_types"], empty,
                                                empty, feat, el, er, s, ex, ret, gout, gf, gel, ger, 0.2)
        else:
            K.backward_relational_fused_gat_separate_coo(sc["eids"], sc["rel_ptrs"], sc["row_indices"],
                                                         sc["col_indices"], 0, {}, feat, el, er, s, ex, ret, gout, gf,
                                                         gel, ger, 0.2)
        close(gf, gf_ref, f"grad_feat/{variant}")
        close(gel, gel_ref, f"grad_el/{variant}")
        close(ger, ger_ref, f"grad_er/{variant}")


@pytest.mark.parametrize("n,e,r,H,F", CASES)
@pytest.mark.parametrize("hub", [False, True])
def test_hgt_edge_ops(cuda, n, e, r, H, F, hub):
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=2 * e + H, hub=hub)
    ic, oc = g.get_in_csr(), g.get_out_csr()
    torch.manual_seed(e + 1)
    attn = torch.randn(e, H, device=cuda)
    mu = torch.rand(r, H, device=cuda) + 0.5
    msg = torch.randn(e, H, F, device=cuda)
    dst_in = R._rows_of(ic["row_ptrs"].cpu())
    s_ref, t_ref, norm_ref = R.hgt_edge_softmax(dst_in, ic["eids"].cpu(), ic["rel_types"].cpu(), n, attn.cpu(), mu.cpu())
    s = torch.full((n, H), float("nan"), device=cuda)
    t = torch.full((e, H), float("nan"), device=cuda)
    norm = torch.full((e, H), float("nan"), device=cuda)
    K.hgt_full_graph_edge_softmax_ops_csr(ic["row_ptrs"], ic["col_indices"], ic["eids"], ic["rel_types"], attn, mu, s,
                                          t, norm)
    close(s, s_ref, "sum")
    close(t, t_ref, "mu_applied")
    close(norm, norm_ref, "norm")
    # softmax rows sum to one wherever a destination has in-edges
    tot = torch.zeros(n, H, dtype=torch.float64).index_add_(0, dst_in, norm.double().cpu()[ic["eids"].cpu()])
    has = (ic["row_ptrs"][1:] > ic["row_ptrs"][:-1]).cpu()
    assert torch.allclose(tot[has], torch.ones_like(tot[has]), atol=1e-5)
    out = torch.full((n, H, F), float("nan"), device=cuda)
    K.hgt_full_graph_message_mean_aggregation_csr(ic["row_ptrs"], ic["col_indices"], ic["rel_types"], ic["eids"], msg,
                                                  t, s, mu, out)
    out_ref = R.hgt_aggregate(dst_in, ic["eids"].cpu(), n, msg.cpu(), t_ref, s_ref)
    close(out, out_ref, "aggregate")
    gout = torch.randn(n, H, F, device=cuda)
    gm = torch.zeros_like(msg)
    K.backward_hgt_full_graph_message_mean_aggregation_csr(oc["row_ptrs"], oc["col_indices"], oc["rel_types"],
                                                           oc["eids"], s, norm, gout, gm)
    close(gm, R.hgt_aggregate_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), norm_ref, gout.cpu()), "grad_msg")
    ga, gmu = torch.zeros_like(attn), torch.zeros_like(mu)
    K.backward_hgt_full_graph_edge_softmax_ops_csr(oc["row_ptrs"], oc["col_indices"], oc["eids"], oc["rel_types"], msg,
                                                   attn, norm, out, gout, mu, ga, gmu)
    ga_ref, gmu_ref = R.hgt_edge_softmax_bwd(oc["col_indices"].cpu(), oc["eids"].cpu(), oc["rel_types"].cpu(),
                                             msg.cpu(), attn.cpu(), norm_ref, out_ref, gout.cpu(), mu.cpu())
    close(ga, ga_ref, "grad_attn")
    close(gmu, gmu_ref, "grad_mu", rtol=5e-4, atol=5e-5)


@pytest.mark.parametrize("n,e,r,H,F", CASES)
def test_inner_product(cuda, n, e, r, H, F):
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=3 * e + H)
    sc = g.get_separate_coo_original()
    torch.manual_seed(e + 2)
    left = torch.randn(e, H, F, device=cuda)
    right = torch.randn(n, H, F, device=cuda)
    out = torch.full((e, H), float("nan"), device=cuda)
    K.rgnn_inner_product_right_node_separatecoo({}, 0, sc["rel_ptrs"], sc["eids"], sc["row_indices"],
                                                sc["col_indices"], left, right, out)
    close(out, R.inner_product(sc["eids"].cpu(), sc["row_indices"].cpu(), left.cpu(), right.cpu()), "inner")
    gout = torch.randn(e, H, device=cuda)
    gl, gr = torch.zeros_like(left), torch.zeros_like(right)
    K.backward_inner_product_right_node_separatecoo({}, 0, sc["rel_ptrs"], sc["eids"], sc["row_indices"],
                                                    sc["col_indices"], left, right, gout, gl, gr)
    gl_ref, gr_ref = R.inner_product_bwd(sc["eids"].cpu(), sc["row_indices"].cpu(), left.cpu(), right.cpu(), gout.cpu())
    close(gl, gl_ref, "grad_left")
    close(gr, gr_ref, "grad_right")


@pytest.mark.gpu
@pytest.mark.parametrize("n,c", [(1, 64), (1000, 64), (4097, 8), (300, 128)])
def test_softmax_cross_entropy(cuda, n, c):
    """Fused loss of the training step vs torch's nll_loss(log_softmax) in fp64 on the CPU."""
    import torch.nn.functional as Fn
    from het_b200 import backend as B
    torch.manual_seed(n + c)
    x = (torch.randn(n, c, device=cuda) * 3).requires_grad_(True)
    y = torch.randint(0, c, (n,), device=cuda)
    loss = B.softmax_cross_entropy(x, y, n + 7)
    loss.backward()
    x64 = x.detach().double().cpu().requires_grad_(True)
    ref = Fn.nll_loss(Fn.log_softmax(x64, dim=-1), y.cpu(), reduction="sum") / (n + 7)
    ref.backward()
    close(loss, ref, "loss")
    close(x.grad, x64.grad, "grad logits", rtol=1e-4, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("rows,cols", [(0, 64), (1, 64), (7300, 64), (94_001, 64), (1000, 16), (513, 128)])
def test_column_sum(cuda, rows, cols):
    """hetb200_column_sum (the bias gradient) against torch in fp64."""
    from het_b200.kernels import K
    torch.manual_seed(0)
    x = torch.randn(rows, cols, device=cuda)
    out = torch.full((cols,), 7.0, device=cuda)  # must be overwritten, not accumulated into
    K.hetb200_column_sum(x, out)
    close(out, x.double().sum(0).cpu(), "column sum", rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1000, 64), (7,), (3, 1, 64, 64), (4097,)])
def test_hetb_adam_matches_torch_adam(cuda, shape):
    """HetbAdam (csrc/optim.cu) against torch.optim.Adam over several steps with changing gradients."""
    from het_b200.train import HetbAdam
    torch.manual_seed(0)
    p0 = torch.randn(*shape, device=cuda)
    a = torch.nn.Parameter(p0.clone())
    b = torch.nn.Parameter(p0.double().clone())
    oa, ob = HetbAdam([a], lr=1e-2), torch.optim.Adam([b], lr=1e-2)
    for step in range(5):
        g = torch.randn(*shape, device=cuda) * (1 + step)
        a.grad, b.grad = g.clone(), g.double()
        oa.step()
        ob.step()
        close(a, b, f"parameter after step {step + 1}", rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("T,R,H,dk", [(4, 8, 4, 16), (3, 5, 1, 64), (2, 7, 2, 32)])
def test_compose_weights(cuda, T, R, H, dk):
    """hetb200_compose_weights (+ backward) against the einsum it replaces, in fp64."""
    from het_b200 import backend as B
    torch.manual_seed(R)
    D = H * dk
    lin = torch.randn(T, 1, 64, D, device=cuda, requires_grad=True)
    rel = torch.randn(R, H, dk, dk, device=cuda, requires_grad=True)
    types = torch.randint(0, T, (R,), device=cuda)
    out = B.compose_weights(lin, types, rel)
    gout = torch.randn_like(out)
    out.backward(gout)
    l64, r64 = lin.detach().double().cpu().requires_grad_(True), rel.detach().double().cpu().requires_grad_(True)
    ref = torch.einsum("rihd,rhde->rihe", l64[types.cpu(), 0].view(-1, 64, H, dk), r64).reshape(-1, 1, 64, D)
    ref.backward(gout.double().cpu())
    close(out, ref, "composed weights")
    close(lin.grad, l64.grad, "grad lin")
    close(rel.grad, r64.grad, "grad rel")
