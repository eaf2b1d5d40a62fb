"""Parity of the relation-typed linear layer (segment MM, gather-MM-scatter) and
its backward, through the torch_hrt ops, against the fp64 CPU oracle."""
import pytest
import torch

from oracle import het_ops_ref as R
from tests.test_gpu_ops import build, close

pytestmark = pytest.mark.gpu

# (nodes, edges, relations, heads, K, N, input_num_head_one)
SHAPES = [
    (60, 500, 3, 1, 64, 64, True),     # RGAT conv_weights, 1 head
    (60, 500, 3, 4, 64, 16, True),     # RGAT conv_weights, 4 heads share the input row
    (80, 900, 8, 4, 16, 16, False),    # HGT relation_att / relation_msg
    (50, 400, 37, 1, 32, 128, True),   # many tiny segments, wider output
    (50, 700, 4, 2, 32, 1, False),     # attention logits el/er (N == 1 row-dot path)
    (50, 300, 5, 1, 64, 1, False),
    (30, 200, 3, 1, 5, 3, True),       # odd sizes -> generic fallback
    (40, 1, 2, 1, 64, 64, True),       # single edge
]


@pytest.mark.parametrize("n,e,r,H,Kd,Nd,head_one", SHAPES)
def test_relational_matmul_fwd_bwd(cuda, n, e, r, H, Kd, Nd, head_one):
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=e + Kd)
    sc = g.get_separate_coo_original()
    torch.manual_seed(Kd * Nd + e)
    a_heads = 1 if head_one else H
    W = torch.randn(r, H, Kd, Nd, device=cuda) / Kd ** 0.5
    for gather_name in ("row_indices", "eids"):
        gather = sc[gather_name]
        rows_in = n if gather_name == "row_indices" else e
        X = torch.randn(rows_in, a_heads, Kd, device=cuda)
        args = {"separate_coo_rel_ptrs": sc["rel_ptrs"], "separate_coo_node_indices": gather,
                "separate_coo_eids": sc["eids"]}
        ret = torch.full((e, H, Nd), float("nan"), device=cuda)
        K.rgnn_relational_matmul(args, 0, W, X, ret, head_one)
        want = R.segment_mm(X.cpu(), gather.cpu(), W.cpu(), sc["eids"].cpu(), sc["rel_ptrs"].cpu(), e, a_heads)
        close(ret, want, f"fwd/{gather_name}")
        gout = torch.randn(e, H, Nd, device=cuda)
        gX, gW = torch.zeros_like(X), torch.zeros_like(W)
        K.backward_rgnn_relational_matmul(args, 0, W.transpose(2, 3).contiguous(), X, gout, gX, gW, head_one)
        gX_ref, gW_ref = R.segment_mm_bwd(X.cpu(), gather.cpu(), W.cpu(), sc["eids"].cpu(), sc["rel_ptrs"].cpu(),
                                          gout.cpu(), a_heads)
        close(gX, gX_ref, f"grad_input/{gather_name}")
        close(gW, gW_ref, f"grad_weight/{gather_name}", rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize("sizes,Kd,Nd", [([40, 0, 300, 7], 64, 64), ([129], 64, 64), ([5, 5, 5], 32, 16),
                                          ([1000, 3], 128, 128), ([17, 60], 12, 20)])
def test_no_scatter_gather_list_fwd_bwd(cuda, sizes, Kd, Nd):
    """rows sorted by node type; weights [T,1,K,N] (HGT k/q/v/a linears)."""
    from het_b200.kernels import K
    T, M = len(sizes), sum(sizes)
    offs = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int64)
    torch.manual_seed(M + Kd)
    W = torch.randn(T, 1, Kd, Nd, device=cuda) / Kd ** 0.5
    X = torch.randn(M, Kd, device=cuda)
    ret = torch.full((M, Nd), float("nan"), device=cuda)
    offs_d = offs.to(cuda)
    K.rgnn_relational_matmul_no_scatter_gather_list(offs_d, W, X, ret)
    want = R.segment_mm(X.cpu(), None, W.cpu(), None, offs, M, 1)
    close(ret, want, "fwd")
    gout = torch.randn(M, Nd, device=cuda)
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    K.backward_rgnn_relational_matmul_no_scatter_gather_list(offs_d, W.transpose(2, 3).contiguous(), X, gout, gX, gW)
    gX_ref, gW_ref = R.segment_mm_bwd(X.cpu(), None, W.cpu(), None, offs, gout.cpu(), 1)
    close(gX, gX_ref, "grad_input")
    close(gW, gW_ref, "grad_weight", rtol=2e-4, atol=2e-5)


def test_ops_reject_cpu_tensors():
    """No CPU fallback: compute ops fail loudly on CPU tensors."""
    from het_b200.kernels import K
    offs = torch.tensor([0, 4])
    with pytest.raises(RuntimeError, match="CUDA"):
        K.rgnn_relational_matmul_no_scatter_gather_list(offs, torch.zeros(1, 1, 4, 4), torch.zeros(4, 4),
                                                        torch.zeros(4, 4))


# ---------------------------------------------------------------------------
# tcgen05 tensor-core path (csrc/segment_mm_tc.cu), forced with impl=2 so that small
# test sizes do not fall back to the SIMT kernels. Same fp32 tolerance (3xTF32 split).
# ---------------------------------------------------------------------------
TC_SHAPES = [
    # (nodes, edges, relations, heads, K, N, input_num_head_one)
    (300, 5000, 3, 1, 64, 64, True),      # dense 64x64, tiles straddle segments
    (300, 4097, 40, 1, 64, 64, True),     # many small segments, ragged last tile
    (200, 3000, 5, 4, 64, 16, True),      # RGAT multi-head: one input row, 4 heads out
    (200, 3000, 8, 4, 16, 16, False),     # HGT per-head 16x16 -> block-diagonal B
    (100, 130, 2, 2, 32, 32, False),      # two heads, barely more than one tile
    (64, 7, 3, 1, 64, 64, True),          # fewer rows than a tile
]


@pytest.fixture(params=[1, 0], ids=["tma", "registers"])
def force_tc(request):
    """Tensor-core kernels at every size, once with the A tile delivered by TMA (segment_mm_tc_tma.cu, the
    default) and once through the producer warps' registers (segment_mm_tc.cu)."""
    from het_b200.kernels import K
    K.hetb200_set_segment_mm_impl(2)
    K.hetb200_set_segment_mm_tma(request.param)
    yield
    K.hetb200_set_segment_mm_tma(1)
    K.hetb200_set_segment_mm_impl(0)


@pytest.mark.parametrize("n,e,r,H,Kd,Nd,head_one", TC_SHAPES)
def test_tensor_core_segment_mm(cuda, force_tc, n, e, r, H, Kd, Nd, head_one):
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=e + Kd + 1)
    sc = g.get_separate_coo_original()
    torch.manual_seed(Kd * Nd + e)
    a_heads = 1 if head_one else H
    W = torch.randn(r, H, Kd, Nd, device=cuda) / Kd ** 0.5
    for gather_name in ("row_indices", "eids"):
        gather = sc[gather_name]
        rows_in = n if gather_name == "row_indices" else e
        X = torch.randn(rows_in, a_heads, Kd, device=cuda)
        args = {"separate_coo_rel_ptrs": sc["rel_ptrs"], "separate_coo_node_indices": gather,
                "separate_coo_eids": sc["eids"]}
        ret = torch.full((e, H, Nd), float("nan"), device=cuda)
        K.rgnn_relational_matmul(args, 0, W, X, ret, head_one)
        want = R.segment_mm(X.cpu(), gather.cpu(), W.cpu(), sc["eids"].cpu(), sc["rel_ptrs"].cpu(), e, a_heads)
        close(ret, want, f"tc fwd/{gather_name}")
        # delta-input half of the backward runs the same kernel (accumulating, stacked B)
        gout = torch.randn(e, H, Nd, device=cuda)
        gX, gW = torch.zeros_like(X), torch.zeros_like(W)
        K.backward_rgnn_relational_matmul(args, 0, W.transpose(2, 3).contiguous(), X, gout, gX, gW, head_one)
        gX_ref, gW_ref = R.segment_mm_bwd(X.cpu(), gather.cpu(), W.cpu(), sc["eids"].cpu(), sc["rel_ptrs"].cpu(),
                                          gout.cpu(), a_heads)
        close(gX, gX_ref, f"tc grad_input/{gather_name}")
        close(gW, gW_ref, f"tc grad_weight/{gather_name}", rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize("gather,scatter,mode", [(False, False, "dense"), (True, False, "blockdiag"),
                                                 (True, False, "heads_out"), (False, True, "dense"),
                                                 (True, False, "bf16")])
def test_tma_and_register_paths_give_the_same_bits(cuda, gather, scatter, mode):
    """The TMA-fed kernel only changes how the A tile reaches shared memory: products, accumulation order and
    stores are those of the register-path kernel, so plain stores agree bit for bit (C ABI, ragged tail,
    segments cut inside tiles, rows past the end of the table in the last tile)."""
    import ctypes
    from het_b200.kernels import KERNEL_LIB
    lib = ctypes.CDLL(str(KERNEL_LIB))
    vp, i64, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    lib.hetb_segment_mm.argtypes = [vp, vp, ci, vp, vp, vp, ci, vp, i64, i64, ci, ci, ci, ci, vp]
    lib.hetb_segment_mm_bf16out.argtypes = [vp, vp, ci, vp, vp, vp, vp, i64, i64, ci, ci, ci, vp]
    g = torch.Generator().manual_seed(17)
    table, rows = 3000, (20003 if gather else 3000 - 1)
    sizes = torch.tensor([rows // 2 - 77, 0, 130, rows - (rows // 2 - 77) - 130])
    seg = torch.cat([torch.zeros(1, dtype=torch.int64), sizes.cumsum(0)]).to(cuda)
    A = torch.randn(table, 64, device=cuda)
    a_idx = torch.randint(0, table, (rows,), generator=g).to(cuda) if gather else None
    c_idx = torch.randperm(table, generator=g)[:rows].to(cuda) if scatter else None
    H, K_, N_, ah, ch = {"dense": (1, 64, 64, 1, 1), "blockdiag": (4, 16, 16, 4, 4), "bf16": (4, 16, 16, 4, 4),
                         "heads_out": (4, 64, 16, 1, 4)}[mode]
    W = torch.randn(4, H, K_, N_, device=cuda) / K_ ** 0.5
    st = torch.cuda.current_stream().cuda_stream
    ptr = lambda t: None if t is None else t.data_ptr()
    outs = []
    lib.hetb_set_segment_mm_impl(2)
    try:
        for tma in (0, 1):
            lib.hetb_set_segment_mm_tma(tma)
            Y = torch.full((table if scatter else rows, 64), float("nan"), device=cuda,
                           dtype=torch.bfloat16 if mode == "bf16" else torch.float32)
            if mode == "bf16":
                rc = lib.hetb_segment_mm_bf16out(ptr(A), ptr(a_idx), ah, ptr(W), ptr(Y), ptr(c_idx), ptr(seg), 4, rows,
                                                 H, K_, N_, st)
            else:
                rc = lib.hetb_segment_mm(ptr(A), ptr(a_idx), ah, ptr(W), ptr(Y), ptr(c_idx), ch, ptr(seg), 4, rows,
                                         H, K_, N_, 0, st)
            assert rc == 0
            torch.cuda.synchronize()
            outs.append(torch.nan_to_num(Y.float(), nan=-7.0))
    finally:
        lib.hetb_set_segment_mm_tma(1)
        lib.hetb_set_segment_mm_impl(0)
    assert torch.equal(outs[0], outs[1])
    want = R.segment_mm(A.cpu(), a_idx.cpu() if gather else None, W.cpu(), None, seg.cpu(), rows, ah)
    got = outs[1][c_idx] if scatter else outs[1]
    close(got.view(rows, H, N_), want, "tma vs oracle", **(dict(rtol=2e-2, atol=2e-2) if mode == "bf16" else {}))


def test_tensor_core_matches_simt_bitwise_close(cuda):
    """A/B: the tcgen05 kernel and the SIMT kernel agree to fp32 rounding on a graph big
    enough for the automatic dispatch to pick the tensor-core path."""
    from het_b200.kernels import K
    g = build(cuda, 5000, 60000, 11, seed=5)
    sc = g.get_separate_coo_original()
    torch.manual_seed(3)
    W = torch.randn(11, 1, 64, 64, device=cuda) / 8
    X = torch.randn(5000, 1, 64, device=cuda)
    args = {"separate_coo_rel_ptrs": sc["rel_ptrs"], "separate_coo_node_indices": sc["row_indices"],
            "separate_coo_eids": sc["eids"]}
    outs = []
    for impl in (1, 2, 0):
        K.hetb200_set_segment_mm_impl(impl)
        ret = torch.full((60000, 1, 64), float("nan"), device=cuda)
        K.rgnn_relational_matmul(args, 0, W, X, ret, True)
        outs.append(ret)
    K.hetb200_set_segment_mm_impl(0)
    close(outs[1], outs[0], "tc vs simt", rtol=1e-5, atol=1e-6)
    assert torch.equal(outs[2], outs[1]), "automatic dispatch should have used the tensor-core kernel"


# ---------------------------------------------------------------------------
# tcgen05 weight gradient (csrc/segment_mm_gradw_tc.cu): contraction over the rows with
# MN-major operands; many CTAs, empty segments, per-head block extraction.
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("sizes,H,Kd,Nd,a_heads", [
    ([40, 0, 300, 7], 1, 64, 64, 1),            # one CTA, empty segment, ragged stages
    ([30000, 0, 1, 12345, 31, 33, 20000], 1, 64, 64, 1),   # 148 CTAs, segment runs spanning CTAs
    ([5000, 7000, 100, 9000], 4, 16, 16, 4),    # HGT per-head: only the diagonal 16x16 blocks are flushed
    ([6000, 2000, 3000], 4, 64, 16, 1),         # RGAT multi-head: one input row, 4 heads of gradient
    ([3] * 700, 1, 64, 64, 1),                  # hundreds of tiny segments (wikikg2-like tail)
])
def test_tensor_core_grad_w(cuda, force_tc, sizes, H, Kd, Nd, a_heads):
    from het_b200.kernels import K
    S, M = len(sizes), sum(sizes)
    ptrs = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int64)
    torch.manual_seed(M + Kd + H)
    nodes = max(8, M // 3)
    gather = torch.randint(0, nodes, (M,))
    eids = torch.randperm(M)
    X = torch.randn(nodes, a_heads, Kd, device=cuda)
    W = torch.randn(S, H, Kd, Nd, device=cuda) / Kd ** 0.5
    gout = torch.randn(M, H, Nd, device=cuda)
    args = {"separate_coo_rel_ptrs": ptrs.to(cuda), "separate_coo_node_indices": gather.to(cuda),
            "separate_coo_eids": eids.to(cuda)}
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    K.backward_rgnn_relational_matmul(args, 0, W.transpose(2, 3).contiguous(), X, gout, gX, gW, a_heads == 1)
    gX_ref, gW_ref = R.segment_mm_bwd(X.cpu(), gather, W.cpu(), eids, ptrs, gout.cpu(), a_heads)
    close(gX, gX_ref, "tc grad_input")
    close(gW, gW_ref, "tc grad_weight", rtol=2e-4, atol=2e-5)
    # SIMT kernel on the same inputs: the two implementations agree far inside the oracle tolerance
    K.hetb200_set_segment_mm_impl(1)
    gX1, gW1 = torch.zeros_like(X), torch.zeros_like(W)
    K.backward_rgnn_relational_matmul(args, 0, W.transpose(2, 3).contiguous(), X, gout, gX1, gW1, a_heads == 1)
    K.hetb200_set_segment_mm_impl(2)
    close(gW, gW1.cpu(), "tc vs simt grad_weight", rtol=2e-5, atol=2e-6)


def test_tensor_core_grad_w_no_gather(cuda, force_tc):
    """HGT k/q/v/a linears: rows sorted by node type, no index lists."""
    from het_b200.kernels import K
    sizes = [7000, 11000, 90, 600]
    T, M = len(sizes), sum(sizes)
    offs = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int64)
    torch.manual_seed(11)
    W = torch.randn(T, 1, 64, 64, device=cuda) / 8
    X = torch.randn(M, 64, device=cuda)
    gout = torch.randn(M, 64, device=cuda)
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    K.backward_rgnn_relational_matmul_no_scatter_gather_list(offs.to(cuda), W.transpose(2, 3).contiguous(), X, gout,
                                                             gX, gW)
    gX_ref, gW_ref = R.segment_mm_bwd(X.cpu(), None, W.cpu(), None, offs, gout.cpu(), 1)
    close(gX, gX_ref, "grad_input")
    close(gW, gW_ref, "grad_weight", rtol=2e-4, atol=2e-5)


# ---------------------------------------------------------------------------
# tcgen05 path for hidden sizes above 64 (csrc/segment_mm_tc_wide.cu): both operands streamed,
# 128 x 256 / 128 / 64 accumulator tiles. Forced with impl=2; 3xTF32, same fp32 tolerance.
# ---------------------------------------------------------------------------
WIDE_SHAPES = [
    # (nodes, edges, relations, heads, K, N, input_num_head_one)
    (300, 5000, 3, 1, 128, 128, True),     # NT = 128, 4 K panels, tiles straddle segments
    (300, 4097, 40, 1, 256, 256, True),    # NT = 256, many small segments, ragged last tile
    (200, 3000, 5, 1, 512, 512, True),     # two N tiles of 256, 16 K panels
    (200, 3000, 5, 4, 128, 64, True),      # RGAT multi-head: one 128-wide input row, 4 x 64 out (NT = 256)
    (100, 1500, 2, 1, 64, 192, True),      # NT = 64 (192 = 3 x 64), K = 64
    (100, 1500, 2, 1, 96, 64, True),       # K = 96 (3 panels), N = 64
    (64, 7, 3, 1, 256, 128, True),         # fewer rows than a tile
]


@pytest.mark.parametrize("n,e,r,H,Kd,Nd,head_one", WIDE_SHAPES)
def test_wide_tensor_core_segment_mm(cuda, force_tc, n, e, r, H, Kd, Nd, head_one):
    from het_b200.kernels import K
    g = build(cuda, n, e, r, seed=e + Kd + 2)
    sc = g.get_separate_coo_original()
    torch.manual_seed(Kd * Nd + e)
    W = torch.randn(r, H, Kd, Nd, device=cuda) / Kd ** 0.5
    for gather_name in ("row_indices", "eids"):
        gather = sc[gather_name]
        rows_in = n if gather_name == "row_indices" else e
        X = torch.randn(rows_in, 1, Kd, device=cuda)
        args = {"separate_coo_rel_ptrs": sc["rel_ptrs"], "separate_coo_node_indices": gather,
                "separate_coo_eids": sc["eids"]}
        ret = torch.full((e, H, Nd), float("nan"), device=cuda)
        K.rgnn_relational_matmul(args, 0, W, X, ret, head_one)
        want = R.segment_mm(X.cpu(), gather.cpu(), W.cpu(), sc["eids"].cpu(), sc["rel_ptrs"].cpu(), e, 1)
        close(ret, want, f"wide fwd/{gather_name}")


@pytest.mark.parametrize("sizes,Kd,Nd", [([1000, 0, 3000, 77], 128, 128), ([4099], 256, 256), ([700, 900], 512, 512),
                                          ([300, 300, 300], 256, 64)])
def test_wide_typed_linear_fwd_bwd(cuda, force_tc, sizes, Kd, Nd):
    """The a3 op (rows sorted by node type) at hidden 128 / 256 / 512: forward and delta-input run the wide
    tcgen05 kernel (the latter accumulating), delta-W the SIMT kernel."""
    from het_b200.kernels import K
    T, M = len(sizes), sum(sizes)
    offs = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int64)
    torch.manual_seed(M + Kd)
    W = torch.randn(T, 1, Kd, Nd, device=cuda) / Kd ** 0.5
    X = torch.randn(M, Kd, device=cuda)
    ret = torch.full((M, Nd), float("nan"), device=cuda)
    offs_d = offs.to(cuda)
    K.rgnn_relational_matmul_no_scatter_gather_list(offs_d, W, X, ret)
    want = R.segment_mm(X.cpu(), None, W.cpu(), None, offs, M, 1)
    close(ret, want, "wide fwd")
    gout = torch.randn(M, Nd, device=cuda)
    gX, gW = torch.zeros_like(X), torch.zeros_like(W)
    K.hetb200_set_segment_mm_impl(0)  # delta-W has no wide tensor-core kernel: automatic dispatch
    K.backward_rgnn_relational_matmul_no_scatter_gather_list(offs_d, W.transpose(2, 3).contiguous(), X, gout, gX, gW)
    gX_ref, gW_ref = R.segment_mm_bwd(X.cpu(), None, W.cpu(), None, offs, gout.cpu(), 1)
    close(gX, gX_ref, "wide grad_input")
    close(gW, gW_ref, "wide grad_weight", rtol=2e-4, atol=2e-5)
