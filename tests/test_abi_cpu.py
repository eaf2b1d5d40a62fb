"""CPU: the C-ABI library loads and exports every symbol include/hetb200.h declares; the torch
shim registers the reference's op names. No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "het_b200", "lib", "libhetb200.so")
HDR = os.path.join(ROOT, "include", "hetb200.h")


def declared_symbols():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hetb_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import __graft_entry__
        __graft_entry__.build()
    return ctypes.CDLL(LIB)


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for must in ("hetb_segment_mm", "hetb_segment_mm_grad_w", "hetb_rgat_fwd_csr", "hetb_rgat_bwd",
                 "hetb_hgt_edge_softmax_fwd_csr", "hetb_hgt_aggregate_fwd_csr", "hetb_inner_product_fwd",
                 "hetb_stable_sort_by_key"):
        assert must in names


def test_every_declared_symbol_is_exported(lib):
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, f"declared in include/hetb200.h but not exported: {missing}"


def test_abi_version_and_build_info(lib):
    lib.hetb_abi_version.restype = ctypes.c_int
    assert lib.hetb_abi_version() >= 1
    lib.hetb_build_info.restype = ctypes.c_char_p
    info = lib.hetb_build_info().decode()
    assert "sm_100a" in info
    lib.hetb_launch_count.restype = ctypes.c_int64
    assert lib.hetb_launch_count() == 0  # nothing launched on a CPU box


def test_argument_validation_without_gpu(lib):
    # bad sizes are rejected before any CUDA call (HETB_EINVAL = -1)
    lib.hetb_segment_mm.restype = ctypes.c_int
    i64 = ctypes.c_int64
    rc = lib.hetb_segment_mm(None, None, 1, None, None, None, 1, None, i64(1), i64(-5), 1, 4, 4, 0, None)
    assert rc == -1
    # zero rows is a no-op success
    rc = lib.hetb_segment_mm(None, None, 1, None, None, None, 1, None, i64(0), i64(0), 1, 4, 4, 0, None)
    assert rc == 0


REFERENCE_OPS = [
    "build_debug_info", "rgnn_relational_matmul", "backward_rgnn_relational_matmul",
    "rgnn_relational_matmul_no_scatter_gather_list", "backward_rgnn_relational_matmul_no_scatter_gather_list",
    "rgnn_inner_product_right_node_separatecoo", "backward_inner_product_right_node_separatecoo",
    "relational_fused_gat_csr", "backward_relational_fused_gat_csr", "relational_fused_gat_separate_coo",
    "backward_relational_fused_gat_separate_coo", "hgt_full_graph_edge_softmax_ops_csr",
    "hgt_full_graph_message_mean_aggregation_csr", "backward_hgt_full_graph_message_mean_aggregation_csr",
    "backward_hgt_full_graph_edge_softmax_ops_csr", "transpose_csr", "convert_integrated_csr_to_separate_coo",
    "convert_integrated_coo_to_separate_coo", "convert_integrated_csr_to_separate_csr",
    "convert_integrated_coo_to_separate_csr",
]


def test_torch_shim_registers_reference_op_names(lib):
    from het_b200.kernels import K
    for name in REFERENCE_OPS:
        assert hasattr(K, name), f"torch.ops.torch_hrt.{name} not registered"


def test_op_schemas_match_reference_argument_order(lib):
    import torch
    from het_b200.kernels import K  # noqa: F401
    s = str(torch._C._get_schema("torch_hrt::relational_fused_gat_csr", ""))
    assert s.count("Tensor(") >= 12 and "float" in s and "bool" in s  # 12 in/out tensors, slope, flag
    s = str(torch._C._get_schema("torch_hrt::rgnn_relational_matmul", ""))
    assert s.startswith("torch_hrt::rgnn_relational_matmul(Dict(str, Tensor)") and s.rstrip().endswith("-> ()")
