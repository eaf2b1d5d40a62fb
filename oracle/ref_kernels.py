"""TEST INFRASTRUCTURE ONLY — ctypes front end of ``oracle/_ref/libhet_ref_kernels.so``.

The library holds the REFERENCE's own CUDA kernel templates (hrt/include/DGLHackKernel/**),
instantiated and launched exactly as the reference's torch op wrappers do
(oracle/ref_kernels_shim.cu cites file:line per launcher) and compiled for sm_100a by
``oracle/build_ref.py: build_kernels`` in the build container. On the GPU box the ``-m gpu`` tests
use it to pin the fp64 oracle (oracle/het_ops_ref.py) and the product to reference outputs.
Nothing under ``het_b200/`` imports this module.

Every wrapper takes CUDA torch tensors (int64 indices, contiguous fp32 data), allocates the
outputs the way the reference's Python does — except that buffers the reference kernels
accumulate into are zero-filled here (the reference hands some of them over uninitialised,
SURVEY.md §8a a7/a8) — and returns them.
"""
from __future__ import annotations

import ctypes
from pathlib import Path

import torch

LIB_PATH = Path(__file__).resolve().parent / "_ref" / "libhet_ref_kernels.so"
_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(str(LIB_PATH))
    return _lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _i(v):
    return ctypes.c_int64(int(v))


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"reference kernel {what} failed: cudaError {rc}")


def rgat_fwd_csr(csr, feat, el, er, slope):
    n = csr["row_ptrs"].numel() - 1
    e, h = feat.shape[0], feat.shape[1]
    d = feat.shape[1] * feat.shape[2]
    s = torch.zeros(n, h, device=feat.device)
    ex = torch.zeros(e, h, device=feat.device)
    ret = torch.zeros(n, h, feat.shape[2], device=feat.device)
    _check(lib().ref_rgat_fwd_csr(_p(csr["row_ptrs"]), _p(csr["col_indices"]), _p(csr["eids"]), _p(csr["rel_types"]),
                                  _i(n), _p(feat), _p(el), _p(er), _p(s), _p(ex), _p(ret), _i(h), _i(d),
                                  ctypes.c_float(slope), _stream()), "rgat_fwd_csr")
    return s, ex, ret


def rgat_bwd_csr(csr, feat, el, er, s, ex, ret, gout, slope):
    n = csr["row_ptrs"].numel() - 1
    h, d = feat.shape[1], feat.shape[1] * feat.shape[2]
    gf, gel, ger = torch.zeros_like(feat), torch.zeros_like(el), torch.zeros_like(er)
    _check(lib().ref_rgat_bwd_csr(_p(csr["row_ptrs"]), _p(csr["col_indices"]), _p(csr["eids"]), _p(csr["rel_types"]),
                                  _i(n), _p(feat), _p(el), _p(er), _p(s), _p(ex), _p(ret), _p(gout), _p(gf), _p(gel),
                                  _p(ger), _i(h), _i(d), ctypes.c_float(slope), _stream()), "rgat_bwd_csr")
    return gf, gel, ger


def hgt_edge_softmax_fwd_csr(csr, attn, mu):
    n = csr["row_ptrs"].numel() - 1
    h = attn.shape[1]
    s = torch.zeros(n, h, device=attn.device)
    t, norm = torch.zeros_like(attn), torch.zeros_like(attn)
    _check(lib().ref_hgt_edge_softmax_fwd_csr(_p(csr["row_ptrs"]), _p(csr["col_indices"]), _p(csr["eids"]),
                                              _p(csr["rel_types"]), _i(n), _p(attn), _p(mu), _p(s), _p(t), _p(norm),
                                              _i(h), _stream()), "hgt_edge_softmax_fwd_csr")
    return s, t, norm


def hgt_aggregate_fwd_csr(csr, msg, score, s, use_switch):
    n = csr["row_ptrs"].numel() - 1
    h, d = msg.shape[1], msg.shape[1] * msg.shape[2]
    ret = torch.zeros(n, h, msg.shape[2], device=msg.device)
    _check(lib().ref_hgt_aggregate_fwd_csr(_p(csr["row_ptrs"]), _p(csr["col_indices"]), _p(csr["rel_types"]),
                                           _p(csr["eids"]), _i(n), _p(msg), _p(score), _p(s), _p(ret), _i(h), _i(d),
                                           ctypes.c_int(use_switch), _stream()), "hgt_aggregate_fwd_csr")
    return ret


def hgt_aggregate_bwd_csr(csr, norm, gout, msg_shape):
    n = csr["row_ptrs"].numel() - 1
    h, d = msg_shape[1], msg_shape[1] * msg_shape[2]
    gm = torch.zeros(msg_shape, device=gout.device)
    _check(lib().ref_hgt_aggregate_bwd_csr(_p(csr["row_ptrs"]), _p(csr["col_indices"]), _p(csr["rel_types"]),
                                           _p(csr["eids"]), _i(n), _p(norm), _p(gout), _p(gm), _i(h), _i(d),
                                           _stream()), "hgt_aggregate_bwd_csr")
    return gm


def hgt_edge_softmax_bwd_csr(csr, msg, attn, norm, out, gout, mu):
    n = csr["row_ptrs"].numel() - 1
    h, d = msg.shape[1], msg.shape[1] * msg.shape[2]
    ga, gmu = torch.zeros_like(attn), torch.zeros_like(mu)
    _check(lib().ref_hgt_edge_softmax_bwd_csr(_p(csr["row_ptrs"]), _p(csr["col_indices"]), _p(csr["eids"]),
                                              _p(csr["rel_types"]), _i(n), _p(msg), _p(attn), _p(norm), _p(out),
                                              _p(gout), _p(mu), _p(ga), _p(gmu), _i(h), _i(d), _stream()),
           "hgt_edge_softmax_bwd_csr")
    return ga, gmu


def inner_product_fwd_coo(coo, left, right):
    e, h, d = left.shape[0], left.shape[1], left.shape[1] * left.shape[2]
    out = torch.empty(e, h, device=left.device)
    _check(lib().ref_inner_product_fwd_coo(_p(coo["rel_ptrs"]), _p(coo["eids"]), _p(coo["row_indices"]),
                                           _p(coo["col_indices"]), _i(e), _i(coo["rel_ptrs"].numel() - 1), _p(left),
                                           _p(right), _p(out), _i(h), _i(d), _stream()), "inner_product_fwd_coo")
    return out


def inner_product_bwd_coo(coo, left, right, gout):
    e, h, d = left.shape[0], left.shape[1], left.shape[1] * left.shape[2]
    gl, gr = torch.zeros_like(left), torch.zeros_like(right)
    _check(lib().ref_inner_product_bwd_coo(_p(coo["rel_ptrs"]), _p(coo["eids"]), _p(coo["row_indices"]),
                                           _p(coo["col_indices"]), _i(e), _i(coo["rel_ptrs"].numel() - 1), _p(left),
                                           _p(right), _p(gout), _p(gl), _p(gr), _i(h), _i(d), _stream()),
           "inner_product_bwd_coo")
    return gl, gr


def relational_matmul_fwd(rel_ptrs, gather, eids, weights, node_feat, input_num_head_one):
    r, h, k, n = weights.shape
    e = eids.numel()
    ret = torch.zeros(e, h, n, device=weights.device)
    _check(lib().ref_relational_matmul_fwd(_p(rel_ptrs), _p(gather), _p(eids), _i(r), _i(e), _p(weights),
                                           _p(node_feat), _p(ret), _i(h), _i(k), _i(n),
                                           ctypes.c_int(int(input_num_head_one)), _stream()), "relational_matmul_fwd")
    return ret


def relational_matmul_bwd(rel_ptrs, gather, eids, weights, node_feat, gradout, input_num_head_one):
    r, h, k, n = weights.shape
    e = eids.numel()
    wt = weights.transpose(2, 3).contiguous()
    gx, gw = torch.zeros_like(node_feat), torch.zeros_like(weights)
    _check(lib().ref_relational_matmul_bwd(_p(rel_ptrs), _p(gather), _p(eids), _i(r), _i(e), _p(wt), _p(node_feat),
                                           _p(gradout), _p(gx), _p(gw), _i(h), _i(k), _i(n),
                                           ctypes.c_int(int(input_num_head_one)), _stream()), "relational_matmul_bwd")
    return gx, gw


def matmul_no_scatter_gather_fwd(offsets, weights, inputs):
    t, h, k, n = weights.shape
    rows = inputs.shape[0]
    ret = torch.zeros(rows, h * n, device=weights.device)
    _check(lib().ref_matmul_no_scatter_gather_fwd(_p(offsets), _i(t), _i(rows), _p(weights), _p(inputs), _p(ret),
                                                  _i(h), _i(k), _i(n), _stream()), "matmul_no_scatter_gather_fwd")
    return ret


def matmul_no_scatter_gather_bwd(offsets, weights, inputs, gradout):
    t, h, k, n = weights.shape
    rows = inputs.shape[0]
    wt = weights.transpose(2, 3).contiguous()
    gx, gw = torch.zeros_like(inputs), torch.zeros_like(weights)
    _check(lib().ref_matmul_no_scatter_gather_bwd(_p(offsets), _i(t), _i(rows), _p(wt), _p(inputs), _p(gradout),
                                                  _p(gx), _p(gw), _i(h), _i(k), _i(n), _stream()),
           "matmul_no_scatter_gather_bwd")
    return gx, gw
