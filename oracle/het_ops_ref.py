"""CPU oracle for the HET/Hector hot path — TEST INFRASTRUCTURE ONLY.

A plain torch (CPU, float64 by default) restatement of what each reference op
computes, written from the reference kernels cited per function (paths relative
to the reference repo root). Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this
package; the product (``het_b200``) never does.

Parity status: PINNED to the reference's own CUDA kernels. The reference ships no golden
vectors for these ops and its CMake build cannot run here (glog, g++-9, cusp vs CUDA 12.9), but
its kernel templates are header-only: ``oracle/ref_kernels_shim.cu`` instantiates them exactly as
the reference's op wrappers do, ``oracle/build_ref.py`` compiles that for sm_100a into
``oracle/_ref/libhet_ref_kernels.so``, and ``tests/test_gpu_reference_kernels.py`` checks on the
B200 that reference kernel == this oracle == hetb200 for every function below (with the
reference's defects — found by running it — either modelled or avoided, and named in the tests).
The integer converters are pinned the same way: ``oracle/converters_ref.py`` and ``oracle/_ref``
(reference MyHyb.h compiled).

Documented deviations from reference bugs (oracle follows the maths):
  * inner-product backward sums into grad_right (reference stores, last edge wins:
    RGNN/InnerProductEdgeParallel.cu.h:180-183);
  * HGT softmax backward uses each edge's relation for mu / grad_mu (reference
    reads relation 0 in the non-compact path: HGT/HGTBackwardKernels.cu.h:533);
  * multi-head inputs use their own head (reference forward reads head 0 for all
    heads through MySGEMMNumHeadKind::AssertANumIsOne:
    RGNN/my_shmem_sgemm_func.cu.h:519 — identical for the 1-head benchmark matrix).
"""
from __future__ import annotations

import torch

F64 = torch.float64


def _seg_ids(seg_ptrs: torch.Tensor, m: int) -> torch.Tensor:
    """segment id of each row 0..m-1 for offsets seg_ptrs[S+1]."""
    counts = (seg_ptrs[1:] - seg_ptrs[:-1]).to(torch.int64)
    return torch.repeat_interleave(torch.arange(counts.numel()), counts)[:m]


# --------------------------------------------------------------------------
# segment MM — RGNN/my_shmem_sgemm_func.cu.h:504-531 (fwd), :711-776 (bwd);
# index roles from OpExport/RGNNOps.inc.h:95-236.
# --------------------------------------------------------------------------
def segment_mm(A, a_idx, W, c_idx, seg_ptrs, out_rows, a_heads, dtype=F64):
    """Y[c_idx[i],h,:] = A[a_idx[i], (0|h), :] @ W[seg(i),h]   -> [out_rows,H,N]"""
    S, H, K, N = W.shape
    m = int(seg_ptrs[-1])
    seg = _seg_ids(seg_ptrs, m)
    ar = a_idx[:m] if a_idx is not None else torch.arange(m)
    cr = c_idx[:m] if c_idx is not None else torch.arange(m)
    a = A.to(dtype).reshape(A.shape[0], a_heads, K)[ar]  # [m,a_heads,K]
    if a_heads == 1:
        a = a.expand(m, H, K)
    w = W.to(dtype)[seg]  # [m,H,K,N]
    y = torch.einsum("mhk,mhkn->mhn", a, w)
    out = torch.zeros(out_rows, H, N, dtype=dtype)
    out[cr] = y
    return out


def segment_mm_bwd(A, a_idx, W, c_idx, seg_ptrs, gout, a_heads, dtype=F64):
    """returns (dA like A viewed [rows,a_heads,K], dW like W)."""
    S, H, K, N = W.shape
    m = int(seg_ptrs[-1])
    seg = _seg_ids(seg_ptrs, m)
    ar = a_idx[:m] if a_idx is not None else torch.arange(m)
    cr = c_idx[:m] if c_idx is not None else torch.arange(m)
    a = A.to(dtype).reshape(A.shape[0], a_heads, K)[ar]
    if a_heads == 1:
        a = a.expand(m, H, K)
    g = gout.to(dtype).reshape(gout.shape[0], H, N)[cr]  # [m,H,N]
    w = W.to(dtype)[seg]
    da_rows = torch.einsum("mhn,mhkn->mhk", g, w)  # [m,H,K]
    if a_heads == 1:
        da_rows = da_rows.sum(1, keepdim=True)
    dA = torch.zeros(A.shape[0], a_heads, K, dtype=dtype)
    dA.index_add_(0, ar, da_rows)
    dW = torch.zeros(S, H, K, N, dtype=dtype)
    dW.index_add_(0, seg, torch.einsum("mhk,mhn->mhkn", a, g))
    return dA, dW


# --------------------------------------------------------------------------
# inner product — RGNN/InnerProductEdgeParallel.cu.h:22-75 (left by eid, right
# by row_indices)
# --------------------------------------------------------------------------
def inner_product(eids, row_idx, left, right, dtype=F64):
    out = torch.zeros(left.shape[0], left.shape[1], dtype=dtype)
    out[eids] = (left.to(dtype)[eids] * right.to(dtype)[row_idx]).sum(-1)
    return out


def inner_product_bwd(eids, row_idx, left, right, gout, dtype=F64):
    g = gout.to(dtype)[eids].unsqueeze(-1)
    gl = torch.zeros_like(left, dtype=dtype)
    gl[eids] = g * right.to(dtype)[row_idx]
    gr = torch.zeros_like(right, dtype=dtype)
    gr.index_add_(0, row_idx, g * left.to(dtype)[eids])
    return gl, gr


def _rows_of(row_ptr: torch.Tensor) -> torch.Tensor:
    counts = row_ptr[1:] - row_ptr[:-1]
    return torch.repeat_interleave(torch.arange(counts.numel()), counts)


# --------------------------------------------------------------------------
# RGAT — GAT/FusedGAT.cu.h:123-211 (exp/sum), :33-105 (aggregate);
# gradients: GAT/FusedGATBackward.cu.h:245-266 (closed form), :31-148.
# --------------------------------------------------------------------------
def rgat_fwd(dst_of_pos, eids, num_nodes, feat, el, er, slope, dtype=F64):
    """dst_of_pos[p] = destination of CSR/COO position p, eids[p] = edge row.
    returns (sum[N,H], exp[E,H], ret[N,H,F])."""
    E, H, Fh = feat.shape
    x = el.to(dtype).reshape(E, H) + er.to(dtype).reshape(E, H)
    ex = torch.exp(torch.where(x > 0, x, slope * x))  # indexed by edge id
    s = torch.zeros(num_nodes, H, dtype=dtype)
    s.index_add_(0, dst_of_pos, ex[eids])
    ret = torch.zeros(num_nodes, H, Fh, dtype=dtype)
    a = ex[eids] / s[dst_of_pos]
    ret.index_add_(0, dst_of_pos, a.unsqueeze(-1) * feat.to(dtype)[eids])
    return s, ex, ret


def rgat_bwd(dst_of_pos, eids, feat, el, er, s, ex, ret, gout, slope, dtype=F64):
    E, H, Fh = feat.shape
    g = gout.to(dtype)[dst_of_pos]
    a = ex.to(dtype)[eids] / s.to(dtype)[dst_of_pos]
    gfeat = torch.zeros(E, H, Fh, dtype=dtype)
    gfeat[eids] = a.unsqueeze(-1) * g
    x = (el.to(dtype).reshape(E, H) + er.to(dtype).reshape(E, H))[eids]
    t = (g * (feat.to(dtype)[eids] - ret.to(dtype)[dst_of_pos])).sum(-1)
    t = t * a * torch.where(x > 0, torch.ones_like(x), torch.full_like(x, slope))
    gel = torch.zeros(E, H, dtype=dtype)
    gel[eids] = t
    return gfeat, gel, gel.clone()


# --------------------------------------------------------------------------
# HGT — HGT/HGTForwardKernels.cu.h:425-587 (softmax), :244-369 (aggregate);
# gradients HGT/HGTBackwardKernels.cu.h:160-170, 343-440, 499-627.
# --------------------------------------------------------------------------
def hgt_edge_softmax(dst_of_pos, eids, rel_of_pos, num_nodes, attn, mu, dtype=F64):
    """returns (sum[N,H], t[E,H] = exp(attn*mu), norm[E,H])."""
    E, H = attn.shape
    t_pos = torch.exp(attn.to(dtype)[eids] * mu.to(dtype)[rel_of_pos])
    s = torch.zeros(num_nodes, H, dtype=dtype)
    s.index_add_(0, dst_of_pos, t_pos)
    t = torch.zeros(E, H, dtype=dtype)
    t[eids] = t_pos
    norm = torch.zeros(E, H, dtype=dtype)
    norm[eids] = t_pos / s[dst_of_pos]
    return s, t, norm


def hgt_aggregate(dst_of_pos, eids, num_nodes, msg, score, s=None, dtype=F64):
    """ret[d] = sum_p w * msg[e]; w = score/sum[d] if s is given (the reference
    kernel's switch 1, HGTForwardKernels.cu.h:304-309) else score."""
    E, H, Fh = msg.shape
    w = score.to(dtype)[eids]
    if s is not None:
        w = w / s.to(dtype)[dst_of_pos]
    ret = torch.zeros(num_nodes, H, Fh, dtype=dtype)
    ret.index_add_(0, dst_of_pos, w.unsqueeze(-1) * msg.to(dtype)[eids])
    return ret


def hgt_aggregate_bwd(dst_of_pos, eids, norm, gout, dtype=F64):
    E, H = norm.shape
    g = gout.to(dtype)[dst_of_pos]
    gm = torch.zeros(E, H, g.shape[-1], dtype=dtype)
    gm[eids] = norm.to(dtype)[eids].unsqueeze(-1) * g
    return gm


def hgt_edge_softmax_bwd(dst_of_pos, eids, rel_of_pos, msg, attn, norm, out, gout,
                         mu, dtype=F64):
    E, H = attn.shape
    g = gout.to(dtype)[dst_of_pos]
    sv = (g * (msg.to(dtype)[eids] - out.to(dtype)[dst_of_pos])).sum(-1)
    sv = sv * norm.to(dtype)[eids]
    ga = torch.zeros(E, H, dtype=dtype)
    ga[eids] = sv * mu.to(dtype)[rel_of_pos]
    gmu = torch.zeros_like(mu, dtype=dtype)
    gmu.index_add_(0, rel_of_pos, sv * attn.to(dtype)[eids])
    return ga, gmu
