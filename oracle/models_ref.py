"""Layer-level CPU oracle (plain differentiable torch) — TEST INFRASTRUCTURE ONLY.

Restates, on an integrated COO (src, dst, etype) and with ordinary torch ops so
that torch.autograd supplies the gradients, what one HET layer computes:

* ``rgat_layer``  hrt/python/RGAT/models.py:265-385 with the op semantics of
  GAT/FusedGAT.cu.h:123-211,33-105 (softmax over ALL in-edges of a destination,
  across relations — SURVEY.md F4).
* ``hgt_layer``   hrt/python/HGT/models.py:156-286 with HGT/HGTForwardKernels.cu.h
  :425-587,244-369 (single softmax normalisation, see het_ops_ref docstring).
* ``dgl_*``       the repo's DGL baselines restated without DGL
  (hrt/python/RGAT/models_dgl.py:64-150: per-relation GATConv summed over
  relations; hrt/python/HGT/models_dgl.py:72-140: per-relation edge_softmax,
  cross-relation mean; graphiler RGCN_DGL.py:9-19) — the CPU baseline that
  bench.py times. DGL itself is not installable here ("unavailable on this box").
"""
from __future__ import annotations

import torch


def _segment_softmax_sum(w, dst, n):
    s = torch.zeros(n, w.shape[1], dtype=w.dtype).index_add_(0, dst, w)
    return w / s[dst]


def _rel_mm(rows, w, etype):
    """rows [E,(H,)K], w [R,H,K,F] -> [E,H,F]; per-relation loop keeps memory O(E*H*F)."""
    E, (R, H, K, Fh) = rows.shape[0], w.shape
    out = torch.zeros(E, H, Fh, dtype=rows.dtype)
    pieces, idxs = [], []
    for r in range(R):
        idx = torch.nonzero(etype == r).squeeze(1)
        if idx.numel() == 0:
            continue
        a = rows[idx]
        y = torch.einsum("ek,hkf->ehf", a, w[r]) if a.dim() == 2 else torch.einsum("ehk,hkf->ehf", a, w[r])
        pieces.append(y)
        idxs.append(idx)
    if pieces:
        out = out.index_add(0, torch.cat(idxs), torch.cat(pieces))
    return out


def rgat_layer(src, dst, etype, x, conv_w, attn_l, attn_r, slope, bias=None, branch=None, kink_eps=2e-5):
    """x [N,Din]; conv_w [R,H,Din,F]; attn_* [R,H,F] -> [N,H*F].

    ``branch`` ([E,H] bool, edge order of src/dst): leaky-ReLU is not differentiable at 0, and on a
    graph of a million edges a few logits el + er land within fp32 rounding of it, where a correct
    fp32 kernel may take either one-sided derivative (both are valid subgradients, and weight
    gradients change by ~1e-3 with the choice). Where |el + er| < kink_eps the oracle takes the
    branch given here — the sign the implementation under test computed for that edge
    (HET_RGATLayer.logit_signs) — everywhere else its own; the forward value moves by
    < 0.8 * kink_eps. None = plain leaky-ReLU."""
    n = x.shape[0]
    fs = _rel_mm(x[src], conv_w, etype)
    fd = _rel_mm(x[dst], conv_w, etype)
    el = (fs * attn_l[etype]).sum(-1)
    er = (fd * attn_r[etype]).sum(-1)
    pre = el + er
    pos = pre > 0
    if branch is not None:
        pos = torch.where(pre.detach().abs() < kink_eps, branch, pos)
    a = _segment_softmax_sum(torch.exp(torch.where(pos, pre, slope * pre)), dst, n)
    out = torch.zeros(n, fs.shape[1], fs.shape[2], dtype=x.dtype).index_add_(0, dst, a.unsqueeze(-1) * fs)
    out = out.reshape(n, -1)
    return out + bias if bias is not None else out


def _typed_linear(x, w, offs):
    """w [T,1,Din,Dout], rows of type t are offs[t]:offs[t+1]."""
    outs = [x[int(offs[t]):int(offs[t + 1])] @ w[t, 0] for t in range(w.shape[0])]
    return torch.cat(outs, 0)


def hgt_layer(src, dst, etype, offs, h, k_w, q_w, v_w, a_w, rel_pri, rel_att, rel_msg, skip):
    n, H, dk = h.shape[0], rel_att.shape[1], rel_att.shape[2]
    k = _typed_linear(h, k_w, offs).view(n, H, dk)
    q = _typed_linear(h, q_w, offs).view(n, H, dk)
    v = _typed_linear(h, v_w, offs).view(n, H, dk)
    wq = _rel_mm(q[dst], rel_att, etype)
    score = (wq * k[src]).sum(-1)  # [E,H]
    msg = _rel_mm(v[src], rel_msg, etype)
    mu = rel_pri / (dk ** 0.5)
    a = _segment_softmax_sum(torch.exp(score * mu[etype]), dst, n)
    new_h = torch.zeros(n, H, dk, dtype=h.dtype).index_add_(0, dst, a.unsqueeze(-1) * msg)
    return _typed_linear(new_h.reshape(n, H * dk), torch.sigmoid(skip) * a_w, offs)


# ------------------------- DGL baselines, restated ------------------------
def dgl_rgat_layer(src, dst, etype, x, fc_w, attn_l, attn_r, slope, num_rels, ranges=None):
    """HeteroGraphConv({rel: GATConv}, aggregate='sum'): softmax per (dst, relation).
    fc_w [R,Din,H*F]; attn_* [R,H,F]. ``ranges[r] = (s_lo, s_hi, d_lo, d_hi)``: node-id
    ranges of relation r's source / destination node types — GATConv's fc runs on the nodes
    of those types only (DGL passes per-type feature dicts); None = all nodes."""
    n, H, Fh = x.shape[0], attn_l.shape[1], attn_l.shape[2]
    out = torch.zeros(n, H, Fh, dtype=x.dtype)
    for r in range(num_rels):
        idx = torch.nonzero(etype == r).squeeze(1)
        if idx.numel() == 0:
            continue
        s, d = src[idx], dst[idx]
        s_lo, s_hi, d_lo, d_hi = ranges[r] if ranges is not None else (0, n, 0, n)
        fs = (x[s_lo:s_hi] @ fc_w[r]).view(-1, H, Fh)
        fd = fs if (s_lo, s_hi) == (d_lo, d_hi) else (x[d_lo:d_hi] @ fc_w[r]).view(-1, H, Fh)
        e = torch.nn.functional.leaky_relu(
            (fs * attn_l[r]).sum(-1)[s - s_lo] + (fd * attn_r[r]).sum(-1)[d - d_lo], slope)
        e = torch.exp(e - e.max())
        a = _segment_softmax_sum(e, d, n)
        out = out.index_add(0, d, a.unsqueeze(-1) * fs[s - s_lo])
    return out.reshape(n, -1)


def dgl_hgt_layer(src, dst, etype, offs, ntype_of_rel_dst, h, k_w, q_w, v_w, a_w, rel_pri,
                  rel_att, rel_msg, skip, num_rels, ranges=None):
    """HGT/models_dgl.py: per-relation edge_softmax, cross_reducer='mean'.
    ``ranges[r] = (s_lo, s_hi, d_lo, d_hi)`` as in dgl_rgat_layer (the relation's K/V
    projections run on its source node type only)."""
    n, H, dk = h.shape[0], rel_att.shape[1], rel_att.shape[2]
    k = _typed_linear(h, k_w, offs).view(n, H, dk)
    q = _typed_linear(h, q_w, offs).view(n, H, dk)
    v = _typed_linear(h, v_w, offs).view(n, H, dk)
    acc = torch.zeros(n, H, dk, dtype=h.dtype)
    cnt = torch.zeros(n, 1, 1, dtype=h.dtype)
    for r in range(num_rels):
        m = etype == r
        if not bool(m.any()):
            continue
        s, d = src[m], dst[m]
        s_lo, s_hi = (ranges[r][0], ranges[r][1]) if ranges is not None else (0, n)
        kk = torch.einsum("nhk,hkf->nhf", k[s_lo:s_hi], rel_att[r])
        vv = torch.einsum("nhk,hkf->nhf", v[s_lo:s_hi], rel_msg[r])
        t = (q[d] * kk[s - s_lo]).sum(-1) * rel_pri[r] / (dk ** 0.5)
        t = torch.exp(t - t.max())
        a = _segment_softmax_sum(t, d, n)
        acc = acc.index_add(0, d, a.unsqueeze(-1) * vv[s - s_lo])
        t_d = int(ntype_of_rel_dst[r])
        cnt[int(offs[t_d]):int(offs[t_d + 1])] += 1
    acc = acc / cnt.clamp(min=1)
    return _typed_linear(acc.reshape(n, H * dk), torch.sigmoid(skip) * a_w, offs)


def dgl_rgcn_layer(src, dst, etype, x, w, norm):
    """RelGraphConv: sum_r sum_{e in r} norm_e * x[src] @ W_r  (RGCN_DGL.py:9-19)."""
    n = x.shape[0]
    out = torch.zeros(n, w.shape[2], dtype=x.dtype)
    for r in range(w.shape[0]):
        m = etype == r
        if bool(m.any()):
            out.index_add_(0, dst[m], (x[src[m]] @ w[r]) * norm[m].unsqueeze(-1))
    return out
