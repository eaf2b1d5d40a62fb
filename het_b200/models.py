"""RGAT / HGT layers with the reference's op sequence (default, non-compact flags).

Host-side mirror of hrt/python/RGAT/models.py:16-385 (``HET_RGATLayer``) and
hrt/python/HGT/models.py:15-286 (``HET_HGTLayerHetero``): same parameters, same
shapes, same ``B.*`` call order, so the parity tests can
compare layer outputs and every parameter gradient with the oracle's restatement.
"""
from __future__ import annotations

import math

import torch as th
import torch.nn as nn

from . import backend as B


class HET_RelGraphEmbed(nn.Module):
    """Learnable [N, D] node features (hrt/python/RGNNUtils/RGNNUtils.py:78-121)."""

    def __init__(self, num_nodes, embed_size):
        super().__init__()
        self.embeds = nn.Parameter(th.empty(num_nodes, embed_size))
        nn.init.uniform_(self.embeds, -1.0, 1.0)

    def forward(self, _=None):
        return self.embeds


class HET_RGATLayer(nn.Module):
    def __init__(self, in_feat, out_feat, num_rels, num_heads, *, bias=True, activation=None,
                 self_loop=False, gat_edge_parallel_flag=False, dropout=0.0, leaky_relu_slope=0.2,
                 compact_as_of_node_flag=False, linear_reorder_flag=False):
        super().__init__()
        self.compact_as_of_node_flag = compact_as_of_node_flag
        # Hector's linear operator reordering on the destination side (compact path only):
        # er = (X.W_r).a_r = X.(W_r.a_r), so the (relation, destination) rows need one
        # row-dot with a per-relation vector instead of a 64-wide typed linear layer.
        self.linear_reorder_flag = linear_reorder_flag
        assert out_feat % num_heads == 0, "out_feat must be a multiple of num_heads"
        self.in_feat, self.out_feat = in_feat, out_feat
        self.num_rels, self.num_heads = num_rels, num_heads
        self.bias, self.activation, self.self_loop = bias, activation, self_loop
        self.gat_edge_parallel_flag = gat_edge_parallel_flag
        self.leaky_relu_slope = leaky_relu_slope
        self.conv_weights = nn.Parameter(th.empty(num_rels, num_heads, in_feat, out_feat // num_heads))
        self.attn_l = nn.Parameter(th.empty(num_rels, num_heads, out_feat // num_heads))
        self.attn_r = nn.Parameter(th.empty(num_rels, num_heads, out_feat // num_heads))
        if bias:
            self.h_bias = nn.Parameter(th.empty(out_feat))
        if self_loop:
            self.loop_weight = nn.Parameter(th.empty(in_feat, out_feat))
        self.dropout = nn.Dropout(dropout)
        self.reset_parameters()

    def reset_parameters(self):
        gain = nn.init.calculate_gain("relu")
        if self.bias:
            nn.init.zeros_(self.h_bias)
        if self.self_loop:
            nn.init.xavier_uniform_(self.loop_weight, gain=gain)
        nn.init.xavier_uniform_(self.conv_weights, gain=gain)
        nn.init.xavier_uniform_(self.attn_l, gain=gain)
        nn.init.xavier_uniform_(self.attn_r, gain=gain)

    def forward(self, g, inputs, inputs_dst=None):
        """``inputs`` rows are indexed by source ids; ``inputs_dst`` (default: the same
        tensor) by destination ids — they differ only on a destination-range shard."""
        if inputs_dst is None:
            inputs_dst = inputs
        if self.compact_as_of_node_flag:
            h = self._forward_compact(g, inputs, inputs_dst)
            return self._epilogue(h, inputs_dst)
        coo = g.get_separate_coo_original()

        def mm(gather, w, x, head_one):
            return B.rgnn_relational_matmul(
                {"separate_coo_rel_ptrs": coo["rel_ptrs"], "separate_coo_node_indices": gather,
                 "separate_coo_eids": coo["eids"]}, w, x, head_one, 0)

        feat_src_per_edge = mm(coo["row_indices"], self.conv_weights, inputs, True)
        el = mm(coo["eids"], self.attn_l.unsqueeze(-1), feat_src_per_edge, False)
        feat_dst_per_edge = mm(coo["col_indices"], self.conv_weights, inputs_dst, True)
        er = mm(coo["eids"], self.attn_r.unsqueeze(-1), feat_dst_per_edge, False)
        if self.gat_edge_parallel_flag:
            h = B.relational_fused_gat_separate_coo(g, feat_src_per_edge, el, er, self.leaky_relu_slope)
        else:
            h = B.relational_fused_gat_csr(g, feat_src_per_edge, el, er, self.leaky_relu_slope)
        return self._epilogue(h, inputs_dst)

    def _forward_compact(self, g, inputs, inputs_dst):
        """compact_as_of_node_flag (hrt/python/RGAT/models.py:151-263): X·W_r is computed once per
        unique (relation, source) and once per unique (relation, destination) pair instead of once
        per edge; the logits are one row-dot per compact row; the edge phase reads the compact rows
        through the direct-indexing maps. The reference runs its edge-parallel COO kernel here
        (the in-CSR compact kernel "not implemented the singe side unique node idex", :254-257);
        this uses the in-CSR walker with the same maps."""
        u = g.get_separate_unique_node_indices_single_sided()
        row = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"],
               "unique_srcs_and_dests_node_indices": u["node_indices_row"]}
        col = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_col"],
               "unique_srcs_and_dests_node_indices": u["node_indices_col"]}
        er_compact = None
        if self.linear_reorder_flag:
            w_r = th.einsum("rhkf,rhf->rhk", self.conv_weights, self.attn_r).unsqueeze(-1)  # [R, H, in, 1]
            if inputs_dst is inputs:
                # both compact products read the same X: one autograd node, one grad_X buffer
                feat_compact, er_compact = B.relational_matmul_compact_pair(g, inputs, self.conv_weights, w_r)
            else:
                feat_compact = B.rgnn_relational_matmul(row, self.conv_weights, inputs, True, 1)
                er_compact = B.rgnn_relational_matmul(col, w_r, inputs_dst, True, 1)
        else:
            feat_compact = B.rgnn_relational_matmul(row, self.conv_weights, inputs, True, 1)
            feat_compact_dst = B.rgnn_relational_matmul(col, self.conv_weights, inputs_dst, True, 1)
            er_compact = B.rgnn_relational_matmul_no_scatter_gather_list(
                u["rel_ptrs_col"], self.attn_r.unsqueeze(-1),
                feat_compact_dst.view(feat_compact_dst.size(0), -1))
        if self.linear_reorder_flag:
            # el row-dot + edge pass as one autograd node (one grad_feat buffer)
            fn = B.relational_fused_gat_plan_with_logits if B.USE_EDGE_PLAN else B.relational_fused_gat_csr_compact_with_logits
            return fn(g, feat_compact, er_compact.view(-1, self.num_heads, 1), self.attn_l, self.leaky_relu_slope)
        el_compact = B.rgnn_relational_matmul_no_scatter_gather_list(
            u["rel_ptrs_row"], self.attn_l.unsqueeze(-1), feat_compact.view(feat_compact.size(0), -1))
        fn = B.relational_fused_gat_plan if B.USE_EDGE_PLAN else B.relational_fused_gat_csr_compact
        return fn(g, feat_compact, el_compact.view(-1, self.num_heads, 1), er_compact.view(-1, self.num_heads, 1),
                  self.leaky_relu_slope)

    @th.no_grad()
    def logit_signs(self, g, inputs, inputs_dst=None):
        """Diagnostic: sign of the attention logit el + er of every edge ([E, H] bool, indexed by
        canonical edge id), recomputed with the same typed-linear kernels the forward uses. Tests
        hand it to the fp64 oracle so that both take the same side of the leaky-ReLU kink for the
        few edges whose logit is within fp32 rounding of zero."""
        if inputs_dst is None:
            inputs_dst = inputs
        u = g.get_separate_unique_node_indices_single_sided()
        row = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"], "unique_srcs_and_dests_node_indices": u["node_indices_row"]}
        col = {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_col"], "unique_srcs_and_dests_node_indices": u["node_indices_col"]}
        feat_c = B.rgnn_relational_matmul(row, self.conv_weights, inputs, True, 1)
        el = B.rgnn_relational_matmul_no_scatter_gather_list(u["rel_ptrs_row"], self.attn_l.unsqueeze(-1),
                                                             feat_c.view(feat_c.size(0), -1))
        if self.compact_as_of_node_flag and self.linear_reorder_flag:
            w_r = th.einsum("rhkf,rhf->rhk", self.conv_weights, self.attn_r).unsqueeze(-1)
            er = B.rgnn_relational_matmul(col, w_r, inputs_dst, True, 1).view(-1, self.num_heads)
        else:
            feat_d = B.rgnn_relational_matmul(col, self.conv_weights, inputs_dst, True, 1)
            er = B.rgnn_relational_matmul_no_scatter_gather_list(u["rel_ptrs_col"], self.attn_r.unsqueeze(-1),
                                                                 feat_d.view(feat_d.size(0), -1))
        return (el.view(-1, self.num_heads)[u["inverse_idx_row"]] + er.view(-1, self.num_heads)[u["inverse_idx_col"]]) > 0

    def _epilogue(self, h, inputs_dst):
        h = h.view(-1, self.out_feat)
        if self.self_loop:
            h = h + th.matmul(inputs_dst, self.loop_weight)
        if self.bias:
            h = B.add_bias(h, self.h_bias)
        if self.activation:
            h = self.activation(h)
        return self.dropout(h)


class HET_RGATModel(nn.Module):
    """Stack of RGAT layers (hrt/python/RGAT/models.py:388-512)."""

    def __init__(self, num_rels, in_dim, h_dim, out_dim, num_heads=1, num_layers=1, **kw):
        super().__init__()
        dims = [in_dim] + [h_dim] * (num_layers - 1) + [out_dim]
        self.layers = nn.ModuleList(
            HET_RGATLayer(dims[i], dims[i + 1], num_rels, num_heads,
                          activation=th.relu if i + 1 < num_layers else None, **kw)
            for i in range(num_layers))

    def forward(self, g, h, gather=None):
        """``gather`` (multi-GPU): maps owned rows to the all-gathered source matrix."""
        for layer in self.layers:
            h = layer(g, h) if gather is None else layer(g, gather(h), h)
        return h


class HET_EglRelGraphConv_EdgeParallel(nn.Module):
    """RGCN layer (hrt/python/RGCN/RGCN.py:194-350): out[dst] = sum_e norm[e] * x[src] @ W[rel(e)]
    (+ bias, activation, dropout). ``compact_as_of_node_flag``: x @ W_r once per unique (relation,
    source) pair, then the aggregation reads the compact rows by direct indexing;
    ``csr_aggregation_flag`` (hetb200): that aggregation as the in-CSR walker instead of the
    reference's edge-parallel COO kernel. Basis regularizer as in the reference."""

    def __init__(self, in_feat, out_feat, num_rels, num_edges=None, regularizer="basis", num_bases=None,
                 bias=True, activation=None, self_loop=False, compact_as_of_node_flag=False,
                 compact_direct_indexing_flag=False, dropout=0.0, layer_type=1, csr_aggregation_flag=False):
        super().__init__()
        if self_loop:
            raise NotImplementedError
        if regularizer != "basis":
            raise ValueError("Regularizer must be either 'basis' or 'bdd'")
        self.in_feat, self.out_feat, self.num_rels = in_feat, out_feat, num_rels
        self.compact_as_of_node_flag = compact_as_of_node_flag
        self.compact_direct_indexing_flag = compact_direct_indexing_flag
        self.csr_aggregation_flag = csr_aggregation_flag
        self.num_bases = num_bases if (num_bases is not None and 0 < num_bases <= num_rels) else num_rels
        self.bias, self.activation, self.layer_type = bias, activation, layer_type
        self.weight = nn.Parameter(th.empty(self.num_bases, in_feat, out_feat))
        nn.init.xavier_uniform_(self.weight, gain=nn.init.calculate_gain("relu"))
        if self.num_bases < num_rels:
            self.w_comp = nn.Parameter(th.empty(num_rels, self.num_bases))
            nn.init.xavier_uniform_(self.w_comp, gain=nn.init.calculate_gain("relu"))
        if bias:
            self.h_bias = nn.Parameter(th.zeros(out_feat))
        self.dropout = nn.Dropout(dropout)

    def forward(self, g, x, norm=None):
        if self.layer_type == 0:
            raise NotImplementedError
        if self.num_bases < self.num_rels:
            weight = th.matmul(self.w_comp, self.weight.view(self.num_bases, -1)).view(
                self.num_rels, self.in_feat, self.out_feat)
        else:
            weight = self.weight
        if norm is None:
            norm = th.ones(g.get_num_edges(), 1, dtype=x.dtype, device=x.device)
        if self.compact_as_of_node_flag:
            u = g.get_separate_unique_node_indices_single_sided()
            feat_compact_src = B.rgnn_relational_matmul(
                {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"],
                 "unique_srcs_and_dests_node_indices": u["node_indices_row"]}, weight.unsqueeze(1), x, True, 1)
            feat_compact_src = feat_compact_src.squeeze(1)
            if self.csr_aggregation_flag:
                node_repr = B.rgcn_node_aggregation_csr_compact(g, feat_compact_src, norm)
            else:
                node_repr = B.rgcn_node_mean_aggregation_compact_as_of_node_separate_coo_single_sided(
                    g, feat_compact_src, norm, self.compact_direct_indexing_flag)
        else:
            node_repr = B.rgcn_layer1_separate_coo(g, x, weight, norm)
        if self.bias:
            node_repr = B.add_bias(node_repr, self.h_bias)
        if self.activation:
            node_repr = self.activation(node_repr)
        return self.dropout(node_repr)


class HET_EGLRGCNSingleLayerModel(nn.Module):
    """hrt/python/RGCN/RGCN.py:353-405 (separate_coo sparse format only)."""

    def __init__(self, n_infeat, out_dim, num_rels, num_nodes=None, num_edges=None, sparse_format="separate_coo",
                 num_bases=None, dropout=0.0, activation=None, compact_as_of_node_flag=False,
                 compact_direct_indexing_flag=False, **kw):
        super().__init__()
        if sparse_format != "separate_coo":
            raise NotImplementedError("seastar CSR/COO RGCN kernels are out of scope (DESIGN.md section 7)")
        self.num_nodes = num_nodes
        self.layer2 = HET_EglRelGraphConv_EdgeParallel(
            n_infeat, out_dim, num_rels, num_edges, num_bases=num_bases, dropout=dropout, activation=activation,
            layer_type=1, compact_as_of_node_flag=compact_as_of_node_flag,
            compact_direct_indexing_flag=compact_direct_indexing_flag, **kw)

    def forward(self, mydglgraph, feats, edge_norm):
        return self.layer2(mydglgraph, feats, edge_norm)


class HET_RGCNModel(nn.Module):
    """Stack of RGCN layers, ReLU between them (the BASELINE "RGCN 2-layer" config; the reference
    wires its HET RGCN as a single layer, RGCN.py:353-405, and its DGL baseline as two)."""

    def __init__(self, num_rels, in_dim, h_dim, out_dim, num_layers=2, **kw):
        super().__init__()
        dims = [in_dim] + [h_dim] * (num_layers - 1) + [out_dim]
        self.layers = nn.ModuleList(
            HET_EglRelGraphConv_EdgeParallel(dims[i], dims[i + 1], num_rels,
                                             activation=th.relu if i + 1 < num_layers else None, **kw)
            for i in range(num_layers))

    def forward(self, g, h, norm, gather=None):
        for layer in self.layers:
            h = layer(g, h if gather is None else gather(h), norm)
        return h


class HET_HGTLayerHetero(nn.Module):
    def __init__(self, num_ntypes, num_rels, in_dim, out_dim, num_heads=1, dropout=0.0,
                 compact_as_of_node_flag=False, compact_direct_indexing_flag=True,
                 fused_attention_aggregation_flag=False, linear_reorder_flag=False, bf16_tables_flag=False,
                 hgt_fused_attn_score_flag=False, fused_message_mean_aggregation_flag=False,
                 multiply_among_weights_first_flag=False):
        super().__init__()
        # the reference's separate-COO op slicings (hrt/python/HGT/models.py:170-174, 251-259): the
        # attention score as one op (k[src].W_att[r] and the inner product with q[dst]; note this op
        # contracts W_att as k.W.q^T where the two-op path computes q.W.k^T), and edge softmax + typed
        # linear + aggregation over the COO. Per-edge (non-compact) paths.
        self.hgt_fused_attn_score_flag = hgt_fused_attn_score_flag
        self.fused_message_mean_aggregation_flag = fused_message_mean_aggregation_flag
        # multiply_among_weights_first_flag (hrt/python/HGT/models.py:124-155): compose the node-type and
        # relation weights once per step so that Q and V are never materialised. The reference composes
        # them with view() calls that "do not transpose" (its own FIXME, :137) and takes v_linears of the
        # destination type; hetb200 implements the optimisation's maths as linear_reorder_flag (compact
        # path), which this flag turns on.
        if multiply_among_weights_first_flag:
            linear_reorder_flag = True
            compact_as_of_node_flag = True
        # bf16 STORAGE of the gathered tables (k, W_att.q, W_msg.v): needs the compact + fused + reordered
        # path on one GPU; arithmetic, outputs and gradients stay fp32 (backend.HGTLayerCoreBF16Tables)
        self.bf16_tables_flag = bf16_tables_flag
        # Hector's linear operator reordering (compact path): (h.W_q[type]).W_att[r] = h.(W_q[type].W_att[r])
        # and the same for W_v / W_msg, so Q and V are never materialised; the composed [R, in, out]
        # weights cost R small products per step. Needs one (src type, dst type) pair per relation.
        self.linear_reorder_flag = linear_reorder_flag
        # compact_as_of_node_flag (hrt/python/HGT/models.py:177-208): W_att·Q is computed once per
        # unique (relation, destination) pair; here the same is also done for W_msg·V per unique
        # (relation, source) pair (a hetb200 extension, same maths). Only direct indexing is built.
        self.compact_as_of_node_flag = compact_as_of_node_flag
        self.compact_direct_indexing_flag = compact_direct_indexing_flag
        # hetb200 fusion (needs the compact flag): score, softmax and aggregation in one in-CSR pass
        self.fused_attention_aggregation_flag = fused_attention_aggregation_flag
        self.in_dim, self.out_dim = in_dim, out_dim
        self.num_ntypes, self.num_relations = num_ntypes, num_rels
        self.num_heads = num_heads
        self.d_k = out_dim // num_heads
        self.sqrt_dk = math.sqrt(self.d_k)
        self.k_linears = nn.Parameter(th.empty(num_ntypes, 1, in_dim, out_dim))
        self.q_linears = nn.Parameter(th.empty(num_ntypes, 1, in_dim, out_dim))
        self.v_linears = nn.Parameter(th.empty(num_ntypes, 1, in_dim, out_dim))
        self.a_linears = nn.Parameter(th.empty(num_ntypes, 1, out_dim, out_dim))
        self.relation_pri = nn.Parameter(th.ones(num_rels, num_heads))
        self.relation_att = nn.Parameter(th.empty(num_rels, num_heads, self.d_k, self.d_k))
        self.relation_msg = nn.Parameter(th.empty(num_rels, num_heads, self.d_k, self.d_k))
        self.skip = nn.Parameter(th.ones(num_ntypes, 1, 1, 1))
        self.drop = nn.Dropout(dropout)
        self.reset_parameters()

    def reset_parameters(self):
        for p in (self.relation_att, self.relation_msg, self.k_linears, self.q_linears,
                  self.v_linears, self.a_linears):
            nn.init.xavier_uniform_(p)

    def forward(self, G, h, gather=None):
        """``gather`` (multi-GPU): all-gathers K and V rows of the owned destinations into
        the source-indexed matrices; Q stays local (it is gathered by destination)."""
        offs = G.get_original_node_type_offsets()
        reorder = (self.linear_reorder_flag and self.compact_as_of_node_flag
                   and G.get_relation_node_types() is not None)
        combined = reorder  # k comes out of hgt_reordered_projections[_sharded] below
        k = None if combined else B.rgnn_relational_matmul_no_scatter_gather_list(
            offs, self.k_linears, h).view(-1, self.num_heads, self.d_k)
        if reorder:
            rel_src_type, rel_dst_type = G.get_relation_node_types()
            H, dk = self.num_heads, self.d_k

            def compose(lin, rel_w, types):  # [R, 1, in, out]: lin[type(r)] . blockdiag_h(rel_w[r, h])
                if lin.is_cuda:
                    return B.compose_weights(lin, types, rel_w)
                w = th.einsum("rihd,rhde->rihe", lin[types, 0].view(-1, self.in_dim, H, dk), rel_w)
                return w.reshape(-1, 1, self.in_dim, self.out_dim)

            q = h  # stands for Q below: the composed weight carries W_q
            v = h
            w_att = compose(self.q_linears, self.relation_att, rel_dst_type)
            w_msg = compose(self.v_linears, self.relation_msg, rel_src_type)
        else:
            q = B.rgnn_relational_matmul_no_scatter_gather_list(offs, self.q_linears, h).view(-1, self.num_heads, self.d_k)
            v = B.rgnn_relational_matmul_no_scatter_gather_list(offs, self.v_linears, h).view(-1, self.num_heads, self.d_k)
            w_att, w_msg = self.relation_att, self.relation_msg
        if gather is not None and not reorder:
            k, v = gather(k), gather(v)
        coo = G.get_separate_coo_original()

        def mm(idx, w, x):
            return B.rgnn_relational_matmul(
                {"separate_coo_rel_ptrs": coo["rel_ptrs"], "separate_coo_node_indices": idx,
                 "separate_coo_eids": coo["eids"]}, w, x, False, 0)

        if self.compact_as_of_node_flag:
            u = G.get_separate_unique_node_indices_single_sided()
            if self.bf16_tables_flag:
                if not (reorder and gather is None and self.fused_attention_aggregation_flag and B.USE_EDGE_PLAN):
                    raise NotImplementedError("bf16_tables_flag needs compact + reorder + fused flags on one GPU")
                new_h = B.hgt_layer_core_bf16_tables(G, h, self.k_linears, w_att, w_msg,
                                                     self.relation_pri / self.sqrt_dk)
                return B.rgnn_relational_matmul_no_scatter_gather_list(
                    offs, th.sigmoid(self.skip) * self.a_linears, new_h.view(-1, self.out_dim))
            if reorder and gather is None:
                # k and the two compact products read the same h: one autograd node, one grad_h buffer
                k, attn_weight_dst_product_compact, message_compact = B.hgt_reordered_projections(
                    G, h, self.k_linears, w_att, w_msg)
                k = k.view(-1, self.num_heads, self.d_k)
            elif reorder:
                # destination shard: ONE exchange (the layer input); k is computed for the gathered rows here
                k, attn_weight_dst_product_compact, message_compact = B.hgt_reordered_projections_sharded(
                    G, h, gather, self.k_linears, w_att, w_msg)
                k = k.view(-1, self.num_heads, self.d_k)
            else:
                attn_weight_dst_product_compact = B.rgnn_relational_matmul(
                    {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_col"],
                     "unique_srcs_and_dests_node_indices": u["node_indices_col"]}, w_att, q, reorder, 1)
                message_compact = B.rgnn_relational_matmul(
                    {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"],
                     "unique_srcs_and_dests_node_indices": u["node_indices_row"]}, w_msg, v, reorder, 1)
            if reorder:
                attn_weight_dst_product_compact = attn_weight_dst_product_compact.view(-1, self.num_heads, self.d_k)
                message_compact = message_compact.view(-1, self.num_heads, self.d_k)
            if self.fused_attention_aggregation_flag:
                fn = (B.hgt_fused_attention_and_aggregation_plan if B.USE_EDGE_PLAN
                      else B.hgt_fused_attention_and_aggregation_csr_compact)
                new_h = fn(G, attn_weight_dst_product_compact, k, message_compact, self.relation_pri / self.sqrt_dk)
            else:
                # CompactAsOfNodeKind 2 (direct indexing) or 1 (binary search in the unique lists)
                attn_score = B.rgnn_inner_product_right_node(G, attn_weight_dst_product_compact, k,
                                                             2 if self.compact_direct_indexing_flag else 1, "_col")
                new_h = B.hgt_full_graph_edge_softmax_and_message_mean_aggregation_csr_compact(
                    G, message_compact, attn_score, self.relation_pri / self.sqrt_dk)
        else:
            if self.hgt_fused_attn_score_flag:
                attn_score = B.hgt_full_graph_hetero_attention_ops_coo(G, self.relation_att, k, q)
            else:
                attn_weight_dst_product_per_edge = mm(coo["col_indices"], self.relation_att, q)
                attn_score = B.rgnn_inner_product_right_node(G, attn_weight_dst_product_per_edge, k, 0, "_col")
            if self.fused_message_mean_aggregation_flag:
                new_h = B.hgt_full_graph_message_calc_edge_softmax_and_message_mean_aggregation_coo(
                    self.relation_msg, v, G, self.relation_pri / self.sqrt_dk, attn_score)
            else:
                message_per_edge = mm(coo["row_indices"], self.relation_msg, v)
                new_h = B.hgt_full_graph_edge_softmax_and_message_mean_aggregation_csr(
                    G, message_per_edge, attn_score, self.relation_pri / self.sqrt_dk)
        new_h = B.rgnn_relational_matmul_no_scatter_gather_list(
            offs, th.sigmoid(self.skip) * self.a_linears, new_h.view(-1, self.out_dim))
        return new_h


class HET_HGT_DGLHetero(nn.Module):
    """Stack of HGT layers. The reference model wires a single layer
    (hrt/python/HGT/models.py:312-330, second layer commented out); the
    BASELINE "HGT 2-layer" config stacks it with ``num_layers=2``."""

    def __init__(self, num_ntypes, num_rels, in_dim, out_dim, num_heads=1, num_layers=1, dropout=0.0, **kw):
        super().__init__()
        dims = [in_dim] + [out_dim] * num_layers
        self.layers = nn.ModuleList(
            HET_HGTLayerHetero(num_ntypes, num_rels, dims[i], dims[i + 1], num_heads, dropout, **kw)
            for i in range(num_layers))

    def forward(self, G, h, gather=None):
        for layer in self.layers:
            h = layer(G, h, gather)
        return h
