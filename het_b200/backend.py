"""autograd wrappers around ``torch.ops.torch_hrt`` with the reference's names.

Host-side mirror of hrt/python/backend/{rgnn,rgat,hgt}_layers_and_funcs.py: each
function allocates the outputs/intermediates the op contract expects, calls one
or two ``K.<op>`` in forward and the matching ``K.backward_<op>`` in backward.
Same function names, argument meaning and op call order as the reference, so the
models in ``het_b200.models`` read like the reference's.
"""
from __future__ import annotations

import torch as th

from .kernels import K


# ----- rgnn_layers_and_funcs.py:8-73 -------------------------------------
class RgnnRelationalMatmul(th.autograd.Function):
    @staticmethod
    def forward(ctx, relptrs, node_indices, eids, weights, inputs, ret, input_num_head_one_flag):
        ctx.save_for_backward(relptrs, node_indices, eids, weights, inputs)
        ctx.input_num_head_one_flag = input_num_head_one_flag
        ctx.mark_dirty(ret)
        K.rgnn_relational_matmul(
            {"separate_coo_rel_ptrs": relptrs, "separate_coo_node_indices": node_indices,
             "separate_coo_eids": eids},
            0, weights, inputs, ret, input_num_head_one_flag)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        relptrs, node_indices, eids, weights, inputs = ctx.saved_tensors
        grad_weight = th.zeros_like(weights, memory_format=th.contiguous_format)
        grad_input = th.zeros_like(inputs, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"separate_coo_rel_ptrs": relptrs, "separate_coo_node_indices": node_indices,
             "separate_coo_eids": eids},
            0, th.transpose(weights, 2, 3).contiguous(), inputs, gradout.contiguous(),
            grad_input, grad_weight, ctx.input_num_head_one_flag)
        return None, None, None, grad_weight, grad_input, None, None


# ----- rgnn_layers_and_funcs.py:76-120 -----------------------------------
class RgnnRelationalMatmulNoScatterGatherList(th.autograd.Function):
    @staticmethod
    def forward(ctx, ntype_offset_ptrs, weights, inputs, ret):
        ctx.mark_dirty(ret)
        K.rgnn_relational_matmul_no_scatter_gather_list(ntype_offset_ptrs, weights, inputs, ret)
        ctx.save_for_backward(ntype_offset_ptrs, weights, inputs)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        ntype_offset_ptrs, weights, inputs = ctx.saved_tensors
        grad_weight = th.zeros_like(weights, memory_format=th.contiguous_format)
        grad_input = th.zeros_like(inputs, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul_no_scatter_gather_list(
            ntype_offset_ptrs, th.transpose(weights, 2, 3).contiguous(), inputs,
            gradout.contiguous(), grad_input, grad_weight)
        return None, grad_weight, grad_input, None


# ----- rgnn_layers_and_funcs.py:350-416 ----------------------------------
class RgnnInnerProductEdgeAndNode(th.autograd.Function):
    @staticmethod
    def forward(ctx, rel_ptrs, eids, row_indices, col_indices, left_edge_data, right_node_vectors, ret):
        ctx.save_for_backward(rel_ptrs, eids, row_indices, col_indices, left_edge_data, right_node_vectors)
        ctx.mark_dirty(ret)
        K.rgnn_inner_product_right_node_separatecoo(
            dict(), 0, rel_ptrs, eids, row_indices, col_indices, left_edge_data,
            right_node_vectors, ret)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        rel_ptrs, eids, row_indices, col_indices, left, right = ctx.saved_tensors
        grad_left = th.zeros_like(left, memory_format=th.contiguous_format)
        grad_right = th.zeros_like(right, memory_format=th.contiguous_format)
        K.backward_inner_product_right_node_separatecoo(
            dict(), 0, rel_ptrs, eids, row_indices, col_indices, left, right,
            gradout.contiguous(), grad_left, grad_right)
        return None, None, None, None, grad_left, grad_right, None


# ----- rgnn_layers_and_funcs.py:122-200 (compact-as-of-node: one output row per unique (rel, node)) -----
class RgnnRelationalMatmulCompactAsOfNode(th.autograd.Function):
    @staticmethod
    def forward(ctx, unique_rel_ptrs, unique_node_indices, weights, inputs, ret, input_num_head_one_flag):
        ctx.save_for_backward(unique_rel_ptrs, unique_node_indices, weights, inputs)
        ctx.input_num_head_one_flag = input_num_head_one_flag
        ctx.mark_dirty(ret)
        K.rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": unique_rel_ptrs,
             "unique_srcs_and_dests_node_indices": unique_node_indices},
            1, weights, inputs, ret, input_num_head_one_flag)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        unique_rel_ptrs, unique_node_indices, weights, inputs = ctx.saved_tensors
        grad_weight = th.zeros_like(weights, memory_format=th.contiguous_format)
        grad_input = th.zeros_like(inputs, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": unique_rel_ptrs,
             "unique_srcs_and_dests_node_indices": unique_node_indices},
            1, th.transpose(weights, 2, 3).contiguous(), inputs, gradout.contiguous(),
            grad_input, grad_weight, ctx.input_num_head_one_flag)
        return None, None, grad_weight, grad_input, None, None


def rgnn_relational_matmul(arg_tensor_dict, weights, inputs, input_num_head_one_flag, compact_as_of_node_kind):
    """rgnn_layers_and_funcs.py:420-470. Kind 0: one output row per edge; kind 1
    (CompactAsOfNodeKind::Enabled): one output row per unique (relation, node) pair."""
    if compact_as_of_node_kind == 1:
        ret = th.empty(
            (arg_tensor_dict["unique_srcs_and_dests_node_indices"].numel(), weights.size(1), weights.size(3)),
            dtype=weights.dtype, device=weights.device)
        return RgnnRelationalMatmulCompactAsOfNode.apply(
            arg_tensor_dict["unique_srcs_and_dests_rel_ptrs"],
            arg_tensor_dict["unique_srcs_and_dests_node_indices"], weights.contiguous(), inputs.contiguous(),
            ret, input_num_head_one_flag)
    if compact_as_of_node_kind != 0:
        raise NotImplementedError("CompactAsOfNodeKind > 1 for the relational matmul")
    ret = th.empty(
        (arg_tensor_dict["separate_coo_node_indices"].numel(), weights.size(1), weights.size(3)),
        dtype=weights.dtype, device=weights.device)
    return RgnnRelationalMatmul.apply(
        arg_tensor_dict["separate_coo_rel_ptrs"], arg_tensor_dict["separate_coo_node_indices"],
        arg_tensor_dict["separate_coo_eids"], weights.contiguous(), inputs.contiguous(), ret,
        input_num_head_one_flag)


def rgnn_relational_matmul_no_scatter_gather_list(ntype_offset_ptrs, weights, inputs):
    """rgnn_layers_and_funcs.py:473-492"""
    ret = th.empty((inputs.size(0), weights.size(1) * weights.size(3)), dtype=weights.dtype,
                   device=weights.device)
    return RgnnRelationalMatmulNoScatterGatherList.apply(
        ntype_offset_ptrs, weights.contiguous(), inputs.contiguous(), ret)


# ----- rgnn_layers_and_funcs.py:202-348 (left data compact, addressed by direct indexing) -----
class RgnnInnerProductNodeCompactAndNode(th.autograd.Function):
    @staticmethod
    def forward(ctx, edata_idx_to_inverse_idx, rel_ptrs, eids, row_indices, col_indices, left_node_compact_data,
                right_node_vectors, ret):
        ctx.save_for_backward(edata_idx_to_inverse_idx, rel_ptrs, eids, row_indices, col_indices,
                              left_node_compact_data, right_node_vectors)
        ctx.mark_dirty(ret)
        K.rgnn_inner_product_right_node_separatecoo(
            {"edata_idx_to_inverse_idx": edata_idx_to_inverse_idx}, 2, rel_ptrs, eids, row_indices, col_indices,
            left_node_compact_data, right_node_vectors, ret)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        inv, rel_ptrs, eids, row_indices, col_indices, left, right = ctx.saved_tensors
        grad_left = th.zeros_like(left, memory_format=th.contiguous_format)
        grad_right = th.zeros_like(right, memory_format=th.contiguous_format)
        K.backward_inner_product_right_node_separatecoo(
            {"edata_idx_to_inverse_idx": inv}, 2, rel_ptrs, eids, row_indices, col_indices, left, right,
            gradout.contiguous(), grad_left, grad_right)
        return None, None, None, None, None, grad_left, grad_right, None


class RgnnInnerProductNodeCompactAndNodeBinarySearch(th.autograd.Function):
    """rgnn_layers_and_funcs.py:202-348 with CompactAsOfNodeKind::Enabled (1): the compact left row of
    an edge is found by binary search of its col node in the relation's sorted unique list."""

    @staticmethod
    def forward(ctx, unique_rel_ptrs, unique_node_indices, rel_ptrs, eids, row_indices, col_indices,
                left_node_compact_data, right_node_vectors, ret):
        ctx.save_for_backward(unique_rel_ptrs, unique_node_indices, rel_ptrs, eids, row_indices, col_indices,
                              left_node_compact_data, right_node_vectors)
        ctx.mark_dirty(ret)
        K.rgnn_inner_product_right_node_separatecoo(
            {"unique_srcs_and_dests_rel_ptrs": unique_rel_ptrs, "unique_srcs_and_dests_node_indices": unique_node_indices},
            1, rel_ptrs, eids, row_indices, col_indices, left_node_compact_data, right_node_vectors, ret)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        up, un, rel_ptrs, eids, row_indices, col_indices, left, right = ctx.saved_tensors
        grad_left = th.zeros_like(left, memory_format=th.contiguous_format)
        grad_right = th.zeros_like(right, memory_format=th.contiguous_format)
        K.backward_inner_product_right_node_separatecoo(
            {"unique_srcs_and_dests_rel_ptrs": up, "unique_srcs_and_dests_node_indices": un}, 1, rel_ptrs, eids,
            row_indices, col_indices, left, right, gradout.contiguous(), grad_left, grad_right)
        return None, None, None, None, None, None, grad_left, grad_right, None


def rgnn_inner_product_right_node(graph, left_side_data, right_node_vectors, compact_as_of_node_kind,
                                  left_mapper_suffix=None):
    """rgnn_layers_and_funcs.py:498-530. Kind 0: left is per-edge; kind 2: left is compact
    ([unique (rel, node), H, F]) and addressed through inverse_idx<left_mapper_suffix>. Kind 1
    (binary search in the unique lists) is not built."""
    coo = graph.get_separate_coo_original()
    if compact_as_of_node_kind == 2:
        uniq = graph.get_separate_unique_node_indices_single_sided()
        ret = th.empty([coo["eids"].numel(), right_node_vectors.size(1)], dtype=right_node_vectors.dtype,
                       device=right_node_vectors.device)
        return RgnnInnerProductNodeCompactAndNode.apply(
            uniq["inverse_idx" + (left_mapper_suffix or "_col")], coo["rel_ptrs"], coo["eids"],
            coo["row_indices"], coo["col_indices"], left_side_data.contiguous(),
            right_node_vectors.contiguous(), ret)
    if compact_as_of_node_kind == 1:
        uniq = graph.get_separate_unique_node_indices_single_sided()
        sfx = left_mapper_suffix or "_col"
        ret = th.empty([coo["eids"].numel(), right_node_vectors.size(1)], dtype=right_node_vectors.dtype,
                       device=right_node_vectors.device)
        return RgnnInnerProductNodeCompactAndNodeBinarySearch.apply(
            uniq["rel_ptrs" + sfx], uniq["node_indices" + sfx], coo["rel_ptrs"], coo["eids"], coo["row_indices"],
            coo["col_indices"], left_side_data.contiguous(), right_node_vectors.contiguous(), ret)
    if compact_as_of_node_kind != 0:
        raise NotImplementedError("CompactAsOfNodeKind > 2 for the inner product")
    ret = th.empty([coo["eids"].numel(), right_node_vectors.size(1)], dtype=right_node_vectors.dtype,
                   device=right_node_vectors.device)
    return RgnnInnerProductEdgeAndNode.apply(
        coo["rel_ptrs"], coo["eids"], coo["row_indices"], coo["col_indices"],
        left_side_data.contiguous(), right_node_vectors.contiguous(), ret)


# ----- rgat_layers_and_funcs.py:8-117, 651-684 ---------------------------
class RelationalFusedGatCSR(th.autograd.Function):
    @staticmethod
    def forward(ctx, incsr_row_ptr, incsr_col_indices, incsr_eids, incsr_reltypes,
                outcsr_row_ptr, outcsr_col_indices, outcsr_eids, outcsr_reltypes,
                unique_rel_ptr, unique_node_indices, feat_src, el, er, s, exp, ret, slope):
        ctx.mark_dirty(ret)
        K.relational_fused_gat_csr(
            incsr_row_ptr, incsr_col_indices, incsr_eids, incsr_reltypes, unique_rel_ptr,
            unique_node_indices, feat_src, el, er, s, exp, ret, slope)
        ctx.save_for_backward(outcsr_row_ptr, outcsr_col_indices, outcsr_eids, outcsr_reltypes,
                              unique_rel_ptr, unique_node_indices, feat_src, el, er, s, exp, ret)
        ctx.slope = slope
        return ret

    @staticmethod
    def backward(ctx, gradout):
        (outcsr_row_ptr, outcsr_col_indices, outcsr_eids, outcsr_reltypes, unique_rel_ptr,
         unique_node_indices, feat_src, el, er, s, exp, ret) = ctx.saved_tensors
        grad_el = th.empty_like(el, memory_format=th.contiguous_format)
        grad_er = th.empty_like(er, memory_format=th.contiguous_format)
        grad_feat_src = th.empty_like(feat_src, memory_format=th.contiguous_format)
        K.backward_relational_fused_gat_csr(
            outcsr_row_ptr, outcsr_col_indices, outcsr_eids, outcsr_reltypes, unique_rel_ptr,
            unique_node_indices, feat_src, el, er, s, exp, ret, gradout.contiguous(),
            grad_feat_src, grad_el, grad_er, ctx.slope)
        return (None,) * 10 + (grad_feat_src, grad_el, grad_er, None, None, None, None)


def relational_fused_gat_csr(graph, feat_src, el, er, slope):
    incsr, outcsr = graph.get_in_csr(), graph.get_out_csr()
    uniq = graph.get_separate_unique_node_indices()
    exp = el.new_empty([graph.get_num_edges()] + list(el.size()[1:]))
    s = el.new_empty([graph.get_num_nodes(), el.size(1)])
    ret = el.new_empty([graph.get_num_nodes()] + list(feat_src.size()[1:]))
    return RelationalFusedGatCSR.apply(
        incsr["row_ptrs"], incsr["col_indices"], incsr["eids"], incsr["rel_types"],
        outcsr["row_ptrs"], outcsr["col_indices"], outcsr["eids"], outcsr["rel_types"],
        uniq["rel_ptrs"], uniq["node_indices"], feat_src.contiguous(), el.contiguous(),
        er.contiguous(), s, exp, ret, slope)


# ----- rgat_layers_and_funcs.py:233-322, 725-750 -------------------------
class RelationalFusedGatSeparateCOO(th.autograd.Function):
    @staticmethod
    def forward(ctx, eids, rel_ptrs, row_indices, col_indices, feat_src, el, er, s, exp, ret, slope):
        ctx.mark_dirty(ret)
        K.relational_fused_gat_separate_coo(
            eids, rel_ptrs, row_indices, col_indices, 0, dict(), feat_src, el, er, s, exp, ret, slope)
        ctx.save_for_backward(eids, rel_ptrs, row_indices, col_indices, feat_src, el, er, s, exp, ret)
        ctx.slope = slope
        return ret

    @staticmethod
    def backward(ctx, gradout):
        eids, rel_ptrs, row_indices, col_indices, feat_src, el, er, s, exp, ret = ctx.saved_tensors
        grad_el = th.empty_like(el, memory_format=th.contiguous_format)
        grad_er = th.empty_like(er, memory_format=th.contiguous_format)
        grad_feat_src = th.empty_like(feat_src, memory_format=th.contiguous_format)
        K.backward_relational_fused_gat_separate_coo(
            eids, rel_ptrs, row_indices, col_indices, 0, dict(), feat_src, el, er, s, exp, ret,
            gradout.contiguous(), grad_feat_src, grad_el, grad_er, ctx.slope)
        return None, None, None, None, grad_feat_src, grad_el, grad_er, None, None, None, None


def relational_fused_gat_separate_coo(g, feat, el, er, negative_slope):
    coo = g.get_separate_coo_original()
    exp = el.new_empty([g.get_num_edges()] + list(el.size()[1:]))
    s = el.new_empty([g.get_num_nodes(), el.size(1)])
    ret = el.new_empty([g.get_num_nodes()] + list(feat.size()[1:]))
    return RelationalFusedGatSeparateCOO.apply(
        coo["eids"], coo["rel_ptrs"], coo["row_indices"], coo["col_indices"], feat.contiguous(),
        el.contiguous(), er.contiguous(), s, exp, ret, negative_slope)


# ----- hgt_layers_and_funcs.py:296-415 -----------------------------------
class HGTFullGraphEdgeSoftmaxAndMessageMeanAggregationOpsCSR(th.autograd.Function):
    @staticmethod
    def forward(ctx, incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes,
                outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes,
                unnormalized_attn_score, mu, edgesoftmax_sum_per_node,
                mu_softmax_applied_unnormalized_attn_score, normalized_attn_score,
                message_per_edge, new_h):
        ctx.mark_dirty(new_h)
        K.hgt_full_graph_edge_softmax_ops_csr(
            incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, unnormalized_attn_score,
            mu, edgesoftmax_sum_per_node, mu_softmax_applied_unnormalized_attn_score,
            normalized_attn_score)
        # The aggregation op divides its score argument by the per-node sum
        # (as the reference kernel does), so it is given exp(mu*score), not the
        # already-normalised score the reference's Python passes
        # (hgt_layers_and_funcs.py:332-342, which double-normalises).
        K.hgt_full_graph_message_mean_aggregation_csr(
            incsr_row_ptrs, incsr_col_indices, incsr_reltypes, incsr_eids, message_per_edge,
            mu_softmax_applied_unnormalized_attn_score, edgesoftmax_sum_per_node, mu, new_h)
        ctx.save_for_backward(outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes,
                              unnormalized_attn_score, mu, edgesoftmax_sum_per_node,
                              normalized_attn_score, message_per_edge, new_h)
        return new_h

    @staticmethod
    def backward(ctx, gradout):
        (outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes, unnormalized_attn_score,
         mu, edgesoftmax_sum_per_node, normalized_attn_score, message_per_edge,
         new_h) = ctx.saved_tensors
        gradout = gradout.contiguous()
        grad_message = th.empty_like(message_per_edge, memory_format=th.contiguous_format)
        K.backward_hgt_full_graph_message_mean_aggregation_csr(
            outcsr_row_ptrs, outcsr_col_indices, outcsr_reltypes, outcsr_eids,
            edgesoftmax_sum_per_node, normalized_attn_score, gradout, grad_message)
        grad_attn_score = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
        grad_mu = th.zeros_like(mu, memory_format=th.contiguous_format)
        K.backward_hgt_full_graph_edge_softmax_ops_csr(
            outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes, message_per_edge,
            unnormalized_attn_score, normalized_attn_score, new_h, gradout, mu, grad_attn_score,
            grad_mu)
        return (None,) * 8 + (grad_attn_score, grad_mu, None, None, None, grad_message, None)


def hgt_full_graph_edge_softmax_and_message_mean_aggregation_csr(graph, message_per_edge,
                                                                 unnormalized_attn_score, mu):
    """hgt_layers_and_funcs.py:520-560"""
    incsr, outcsr = graph.get_in_csr(), graph.get_out_csr()
    n, h = graph.get_num_nodes(), mu.size(1)
    new_h = message_per_edge.new_empty(n, h, message_per_edge.size(2))
    ssum = message_per_edge.new_empty(n, h)
    t = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
    norm = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
    return HGTFullGraphEdgeSoftmaxAndMessageMeanAggregationOpsCSR.apply(
        incsr["row_ptrs"], incsr["col_indices"], incsr["eids"], incsr["rel_types"],
        outcsr["row_ptrs"], outcsr["col_indices"], outcsr["eids"], outcsr["rel_types"],
        unnormalized_attn_score.contiguous(), mu.contiguous(), ssum, t, norm,
        message_per_edge.contiguous(), new_h)


# ----- hgt_layers_and_funcs.py:9-295, 430-520: the separate-COO (edge-parallel) HGT ops (P1) -----
class HGTFullGraphHeteroAttentionOps(th.autograd.Function):
    @staticmethod
    def forward(ctx, incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, separate_coo_row_indices,
                separate_coo_col_indices, separate_coo_eids, separate_coo_relptrs, applied_klinear_node_features,
                applied_qlinear_node_features, attn_score_weight):
        e = separate_coo_row_indices.numel()
        unnormalized_attn_score = th.empty((e, attn_score_weight.size(1)), dtype=attn_score_weight.dtype,
                                           device=attn_score_weight.device)
        attn_score_inner_product = th.empty((e, attn_score_weight.size(1), attn_score_weight.size(3)),
                                            dtype=attn_score_weight.dtype, device=attn_score_weight.device)
        K.hgt_full_graph_hetero_attention_ops_coo(
            separate_coo_row_indices, separate_coo_col_indices, separate_coo_eids, separate_coo_relptrs,
            applied_klinear_node_features, applied_qlinear_node_features, attn_score_weight, attn_score_inner_product,
            unnormalized_attn_score)
        ctx.save_for_backward(incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, separate_coo_row_indices,
                              separate_coo_col_indices, separate_coo_eids, separate_coo_relptrs,
                              applied_klinear_node_features, applied_qlinear_node_features, attn_score_inner_product,
                              attn_score_weight)
        return unnormalized_attn_score

    @staticmethod
    def backward(ctx, grad_unnorm_attn_score):
        (incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, row, col, eids, relptrs, k, q, inner,
         w) = ctx.saved_tensors
        grad_w = th.zeros_like(w, memory_format=th.contiguous_format)
        grad_k = th.zeros_like(k, memory_format=th.contiguous_format)
        grad_q = th.zeros_like(q, memory_format=th.contiguous_format)
        K.backward_hgt_full_graph_hetero_attention_ops_coo(
            incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, row, col, eids, relptrs, grad_w,
            th.transpose(w, 2, 3).contiguous(), k, q, inner, grad_unnorm_attn_score.contiguous(), grad_k, grad_q)
        return (None,) * 8 + (grad_k, grad_q, grad_w)


def hgt_full_graph_hetero_attention_ops_coo(graph, weight, applied_klinear_node_features,
                                            applied_qlinear_node_features):
    """hgt_layers_and_funcs.py:430-455: score[e,h] = <k[src]·W[rel,h], q[dst]> in one op."""
    incsr, coo = graph.get_in_csr(), graph.get_separate_coo_original()
    return HGTFullGraphHeteroAttentionOps.apply(
        incsr["row_ptrs"], incsr["col_indices"], incsr["eids"], incsr["rel_types"], coo["row_indices"],
        coo["col_indices"], coo["eids"], coo["rel_ptrs"], applied_klinear_node_features.contiguous(),
        applied_qlinear_node_features.contiguous(), weight.contiguous())


class HGTFullGraphMessageCalcEdgeSoftmaxAndMessageMeanAggregationCOO(th.autograd.Function):
    @staticmethod
    def forward(ctx, incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, separate_coo_relptrs,
                separate_coo_row_indices, separate_coo_col_indices, separate_coo_eids, message_generation_weights,
                inputs, unnormalized_attn_score, edgesoftmax_sum_per_node, mu,
                mu_softmax_applied_unnormalized_attn_score, normalized_attn_score, new_h):
        ctx.mark_dirty(new_h)
        K.hgt_full_graph_edge_softmax_ops_separate_coo(
            separate_coo_row_indices, separate_coo_col_indices, separate_coo_eids, separate_coo_relptrs,
            unnormalized_attn_score, mu, edgesoftmax_sum_per_node, mu_softmax_applied_unnormalized_attn_score,
            normalized_attn_score)
        K.hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo(
            separate_coo_relptrs, separate_coo_eids, separate_coo_row_indices, separate_coo_col_indices, inputs,
            message_generation_weights, normalized_attn_score, new_h)
        ctx.save_for_backward(incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, separate_coo_relptrs,
                              separate_coo_row_indices, separate_coo_col_indices, separate_coo_eids,
                              message_generation_weights, inputs, unnormalized_attn_score, mu, normalized_attn_score,
                              new_h)
        return new_h

    USE_CSR_ENORM_BACKWARD = False  # the reference exposes both (`if 0:` picks the COO one, :253-290)

    @staticmethod
    def backward(ctx, gradout):
        (incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, relptrs, row, col, eids, w, inputs, attn, mu,
         norm, new_h) = ctx.saved_tensors
        grad_w = th.zeros_like(w, memory_format=th.contiguous_format)
        grad_input = th.zeros_like(inputs, memory_format=th.contiguous_format)
        grad_norm = th.zeros_like(norm, memory_format=th.contiguous_format)
        grad_attn = th.zeros_like(attn, memory_format=th.contiguous_format)
        grad_mu = th.zeros_like(mu, memory_format=th.contiguous_format)
        K.backward_hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo(
            relptrs, eids, row, col, inputs, th.transpose(w, 2, 3).contiguous(), norm, new_h, grad_input, grad_w,
            grad_norm, gradout.contiguous())
        if HGTFullGraphMessageCalcEdgeSoftmaxAndMessageMeanAggregationCOO.USE_CSR_ENORM_BACKWARD:
            K.backward_hgt_full_graph_enorm_to_unnormalized_attn_score_csr(
                incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, attn, norm, grad_norm, mu, grad_attn,
                grad_mu)
        else:
            sumprod = th.zeros([incsr_row_ptrs.shape[0] - 1, attn.shape[1]], dtype=norm.dtype, device=norm.device)
            K.backward_hgt_full_graph_enorm_to_unnormalized_attn_score_separate_coo(
                row, col, eids, relptrs, attn, norm, grad_norm, mu, grad_attn, grad_mu, sumprod)
        return (None,) * 8 + (grad_w, grad_input, grad_attn, None, grad_mu, None, None, None)


def hgt_full_graph_message_calc_edge_softmax_and_message_mean_aggregation_coo(
        message_generation_weights, inputs, graph, mu, unnormalized_attn_score):
    """hgt_layers_and_funcs.py:457-518: edge softmax over the COO + (typed linear fused with the
    attention-weighted aggregation)."""
    incsr, coo = graph.get_in_csr(), graph.get_separate_coo_original()
    n, h = graph.get_num_nodes(), mu.size(1)
    new_h = th.empty((n, h, message_generation_weights.size(3)), dtype=inputs.dtype, device=inputs.device)
    ssum = th.empty((n, h), dtype=inputs.dtype, device=inputs.device)
    t = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
    norm = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
    return HGTFullGraphMessageCalcEdgeSoftmaxAndMessageMeanAggregationCOO.apply(
        incsr["row_ptrs"], incsr["col_indices"], incsr["eids"], incsr["rel_types"], coo["rel_ptrs"],
        coo["row_indices"], coo["col_indices"], coo["eids"], message_generation_weights.contiguous(),
        inputs.contiguous(), unnormalized_attn_score.contiguous(), ssum, mu.contiguous(), t, norm, new_h)


# ----- hetb200 extensions: compact-as-of-node (direct indexing) forms of the in-CSR ops -----
class RelationalFusedGatCSRCompact(th.autograd.Function):
    """relational_fused_gat_csr with feat_src / el compact by (relation, source) and er compact by
    (relation, destination); the reference offers this only for its edge-parallel COO kernel
    (rgat_layers_and_funcs.py:325-470, relational_fused_gat_compact_as_of_node_separate_coo_single_sided)."""

    @staticmethod
    def forward(ctx, incsr_row_ptr, incsr_eids, outcsr_col_indices, outcsr_eids, in_inverse_row, in_inverse_col,
                inverse_row, inverse_col, feat_compact, el_compact, er_compact, s, exp, ret, slope, by_rows=False):
        # in_inverse_* are aligned with incsr_eids, inverse_* with outcsr_eids. by_rows: the backward
        # walks the edges grouped by compact source row instead of the out-CSR, and the four
        # backward arguments are (urow_dst, urow_eids, urow_ptrs, urow_inverse_idx_col)
        ctx.by_rows = by_rows
        ctx.mark_dirty(ret)
        K.hetb200_relational_fused_gat_csr_compact(
            incsr_row_ptr, incsr_eids, in_inverse_row, in_inverse_col, feat_compact, el_compact, er_compact, s,
            exp, ret, slope)
        ctx.save_for_backward(outcsr_col_indices, outcsr_eids, inverse_row, inverse_col, feat_compact,
                              el_compact, er_compact, s, exp, ret)
        ctx.slope = slope
        return ret

    @staticmethod
    def backward(ctx, gradout):
        (outcsr_col_indices, outcsr_eids, inverse_row, inverse_col, feat, el, er, s, exp, ret) = ctx.saved_tensors
        grad_el = th.zeros_like(el, memory_format=th.contiguous_format)
        grad_er = th.zeros_like(er, memory_format=th.contiguous_format)
        grad_feat = th.zeros_like(feat, memory_format=th.contiguous_format)
        if ctx.by_rows:
            K.hetb200_backward_relational_fused_gat_compact_rows(
                inverse_row, outcsr_col_indices, outcsr_eids, inverse_col, feat, el, er, s, exp, ret,
                gradout.contiguous(), grad_feat, grad_el, grad_er, ctx.slope)
        else:
            K.hetb200_backward_relational_fused_gat_csr_compact(
                outcsr_col_indices, outcsr_eids, inverse_row, inverse_col, feat, el, er, s, exp, ret,
                gradout.contiguous(), grad_feat, grad_el, grad_er, ctx.slope)
        return (None,) * 8 + (grad_feat, grad_el, grad_er, None, None, None, None, None)


RGAT_COMPACT_BACKWARD_BY_ROWS = True  # False: edge-parallel out-CSR backward with one reduction per edge


def relational_fused_gat_csr_compact(graph, feat_compact, el_compact, er_compact, slope):
    incsr, outcsr = graph.get_in_csr(), graph.get_out_csr()
    uniq = graph.get_separate_unique_node_indices_single_sided()
    h = el_compact.size(1)
    exp = el_compact.new_empty([graph.get_num_edges(), h, 1])
    s = el_compact.new_empty([graph.get_num_nodes(), h])
    ret = el_compact.new_empty([graph.get_num_nodes()] + list(feat_compact.size()[1:]))
    if RGAT_COMPACT_BACKWARD_BY_ROWS:
        bwd = (uniq["urow_dst"], uniq["urow_eids"], uniq["urow_ptrs"], uniq["urow_inverse_idx_col"])
    else:
        bwd = (outcsr["col_indices"], outcsr["eids"], uniq["outcsr_inverse_idx_row"], uniq["outcsr_inverse_idx_col"])
    return RelationalFusedGatCSRCompact.apply(
        incsr["row_ptrs"], incsr["eids"], bwd[0], bwd[1], uniq["incsr_inverse_idx_row"],
        uniq["incsr_inverse_idx_col"], bwd[2], bwd[3],
        feat_compact.contiguous(), el_compact.contiguous(), er_compact.contiguous(), s, exp, ret, slope,
        RGAT_COMPACT_BACKWARD_BY_ROWS)


class HGTFullGraphEdgeSoftmaxAndMessageMeanAggregationOpsCSRCompact(th.autograd.Function):
    """hgt_full_graph_edge_softmax_and_message_mean_aggregation_csr with the messages compact by
    (relation, source): V[src]·W_msg[r] is computed once per unique pair, not once per edge."""

    @staticmethod
    def forward(ctx, incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, outcsr_col_indices,
                outcsr_eids, outcsr_reltypes, in_msg_idx, msg_idx, unnormalized_attn_score, mu, ssum, t, norm,
                message_compact, new_h):
        # in_msg_idx is aligned with incsr_eids, msg_idx with outcsr_eids
        ctx.mark_dirty(new_h)
        K.hgt_full_graph_edge_softmax_ops_csr(
            incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, unnormalized_attn_score, mu, ssum, t,
            norm)
        K.hetb200_hgt_message_aggregation_csr_compact(
            incsr_row_ptrs, incsr_eids, in_msg_idx, message_compact, t, ssum, new_h)
        ctx.save_for_backward(outcsr_col_indices, outcsr_eids, outcsr_reltypes, msg_idx,
                              unnormalized_attn_score, mu, norm, message_compact, new_h)
        return new_h

    @staticmethod
    def backward(ctx, gradout):
        (outcsr_col_indices, outcsr_eids, outcsr_reltypes, msg_idx, unnormalized_attn_score, mu, norm,
         message_compact, new_h) = ctx.saved_tensors
        gradout = gradout.contiguous()
        grad_message = th.zeros_like(message_compact, memory_format=th.contiguous_format)
        K.hetb200_backward_hgt_message_aggregation_csr_compact(
            outcsr_col_indices, outcsr_eids, msg_idx, norm, gradout, grad_message)
        grad_attn_score = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
        grad_mu = th.zeros_like(mu, memory_format=th.contiguous_format)
        K.hetb200_backward_hgt_edge_softmax_ops_csr_compact(
            outcsr_col_indices, outcsr_eids, outcsr_reltypes, msg_idx, message_compact, unnormalized_attn_score,
            norm, new_h, gradout, mu, grad_attn_score, grad_mu)
        return (None,) * 9 + (grad_attn_score, grad_mu, None, None, None, grad_message, None)


def hgt_full_graph_edge_softmax_and_message_mean_aggregation_csr_compact(graph, message_compact,
                                                                         unnormalized_attn_score, mu):
    incsr, outcsr = graph.get_in_csr(), graph.get_out_csr()
    uniq = graph.get_separate_unique_node_indices_single_sided()
    n, h = graph.get_num_nodes(), mu.size(1)
    new_h = message_compact.new_empty(n, h, message_compact.size(2))
    ssum = message_compact.new_empty(n, h)
    t = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
    norm = th.empty_like(unnormalized_attn_score, memory_format=th.contiguous_format)
    return HGTFullGraphEdgeSoftmaxAndMessageMeanAggregationOpsCSRCompact.apply(
        incsr["row_ptrs"], incsr["col_indices"], incsr["eids"], incsr["rel_types"], outcsr["col_indices"],
        outcsr["eids"], outcsr["rel_types"], uniq["incsr_inverse_idx_row"], uniq["outcsr_inverse_idx_row"],
        unnormalized_attn_score.contiguous(), mu.contiguous(), ssum, t, norm, message_compact.contiguous(), new_h)


class HGTFusedAttentionAggregationCSRCompact(th.autograd.Function):
    """hetb200 fusion of rgnn_inner_product_right_node (kind 2) + the edge softmax + the message
    aggregation: one in-CSR pass computes the score, exponentiates it and accumulates the weighted
    compact messages, so the score never leaves registers on its way to the aggregation. The
    backward is one out-CSR pass as well (aggregation-, softmax- and inner-product-backward fused)."""

    @staticmethod
    def forward(ctx, incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, outcsr_row_ptrs,
                outcsr_col_indices, outcsr_eids, outcsr_reltypes, in_left_idx, in_msg_idx, left_idx, msg_idx,
                left_compact, k, message_compact, mu, attn, t, ssum, new_h):
        # in_*_idx are aligned with incsr_eids, *_idx with outcsr_eids
        ctx.mark_dirty(new_h)
        K.hetb200_hgt_fused_attention_aggregation_csr(
            incsr_row_ptrs, incsr_col_indices, incsr_eids, incsr_reltypes, in_left_idx, in_msg_idx, left_compact,
            k, message_compact, mu, attn, t, ssum, new_h)
        ctx.save_for_backward(outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes, left_idx,
                              msg_idx, left_compact, k, message_compact, mu, attn, t, ssum, new_h)
        return new_h

    @staticmethod
    def backward(ctx, gradout):
        (outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes, left_idx, msg_idx, left, k, msg,
         mu, attn, t, ssum, new_h) = ctx.saved_tensors
        grad_msg = th.zeros_like(msg, memory_format=th.contiguous_format)
        grad_left = th.zeros_like(left, memory_format=th.contiguous_format)
        grad_k = th.zeros_like(k, memory_format=th.contiguous_format)
        grad_mu = th.zeros_like(mu, memory_format=th.contiguous_format)
        K.hetb200_backward_hgt_fused_attention_aggregation_csr(
            outcsr_row_ptrs, outcsr_col_indices, outcsr_eids, outcsr_reltypes, left_idx, msg_idx, left, k, msg,
            mu, attn, t, ssum, new_h, gradout.contiguous(), grad_left, grad_k, grad_msg, grad_mu)
        return (None,) * 12 + (grad_left, grad_k, grad_msg, grad_mu, None, None, None, None)


def hgt_fused_attention_and_aggregation_csr_compact(graph, left_compact, k, message_compact, mu):
    incsr, outcsr, coo = graph.get_in_csr(), graph.get_out_csr(), graph.get_separate_coo_original()
    uniq = graph.get_separate_unique_node_indices_single_sided()
    n, e, h = graph.get_num_nodes(), graph.get_num_edges(), mu.size(1)
    new_h = message_compact.new_empty(n, h, message_compact.size(2))
    ssum = message_compact.new_empty(n, h)
    attn = message_compact.new_empty(e, h)
    t = message_compact.new_empty(e, h)
    return HGTFusedAttentionAggregationCSRCompact.apply(
        incsr["row_ptrs"], incsr["col_indices"], incsr["eids"], incsr["rel_types"], outcsr["row_ptrs"],
        outcsr["col_indices"], outcsr["eids"], outcsr["rel_types"], uniq["incsr_inverse_idx_col"],
        uniq["incsr_inverse_idx_row"], uniq["outcsr_inverse_idx_col"], uniq["outcsr_inverse_idx_row"],
        left_compact.contiguous(),
        k.contiguous(), message_compact.contiguous(), mu.contiguous(), attn, t, ssum, new_h)


class ComposeWeights(th.autograd.Function):
    """W[r] = lin[type_of_rel[r]] . blockdiag_h(rel[r, h]) -> [R, 1, in, H*dk]: the composed weights of HGT's
    linear operator reordering, one small kernel instead of a batched cuBLAS product per step."""

    @staticmethod
    def forward(ctx, lin, type_of_rel, rel):
        out = th.empty((rel.size(0), 1, lin.size(2), lin.size(3)), dtype=lin.dtype, device=lin.device)
        K.hetb200_compose_weights(lin, type_of_rel, rel, out)
        ctx.save_for_backward(lin, type_of_rel, rel)
        return out

    @staticmethod
    def backward(ctx, gradout):
        lin, type_of_rel, rel = ctx.saved_tensors
        grad_lin = th.zeros_like(lin, memory_format=th.contiguous_format)
        grad_rel = th.empty_like(rel, memory_format=th.contiguous_format)
        K.hetb200_backward_compose_weights(gradout.contiguous(), lin, type_of_rel, rel, grad_lin, grad_rel)
        return grad_lin, None, grad_rel


def compose_weights(lin, type_of_rel, rel):
    return ComposeWeights.apply(lin.contiguous(), type_of_rel.contiguous(), rel.contiguous())


class SoftmaxCrossEntropy(th.autograd.Function):
    """loss = nll_loss(log_softmax(logits), labels, reduction="sum") / denom in one pass
    (the loss of HET_RGNN_train, hrt/python/RGNNUtils/RGNNUtils.py:291-345)."""

    @staticmethod
    def forward(ctx, logits, labels, denom):
        logits = logits.contiguous()
        lse = logits.new_empty(logits.size(0))
        loss = logits.new_empty(())
        K.hetb200_softmax_cross_entropy(logits, labels, float(denom), lse, loss)
        ctx.save_for_backward(logits, labels, lse)
        ctx.denom = float(denom)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        logits, labels, lse = ctx.saved_tensors
        grad = th.empty_like(logits)
        K.hetb200_backward_softmax_cross_entropy(logits, labels, lse, grad_loss.contiguous(), ctx.denom, grad)
        return grad, None, None


def softmax_cross_entropy(logits, labels, denom):
    return SoftmaxCrossEntropy.apply(logits, labels, denom)


# ----- rgcn_layers_and_funcs.py:481-601 (RGCN) --------------------------------
class RgcnLayer1SeparateCoo(th.autograd.Function):
    @staticmethod
    def forward(ctx, separate_coo_rel_ptrs, separate_coo_eids, separate_coo_row_indices,
                separate_coo_col_indices, x, weight, norm, ret):
        ctx.save_for_backward(separate_coo_rel_ptrs, separate_coo_eids, separate_coo_row_indices,
                              separate_coo_col_indices, weight, norm, x)
        ctx.mark_dirty(ret)
        K.rgcn_layer1_separate_coo(separate_coo_rel_ptrs, separate_coo_eids, separate_coo_row_indices,
                                   separate_coo_col_indices, x, weight, norm, ret)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        rel_ptrs, eids, row_indices, col_indices, weight, norm, x = ctx.saved_tensors
        grad_x = th.zeros_like(x, memory_format=th.contiguous_format)
        grad_weight = th.zeros_like(weight, memory_format=th.contiguous_format)
        grad_norm = th.zeros_like(norm, memory_format=th.contiguous_format)
        K.backward_rgcn_layer1_separate_coo(
            rel_ptrs, eids, row_indices, col_indices, x, th.transpose(weight, 1, 2).contiguous(), norm,
            grad_norm, grad_x, gradout.contiguous(), grad_weight)
        return None, None, None, None, grad_x, grad_weight, grad_norm, None


def rgcn_layer1_separate_coo(graph, x, weight, norm):
    """rgcn_layers_and_funcs.py:570-601: out[dst] = sum_e norm[e] * x[src] @ weight[rel(e)]."""
    coo = graph.get_separate_coo_original()
    ret = th.empty((graph.get_num_nodes(), weight.size(2)), dtype=weight.dtype, device=weight.device)
    return RgcnLayer1SeparateCoo.apply(coo["rel_ptrs"], coo["eids"], coo["row_indices"], coo["col_indices"],
                                       x.contiguous(), weight.contiguous(), norm.contiguous(), ret)


# ----- rgcn_layers_and_funcs.py:684-755 (compact sources, direct indexing) -------------------
class RGCNNodeMeanAggregationCompactAsOfNodeDirectIndexingSeparateCOO(th.autograd.Function):
    @staticmethod
    def forward(ctx, separate_coo_eids, separate_coo_rel_ptrs, separate_coo_row_indices,
                separate_coo_col_indices, inverse_indices_row, feat_src, enorm, ret):
        ctx.save_for_backward(separate_coo_eids, separate_coo_rel_ptrs, separate_coo_row_indices,
                              separate_coo_col_indices, inverse_indices_row, feat_src, enorm, ret)
        ctx.mark_dirty(ret)
        K.rgcn_node_mean_aggregation_compact_as_of_node_separate_coo(
            separate_coo_eids, separate_coo_rel_ptrs, separate_coo_row_indices, separate_coo_col_indices,
            {"inverse_indices_row": inverse_indices_row}, feat_src, enorm, ret, True)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        eids, rel_ptrs, row_indices, col_indices, inverse_indices_row, feat_src, enorm, ret = ctx.saved_tensors
        grad_feat_src = th.zeros_like(feat_src, memory_format=th.contiguous_format)
        K.backward_rgcn_node_mean_aggregation_compact_as_of_node_separate_coo(
            eids, rel_ptrs, row_indices, col_indices, {"inverse_indices_row": inverse_indices_row}, feat_src,
            enorm, ret, gradout.contiguous(), grad_feat_src, True)
        return None, None, None, None, None, grad_feat_src, None, None


def rgcn_node_mean_aggregation_compact_as_of_node_separate_coo_single_sided(
        g, feat_compact_src, enorm, compact_direct_indexing_flag):
    """rgcn_layers_and_funcs.py:782-830. Only direct indexing is built (the binary-search kind
    raises, like the other compact ops)."""
    if not compact_direct_indexing_flag:
        raise NotImplementedError("compact_as_of_node without direct indexing (binary search)")
    coo = g.get_separate_coo_original()
    uniq = g.get_separate_unique_node_indices_single_sided()
    ret = th.empty([g.get_num_nodes()] + list(feat_compact_src.size()[1:]), dtype=feat_compact_src.dtype,
                   device=feat_compact_src.device)
    return RGCNNodeMeanAggregationCompactAsOfNodeDirectIndexingSeparateCOO.apply(
        coo["eids"], coo["rel_ptrs"], coo["row_indices"], coo["col_indices"], uniq["inverse_idx_row"],
        feat_compact_src.contiguous(), enorm.contiguous(), ret)


class RgcnNodeAggregationCSRCompact(th.autograd.Function):
    """hetb200 extension: the same aggregation over the in-CSR with the edge-balanced walker (no
    atomics in the forward); backward over the out-CSR into the compact rows."""

    @staticmethod
    def forward(ctx, incsr_row_ptrs, incsr_eids, outcsr_col_indices, outcsr_eids, in_msg_idx, msg_idx,
                feat_compact, enorm, ret):
        ctx.mark_dirty(ret)
        K.hetb200_rgcn_node_aggregation_csr_compact(incsr_row_ptrs, incsr_eids, in_msg_idx, feat_compact, enorm, ret)
        ctx.save_for_backward(outcsr_col_indices, outcsr_eids, msg_idx, feat_compact, enorm)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        outcsr_col_indices, outcsr_eids, msg_idx, feat_compact, enorm = ctx.saved_tensors
        grad_feat = th.zeros_like(feat_compact, memory_format=th.contiguous_format)
        K.hetb200_backward_hgt_message_aggregation_csr_compact(
            outcsr_col_indices, outcsr_eids, msg_idx, enorm, gradout.contiguous(), grad_feat)
        return (None,) * 6 + (grad_feat, None, None)


def rgcn_node_aggregation_csr_compact(g, feat_compact_src, enorm):
    incsr, outcsr = g.get_in_csr(), g.get_out_csr()
    uniq = g.get_separate_unique_node_indices_single_sided()
    ret = th.empty([g.get_num_nodes()] + list(feat_compact_src.size()[1:]), dtype=feat_compact_src.dtype,
                   device=feat_compact_src.device)
    return RgcnNodeAggregationCSRCompact.apply(
        incsr["row_ptrs"], incsr["eids"], outcsr["col_indices"], outcsr["eids"], uniq["incsr_inverse_idx_row"],
        uniq["outcsr_inverse_idx_row"], feat_compact_src.contiguous(), enorm.contiguous(), ret)


class HGTReorderedProjections(th.autograd.Function):
    """The three typed linears of one HGT layer input under linear operator reordering, as ONE
    autograd node:  k = h.W_k[type];  left_c = h[dst].(W_q.W_att)[r] per unique (relation, dst);
    msg_c = h[src].(W_v.W_msg)[r] per unique (relation, src). Same ops as calling
    rgnn_relational_matmul_no_scatter_gather_list + rgnn_relational_matmul (kind 1) twice; the
    point is the backward: the three input gradients land in ONE buffer (the node-typed linear
    stores every row, the two compact ones scatter-add into it), which saves two zero-fills and
    two [N, D] additions per layer that autograd would otherwise do to sum them."""

    @staticmethod
    def forward(ctx, offs, rel_ptrs_col, nodes_col, rel_ptrs_row, nodes_row, h, k_linears, w_att, w_msg):
        n, d_out = h.size(0), k_linears.size(1) * k_linears.size(3)
        k = th.empty((n, d_out), dtype=h.dtype, device=h.device)
        K.rgnn_relational_matmul_no_scatter_gather_list(offs, k_linears, h, k)
        left = th.empty((nodes_col.numel(), w_att.size(1), w_att.size(3)), dtype=h.dtype, device=h.device)
        K.rgnn_relational_matmul({"unique_srcs_and_dests_rel_ptrs": rel_ptrs_col,
                                  "unique_srcs_and_dests_node_indices": nodes_col}, 1, w_att, h, left, True)
        msg = th.empty((nodes_row.numel(), w_msg.size(1), w_msg.size(3)), dtype=h.dtype, device=h.device)
        K.rgnn_relational_matmul({"unique_srcs_and_dests_rel_ptrs": rel_ptrs_row,
                                  "unique_srcs_and_dests_node_indices": nodes_row}, 1, w_msg, h, msg, True)
        ctx.save_for_backward(offs, rel_ptrs_col, nodes_col, rel_ptrs_row, nodes_row, h, k_linears, w_att, w_msg)
        return k, left, msg

    @staticmethod
    def backward(ctx, grad_k, grad_left, grad_msg):
        offs, rel_ptrs_col, nodes_col, rel_ptrs_row, nodes_row, h, k_linears, w_att, w_msg = ctx.saved_tensors
        # the node-type offsets cover every row, so the first op writes all of grad_h
        grad_h = th.empty_like(h, memory_format=th.contiguous_format)
        grad_wk = th.zeros_like(k_linears, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul_no_scatter_gather_list(
            offs, th.transpose(k_linears, 2, 3).contiguous(), h, grad_k.contiguous(), grad_h, grad_wk)
        grad_watt = th.zeros_like(w_att, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": rel_ptrs_col, "unique_srcs_and_dests_node_indices": nodes_col}, 1,
            th.transpose(w_att, 2, 3).contiguous(), h, grad_left.contiguous(), grad_h, grad_watt, True)
        grad_wmsg = th.zeros_like(w_msg, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": rel_ptrs_row, "unique_srcs_and_dests_node_indices": nodes_row}, 1,
            th.transpose(w_msg, 2, 3).contiguous(), h, grad_msg.contiguous(), grad_h, grad_wmsg, True)
        return None, None, None, None, None, grad_h, grad_wk, grad_watt, grad_wmsg


def hgt_reordered_projections(g, h, k_linears, w_att, w_msg):
    u = g.get_separate_unique_node_indices_single_sided()
    return HGTReorderedProjections.apply(
        g.get_original_node_type_offsets(), u["rel_ptrs_col"], u["node_indices_col"], u["rel_ptrs_row"],
        u["node_indices_row"], h.contiguous(), k_linears.contiguous(), w_att.contiguous(), w_msg.contiguous())


class HGTReorderedProjectionsSharded(th.autograd.Function):
    """HGTReorderedProjections on a destination shard (het_b200/parallel.py): ``h`` are the owned rows;
    the layer-input exchange (``xch``: a parallel.RowExchange) happens INSIDE this node so that it can
    overlap: forward = start the exchange, run the destination-side typed linear (left_c reads only the
    local h), finish the exchange, then k = hg.W_k[type] for every received row (one exchange per layer
    instead of gathering k and v; the price is a typed linear over the received rows instead of N/P)
    and msg_c through the (relation, source) lists. Backward = the two input gradients of hg into ONE
    buffer, start its reverse exchange, run the three weight-gradient kernels and the left_c input
    gradient under it, then add the returned rows into grad_h."""

    @staticmethod
    def forward(ctx, xch, gathered_type_ptrs, rel_ptrs_col, nodes_col, rel_ptrs_row, nodes_row, h, k_ext, w_att, w_msg):
        handle = xch.forward_start(h)
        left = th.empty((nodes_col.numel(), w_att.size(1), w_att.size(3)), dtype=h.dtype, device=h.device)
        K.rgnn_relational_matmul({"unique_srcs_and_dests_rel_ptrs": rel_ptrs_col,
                                  "unique_srcs_and_dests_node_indices": nodes_col}, 1, w_att, h, left, True)
        hg = xch.forward_finish(handle)
        ng, d_out = hg.size(0), k_ext.size(1) * k_ext.size(3)
        k = th.empty((ng, d_out), dtype=hg.dtype, device=hg.device)
        K.rgnn_relational_matmul_no_scatter_gather_list(gathered_type_ptrs, k_ext, hg, k)
        msg = th.empty((nodes_row.numel(), w_msg.size(1), w_msg.size(3)), dtype=h.dtype, device=h.device)
        K.rgnn_relational_matmul({"unique_srcs_and_dests_rel_ptrs": rel_ptrs_row,
                                  "unique_srcs_and_dests_node_indices": nodes_row}, 1, w_msg, hg, msg, True)
        ctx.xch = xch
        ctx.save_for_backward(gathered_type_ptrs, rel_ptrs_col, nodes_col, rel_ptrs_row, nodes_row, h, hg, k_ext, w_att,
                              w_msg)
        return k, left, msg

    @staticmethod
    def backward(ctx, grad_k, grad_left, grad_msg):
        gptrs, rel_ptrs_col, nodes_col, rel_ptrs_row, nodes_row, h, hg, k_ext, w_att, w_msg = ctx.saved_tensors
        row = {"unique_srcs_and_dests_rel_ptrs": rel_ptrs_row, "unique_srcs_and_dests_node_indices": nodes_row}
        col = {"unique_srcs_and_dests_rel_ptrs": rel_ptrs_col, "unique_srcs_and_dests_node_indices": nodes_col}
        none = h.new_empty(0)  # "skip this half" for the backward ops
        grad_k, grad_msg = grad_k.contiguous(), grad_msg.contiguous()
        wk_t, wmsg_t = th.transpose(k_ext, 2, 3).contiguous(), th.transpose(w_msg, 2, 3).contiguous()
        # input gradients of the received rows first: their reverse exchange is the long pole
        grad_hg = th.empty_like(hg, memory_format=th.contiguous_format)  # the segments cover every received row
        K.backward_rgnn_relational_matmul_no_scatter_gather_list(gptrs, wk_t, hg, grad_k, grad_hg, none)
        K.backward_rgnn_relational_matmul(row, 1, wmsg_t, hg, grad_msg, grad_hg, none, True)
        handle = ctx.xch.backward_start(grad_hg)
        # under the exchange: the weight gradients and the destination-side input gradient
        grad_wk = th.zeros_like(k_ext, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul_no_scatter_gather_list(gptrs, wk_t, hg, grad_k, none, grad_wk)
        grad_wmsg = th.zeros_like(w_msg, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(row, 1, wmsg_t, hg, grad_msg, none, grad_wmsg, True)
        grad_h = th.zeros_like(h, memory_format=th.contiguous_format)
        grad_watt = th.zeros_like(w_att, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(col, 1, th.transpose(w_att, 2, 3).contiguous(), h, grad_left.contiguous(),
                                          grad_h, grad_watt, True)
        ctx.xch.backward_finish(handle, grad_h)
        return None, None, None, None, None, None, grad_h, grad_wk, grad_watt, grad_wmsg


def hgt_reordered_projections_sharded(g, h, xch, k_linears, w_att, w_msg):
    """k_linears [T, 1, in, out] -> one weight per segment of the received-row layout: per rank the T
    types (and, in the all-gather layout, a zero matrix for its padding rows)."""
    u = g.get_separate_unique_node_indices_single_sided()
    if g.gathered_segments_per_rank == k_linears.size(0) + 1:   # all-gather layout: T types + padding per rank
        k_ext = th.cat([k_linears, th.zeros_like(k_linears[:1])], 0).repeat(g.world, 1, 1, 1)
    else:                                                        # halo table: T types per owner block
        k_ext = k_linears.repeat(g.world, 1, 1, 1)
    return HGTReorderedProjectionsSharded.apply(
        xch, g.gathered_type_ptrs, u["rel_ptrs_col"], u["node_indices_col"], u["rel_ptrs_row"], u["node_indices_row"],
        h.contiguous(), k_ext.contiguous(), w_att.contiguous(), w_msg.contiguous())


class RelationalMatmulCompactPair(th.autograd.Function):
    """Two compact typed linears of the SAME input as one autograd node (RGAT under linear operator
    reordering: X.W_r per unique (relation, source) and X.(W_r.a_r) per unique (relation,
    destination)): both input gradients scatter-add into one zero-filled buffer instead of two that
    autograd then has to sum."""

    @staticmethod
    def forward(ctx, rel_ptrs_a, nodes_a, rel_ptrs_b, nodes_b, x, w_a, w_b):
        out_a = th.empty((nodes_a.numel(), w_a.size(1), w_a.size(3)), dtype=x.dtype, device=x.device)
        K.rgnn_relational_matmul({"unique_srcs_and_dests_rel_ptrs": rel_ptrs_a,
                                  "unique_srcs_and_dests_node_indices": nodes_a}, 1, w_a, x, out_a, True)
        out_b = th.empty((nodes_b.numel(), w_b.size(1), w_b.size(3)), dtype=x.dtype, device=x.device)
        K.rgnn_relational_matmul({"unique_srcs_and_dests_rel_ptrs": rel_ptrs_b,
                                  "unique_srcs_and_dests_node_indices": nodes_b}, 1, w_b, x, out_b, True)
        ctx.save_for_backward(rel_ptrs_a, nodes_a, rel_ptrs_b, nodes_b, x, w_a, w_b)
        return out_a, out_b

    @staticmethod
    def backward(ctx, grad_a, grad_b):
        rel_ptrs_a, nodes_a, rel_ptrs_b, nodes_b, x, w_a, w_b = ctx.saved_tensors
        grad_x = th.zeros_like(x, memory_format=th.contiguous_format)
        grad_wa = th.zeros_like(w_a, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": rel_ptrs_a, "unique_srcs_and_dests_node_indices": nodes_a}, 1,
            th.transpose(w_a, 2, 3).contiguous(), x, grad_a.contiguous(), grad_x, grad_wa, True)
        grad_wb = th.zeros_like(w_b, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": rel_ptrs_b, "unique_srcs_and_dests_node_indices": nodes_b}, 1,
            th.transpose(w_b, 2, 3).contiguous(), x, grad_b.contiguous(), grad_x, grad_wb, True)
        return None, None, None, None, grad_x, grad_wa, grad_wb


def relational_matmul_compact_pair(g, x, w_row, w_col):
    u = g.get_separate_unique_node_indices_single_sided()
    return RelationalMatmulCompactPair.apply(u["rel_ptrs_row"], u["node_indices_row"], u["rel_ptrs_col"],
                                             u["node_indices_col"], x.contiguous(), w_row.contiguous(),
                                             w_col.contiguous())


class AddBias(th.autograd.Function):
    """h + bias (the layer epilogue, RGAT/models.py:264-278); the bias gradient is one column-sum
    kernel over the incoming gradient and the input gradient is the incoming gradient itself."""

    @staticmethod
    def forward(ctx, h, bias):
        return h + bias

    @staticmethod
    def backward(ctx, gradout):
        gradout = gradout.contiguous()
        grad_bias = th.empty(gradout.size(-1), dtype=gradout.dtype, device=gradout.device)
        K.hetb200_column_sum(gradout.view(-1, gradout.size(-1)), grad_bias)
        return gradout, grad_bias


def add_bias(h, bias):
    if h.is_cuda and h.dim() == 2 and h.size(1) % 4 == 0 and 256 % (h.size(1) // 4) == 0 and h.dtype == th.float32:
        return AddBias.apply(h, bias)
    return h + bias


class RelationalFusedGatCSRCompactWithLogits(th.autograd.Function):
    """el = <feat_compact, attn_l[r]> followed by relational_fused_gat_csr_compact as ONE autograd node:
    the same two ops (rgnn_relational_matmul_no_scatter_gather_list + the compact in-CSR RGAT op),
    but in the backward the el path adds its part of grad_feat straight into the buffer the edge
    backward produced, instead of a second [U, D] tensor that autograd then sums with it."""

    @staticmethod
    def forward(ctx, rel_ptrs_row, incsr_row_ptr, incsr_eids, in_inverse_row, in_inverse_col, bwd0, bwd1, bwd2, bwd3,
                feat_compact, er_compact, attn_l, s, exp, ret, slope, by_rows):
        u, h = feat_compact.size(0), attn_l.size(1)
        el = th.empty((u, h), dtype=feat_compact.dtype, device=feat_compact.device)
        w_l = attn_l.unsqueeze(-1).contiguous()
        K.rgnn_relational_matmul_no_scatter_gather_list(rel_ptrs_row, w_l, feat_compact.view(u, -1), el)
        el = el.view(u, h, 1)
        ctx.mark_dirty(ret)
        K.hetb200_relational_fused_gat_csr_compact(
            incsr_row_ptr, incsr_eids, in_inverse_row, in_inverse_col, feat_compact, el, er_compact, s, exp, ret,
            slope)
        ctx.save_for_backward(rel_ptrs_row, bwd0, bwd1, bwd2, bwd3, feat_compact, el, er_compact, w_l, s, exp, ret)
        ctx.slope, ctx.by_rows = slope, by_rows
        return ret

    @staticmethod
    def backward(ctx, gradout):
        rel_ptrs_row, bwd0, bwd1, bwd2, bwd3, feat, el, er, w_l, s, exp, ret = ctx.saved_tensors
        grad_el = th.zeros_like(el, memory_format=th.contiguous_format)
        grad_er = th.zeros_like(er, memory_format=th.contiguous_format)
        grad_feat = th.zeros_like(feat, memory_format=th.contiguous_format)
        if ctx.by_rows:  # (urow_dst, urow_eids, urow_ptrs, urow_inverse_idx_col)
            K.hetb200_backward_relational_fused_gat_compact_rows(
                bwd2, bwd0, bwd1, bwd3, feat, el, er, s, exp, ret, gradout.contiguous(), grad_feat, grad_el,
                grad_er, ctx.slope)
        else:            # (outcsr_col_indices, outcsr_eids, outcsr_inverse_idx_row, outcsr_inverse_idx_col)
            K.hetb200_backward_relational_fused_gat_csr_compact(
                bwd0, bwd1, bwd2, bwd3, feat, el, er, s, exp, ret, gradout.contiguous(), grad_feat, grad_el,
                grad_er, ctx.slope)
        grad_wl = th.zeros_like(w_l, memory_format=th.contiguous_format)
        u = feat.size(0)
        K.hetb200_backward_rgnn_relational_matmul_no_scatter_gather_list_accumulate(
            rel_ptrs_row, th.transpose(w_l, 2, 3).contiguous(), feat.view(u, -1), grad_el.view(u, -1),
            grad_feat.view(u, -1), grad_wl)
        return (None,) * 9 + (grad_feat, grad_er, grad_wl.squeeze(-1), None, None, None, None, None)


def relational_fused_gat_csr_compact_with_logits(graph, feat_compact, er_compact, attn_l, slope):
    incsr, outcsr = graph.get_in_csr(), graph.get_out_csr()
    uniq = graph.get_separate_unique_node_indices_single_sided()
    h = attn_l.size(1)
    exp = feat_compact.new_empty([graph.get_num_edges(), h, 1])
    s = feat_compact.new_empty([graph.get_num_nodes(), h])
    ret = feat_compact.new_empty([graph.get_num_nodes()] + list(feat_compact.size()[1:]))
    if RGAT_COMPACT_BACKWARD_BY_ROWS:
        bwd = (uniq["urow_dst"], uniq["urow_eids"], uniq["urow_ptrs"], uniq["urow_inverse_idx_col"])
    else:
        bwd = (outcsr["col_indices"], outcsr["eids"], uniq["outcsr_inverse_idx_row"], uniq["outcsr_inverse_idx_col"])
    return RelationalFusedGatCSRCompactWithLogits.apply(
        uniq["rel_ptrs_row"], incsr["row_ptrs"], incsr["eids"], uniq["incsr_inverse_idx_row"],
        uniq["incsr_inverse_idx_col"], bwd[0], bwd[1], bwd[2], bwd[3], feat_compact.contiguous(),
        er_compact.contiguous(), attn_l.contiguous(), s, exp, ret, slope, RGAT_COMPACT_BACKWARD_BY_ROWS)


# ----- plan-based fused edge passes (csrc/edge_plan.cu): the bench default -----------------------
USE_EDGE_PLAN = True  # False: the compact in-CSR / out-CSR walkers of csrc/edge_ops.cu (int64 indices, no max-subtraction)


# Variants of the HGT plan backward (A/B partners of the default, measured in tools/bench_edge_plan.py and
# covered by tests/test_gpu_fused_ops.py). Default: ONE walk that recomputes the raw score from the two rows it
# holds anyway; the forward saves nothing per edge.
# SAVE_HGT_SCORES: the forward stores the raw score per edge and the backward reads it back (one more random
#   access per edge: 3.99 ms against the default at mag size).
# TWO_PASS_HGT_BACKWARD: hetb_hgt_fused_bwd_plan2 — the destination-side gradient is owned by a second,
#   destination-major walk instead of one vector reduction per edge: no atomics on grad_left, run-to-run
#   identical, but 4.75 ms (needs the saved scores: it overwrites them with their gradient).
SAVE_HGT_SCORES = False
TWO_PASS_HGT_BACKWARD = False


class HGTFusedAttentionAggregationPlan(th.autograd.Function):
    """Same op as HGTFusedAttentionAggregationCSRCompact (inner product + edge softmax + message
    aggregation, compact inputs) on the graph's walk plans: int32 indices, (destination, relation)-
    sorted forward / (source, relation)-sorted backward walks, max-subtracted online softmax. Saved
    for the backward: the raw score per edge and the log-sum-exp per (destination, head)."""

    @staticmethod
    def forward(ctx, graph, left_compact, k, message_compact, mu):
        pl = graph.get_edge_plan()
        n, e, h = graph.get_num_nodes(), graph.get_num_edges(), mu.size(1)
        f = message_compact.size(2)
        sc = graph.plan_scratch(h, h * f)
        dev = mu.device
        new_h = th.empty((n, h, f), dtype=th.float32, device=dev)
        lse = th.empty((n, h), dtype=th.float32, device=dev)
        attn = th.empty((e, h) if (SAVE_HGT_SCORES or TWO_PASS_HGT_BACKWARD) else (0, h), dtype=th.float32, device=dev)
        K.hetb200_hgt_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, left_compact, k, message_compact, mu, attn, lse,
                                     new_h, sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"])
        ctx.graph = graph
        ctx.save_for_backward(left_compact, k, message_compact, mu, attn, lse, new_h)
        return new_h

    @staticmethod
    def backward(ctx, gradout):
        left, k, msg, mu, attn, lse, new_h = ctx.saved_tensors
        graph = ctx.graph
        pl = graph.get_edge_plan()
        h, f = mu.size(1), msg.size(2)
        sc = graph.plan_scratch(h, h * f)
        f32 = dict(dtype=th.float32, device=mu.device)
        grad_mu = th.zeros(mu.shape, **f32)
        grad_k = th.empty(k.shape, **f32)
        grad_msg = th.empty(msg.shape, **f32)
        if TWO_PASS_HGT_BACKWARD and attn.numel():
            # attn is this node's own buffer and is consumed here (score in, gradient of the score out)
            assert not getattr(ctx, "attn_consumed", False), "this node's backward overwrites its saved scores: no second pass"
            ctx.attn_consumed = True
            grad_left = th.empty(left.shape, **f32)
            K.hetb200_hgt_fused_bwd_plan2(pl["bwd"].blob, pl["bwd"].meta, pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu,
                                          attn, lse, new_h, gradout.contiguous(), sc["gdot"], grad_left, grad_k, grad_msg,
                                          grad_mu)
        else:
            grad_left = th.zeros(left.shape, **f32)
            K.hetb200_hgt_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, left, k, msg, mu, attn, lse, new_h,
                                         gradout.contiguous(), sc["gdot"], grad_left, grad_k, grad_msg, grad_mu)
        return None, grad_left, grad_k, grad_msg, grad_mu


def hgt_fused_attention_and_aggregation_plan(graph, left_compact, k, message_compact, mu):
    return HGTFusedAttentionAggregationPlan.apply(graph, left_compact.contiguous(), k.contiguous(),
                                                  message_compact.contiguous(), mu.contiguous())


class HGTLayerCoreBF16Tables(th.autograd.Function):
    """bf16 STORAGE mode of the HGT layer core (north star: "rtol 2e-2 in bf16"): the three typed
    linears of HGTReorderedProjections write k, W_att.q and W_msg.v as bf16 straight from the
    tcgen05 epilogue (fp32 accumulate, one rounding at the store), the plan-based edge passes gather
    the bf16 rows (half the bytes of the dominant gathers) and accumulate in fp32. Everything else —
    layer input, weights, scores, log-sum-exp, outputs and EVERY gradient — stays fp32. One autograd
    node, so the fp32 gradients of the bf16 tables never pass through autograd's dtype cast."""

    @staticmethod
    def forward(ctx, graph, h, k_linears, w_att, w_msg, mu):
        u = graph.get_separate_unique_node_indices_single_sided()
        offs = graph.get_original_node_type_offsets()
        pl = graph.get_edge_plan()
        n, e, hh = graph.get_num_nodes(), graph.get_num_edges(), mu.size(1)
        d = k_linears.size(1) * k_linears.size(3)
        f = d // hh
        dev = h.device
        none = th.empty(0, dtype=th.int64, device=dev)
        bf = dict(dtype=th.bfloat16, device=dev)
        k = th.empty((h.size(0), hh, f), **bf)
        K.hetb200_typed_linear_bf16out(offs, none, k_linears, h, k)
        left = th.empty((u["node_indices_col"].numel(), hh, f), **bf)
        K.hetb200_typed_linear_bf16out(u["rel_ptrs_col"], u["node_indices_col"], w_att, h, left)
        msg = th.empty((u["node_indices_row"].numel(), hh, f), **bf)
        K.hetb200_typed_linear_bf16out(u["rel_ptrs_row"], u["node_indices_row"], w_msg, h, msg)
        sc = graph.plan_scratch(hh, d)
        new_h = th.empty((n, hh, f), dtype=th.float32, device=dev)
        lse = th.empty((n, hh), dtype=th.float32, device=dev)
        attn = th.empty((e, hh) if (SAVE_HGT_SCORES or TWO_PASS_HGT_BACKWARD) else (0, hh), dtype=th.float32, device=dev)
        K.hetb200_hgt_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu, attn, lse, new_h, sc["part_m"],
                                     sc["part_s"], sc["part_acc"], sc["ssum"])
        ctx.graph = graph
        ctx.save_for_backward(h, k_linears, w_att, w_msg, mu, k, left, msg, attn, lse, new_h)
        return new_h

    @staticmethod
    def backward(ctx, gradout):
        h, k_linears, w_att, w_msg, mu, k, left, msg, attn, lse, new_h = ctx.saved_tensors
        graph = ctx.graph
        u = graph.get_separate_unique_node_indices_single_sided()
        offs = graph.get_original_node_type_offsets()
        pl = graph.get_edge_plan()
        hh, f = mu.size(1), msg.size(2)
        sc = graph.plan_scratch(hh, hh * f)
        f32 = dict(dtype=th.float32, device=mu.device)
        grad_mu = th.zeros(mu.shape, **f32)
        grad_k, grad_msg = th.empty(k.shape, **f32), th.empty(msg.shape, **f32)
        if TWO_PASS_HGT_BACKWARD and attn.numel():
            assert not getattr(ctx, "attn_consumed", False), "this node's backward overwrites its saved scores: no second pass"
            ctx.attn_consumed = True
            grad_left = th.empty(left.shape, **f32)
            K.hetb200_hgt_fused_bwd_plan2(pl["bwd"].blob, pl["bwd"].meta, pl["fwd"].blob, pl["fwd"].meta, left, k, msg, mu,
                                          attn, lse, new_h, gradout.contiguous(), sc["gdot"], grad_left, grad_k, grad_msg,
                                          grad_mu)
        else:
            grad_left = th.zeros(left.shape, **f32)
            K.hetb200_hgt_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, left, k, msg, mu, attn, lse, new_h,
                                         gradout.contiguous(), sc["gdot"], grad_left, grad_k, grad_msg, grad_mu)
        grad_h = th.empty_like(h, memory_format=th.contiguous_format)
        grad_wk = th.zeros_like(k_linears, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul_no_scatter_gather_list(
            offs, th.transpose(k_linears, 2, 3).contiguous(), h, grad_k.view(h.size(0), -1), grad_h, grad_wk)
        grad_watt = th.zeros_like(w_att, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_col"], "unique_srcs_and_dests_node_indices": u["node_indices_col"]},
            1, th.transpose(w_att, 2, 3).contiguous(), h, grad_left.view(grad_left.size(0), 1, -1), grad_h, grad_watt, True)
        grad_wmsg = th.zeros_like(w_msg, memory_format=th.contiguous_format)
        K.backward_rgnn_relational_matmul(
            {"unique_srcs_and_dests_rel_ptrs": u["rel_ptrs_row"], "unique_srcs_and_dests_node_indices": u["node_indices_row"]},
            1, th.transpose(w_msg, 2, 3).contiguous(), h, grad_msg.view(grad_msg.size(0), 1, -1), grad_h, grad_wmsg, True)
        return None, grad_h, grad_wk, grad_watt, grad_wmsg, grad_mu


def hgt_layer_core_bf16_tables(graph, h, k_linears, w_att, w_msg, mu):
    return HGTLayerCoreBF16Tables.apply(graph, h.contiguous(), k_linears.contiguous(), w_att.contiguous(),
                                        w_msg.contiguous(), mu.contiguous())


class RelationalFusedGatPlan(th.autograd.Function):
    """relational_fused_gat_csr on compact inputs over the walk plans; with ``attn_l`` given the
    el = <feat_compact, attn_l[r]> row-dot and its backward are part of the node (as in
    RelationalFusedGatCSRCompactWithLogits). Saves only lse [N, H] and ret: no per-edge tensor."""

    @staticmethod
    def forward(ctx, graph, feat_compact, el_compact, er_compact, attn_l, slope):
        pl = graph.get_edge_plan()
        uniq = graph.get_separate_unique_node_indices_single_sided()
        n = graph.get_num_nodes()
        u, h, f = feat_compact.size(0), feat_compact.size(1), feat_compact.size(2)
        w_l = None
        if attn_l is not None:
            el_compact = th.empty((u, h), dtype=th.float32, device=feat_compact.device)
            w_l = attn_l.unsqueeze(-1).contiguous()
            K.rgnn_relational_matmul_no_scatter_gather_list(uniq["rel_ptrs_row"], w_l, feat_compact.view(u, -1),
                                                            el_compact)
        el_compact = el_compact.view(u, h).contiguous()
        er_compact = er_compact.reshape(-1, h).contiguous()
        sc = graph.plan_scratch(h, h * f)
        ret = th.empty((n, h, f), dtype=th.float32, device=feat_compact.device)
        lse = th.empty((n, h), dtype=th.float32, device=feat_compact.device)
        K.hetb200_rgat_fused_fwd_plan(pl["fwd"].blob, pl["fwd"].meta, feat_compact, el_compact, er_compact, slope, lse,
                                      ret, sc["part_m"], sc["part_s"], sc["part_acc"], sc["ssum"])
        ctx.graph, ctx.slope, ctx.with_logits = graph, slope, attn_l is not None
        ctx.save_for_backward(feat_compact, el_compact, er_compact, lse, ret, w_l if w_l is not None else lse)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        feat, el, er, lse, ret, w_l = ctx.saved_tensors
        graph = ctx.graph
        pl = graph.get_edge_plan()
        u, h, f = feat.size(0), feat.size(1), feat.size(2)
        sc = graph.plan_scratch(h, h * f)
        grad_feat = th.empty_like(feat)
        grad_el = th.empty_like(el)
        grad_er = th.zeros_like(er)
        K.hetb200_rgat_fused_bwd_plan(pl["bwd"].blob, pl["bwd"].meta, feat, el, er, lse, ret, gradout.contiguous(),
                                      sc["gdot"], grad_feat, grad_el, grad_er, ctx.slope)
        if not ctx.with_logits:
            return None, grad_feat, grad_el.view(u, h, 1), grad_er.view(-1, h, 1), None, None
        uniq = graph.get_separate_unique_node_indices_single_sided()
        grad_wl = th.zeros_like(w_l)
        K.hetb200_backward_rgnn_relational_matmul_no_scatter_gather_list_accumulate(
            uniq["rel_ptrs_row"], th.transpose(w_l, 2, 3).contiguous(), feat.view(u, -1), grad_el.view(u, -1),
            grad_feat.view(u, -1), grad_wl)
        return None, grad_feat, None, grad_er.view(-1, h, 1), grad_wl.squeeze(-1), None


def relational_fused_gat_plan(graph, feat_compact, el_compact, er_compact, slope):
    return RelationalFusedGatPlan.apply(graph, feat_compact.contiguous(), el_compact.contiguous(),
                                        er_compact.contiguous(), None, slope)


def relational_fused_gat_plan_with_logits(graph, feat_compact, er_compact, attn_l, slope):
    return RelationalFusedGatPlan.apply(graph, feat_compact.contiguous(), None, er_compact.contiguous(),
                                        attn_l.contiguous(), slope)
