// torch_hrt operator surface: registers the SAME op names and argument orders
// that the reference registers from hrt/include/DGLHackKernel/OpExport/*.inc.h
// (TORCH_LIBRARY_FRAGMENT(torch_hrt, ...)), and forwards every call to the
// C ABI of libhetb200.so (include/hetb200.h). No math lives here: this file
// only validates tensors, unpacks pointers/sizes and picks the current stream.
//
// There is deliberately no CPU implementation: every compute op requires CUDA
// tensors and fails loudly otherwise.
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include <cstdio>
#include <string>
#include <vector>

#include "hetb200.h"

namespace {

using at::Tensor;
using TensorDict = c10::Dict<std::string, at::Tensor>;

void *stream_of(const Tensor &t) {
  return (void *)c10::cuda::getCurrentCUDAStream(t.device().index()).stream();
}

void check_rc(int rc, const char *op) {
  if (rc == 0) return;
  const char *why = rc == HETB_EINVAL          ? "invalid argument"
                    : rc == HETB_EUNSUPPORTED ? "shape/kind not supported by hetb200"
                    : rc == HETB_ENOMEM       ? "out of scratch memory"
                                              : "CUDA launch error";
  TORCH_CHECK(false, "torch_hrt::", op, ": ", why, " (code ", rc, ")");
}

const float *f32(const Tensor &t, const char *op, const char *name) {
  TORCH_CHECK(t.is_cuda(), "torch_hrt::", op, ": `", name,
              "` must be a CUDA tensor (hetb200 has no CPU path)");
  TORCH_CHECK(t.scalar_type() == at::kFloat, "torch_hrt::", op, ": `", name,
              "` must be float32");
  TORCH_CHECK(t.is_contiguous(), "torch_hrt::", op, ": `", name,
              "` must be contiguous");
  return t.data_ptr<float>();
}
float *f32w(Tensor &t, const char *op, const char *name) {
  return const_cast<float *>(f32(t, op, name));
}
const int64_t *i64(const Tensor &t, const char *op, const char *name) {
  TORCH_CHECK(t.is_cuda(), "torch_hrt::", op, ": `", name,
              "` must be a CUDA tensor (hetb200 has no CPU path)");
  TORCH_CHECK(t.scalar_type() == at::kLong, "torch_hrt::", op, ": `", name,
              "` must be int64");
  TORCH_CHECK(t.is_contiguous(), "torch_hrt::", op, ": `", name,
              "` must be contiguous");
  return t.data_ptr<int64_t>();
}
// product of all dims after the first (the reference's SeastarComputeXLength)
int64_t xlen(const Tensor &t) {
  int64_t r = 1;
  for (int64_t d = 1; d < t.dim(); ++d) r *= t.size(d);
  return r;
}
void not_implemented(const char *op) {
  TORCH_CHECK(false, "torch_hrt::", op,
              " is registered for interface compatibility but is outside the "
              "hot path implemented by hetb200");
}

Tensor compact_index_by_position(const char *op, const Tensor &node, const int64_t *rel_of_pos, const int64_t *rel_ptrs,
                                 int64_t num_rels, const Tensor &unique_rel_ptrs, const Tensor &unique_nodes);

// row id of every CSR position (CSR -> COO row expansion)
Tensor expand_rows(const char *op, const Tensor &row_ptr, int64_t num_edges) {
  Tensor rows = at::empty({num_edges}, row_ptr.options());
  check_rc(hetb_expand_row_ptr(i64(row_ptr, op, "row_ptr"), row_ptr.numel() - 1, num_edges, rows.data_ptr<int64_t>(),
                               stream_of(row_ptr)),
           op);
  return rows;
}

// ---------------------------------------------------------------- misc ---
void build_debug_info() { std::printf("%s\n", hetb_build_info()); }
int64_t hetb200_launch_count() { return hetb_launch_count(); }

// ------------------------------------------------------------ RGNN ops ---
// OpExport/RGNNOps.inc.h:238-295
void rgnn_relational_matmul(TensorDict args, int64_t kind, Tensor &weights,
                            Tensor &node_feat, Tensor &ret,
                            bool input_num_head_one) {
  const char *op = "rgnn_relational_matmul";
  TORCH_CHECK(weights.dim() == 4, op, ": weights must be [R,H,in,out/H]");
  const int H = (int)weights.size(1), K = (int)weights.size(2),
            N = (int)weights.size(3);
  const int a_heads = input_num_head_one ? 1 : H;
  c10::cuda::CUDAGuard guard(ret.device());
  if (kind == 0) {
    Tensor rel_ptrs = args.at("separate_coo_rel_ptrs");
    Tensor gather = args.at("separate_coo_node_indices");
    Tensor eids = args.at("separate_coo_eids");
    check_rc(hetb_segment_mm(f32(node_feat, op, "node_feat"),
                             i64(gather, op, "separate_coo_node_indices"),
                             a_heads, f32(weights, op, "weights"),
                             f32w(ret, op, "ret"),
                             i64(eids, op, "separate_coo_eids"), H,
                             i64(rel_ptrs, op, "separate_coo_rel_ptrs"),
                             rel_ptrs.numel() - 1, eids.numel(), H, K, N, 0,
                             stream_of(ret)),
             op);
  } else if (kind == 1) {  // CompactAsOfNodeKind::Enabled: rows are (rel, unique node)
    Tensor rel_ptrs = args.at("unique_srcs_and_dests_rel_ptrs");
    Tensor nodes = args.at("unique_srcs_and_dests_node_indices");
    check_rc(hetb_segment_mm(f32(node_feat, op, "node_feat"),
                             i64(nodes, op, "unique_srcs_and_dests_node_indices"),
                             a_heads, f32(weights, op, "weights"),
                             f32w(ret, op, "ret"), nullptr, H,
                             i64(rel_ptrs, op, "unique_srcs_and_dests_rel_ptrs"),
                             rel_ptrs.numel() - 1, nodes.numel(), H, K, N, 0,
                             stream_of(ret)),
             op);
  } else {
    check_rc(HETB_EUNSUPPORTED, op);
  }
}

// OpExport/RGNNOps.inc.h:946-1010. weights_transposed is [R,H,out/H,in].
void backward_rgnn_relational_matmul(TensorDict args, int64_t kind,
                                     Tensor &weights_transposed,
                                     Tensor &node_feat, Tensor &gradout,
                                     Tensor &grad_node_feat,
                                     Tensor &grad_weights,
                                     bool input_num_head_one) {
  const char *op = "backward_rgnn_relational_matmul";
  TORCH_CHECK(weights_transposed.dim() == 4, op, ": weights must be 4-D");
  const int H = (int)weights_transposed.size(1),
            N = (int)weights_transposed.size(2),
            K = (int)weights_transposed.size(3);
  const int a_heads = input_num_head_one ? 1 : H;
  c10::cuda::CUDAGuard guard(gradout.device());
  void *st = stream_of(gradout);
  const int64_t *rel_ptrs, *gather, *scatter;
  int64_t R, M;
  bool identical = false;
  if (kind == 0) {
    Tensor t_rel = args.at("separate_coo_rel_ptrs");
    Tensor t_gather = args.at("separate_coo_node_indices");
    Tensor t_eids = args.at("separate_coo_eids");
    rel_ptrs = i64(t_rel, op, "separate_coo_rel_ptrs");
    gather = i64(t_gather, op, "separate_coo_node_indices");
    scatter = i64(t_eids, op, "separate_coo_eids");
    R = t_rel.numel() - 1;
    M = t_eids.numel();
    // same pointer-equality dispatch as the reference (RGNNOps.inc.h:253):
    // the gather list is the edge-id permutation, so no output row repeats.
    identical = (gather == scatter);
  } else if (kind == 1) {
    Tensor t_rel = args.at("unique_srcs_and_dests_rel_ptrs");
    Tensor t_nodes = args.at("unique_srcs_and_dests_node_indices");
    rel_ptrs = i64(t_rel, op, "unique_srcs_and_dests_rel_ptrs");
    gather = i64(t_nodes, op, "unique_srcs_and_dests_node_indices");
    scatter = nullptr;
    R = t_rel.numel() - 1;
    M = t_nodes.numel();
  } else {
    check_rc(HETB_EUNSUPPORTED, op);
    return;
  }
  const int accumulate = (identical && a_heads == H) ? 0 : 1;
  // hetb200 extension: a zero-element grad_node_feat / grad_weights skips that half, so a caller can put
  // other work (a collective on the input gradient) between the two.
  // dX[gather[i]] += dY[scatter[i]] · W^T[r]
  if (grad_node_feat.numel() > 0)
    check_rc(hetb_segment_mm(f32(gradout, op, "gradout"), scatter, H,
                             f32(weights_transposed, op, "weights_transposed"),
                             f32w(grad_node_feat, op, "grad_node_feat"), gather,
                             a_heads, rel_ptrs, R, M, H, N, K, accumulate, st),
             op);
  // dW[r] += X[gather[i]]^T · dY[scatter[i]]
  if (grad_weights.numel() > 0)
    check_rc(hetb_segment_mm_grad_w(f32(node_feat, op, "node_feat"), gather,
                                    a_heads, f32(gradout, op, "gradout"), scatter,
                                    H, f32w(grad_weights, op, "grad_weights"),
                                    rel_ptrs, R, M, H, K, N, st),
             op);
}

// OpExport/RGNNOps.inc.h:83-88
void rgnn_relational_matmul_no_scatter_gather_list(Tensor &ntype_offset_ptrs,
                                                   Tensor &weights,
                                                   Tensor &inputs,
                                                   Tensor &ret) {
  const char *op = "rgnn_relational_matmul_no_scatter_gather_list";
  TORCH_CHECK(weights.dim() == 4, op, ": weights must be [T,H,in,out]");
  const int H = (int)weights.size(1), K = (int)weights.size(2),
            N = (int)weights.size(3);
  const int64_t M = inputs.size(0);
  const int a_heads = (xlen(inputs) == (int64_t)H * K) ? H : 1;
  TORCH_CHECK(xlen(inputs) == (int64_t)a_heads * K, op, ": inputs row length ",
              xlen(inputs), " does not match weights");
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_segment_mm(f32(inputs, op, "inputs"), nullptr, a_heads,
                           f32(weights, op, "weights"), f32w(ret, op, "ret"),
                           nullptr, H, i64(ntype_offset_ptrs, op, "ntype_offset_ptrs"),
                           ntype_offset_ptrs.numel() - 1, M, H, K, N, 0,
                           stream_of(ret)),
           op);
}

// OpExport/RGNNOps.inc.h:744-753
static void backward_no_scatter_gather_impl(const char *op, bool force_accumulate,
    
    Tensor &ntype_offset_ptrs, Tensor &weights_transposed,
    Tensor &node_feat_input, Tensor &grad_node_feat_output,
    Tensor &grad_node_feat_input, Tensor &grad_weights) {
  TORCH_CHECK(weights_transposed.dim() == 4, op, ": weights must be 4-D");
  const int H = (int)weights_transposed.size(1),
            N = (int)weights_transposed.size(2),
            K = (int)weights_transposed.size(3);
  const int64_t M = node_feat_input.size(0);
  const int a_heads = (xlen(node_feat_input) == (int64_t)H * K) ? H : 1;
  c10::cuda::CUDAGuard guard(grad_node_feat_output.device());
  void *st = stream_of(grad_node_feat_output);
  const int64_t *ptrs = i64(ntype_offset_ptrs, op, "ntype_offset_ptrs");
  const int64_t T = ntype_offset_ptrs.numel() - 1;
  const int accumulate = (force_accumulate || (a_heads == 1 && H > 1)) ? 1 : 0;
  // a zero-element grad_node_feat_input / grad_weights skips that half (hetb200 extension, see above)
  if (grad_node_feat_input.numel() > 0)
    check_rc(hetb_segment_mm(f32(grad_node_feat_output, op, "grad_node_feat_output"),
                             nullptr, H,
                             f32(weights_transposed, op, "weights_transposed"),
                             f32w(grad_node_feat_input, op, "grad_node_feat_input"),
                             nullptr, a_heads, ptrs, T, M, H, N, K, accumulate, st),
             op);
  if (grad_weights.numel() > 0)
    check_rc(hetb_segment_mm_grad_w(
                 f32(node_feat_input, op, "node_feat_input"), nullptr, a_heads,
                 f32(grad_node_feat_output, op, "grad_node_feat_output"), nullptr,
                 H, f32w(grad_weights, op, "grad_weights"), ptrs, T, M, H, K, N, st),
             op);
}

void backward_rgnn_relational_matmul_no_scatter_gather_list(
    Tensor &ntype_offset_ptrs, Tensor &weights_transposed, Tensor &node_feat_input,
    Tensor &grad_node_feat_output, Tensor &grad_node_feat_input, Tensor &grad_weights) {
  backward_no_scatter_gather_impl("backward_rgnn_relational_matmul_no_scatter_gather_list", false, ntype_offset_ptrs,
                                  weights_transposed, node_feat_input, grad_node_feat_output, grad_node_feat_input,
                                  grad_weights);
}

// hetb200 extension: the same backward, but the input gradient is ADDED to grad_node_feat_input
// (which already holds the gradient that reached the same tensor through another consumer).
void hetb200_backward_rgnn_relational_matmul_no_scatter_gather_list_accumulate(
    Tensor &ntype_offset_ptrs, Tensor &weights_transposed, Tensor &node_feat_input,
    Tensor &grad_node_feat_output, Tensor &grad_node_feat_input, Tensor &grad_weights) {
  backward_no_scatter_gather_impl("hetb200_backward_rgnn_relational_matmul_no_scatter_gather_list_accumulate", true,
                                  ntype_offset_ptrs, weights_transposed, node_feat_input, grad_node_feat_output,
                                  grad_node_feat_input, grad_weights);
}

// OpExport/RGNNOps.inc.h:609-658
void rgnn_inner_product_right_node_separatecoo(
    TensorDict args, int64_t kind, Tensor &rel_ptrs, Tensor &eids,
    Tensor &row_indices, Tensor &col_indices, Tensor &left_side_data,
    Tensor &right_node_vectors, Tensor &edge_inner_product) {
  const char *op = "rgnn_inner_product_right_node_separatecoo";
  // kind 0: per-edge left data; kind 2 (EnabledWithDirectIndexing): compact left rows addressed
  // through args["edata_idx_to_inverse_idx"]; kind 1 (Enabled): the compact row of (relation, col
  // node) found by binary search in the unique lists (InnerProductEdgeParallel.cu.h:36-56)
  if (kind != 0 && kind != 1 && kind != 2) check_rc(HETB_EUNSUPPORTED, op);
  const int64_t *left_idx = nullptr;
  Tensor searched;
  if (kind == 2) {
    TORCH_CHECK(args.contains("edata_idx_to_inverse_idx"), op, ": kind 2 needs args[edata_idx_to_inverse_idx]");
    left_idx = i64(args.at("edata_idx_to_inverse_idx"), op, "edata_idx_to_inverse_idx");
  } else if (kind == 1) {
    c10::cuda::CUDAGuard g1(edge_inner_product.device());
    Tensor by_pos = compact_index_by_position(op, col_indices, nullptr, i64(rel_ptrs, op, "rel_ptrs"),
                                              rel_ptrs.numel() - 1, args.at("unique_srcs_and_dests_rel_ptrs"),
                                              args.at("unique_srcs_and_dests_node_indices"));
    searched = at::empty_like(by_pos);
    searched.index_put_({eids}, by_pos);  // the kernel addresses the map by edge id
    left_idx = searched.data_ptr<int64_t>();
  }
  TORCH_CHECK(right_node_vectors.dim() == 3, op,
              ": right_node_vectors must be [N,H,F]");
  const int H = (int)right_node_vectors.size(1), F = (int)right_node_vectors.size(2);
  c10::cuda::CUDAGuard guard(edge_inner_product.device());
  check_rc(hetb_inner_product_fwd(
               i64(eids, op, "eids"), i64(row_indices, op, "row_indices"), left_idx,
               eids.numel(), f32(left_side_data, op, "left_side_data"),
               f32(right_node_vectors, op, "right_node_vectors"),
               f32w(edge_inner_product, op, "edge_inner_product"), H, F,
               stream_of(edge_inner_product)),
           op);
}

// OpExport/RGNNOps.inc.h:1131-1181
void backward_inner_product_right_node_separatecoo(
    TensorDict args, int64_t kind, Tensor &rel_ptrs, Tensor &eids,
    Tensor &row_indices, Tensor &col_indices, Tensor &left_side_data,
    Tensor &right_node_vectors, Tensor &grad_inner_product,
    Tensor &grad_left_side_data, Tensor &grad_right_node_vectors) {
  const char *op = "backward_inner_product_right_node_separatecoo";
  if (kind != 0 && kind != 1 && kind != 2) check_rc(HETB_EUNSUPPORTED, op);
  const int64_t *left_idx = nullptr;
  Tensor searched;
  if (kind == 2) {
    TORCH_CHECK(args.contains("edata_idx_to_inverse_idx"), op, ": kind 2 needs args[edata_idx_to_inverse_idx]");
    left_idx = i64(args.at("edata_idx_to_inverse_idx"), op, "edata_idx_to_inverse_idx");
  } else if (kind == 1) {
    c10::cuda::CUDAGuard g1(grad_inner_product.device());
    Tensor by_pos = compact_index_by_position(op, col_indices, nullptr, i64(rel_ptrs, op, "rel_ptrs"),
                                              rel_ptrs.numel() - 1, args.at("unique_srcs_and_dests_rel_ptrs"),
                                              args.at("unique_srcs_and_dests_node_indices"));
    searched = at::empty_like(by_pos);
    searched.index_put_({eids}, by_pos);
    left_idx = searched.data_ptr<int64_t>();
  }
  const int H = (int)right_node_vectors.size(1), F = (int)right_node_vectors.size(2);
  c10::cuda::CUDAGuard guard(grad_inner_product.device());
  check_rc(hetb_inner_product_bwd(
               i64(eids, op, "eids"), i64(row_indices, op, "row_indices"), left_idx,
               eids.numel(), f32(left_side_data, op, "left_side_data"),
               f32(right_node_vectors, op, "right_node_vectors"),
               f32(grad_inner_product, op, "grad_inner_product"),
               f32w(grad_left_side_data, op, "grad_left_side_data"),
               f32w(grad_right_node_vectors, op, "grad_right_node_vectors"), H,
               F, stream_of(grad_inner_product)),
           op);
}

// ------------------------------------------------------------ RGAT ops ---
// OpExport/RGATOps.inc.h:251-277
void relational_fused_gat_csr(Tensor &incsr_row_ptr, Tensor &incsr_col_indices,
                              Tensor &incsr_eids, Tensor &incsr_reltypes,
                              Tensor &unique_rel_ptrs,
                              Tensor &unique_node_indices, Tensor &feat_src,
                              Tensor &el, Tensor &er, Tensor &sum, Tensor &exp,
                              Tensor &ret, double slope,
                              bool compact_as_of_node) {
  const char *op = "relational_fused_gat_csr";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(ret.device());
  // CompactAsOfNodeFlag (CompactAsOfNodeKind::Enabled, RGATOps.inc.h:251-277): feat_src / el / er hold
  // one row per (relation, unique node) of the double-sided list; the row of a position is found by
  // binary search (kernel_enums.h:96-119), here once per call into position-aligned index arrays.
  hetb_compact_idx ci = {nullptr, nullptr, nullptr};
  Tensor src_idx, dst_idx;
  if (compact_as_of_node) {
    Tensor dst = expand_rows(op, incsr_row_ptr, incsr_eids.numel());
    src_idx = compact_index_by_position(op, incsr_col_indices, i64(incsr_reltypes, op, "incsr_reltypes"), nullptr, 0,
                                        unique_rel_ptrs, unique_node_indices);
    dst_idx = compact_index_by_position(op, dst, i64(incsr_reltypes, op, "incsr_reltypes"), nullptr, 0,
                                        unique_rel_ptrs, unique_node_indices);
    ci.feat_idx = ci.el_idx = src_idx.data_ptr<int64_t>();
    ci.er_idx = dst_idx.data_ptr<int64_t>();
  }
  check_rc(hetb_rgat_fwd_csr(i64(incsr_row_ptr, op, "incsr_row_ptr"),
                             i64(incsr_eids, op, "incsr_eids"),
                             incsr_row_ptr.numel() - 1, incsr_eids.numel(),
                             f32(feat_src, op, "feat_src"), f32(el, op, "el"),
                             f32(er, op, "er"), f32w(sum, op, "sum"),
                             f32w(exp, op, "exp"), f32w(ret, op, "ret"), H, F,
                             (float)slope, compact_as_of_node ? &ci : nullptr, stream_of(ret)),
           op);
}

// hetb200 extension: relational_fused_gat_csr with compact-as-of-node inputs addressed by
// direct indexing (the single-sided lists of RGAT/models.py:152-263, CompactAsOfNodeKind 4):
// feat_src / el hold one row per unique (relation, source), er one per unique (relation,
// destination); inverse_row / inverse_col map an edge id to its row. exp stays per edge.
void hetb200_relational_fused_gat_csr_compact(
    Tensor &incsr_row_ptr, Tensor &incsr_eids, Tensor &inverse_row, Tensor &inverse_col,
    Tensor &feat_src, Tensor &el, Tensor &er, Tensor &sum, Tensor &exp, Tensor &ret, double slope) {
  const char *op = "hetb200_relational_fused_gat_csr_compact";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(ret.device());
  const int64_t *ir = i64(inverse_row, op, "inverse_row"), *ic = i64(inverse_col, op, "inverse_col");
  const hetb_compact_idx ci = {ir, ir, ic};
  check_rc(hetb_rgat_fwd_csr(i64(incsr_row_ptr, op, "incsr_row_ptr"), i64(incsr_eids, op, "incsr_eids"),
                             incsr_row_ptr.numel() - 1, incsr_eids.numel(), f32(feat_src, op, "feat_src"),
                             f32(el, op, "el"), f32(er, op, "er"), f32w(sum, op, "sum"), f32w(exp, op, "exp"),
                             f32w(ret, op, "ret"), H, F, (float)slope, &ci, stream_of(ret)),
           op);
}

// grad_feat_src / grad_el / grad_er are compact and must arrive zero-filled (summed into).
void hetb200_backward_relational_fused_gat_csr_compact(
    Tensor &outcsr_col_indices, Tensor &outcsr_eids, Tensor &inverse_row, Tensor &inverse_col,
    Tensor &feat_src, Tensor &el, Tensor &er, Tensor &sum, Tensor &exp, Tensor &ret, Tensor &gradout,
    Tensor &grad_feat_src, Tensor &grad_el, Tensor &grad_er, double slope) {
  const char *op = "hetb200_backward_relational_fused_gat_csr_compact";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  const int64_t *ir = i64(inverse_row, op, "inverse_row"), *ic = i64(inverse_col, op, "inverse_col");
  const hetb_compact_idx ci = {ir, ir, ic};
  check_rc(hetb_rgat_bwd(i64(outcsr_col_indices, op, "outcsr_col_indices"), i64(outcsr_eids, op, "outcsr_eids"),
                         outcsr_eids.numel(), f32(feat_src, op, "feat_src"), f32(el, op, "el"), f32(er, op, "er"),
                         f32(sum, op, "sum"), f32(exp, op, "exp"), f32(ret, op, "ret"), f32(gradout, op, "gradout"),
                         f32w(grad_feat_src, op, "grad_feat_src"), f32w(grad_el, op, "grad_el"),
                         f32w(grad_er, op, "grad_er"), H, F, (float)slope, &ci, stream_of(gradout)),
           op);
}

// The same backward walking the edges grouped by compact (relation, source) row
// (HetGraph: urow_ptrs / urow_dst / urow_eids / urow_inverse_idx_col).
void hetb200_backward_relational_fused_gat_compact_rows(
    Tensor &urow_ptrs, Tensor &urow_dst, Tensor &urow_eids, Tensor &urow_inverse_col,
    Tensor &feat_src, Tensor &el, Tensor &er, Tensor &sum, Tensor &exp, Tensor &ret, Tensor &gradout,
    Tensor &grad_feat_src, Tensor &grad_el, Tensor &grad_er, double slope) {
  const char *op = "hetb200_backward_relational_fused_gat_compact_rows";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  Tensor scratch = at::empty({sum.numel()}, sum.options());
  check_rc(hetb_rgat_bwd_compact_rows(
               i64(urow_ptrs, op, "urow_ptrs"), i64(urow_dst, op, "urow_dst"), i64(urow_eids, op, "urow_eids"),
               i64(urow_inverse_col, op, "urow_inverse_col"), urow_ptrs.numel() - 1, urow_eids.numel(),
               sum.numel() / H, f32(feat_src, op, "feat_src"), f32(el, op, "el"), f32(er, op, "er"),
               f32(sum, op, "sum"), f32(exp, op, "exp"), f32(ret, op, "ret"), f32(gradout, op, "gradout"),
               f32w(scratch, op, "scratch"), f32w(grad_feat_src, op, "grad_feat_src"), f32w(grad_el, op, "grad_el"),
               f32w(grad_er, op, "grad_er"), H, F, (float)slope, stream_of(gradout)),
           op);
}

// OpExport/RGATOps.inc.h:430-461
void backward_relational_fused_gat_csr(
    Tensor &outcsr_row_ptr, Tensor &outcsr_col_indices, Tensor &outcsr_eids,
    Tensor &outcsr_reltypes, Tensor &unique_rel_ptrs,
    Tensor &unique_node_indices, Tensor &feat_src, Tensor &el, Tensor &er,
    Tensor &sum, Tensor &exp, Tensor &ret, Tensor &gradout,
    Tensor &grad_feat_src, Tensor &grad_el, Tensor &grad_er, double slope,
    bool compact_as_of_node) {
  const char *op = "backward_relational_fused_gat_csr";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  // CompactAsOfNodeFlag: out-CSR rows are sources; the gradients of compact rows are summed over
  // their edges, so grad_feat_src / grad_el / grad_er must arrive zero-filled (as the reference's do)
  hetb_compact_idx ci = {nullptr, nullptr, nullptr};
  Tensor src_idx, dst_idx;
  if (compact_as_of_node) {
    Tensor src = expand_rows(op, outcsr_row_ptr, outcsr_eids.numel());
    src_idx = compact_index_by_position(op, src, i64(outcsr_reltypes, op, "outcsr_reltypes"), nullptr, 0,
                                        unique_rel_ptrs, unique_node_indices);
    dst_idx = compact_index_by_position(op, outcsr_col_indices, i64(outcsr_reltypes, op, "outcsr_reltypes"), nullptr, 0,
                                        unique_rel_ptrs, unique_node_indices);
    ci.feat_idx = ci.el_idx = src_idx.data_ptr<int64_t>();
    ci.er_idx = dst_idx.data_ptr<int64_t>();
  }
  check_rc(hetb_rgat_bwd(i64(outcsr_col_indices, op, "outcsr_col_indices"),
                         i64(outcsr_eids, op, "outcsr_eids"),
                         outcsr_eids.numel(), f32(feat_src, op, "feat_src"),
                         f32(el, op, "el"), f32(er, op, "er"),
                         f32(sum, op, "sum"), f32(exp, op, "exp"),
                         f32(ret, op, "ret"), f32(gradout, op, "gradout"),
                         f32w(grad_feat_src, op, "grad_feat_src"),
                         f32w(grad_el, op, "grad_el"),
                         f32w(grad_er, op, "grad_er"), H, F, (float)slope, compact_as_of_node ? &ci : nullptr,
                         stream_of(gradout)),
           op);
}

// Compact rows of every COO position for the compact kinds of the separate-COO RGAT ops
// (RGATOps.inc.h:170-245 dict keys): kind 1 one double-sided list, binary search; kind 3 dual lists
// (row / col), binary search; kind 4 dual direct maps by edge id; kind 2 one direct map by edge id
// (the reference then addresses source AND destination rows with it: kernel_enums.h:116-118).
struct CooCompact {
  Tensor src, dst;
  hetb_compact_idx ci{nullptr, nullptr, nullptr};
};
CooCompact coo_compact_indices(const char *op, int64_t kind, TensorDict &args, Tensor &eids, Tensor &rel_ptrs,
                               Tensor &row_indices, Tensor &col_indices) {
  CooCompact c;
  if (kind == 0) return c;
  const int64_t *rp = i64(rel_ptrs, op, "rel_ptrs");
  const int64_t R = rel_ptrs.numel() - 1;
  if (kind == 1) {
    Tensor up = args.at("unique_srcs_and_dests_rel_ptrs"), un = args.at("unique_srcs_and_dests_node_indices");
    c.src = compact_index_by_position(op, row_indices, nullptr, rp, R, up, un);
    c.dst = compact_index_by_position(op, col_indices, nullptr, rp, R, up, un);
  } else if (kind == 3) {
    c.src = compact_index_by_position(op, row_indices, nullptr, rp, R, args.at("unique_srcs_and_dests_rel_ptrs"),
                                      args.at("unique_srcs_and_dests_node_indices_row"));
    c.dst = compact_index_by_position(op, col_indices, nullptr, rp, R, args.at("unique_srcs_and_dests_rel_ptrs_col"),
                                      args.at("unique_srcs_and_dests_node_indices_col"));
  } else if (kind == 4) {
    c.src = args.at("edata_idx_to_inverse_idx_row").index({eids}).contiguous();
    c.dst = args.at("edata_idx_to_inverse_idx_col").index({eids}).contiguous();
  } else if (kind == 2) {
    c.src = args.at("edata_idx_to_inverse_idx").index({eids}).contiguous();
    c.dst = c.src;
  } else {
    check_rc(HETB_EUNSUPPORTED, op);
  }
  c.ci.feat_idx = c.ci.el_idx = i64(c.src, op, "compact source index");
  c.ci.er_idx = i64(c.dst, op, "compact destination index");
  return c;
}

// OpExport/RGATOps.inc.h:170-245
void relational_fused_gat_separate_coo(Tensor &eids, Tensor &rel_ptrs,
                                       Tensor &row_indices, Tensor &col_indices,
                                       int64_t kind, TensorDict args,
                                       Tensor &feat_src, Tensor &el, Tensor &er,
                                       Tensor &sum, Tensor &exp, Tensor &ret,
                                       double slope) {
  const char *op = "relational_fused_gat_separate_coo";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(ret.device());
  CooCompact cc = coo_compact_indices(op, kind, args, eids, rel_ptrs, row_indices, col_indices);
  check_rc(hetb_rgat_fwd_coo(i64(col_indices, op, "col_indices"),
                             i64(eids, op, "eids"), ret.size(0), eids.numel(),
                             f32(feat_src, op, "feat_src"), f32(el, op, "el"),
                             f32(er, op, "er"), f32w(sum, op, "sum"),
                             f32w(exp, op, "exp"), f32w(ret, op, "ret"), H, F,
                             (float)slope, kind ? &cc.ci : nullptr, stream_of(ret)),
           op);
}

// OpExport/RGATOps.inc.h:465-551
void backward_relational_fused_gat_separate_coo(
    Tensor &eids, Tensor &rel_ptrs, Tensor &row_indices, Tensor &col_indices,
    int64_t kind, TensorDict args, Tensor &feat_src, Tensor &el, Tensor &er,
    Tensor &sum, Tensor &exp, Tensor &ret, Tensor &gradout,
    Tensor &grad_feat_src, Tensor &grad_el, Tensor &grad_er, double slope) {
  const char *op = "backward_relational_fused_gat_separate_coo";
  const int H = (int)xlen(el);
  const int F = (int)(xlen(feat_src) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  // compact kinds: the gradients of compact rows are summed, grad_* must arrive zero-filled
  CooCompact cc = coo_compact_indices(op, kind, args, eids, rel_ptrs, row_indices, col_indices);
  check_rc(hetb_rgat_bwd(i64(col_indices, op, "col_indices"),
                         i64(eids, op, "eids"), eids.numel(),
                         f32(feat_src, op, "feat_src"), f32(el, op, "el"),
                         f32(er, op, "er"), f32(sum, op, "sum"),
                         f32(exp, op, "exp"), f32(ret, op, "ret"),
                         f32(gradout, op, "gradout"),
                         f32w(grad_feat_src, op, "grad_feat_src"),
                         f32w(grad_el, op, "grad_el"),
                         f32w(grad_er, op, "grad_er"), H, F, (float)slope, kind ? &cc.ci : nullptr,
                         stream_of(gradout)),
           op);
}

// ------------------------------------------------------------- HGT ops ---
// OpExport/HGTOps.inc.h:190-203
void hgt_full_graph_edge_softmax_ops_csr(
    Tensor &row_ptr, Tensor &col_indices, Tensor &eids, Tensor &reltypes,
    Tensor &unnormalized_attn_score, Tensor &mu,
    Tensor &edgesoftmax_sum_per_node,
    Tensor &mu_softmax_applied_unnormalized_attn_score,
    Tensor &normalized_attn_score) {
  const char *op = "hgt_full_graph_edge_softmax_ops_csr";
  (void)col_indices;
  const int H = (int)xlen(unnormalized_attn_score);
  c10::cuda::CUDAGuard guard(normalized_attn_score.device());
  check_rc(hetb_hgt_edge_softmax_fwd_csr(
               i64(row_ptr, op, "row_ptr"), i64(eids, op, "eids"),
               i64(reltypes, op, "reltypes"), row_ptr.numel() - 1, eids.numel(),
               f32(unnormalized_attn_score, op, "unnormalized_attn_score"),
               f32(mu, op, "mu"),
               f32w(edgesoftmax_sum_per_node, op, "edgesoftmax_sum_per_node"),
               f32w(mu_softmax_applied_unnormalized_attn_score, op,
                    "mu_softmax_applied_unnormalized_attn_score"),
               f32w(normalized_attn_score, op, "normalized_attn_score"), H,
               stream_of(normalized_attn_score)),
           op);
}

// OpExport/HGTOps.inc.h:180-188 (reltypes before eids, as in the reference).
// Like the reference kernel (HGTForwardKernels.cu.h:304-309) this divides
// edge_attn_score by edgesoftmax_sum_per_node: pass the UN-normalised
// exp(mu*score) tensor to get a softmax-weighted sum.
void hgt_full_graph_message_mean_aggregation_csr(
    Tensor &incsr_row_ptrs, Tensor &incsr_col_indices, Tensor &incsr_reltypes,
    Tensor &incsr_eids, Tensor &edge_messages, Tensor &edge_attn_score,
    Tensor &edgesoftmax_sum_per_node, Tensor &mu, Tensor &ret) {
  const char *op = "hgt_full_graph_message_mean_aggregation_csr";
  (void)incsr_col_indices; (void)incsr_reltypes; (void)mu;
  const int H = (int)xlen(edge_attn_score);
  const int F = (int)(xlen(edge_messages) / H);
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_hgt_aggregate_fwd_csr(
               i64(incsr_row_ptrs, op, "incsr_row_ptrs"),
               i64(incsr_eids, op, "incsr_eids"), incsr_row_ptrs.numel() - 1,
               incsr_eids.numel(), f32(edge_messages, op, "edge_messages"),
               f32(edge_attn_score, op, "edge_attn_score"),
               f32(edgesoftmax_sum_per_node, op, "edgesoftmax_sum_per_node"),
               f32w(ret, op, "ret"), H, F, 1, nullptr, stream_of(ret)),
           op);
}

// hetb200 extension: the same aggregation with edge_messages holding one row per unique
// (relation, source) pair (compact-as-of-node, direct indexing: msg_idx[eid] is the row).
void hetb200_hgt_message_aggregation_csr_compact(
    Tensor &incsr_row_ptrs, Tensor &incsr_eids, Tensor &msg_idx, Tensor &messages_compact,
    Tensor &edge_attn_score, Tensor &edgesoftmax_sum_per_node, Tensor &ret) {
  const char *op = "hetb200_hgt_message_aggregation_csr_compact";
  const int H = (int)xlen(edge_attn_score);
  const int F = (int)(xlen(messages_compact) / H);
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_hgt_aggregate_fwd_csr(
               i64(incsr_row_ptrs, op, "incsr_row_ptrs"), i64(incsr_eids, op, "incsr_eids"),
               incsr_row_ptrs.numel() - 1, incsr_eids.numel(), f32(messages_compact, op, "messages_compact"),
               f32(edge_attn_score, op, "edge_attn_score"),
               f32(edgesoftmax_sum_per_node, op, "edgesoftmax_sum_per_node"), f32w(ret, op, "ret"), H, F, 1,
               i64(msg_idx, op, "msg_idx"), stream_of(ret)),
           op);
}

// ---------------------------------------------------------------------------
// RGCN. OpExport/RGCNOps.inc.h:84-138: one fused SGEMM with an atomic, norm-scaled
// scatter in the reference; here the typed linear (tcgen05 segment MM, per-edge rows) followed by
// the edge-parallel aggregation. weights [R, in, out], edge_norm [E, 1], node_feat_output [N, out].
// ---------------------------------------------------------------------------
void rgcn_layer1_separate_coo(Tensor &separate_coo_relptrs, Tensor &separate_coo_eids,
                              Tensor &separate_coo_row_indices, Tensor &separate_coo_col_indices,
                              Tensor &node_feat_input, Tensor &weights, Tensor &edge_norm,
                              Tensor &node_feat_output) {
  const char *op = "rgcn_layer1_separate_coo";
  TORCH_CHECK(weights.dim() == 3, op, ": weights must be [R,in,out]");
  const int K = (int)weights.size(1), N = (int)weights.size(2);
  const int64_t E = separate_coo_eids.numel();
  c10::cuda::CUDAGuard guard(node_feat_output.device());
  void *st = stream_of(node_feat_output);
  Tensor per_edge = at::empty({E, N}, node_feat_output.options());
  const int64_t *eids = i64(separate_coo_eids, op, "separate_coo_eids");
  const int64_t *rel_ptrs = i64(separate_coo_relptrs, op, "separate_coo_relptrs");
  check_rc(hetb_segment_mm(f32(node_feat_input, op, "node_feat_input"),
                           i64(separate_coo_row_indices, op, "separate_coo_row_indices"), 1,
                           f32(weights, op, "weights"), f32w(per_edge, op, "per_edge"), eids, 1, rel_ptrs,
                           separate_coo_relptrs.numel() - 1, E, 1, K, N, 0, st),
           op);
  check_rc(hetb_rgcn_aggregate_coo(eids, i64(separate_coo_col_indices, op, "separate_coo_col_indices"), nullptr,
                                   E, node_feat_output.size(0), f32(per_edge, op, "per_edge"),
                                   f32(edge_norm, op, "edge_norm"), f32w(node_feat_output, op, "node_feat_output"),
                                   N, st),
           op);
}

// OpExport/RGCNOps.inc.h:368-470. weights_transposed [R, out, in]; grad_* arrive zero-filled.
//   G[e] = norm[e] * gout[dst];  grad_x[src] += G[e] W_r^T;  grad_W[r] += x[src]^T G[e];
//   grad_norm[e] = <x[src], gout[dst] W_r^T>
void backward_rgcn_layer1_separate_coo(Tensor &separate_coo_relptrs, Tensor &separate_coo_eids,
                                       Tensor &separate_coo_row_indices, Tensor &separate_coo_col_indices,
                                       Tensor &node_feat_input, Tensor &weights_transposed, Tensor &edge_norm,
                                       Tensor &grad_edge_norm, Tensor &delta_node_feat_input,
                                       Tensor &delta_node_feat_output, Tensor &delta_weights) {
  const char *op = "backward_rgcn_layer1_separate_coo";
  TORCH_CHECK(weights_transposed.dim() == 3, op, ": weights_transposed must be [R,out,in]");
  const int N = (int)weights_transposed.size(1), K = (int)weights_transposed.size(2);  // forward: K in, N out
  const int64_t E = separate_coo_eids.numel();
  const int64_t R = separate_coo_relptrs.numel() - 1;
  c10::cuda::CUDAGuard guard(delta_node_feat_output.device());
  void *st = stream_of(delta_node_feat_output);
  const int64_t *eids = i64(separate_coo_eids, op, "separate_coo_eids");
  const int64_t *rel_ptrs = i64(separate_coo_relptrs, op, "separate_coo_relptrs");
  const int64_t *row = i64(separate_coo_row_indices, op, "separate_coo_row_indices");
  const int64_t *col = i64(separate_coo_col_indices, op, "separate_coo_col_indices");
  const float *x = f32(node_feat_input, op, "node_feat_input");
  const float *wt = f32(weights_transposed, op, "weights_transposed");
  const float *gout = f32(delta_node_feat_output, op, "delta_node_feat_output");
  Tensor scaled = at::empty({E, N}, delta_node_feat_output.options());  // G
  check_rc(hetb_hgt_aggregate_bwd(col, eids, E, f32(edge_norm, op, "edge_norm"), gout, f32w(scaled, op, "scaled"), 1,
                                  N, nullptr, nullptr, st),
           op);
  check_rc(hetb_segment_mm(f32(scaled, op, "scaled"), eids, 1, wt, f32w(delta_node_feat_input, op, "grad_x"), row, 1,
                           rel_ptrs, R, E, 1, N, K, 1, st),
           op);
  check_rc(hetb_segment_mm_grad_w(x, row, 1, f32(scaled, op, "scaled"), eids, 1, f32w(delta_weights, op, "grad_w"),
                                  rel_ptrs, R, E, 1, K, N, st),
           op);
  // grad_norm: P[e] = gout[dst] W_r^T (per-edge, [E, in]), then <P[e], x[src]>
  Tensor proj = at::empty({E, K}, delta_node_feat_output.options());
  check_rc(hetb_segment_mm(gout, col, 1, wt, f32w(proj, op, "proj"), eids, 1, rel_ptrs, R, E, 1, N, K, 0, st), op);
  check_rc(hetb_inner_product_fwd(eids, row, nullptr, E, f32(proj, op, "proj"), x,
                                  f32w(grad_edge_norm, op, "grad_edge_norm"), 1, K, st),
           op);
}

// OpExport/RGCNOps.inc.h:23-82: ret[dst] = sum_e enorm[e] * feat_src[inverse_indices_row[e]] (direct indexing).
void rgcn_node_mean_aggregation_compact_as_of_node_separate_coo(
    Tensor &separate_coo_eids, Tensor &separate_coo_rel_ptrs, Tensor &separate_coo_row_indices,
    Tensor &separate_coo_col_indices, TensorDict args_tensor_dict, Tensor &feat_src, Tensor &enorm, Tensor &ret,
    bool direct_index_flag) {
  const char *op = "rgcn_node_mean_aggregation_compact_as_of_node_separate_coo";
  (void)separate_coo_rel_ptrs; (void)separate_coo_row_indices;
  if (!direct_index_flag) check_rc(HETB_EUNSUPPORTED, op);  // binary-search kind: not built
  Tensor inverse = args_tensor_dict.at("inverse_indices_row");
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_rgcn_aggregate_coo(i64(separate_coo_eids, op, "separate_coo_eids"),
                                   i64(separate_coo_col_indices, op, "separate_coo_col_indices"),
                                   i64(inverse, op, "inverse_indices_row"), separate_coo_eids.numel(), ret.size(0),
                                   f32(feat_src, op, "feat_src"), f32(enorm, op, "enorm"), f32w(ret, op, "ret"),
                                   (int)xlen(feat_src), stream_of(ret)),
           op);
}

// OpExport/RGCNOps.inc.h:303-366: grad_feat_src[inverse_indices_row[e]] += enorm[e] * gradout[dst].
// Separate-COO edge ids are canonical (position == edge id), which the index map relies on.
void backward_rgcn_node_mean_aggregation_compact_as_of_node_separate_coo(
    Tensor &separate_coo_eids, Tensor &separate_coo_rel_ptrs, Tensor &separate_coo_row_indices,
    Tensor &separate_coo_col_indices, TensorDict args_tensor_dict, Tensor &feat_src, Tensor &enorm, Tensor &ret,
    Tensor &gradout, Tensor &grad_feat_src, bool direct_index_flag) {
  const char *op = "backward_rgcn_node_mean_aggregation_compact_as_of_node_separate_coo";
  (void)separate_coo_rel_ptrs; (void)separate_coo_row_indices; (void)feat_src; (void)ret;
  if (!direct_index_flag) check_rc(HETB_EUNSUPPORTED, op);
  Tensor inverse = args_tensor_dict.at("inverse_indices_row");
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_aggregate_bwd(i64(separate_coo_col_indices, op, "separate_coo_col_indices"),
                                  i64(separate_coo_eids, op, "separate_coo_eids"), separate_coo_eids.numel(),
                                  f32(enorm, op, "enorm"), f32(gradout, op, "gradout"),
                                  f32w(grad_feat_src, op, "grad_feat_src"), 1, (int)xlen(grad_feat_src),
                                  i64(inverse, op, "inverse_indices_row"), nullptr, stream_of(gradout)),
           op);
}

// hetb200 extension: the same aggregation over the in-CSR (edge-balanced walker, no atomics);
// msg_idx is aligned with incsr_eids. Backward: hetb200_backward_hgt_message_aggregation_csr_compact.
void hetb200_rgcn_node_aggregation_csr_compact(Tensor &incsr_row_ptrs, Tensor &incsr_eids, Tensor &msg_idx,
                                               Tensor &feat_compact, Tensor &enorm, Tensor &ret) {
  const char *op = "hetb200_rgcn_node_aggregation_csr_compact";
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_hgt_aggregate_fwd_csr(i64(incsr_row_ptrs, op, "incsr_row_ptrs"), i64(incsr_eids, op, "incsr_eids"),
                                      incsr_row_ptrs.numel() - 1, incsr_eids.numel(),
                                      f32(feat_compact, op, "feat_compact"), f32(enorm, op, "enorm"), nullptr,
                                      f32w(ret, op, "ret"), 1, (int)xlen(feat_compact), 0,
                                      i64(msg_idx, op, "msg_idx"), stream_of(ret)),
           op);
}

// hetb200 extension: grad of a layer bias = column sum of the incoming gradient.
void hetb200_column_sum(Tensor &x, Tensor &out) {
  const char *op = "hetb200_column_sum";
  TORCH_CHECK(x.dim() == 2 && out.numel() == x.size(1), op, ": x must be [rows, cols] and out [cols]");
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(hetb_column_sum(f32(x, op, "x"), x.size(0), (int)x.size(1), f32w(out, op, "out"), stream_of(x)), op);
}

// OpExport/HGTOps.inc.h:477-487
void backward_hgt_full_graph_message_mean_aggregation_csr(
    Tensor &outcsr_row_ptrs, Tensor &outcsr_col_indices,
    Tensor &outcsr_reltypes, Tensor &outcsr_eids,
    Tensor &edgesoftmax_sum_per_node, Tensor &normalized_attn_score,
    Tensor &gradout, Tensor &grad_message) {
  const char *op = "backward_hgt_full_graph_message_mean_aggregation_csr";
  (void)outcsr_row_ptrs; (void)outcsr_reltypes; (void)edgesoftmax_sum_per_node;
  const int H = (int)xlen(normalized_attn_score);
  const int F = (int)(xlen(grad_message) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_aggregate_bwd(
               i64(outcsr_col_indices, op, "outcsr_col_indices"),
               i64(outcsr_eids, op, "outcsr_eids"), outcsr_eids.numel(),
               f32(normalized_attn_score, op, "normalized_attn_score"),
               f32(gradout, op, "gradout"), f32w(grad_message, op, "grad_message"),
               H, F, nullptr, nullptr, stream_of(gradout)),
           op);
}

// hetb200 extension: grad_message is compact ([unique (rel, src), H, F], zero-filled on entry)
void hetb200_backward_hgt_message_aggregation_csr_compact(
    Tensor &outcsr_col_indices, Tensor &outcsr_eids, Tensor &msg_idx, Tensor &normalized_attn_score,
    Tensor &gradout, Tensor &grad_message_compact) {
  const char *op = "hetb200_backward_hgt_message_aggregation_csr_compact";
  const int H = (int)xlen(normalized_attn_score);
  const int F = (int)(xlen(grad_message_compact) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_aggregate_bwd(
               i64(outcsr_col_indices, op, "outcsr_col_indices"), i64(outcsr_eids, op, "outcsr_eids"),
               outcsr_eids.numel(), f32(normalized_attn_score, op, "normalized_attn_score"),
               f32(gradout, op, "gradout"), f32w(grad_message_compact, op, "grad_message_compact"), H, F,
               i64(msg_idx, op, "msg_idx"), nullptr, stream_of(gradout)),
           op);
}

// OpExport/HGTOps.inc.h:553-566
void backward_hgt_full_graph_edge_softmax_ops_csr(
    Tensor &outcsr_row_ptr, Tensor &outcsr_col_indices, Tensor &outcsr_eids,
    Tensor &outcsr_reltypes, Tensor &message, Tensor &unnormalized_attn_score,
    Tensor &normalized_attn_score, Tensor &out, Tensor &gradout, Tensor &mu,
    Tensor &grad_attn_score, Tensor &grad_mu) {
  const char *op = "backward_hgt_full_graph_edge_softmax_ops_csr";
  (void)outcsr_row_ptr;
  const int H = (int)xlen(unnormalized_attn_score);
  const int F = (int)(xlen(message) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_edge_softmax_bwd(
               i64(outcsr_col_indices, op, "outcsr_col_indices"),
               i64(outcsr_eids, op, "outcsr_eids"),
               i64(outcsr_reltypes, op, "outcsr_reltypes"), outcsr_eids.numel(),
               mu.size(0), f32(message, op, "message"),
               f32(unnormalized_attn_score, op, "unnormalized_attn_score"),
               f32(normalized_attn_score, op, "normalized_attn_score"),
               f32(out, op, "out"), f32(gradout, op, "gradout"),
               f32(mu, op, "mu"), f32w(grad_attn_score, op, "grad_attn_score"),
               f32w(grad_mu, op, "grad_mu"), H, F, nullptr, nullptr, stream_of(gradout)),
           op);
}

// hetb200 extension: message is compact (see hetb200_hgt_message_aggregation_csr_compact)
void hetb200_backward_hgt_edge_softmax_ops_csr_compact(
    Tensor &outcsr_col_indices, Tensor &outcsr_eids, Tensor &outcsr_reltypes, Tensor &msg_idx,
    Tensor &messages_compact, Tensor &unnormalized_attn_score, Tensor &normalized_attn_score, Tensor &out,
    Tensor &gradout, Tensor &mu, Tensor &grad_attn_score, Tensor &grad_mu) {
  const char *op = "hetb200_backward_hgt_edge_softmax_ops_csr_compact";
  const int H = (int)xlen(unnormalized_attn_score);
  const int F = (int)(xlen(messages_compact) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_edge_softmax_bwd(
               i64(outcsr_col_indices, op, "outcsr_col_indices"), i64(outcsr_eids, op, "outcsr_eids"),
               i64(outcsr_reltypes, op, "outcsr_reltypes"), outcsr_eids.numel(), mu.size(0),
               f32(messages_compact, op, "messages_compact"),
               f32(unnormalized_attn_score, op, "unnormalized_attn_score"),
               f32(normalized_attn_score, op, "normalized_attn_score"), f32(out, op, "out"),
               f32(gradout, op, "gradout"), f32(mu, op, "mu"), f32w(grad_attn_score, op, "grad_attn_score"),
               f32w(grad_mu, op, "grad_mu"), H, F, i64(msg_idx, op, "msg_idx"), nullptr, stream_of(gradout)),
           op);
}

// hetb200 fusion: inner product + edge softmax + aggregation in one in-CSR pass, compact inputs.
// attn / t are still written (the backward needs them); sum, ret as in the unfused ops.
void hetb200_hgt_fused_attention_aggregation_csr(
    Tensor &incsr_row_ptrs, Tensor &incsr_col_indices, Tensor &incsr_eids, Tensor &incsr_reltypes,
    Tensor &left_idx, Tensor &msg_idx, Tensor &left_compact, Tensor &k, Tensor &messages_compact, Tensor &mu,
    Tensor &attn, Tensor &t, Tensor &sum, Tensor &ret) {
  const char *op = "hetb200_hgt_fused_attention_aggregation_csr";
  const int H = (int)mu.size(1);
  const int F = (int)(xlen(messages_compact) / H);
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_hgt_fused_fwd_csr(
               i64(incsr_row_ptrs, op, "incsr_row_ptrs"), i64(incsr_eids, op, "incsr_eids"),
               i64(incsr_reltypes, op, "incsr_reltypes"), i64(incsr_col_indices, op, "incsr_col_indices"),
               incsr_row_ptrs.numel() - 1, incsr_eids.numel(), f32(left_compact, op, "left_compact"),
               i64(left_idx, op, "left_idx"), f32(k, op, "k"), f32(messages_compact, op, "messages_compact"),
               i64(msg_idx, op, "msg_idx"), f32(mu, op, "mu"), f32w(attn, op, "attn"), f32w(t, op, "t"),
               f32w(sum, op, "sum"), f32w(ret, op, "ret"), H, F, stream_of(ret)),
           op);
}

// Backward of the softmax + aggregation half with t = exp(mu·score) and the per-node sums instead of
// the normalised scores (norm = t / sum[d] is formed in the kernels).
void hetb200_backward_hgt_fused_softmax_aggregation_csr(
    Tensor &outcsr_col_indices, Tensor &outcsr_eids, Tensor &outcsr_reltypes, Tensor &msg_idx,
    Tensor &messages_compact, Tensor &attn, Tensor &t, Tensor &sum, Tensor &out, Tensor &gradout, Tensor &mu,
    Tensor &grad_message_compact, Tensor &grad_attn, Tensor &grad_mu) {
  const char *op = "hetb200_backward_hgt_fused_softmax_aggregation_csr";
  const int H = (int)mu.size(1);
  const int F = (int)(xlen(messages_compact) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_aggregate_bwd(
               i64(outcsr_col_indices, op, "outcsr_col_indices"), i64(outcsr_eids, op, "outcsr_eids"),
               outcsr_eids.numel(), f32(t, op, "t"), f32(gradout, op, "gradout"),
               f32w(grad_message_compact, op, "grad_message_compact"), H, F, i64(msg_idx, op, "msg_idx"),
               f32(sum, op, "sum"), stream_of(gradout)),
           op);
  check_rc(hetb_hgt_edge_softmax_bwd(
               i64(outcsr_col_indices, op, "outcsr_col_indices"), i64(outcsr_eids, op, "outcsr_eids"),
               i64(outcsr_reltypes, op, "outcsr_reltypes"), outcsr_eids.numel(), mu.size(0),
               f32(messages_compact, op, "messages_compact"), f32(attn, op, "attn"), f32(t, op, "t"),
               f32(out, op, "out"), f32(gradout, op, "gradout"), f32(mu, op, "mu"),
               f32w(grad_attn, op, "grad_attn"), f32w(grad_mu, op, "grad_mu"), H, F,
               i64(msg_idx, op, "msg_idx"), f32(sum, op, "sum"), stream_of(gradout)),
           op);
}

// hetb200 fusion: the whole backward of hetb200_hgt_fused_attention_aggregation_csr in one out-CSR pass.
void hetb200_backward_hgt_fused_attention_aggregation_csr(
    Tensor &outcsr_row_ptrs, Tensor &outcsr_col_indices, Tensor &outcsr_eids, Tensor &outcsr_reltypes,
    Tensor &left_idx, Tensor &msg_idx, Tensor &left_compact, Tensor &k, Tensor &messages_compact, Tensor &mu,
    Tensor &attn, Tensor &t, Tensor &sum, Tensor &out, Tensor &gradout, Tensor &grad_left, Tensor &grad_k,
    Tensor &grad_message_compact, Tensor &grad_mu) {
  const char *op = "hetb200_backward_hgt_fused_attention_aggregation_csr";
  const int H = (int)mu.size(1);
  const int F = (int)(xlen(messages_compact) / H);
  c10::cuda::CUDAGuard guard(gradout.device());
  Tensor scratch = at::empty({sum.numel()}, sum.options());
  check_rc(hetb_hgt_fused_bwd_csr(
               i64(outcsr_row_ptrs, op, "outcsr_row_ptrs"), i64(outcsr_col_indices, op, "outcsr_col_indices"),
               i64(outcsr_eids, op, "outcsr_eids"), i64(outcsr_reltypes, op, "outcsr_reltypes"),
               outcsr_row_ptrs.numel() - 1, outcsr_eids.numel(), mu.size(0), f32(left_compact, op, "left_compact"),
               i64(left_idx, op, "left_idx"), f32(k, op, "k"), f32(messages_compact, op, "messages_compact"),
               i64(msg_idx, op, "msg_idx"), f32(mu, op, "mu"), f32(attn, op, "attn"), f32(t, op, "t"),
               f32(sum, op, "sum"), f32(out, op, "out"), f32(gradout, op, "gradout"), sum.numel() / H,
               f32w(scratch, op, "scratch"), f32w(grad_left, op, "grad_left"), f32w(grad_k, op, "grad_k"),
               f32w(grad_message_compact, op, "grad_message_compact"), f32w(grad_mu, op, "grad_mu"), H, F,
               stream_of(gradout)),
           op);
}

// hetb200 extension: typed linear with the output stored as bf16 (a table the plan-based edge passes
// gather). `gather` (int64 node list) may be an empty tensor: rows are then the segments' own rows
// (the no_scatter_gather_list form). Tensor-core shapes only; fails loudly otherwise.
void hetb200_typed_linear_bf16out(Tensor &seg_ptrs, Tensor &gather, Tensor &weights, Tensor &inputs, Tensor &ret) {
  const char *op = "hetb200_typed_linear_bf16out";
  TORCH_CHECK(weights.dim() == 4, op, ": weights must be [S,H,in,out/H]");
  TORCH_CHECK(ret.is_cuda() && ret.is_contiguous() && ret.scalar_type() == at::kBFloat16, op,
              ": `ret` must be a contiguous bfloat16 CUDA tensor");
  const int H = (int)weights.size(1), K = (int)weights.size(2), N = (int)weights.size(3);
  const int a_heads = (xlen(inputs) == (int64_t)H * K) ? H : 1;
  const bool has_gather = gather.numel() > 0;
  const int64_t M = has_gather ? gather.numel() : inputs.size(0);
  TORCH_CHECK(ret.numel() == M * H * N, op, ": `ret` has the wrong size");
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_segment_mm_bf16out(f32(inputs, op, "inputs"), has_gather ? i64(gather, op, "gather") : nullptr, a_heads,
                                   f32(weights, op, "weights"), ret.data_ptr(), nullptr,
                                   i64(seg_ptrs, op, "seg_ptrs"), seg_ptrs.numel() - 1, M, H, K, N, stream_of(ret)),
           op);
}

// ------------------------------------------------- plan-based fused edge passes ---
// A walk plan travels as ONE int32 device tensor plus a host list of integers
// (het_b200/graph.py: WalkPlan.blob / .meta): counts first, then the element offset of every array
// inside the blob (-1 = absent), in the order of hetb_walk_plan's fields.
constexpr int kPlanMetaLen = 21;

hetb_walk_plan unpack_plan(const Tensor &blob, c10::IntArrayRef meta, const char *op) {
  TORCH_CHECK(blob.is_cuda() && blob.scalar_type() == at::kInt && blob.is_contiguous(), "torch_hrt::", op,
              ": the plan blob must be a contiguous int32 CUDA tensor");
  TORCH_CHECK((int)meta.size() == kPlanMetaLen, "torch_hrt::", op, ": plan meta must hold ", kPlanMetaLen, " integers");
  const int32_t *base = blob.data_ptr<int32_t>();
  const int64_t n = blob.numel();
  auto at_off = [&](int i) -> const int32_t * {
    const int64_t o = meta[i];
    TORCH_CHECK(o >= -1 && o < n, "torch_hrt::", op, ": plan offset out of range");
    return o < 0 ? nullptr : base + o;
  };
  hetb_walk_plan pl;
  pl.num_positions = meta[0];
  pl.num_items = (int32_t)meta[1];
  pl.num_nz_rows = (int32_t)meta[2];
  pl.num_fix_rows = (int32_t)meta[3];
  pl.num_empty_rows = (int32_t)meta[4];
  pl.num_pslots = (int32_t)meta[5];
  pl.num_fix_runs = (int32_t)meta[6];
  pl.item_ptr = at_off(7);
  pl.item_row = at_off(8);
  pl.nz_rows = at_off(9);
  pl.nz_ptr = at_off(10);
  pl.fix_rows = at_off(11);
  pl.empty_rows = at_off(12);
  pl.pslot = at_off(13);
  pl.pslot_row = at_off(14);
  pl.fix_runs = at_off(15);
  pl.idx_a = at_off(16);
  pl.idx_b = at_off(17);
  pl.idx_c = at_off(18);
  pl.idx_d = at_off(19);
  pl.row_rel = at_off(20);
  return pl;
}

// fp32 or bf16 storage of a gathered table; returns the dtype tag of the C ABI
int table_dtype(const Tensor &t, const char *op, const char *name) {
  TORCH_CHECK(t.is_cuda() && t.is_contiguous(), "torch_hrt::", op, ": `", name, "` must be a contiguous CUDA tensor");
  if (t.scalar_type() == at::kFloat) return HETB_DTYPE_F32;
  if (t.scalar_type() == at::kBFloat16) return HETB_DTYPE_BF16;
  TORCH_CHECK(false, "torch_hrt::", op, ": `", name, "` must be float32 or bfloat16");
  return -1;
}

void hetb200_hgt_fused_fwd_plan(Tensor &plan, c10::IntArrayRef meta, Tensor &left_compact, Tensor &k,
                                Tensor &messages_compact, Tensor &mu, Tensor &attn, Tensor &lse, Tensor &ret,
                                Tensor &part_m, Tensor &part_s, Tensor &part_acc, Tensor &ssum) {
  const char *op = "hetb200_hgt_fused_fwd_plan";
  const hetb_walk_plan pl = unpack_plan(plan, meta, op);
  const int H = (int)mu.size(1);
  const int F = (int)(xlen(ret) / H);
  const int dt = table_dtype(left_compact, op, "left_compact");
  TORCH_CHECK(table_dtype(k, op, "k") == dt && table_dtype(messages_compact, op, "messages_compact") == dt,
              "torch_hrt::", op, ": left_compact, k and messages_compact must share one dtype");
  // attn: empty = do not save the raw scores (the default backward recomputes them)
  TORCH_CHECK((attn.numel() == 0 || attn.numel() == pl.num_positions * H) && part_m.numel() >= 2 * (int64_t)pl.num_items * H &&
                  part_acc.numel() >= 2 * (int64_t)pl.num_items * H * F,
              "torch_hrt::", op, ": attn / scratch sizes do not match the plan");
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_hgt_fused_fwd_plan(&pl, left_compact.data_ptr(), k.data_ptr(), messages_compact.data_ptr(), dt,
                                   f32(mu, op, "mu"), attn.numel() ? f32w(attn, op, "attn") : nullptr, f32w(lse, op, "lse"),
                                   f32w(ret, op, "ret"), f32w(part_m, op, "part_m"), f32w(part_s, op, "part_s"),
                                   f32w(part_acc, op, "part_acc"), f32w(ssum, op, "ssum"), H, F, stream_of(ret)),
           op);
}

void hetb200_hgt_fused_bwd_plan(Tensor &plan, c10::IntArrayRef meta, Tensor &left_compact, Tensor &k,
                                Tensor &messages_compact, Tensor &mu, Tensor &attn, Tensor &lse, Tensor &out,
                                Tensor &gradout, Tensor &gdoto_scratch, Tensor &grad_left, Tensor &grad_k,
                                Tensor &grad_message_compact, Tensor &grad_mu) {
  const char *op = "hetb200_hgt_fused_bwd_plan";
  const hetb_walk_plan pl = unpack_plan(plan, meta, op);
  const int H = (int)mu.size(1);
  const int F = (int)(xlen(out) / H);
  const int dt = table_dtype(left_compact, op, "left_compact");
  TORCH_CHECK(table_dtype(k, op, "k") == dt && table_dtype(messages_compact, op, "messages_compact") == dt,
              "torch_hrt::", op, ": left_compact, k and messages_compact must share one dtype");
  TORCH_CHECK(gdoto_scratch.numel() >= 2 * out.size(0) * H, "torch_hrt::", op, ": gdoto_scratch too small ([N, 2H])");
  TORCH_CHECK(attn.numel() == 0 || attn.numel() == pl.num_positions * H, "torch_hrt::", op, ": attn does not match the plan");
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_fused_bwd_plan(&pl, mu.size(0), out.size(0), left_compact.data_ptr(), k.data_ptr(),
                                   messages_compact.data_ptr(), dt, f32(mu, op, "mu"),
                                   attn.numel() ? f32(attn, op, "attn") : nullptr,
                                   f32(lse, op, "lse"), f32(out, op, "out"), f32(gradout, op, "gradout"),
                                   f32w(gdoto_scratch, op, "gdoto_scratch"), f32w(grad_left, op, "grad_left"),
                                   f32w(grad_k, op, "grad_k"), f32w(grad_message_compact, op, "grad_message_compact"),
                                   f32w(grad_mu, op, "grad_mu"), H, F, stream_of(gradout)),
           op);
}

// Two-pass form: attn is overwritten with the gradient of the raw scores; grad_left needs no zero-fill.
void hetb200_hgt_fused_bwd_plan2(Tensor &plan, c10::IntArrayRef meta, Tensor &fwd_plan, c10::IntArrayRef fwd_meta,
                                 Tensor &left_compact, Tensor &k, Tensor &messages_compact, Tensor &mu, Tensor &attn,
                                 Tensor &lse, Tensor &out, Tensor &gradout, Tensor &gdoto_scratch, Tensor &grad_left,
                                 Tensor &grad_k, Tensor &grad_message_compact, Tensor &grad_mu) {
  const char *op = "hetb200_hgt_fused_bwd_plan2";
  const hetb_walk_plan pl = unpack_plan(plan, meta, op);
  const hetb_walk_plan fpl = unpack_plan(fwd_plan, fwd_meta, op);
  const int H = (int)mu.size(1);
  const int F = (int)(xlen(out) / H);
  const int dt = table_dtype(left_compact, op, "left_compact");
  TORCH_CHECK(table_dtype(k, op, "k") == dt && table_dtype(messages_compact, op, "messages_compact") == dt,
              "torch_hrt::", op, ": left_compact, k and messages_compact must share one dtype");
  TORCH_CHECK(gdoto_scratch.numel() >= 2 * out.size(0) * H, "torch_hrt::", op, ": gdoto_scratch too small ([N, 2H])");
  TORCH_CHECK(attn.numel() == pl.num_positions * H, "torch_hrt::", op, ": attn does not match the plan");
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_hgt_fused_bwd_plan2(&pl, &fpl, mu.size(0), out.size(0), left_compact.data_ptr(), k.data_ptr(),
                                    messages_compact.data_ptr(), dt, f32(mu, op, "mu"), f32w(attn, op, "attn"),
                                    f32(lse, op, "lse"), f32(out, op, "out"), f32(gradout, op, "gradout"),
                                    f32w(gdoto_scratch, op, "gdoto_scratch"), f32w(grad_left, op, "grad_left"),
                                    f32w(grad_k, op, "grad_k"), f32w(grad_message_compact, op, "grad_message_compact"),
                                    f32w(grad_mu, op, "grad_mu"), H, F, stream_of(gradout)),
           op);
}

void hetb200_rgat_fused_fwd_plan(Tensor &plan, c10::IntArrayRef meta, Tensor &feat_compact, Tensor &el, Tensor &er,
                                 double slope, Tensor &lse, Tensor &ret, Tensor &part_m, Tensor &part_s,
                                 Tensor &part_acc, Tensor &ssum) {
  const char *op = "hetb200_rgat_fused_fwd_plan";
  const hetb_walk_plan pl = unpack_plan(plan, meta, op);
  const int H = (int)xlen(el);
  const int F = (int)(xlen(ret) / H);
  TORCH_CHECK(part_m.numel() >= 2 * (int64_t)pl.num_items * H && part_acc.numel() >= 2 * (int64_t)pl.num_items * H * F,
              "torch_hrt::", op, ": scratch sizes do not match the plan");
  c10::cuda::CUDAGuard guard(ret.device());
  check_rc(hetb_rgat_fused_fwd_plan(&pl, feat_compact.data_ptr(), table_dtype(feat_compact, op, "feat_compact"),
                                    f32(el, op, "el"), f32(er, op, "er"), (float)slope, f32w(lse, op, "lse"),
                                    f32w(ret, op, "ret"), f32w(part_m, op, "part_m"), f32w(part_s, op, "part_s"),
                                    f32w(part_acc, op, "part_acc"), f32w(ssum, op, "ssum"), H, F, stream_of(ret)),
           op);
}

void hetb200_rgat_fused_bwd_plan(Tensor &plan, c10::IntArrayRef meta, Tensor &feat_compact, Tensor &el, Tensor &er,
                                 Tensor &lse, Tensor &ret, Tensor &gradout, Tensor &gdotr_scratch, Tensor &grad_feat,
                                 Tensor &grad_el, Tensor &grad_er, double slope) {
  const char *op = "hetb200_rgat_fused_bwd_plan";
  const hetb_walk_plan pl = unpack_plan(plan, meta, op);
  const int H = (int)xlen(el);
  const int F = (int)(xlen(ret) / H);
  TORCH_CHECK(gdotr_scratch.numel() >= 2 * ret.size(0) * H, "torch_hrt::", op, ": gdotr_scratch too small ([N, 2H])");
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_rgat_fused_bwd_plan(&pl, ret.size(0), feat_compact.data_ptr(),
                                    table_dtype(feat_compact, op, "feat_compact"), f32(el, op, "el"),
                                    f32(er, op, "er"), f32(lse, op, "lse"), f32(ret, op, "ret"),
                                    f32(gradout, op, "gradout"), f32w(gdotr_scratch, op, "gdotr_scratch"),
                                    f32w(grad_feat, op, "grad_feat"), f32w(grad_el, op, "grad_el"),
                                    f32w(grad_er, op, "grad_er"), H, F, (float)slope, stream_of(gradout)),
           op);
}

// ------------------------------------------------- P1: separate-COO HGT ops ---
// Same maths as the P0 ops in the reference's other slicing; built on the existing kernels
// (typed linear + inner product / weighted scatter). Per-edge temporaries live only inside the call.

// compact row of every COO position by binary search in per-relation sorted unique lists
// (CompactAsOfNodeKind::Enabled / EnabledWithDualList): position-aligned int64 tensor
Tensor compact_index_by_position(const char *op, const Tensor &node, const int64_t *rel_of_pos, const int64_t *rel_ptrs,
                                 int64_t num_rels, const Tensor &unique_rel_ptrs, const Tensor &unique_nodes) {
  Tensor idx = at::empty({node.numel()}, node.options());
  Tensor missing = at::zeros({1}, node.options().dtype(at::kInt));
  check_rc(hetb_compact_index(i64(node, op, "node indices"), rel_of_pos, rel_ptrs, num_rels, node.numel(),
                              i64(unique_rel_ptrs, op, "unique_srcs_and_dests_rel_ptrs"),
                              i64(unique_nodes, op, "unique_srcs_and_dests_node_indices"), idx.data_ptr<int64_t>(),
                              missing.data_ptr<int>(), stream_of(node)),
           op);
  return idx;
}

// OpExport/HGTOpsEdgeParallel.inc.h:95-158: inner[eid] = k[row]·W_att[rel]; score[eid,h] = <inner[eid,h,:], q[col,h,:]>
void hgt_full_graph_hetero_attention_ops_coo(Tensor &row_indices, Tensor &col_indices, Tensor &eids, Tensor &rel_ptrs,
                                             Tensor &applied_klinear_node_features,
                                             Tensor &applied_qlinear_node_features, Tensor &attn_score_weight,
                                             Tensor &attn_score_inner_product, Tensor &unnormalized_attn_score) {
  const char *op = "hgt_full_graph_hetero_attention_ops_coo";
  TORCH_CHECK(attn_score_weight.dim() == 4, op, ": attn_score_weight must be [R,H,in,out]");
  const int H = (int)attn_score_weight.size(1), K = (int)attn_score_weight.size(2), N = (int)attn_score_weight.size(3);
  c10::cuda::CUDAGuard guard(unnormalized_attn_score.device());
  void *st = stream_of(unnormalized_attn_score);
  check_rc(hetb_segment_mm(f32(applied_klinear_node_features, op, "k"), i64(row_indices, op, "row_indices"), H,
                           f32(attn_score_weight, op, "attn_score_weight"),
                           f32w(attn_score_inner_product, op, "attn_score_inner_product"), i64(eids, op, "eids"), H,
                           i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, K, N, 0, st),
           op);
  check_rc(hetb_inner_product_fwd(i64(eids, op, "eids"), i64(col_indices, op, "col_indices"), nullptr, eids.numel(),
                                  f32(attn_score_inner_product, op, "attn_score_inner_product"),
                                  f32(applied_qlinear_node_features, op, "q"),
                                  f32w(unnormalized_attn_score, op, "unnormalized_attn_score"), H, N, st),
           op);
}

// OpExport/HGTOpsEdgeParallel.inc.h:166-293. grad_k / grad_q / grad_attn_score_weight arrive zero-filled.
void backward_hgt_full_graph_hetero_attention_ops_coo(
    Tensor &incsr_row_ptr, Tensor &incsr_col_indices, Tensor &incsr_eids, Tensor &incsr_reltypes,
    Tensor &row_indices, Tensor &col_indices, Tensor &eids, Tensor &rel_ptrs, Tensor &grad_attn_score_weight,
    Tensor &attn_score_weight_transposed, Tensor &applied_klinear_node_features,
    Tensor &applied_qlinear_node_features, Tensor &attn_score_inner_product, Tensor &grad_unnorm_attn_score,
    Tensor &grad_k, Tensor &grad_q) {
  const char *op = "backward_hgt_full_graph_hetero_attention_ops_coo";
  (void)incsr_row_ptr; (void)incsr_col_indices; (void)incsr_eids; (void)incsr_reltypes;
  const int H = (int)attn_score_weight_transposed.size(1), N = (int)attn_score_weight_transposed.size(2),
            K = (int)attn_score_weight_transposed.size(3);
  c10::cuda::CUDAGuard guard(grad_k.device());
  void *st = stream_of(grad_k);
  Tensor grad_inner = at::empty_like(attn_score_inner_product);
  check_rc(hetb_inner_product_bwd(i64(eids, op, "eids"), i64(col_indices, op, "col_indices"), nullptr, eids.numel(),
                                  f32(attn_score_inner_product, op, "attn_score_inner_product"),
                                  f32(applied_qlinear_node_features, op, "q"),
                                  f32(grad_unnorm_attn_score, op, "grad_unnorm_attn_score"),
                                  grad_inner.data_ptr<float>(), f32w(grad_q, op, "grad_q"), H, N, st),
           op);
  check_rc(hetb_segment_mm(grad_inner.data_ptr<float>(), i64(eids, op, "eids"), H,
                           f32(attn_score_weight_transposed, op, "attn_score_weight_transposed"),
                           f32w(grad_k, op, "grad_k"), i64(row_indices, op, "row_indices"), H,
                           i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, N, K, 1, st),
           op);
  check_rc(hetb_segment_mm_grad_w(f32(applied_klinear_node_features, op, "k"), i64(row_indices, op, "row_indices"), H,
                                  grad_inner.data_ptr<float>(), i64(eids, op, "eids"), H,
                                  f32w(grad_attn_score_weight, op, "grad_attn_score_weight"),
                                  i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, K, N, st),
           op);
}

// OpExport/HGTOpsEdgeParallel.inc.h:18-32
void hgt_full_graph_edge_softmax_ops_separate_coo(Tensor &row_indices, Tensor &col_indices, Tensor &eids,
                                                  Tensor &rel_ptrs, Tensor &unnormalized_attn_score, Tensor &mu,
                                                  Tensor &edgesoftmax_sum_per_node,
                                                  Tensor &mu_softmax_applied_unnormalized_attn_score,
                                                  Tensor &normalized_attn_score) {
  const char *op = "hgt_full_graph_edge_softmax_ops_separate_coo";
  (void)row_indices;
  const int H = (int)mu.size(1);
  c10::cuda::CUDAGuard guard(normalized_attn_score.device());
  check_rc(hetb_hgt_edge_softmax_fwd_coo(
               i64(col_indices, op, "col_indices"), i64(eids, op, "eids"), nullptr, i64(rel_ptrs, op, "rel_ptrs"),
               rel_ptrs.numel() - 1, edgesoftmax_sum_per_node.size(0), eids.numel(),
               f32(unnormalized_attn_score, op, "unnormalized_attn_score"), f32(mu, op, "mu"),
               f32w(edgesoftmax_sum_per_node, op, "edgesoftmax_sum_per_node"),
               f32w(mu_softmax_applied_unnormalized_attn_score, op, "mu_softmax_applied_unnormalized_attn_score"),
               f32w(normalized_attn_score, op, "normalized_attn_score"), H, stream_of(normalized_attn_score)),
           op);
}

// OpExport/HGTOpsEdgeParallel.inc.h:33-92: out[col] = sum_e edge_norm[eid,h] * (input[row]·W[rel])
void hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo(
    Tensor &rel_ptrs, Tensor &eids, Tensor &row_indices, Tensor &col_indices, Tensor &node_feat_input,
    Tensor &weights, Tensor &edge_norm, Tensor &node_feat_output) {
  const char *op = "hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo";
  const int H = (int)weights.size(1), K = (int)weights.size(2), N = (int)weights.size(3);
  c10::cuda::CUDAGuard guard(node_feat_output.device());
  void *st = stream_of(node_feat_output);
  Tensor msg = at::empty({eids.numel(), H, N}, node_feat_output.options());
  check_rc(hetb_segment_mm(f32(node_feat_input, op, "node_feat_input"), i64(row_indices, op, "row_indices"), H,
                           f32(weights, op, "weights"), msg.data_ptr<float>(), i64(eids, op, "eids"), H,
                           i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, K, N, 0, st),
           op);
  check_rc(hetb_weighted_scatter_coo(i64(col_indices, op, "col_indices"), i64(eids, op, "eids"), eids.numel(),
                                     node_feat_output.size(0), msg.data_ptr<float>(), f32(edge_norm, op, "edge_norm"),
                                     f32w(node_feat_output, op, "node_feat_output"), H, N, st),
           op);
}

// OpExport/HGTOpsEdgeParallel.inc.h:295-370. grad_* arrive zero-filled.
void backward_hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo(
    Tensor &rel_ptrs, Tensor &eids, Tensor &row_indices, Tensor &col_indices, Tensor &node_feat_input,
    Tensor &weights_transposed, Tensor &edge_norm, Tensor &node_feat_output, Tensor &grad_node_feat_input,
    Tensor &grad_weights, Tensor &grad_edge_norm, Tensor &grad_node_feat_output) {
  const char *op = "backward_hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo";
  (void)node_feat_output;
  const int H = (int)weights_transposed.size(1), N = (int)weights_transposed.size(2), K = (int)weights_transposed.size(3);
  c10::cuda::CUDAGuard guard(grad_node_feat_input.device());
  void *st = stream_of(grad_node_feat_input);
  Tensor weights = weights_transposed.transpose(2, 3).contiguous();
  Tensor msg = at::empty({eids.numel(), H, N}, grad_node_feat_output.options());
  check_rc(hetb_segment_mm(f32(node_feat_input, op, "node_feat_input"), i64(row_indices, op, "row_indices"), H,
                           weights.data_ptr<float>(), msg.data_ptr<float>(), i64(eids, op, "eids"), H,
                           i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, K, N, 0, st),
           op);
  // grad_edge_norm[eid,h] = <gout[col,h,:], msg[eid,h,:]>
  check_rc(hetb_inner_product_fwd(i64(eids, op, "eids"), i64(col_indices, op, "col_indices"), nullptr, eids.numel(),
                                  msg.data_ptr<float>(), f32(grad_node_feat_output, op, "grad_node_feat_output"),
                                  f32w(grad_edge_norm, op, "grad_edge_norm"), H, N, st),
           op);
  // grad_msg[eid] = edge_norm[eid,h] * gout[col]  (overwrites msg)
  check_rc(hetb_hgt_aggregate_bwd(i64(col_indices, op, "col_indices"), i64(eids, op, "eids"), eids.numel(),
                                  f32(edge_norm, op, "edge_norm"), f32(grad_node_feat_output, op, "grad_node_feat_output"),
                                  msg.data_ptr<float>(), H, N, nullptr, nullptr, st),
           op);
  check_rc(hetb_segment_mm(msg.data_ptr<float>(), i64(eids, op, "eids"), H,
                           f32(weights_transposed, op, "weights_transposed"),
                           f32w(grad_node_feat_input, op, "grad_node_feat_input"), i64(row_indices, op, "row_indices"), H,
                           i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, N, K, 1, st),
           op);
  check_rc(hetb_segment_mm_grad_w(f32(node_feat_input, op, "node_feat_input"), i64(row_indices, op, "row_indices"), H,
                                  msg.data_ptr<float>(), i64(eids, op, "eids"), H, f32w(grad_weights, op, "grad_weights"),
                                  i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1, eids.numel(), H, K, N, st),
           op);
}

// OpExport/HGTOps.inc.h:283-340 (in-CSR: rows = destinations)
void backward_hgt_full_graph_enorm_to_unnormalized_attn_score_csr(
    Tensor &incsr_row_ptr, Tensor &incsr_col_indices, Tensor &incsr_eids, Tensor &incsr_reltypes,
    Tensor &unnormalized_attn_score, Tensor &normalized_attn_score, Tensor &grad_normalized_attn_score, Tensor &mu,
    Tensor &grad_unnormalized_attn_score, Tensor &grad_mu) {
  const char *op = "backward_hgt_full_graph_enorm_to_unnormalized_attn_score_csr";
  (void)incsr_col_indices;
  const int H = (int)mu.size(1);
  const int64_t n = incsr_row_ptr.numel() - 1, e = incsr_eids.numel();
  c10::cuda::CUDAGuard guard(grad_mu.device());
  void *st = stream_of(grad_mu);
  Tensor dst = at::empty({e}, incsr_eids.options());
  check_rc(hetb_expand_row_ptr(i64(incsr_row_ptr, op, "incsr_row_ptr"), n, e, dst.data_ptr<int64_t>(), st), op);
  Tensor scratch = at::empty({n * H}, mu.options());
  check_rc(hetb_hgt_enorm_to_unnormalized_bwd(
               dst.data_ptr<int64_t>(), i64(incsr_eids, op, "incsr_eids"), i64(incsr_reltypes, op, "incsr_reltypes"),
               nullptr, mu.size(0), n, e, f32(unnormalized_attn_score, op, "unnormalized_attn_score"),
               f32(normalized_attn_score, op, "normalized_attn_score"),
               f32(grad_normalized_attn_score, op, "grad_normalized_attn_score"), f32(mu, op, "mu"),
               scratch.data_ptr<float>(), f32w(grad_unnormalized_attn_score, op, "grad_unnormalized_attn_score"),
               f32w(grad_mu, op, "grad_mu"), H, st),
           op);
}

// OpExport/HGTOps.inc.h:598-650 (separate COO: col_indices = destinations)
void backward_hgt_full_graph_enorm_to_unnormalized_attn_score_separate_coo(
    Tensor &row_indices, Tensor &col_indices, Tensor &eids, Tensor &rel_ptrs, Tensor &unnormalized_attn_score,
    Tensor &normalized_attn_score, Tensor &grad_normalized_attn_score, Tensor &mu,
    Tensor &grad_unnormalized_attn_score, Tensor &grad_mu, Tensor &sum_incoming_edges_product_softmax_score) {
  const char *op = "backward_hgt_full_graph_enorm_to_unnormalized_attn_score_separate_coo";
  (void)row_indices;
  const int H = (int)mu.size(1);
  c10::cuda::CUDAGuard guard(grad_mu.device());
  check_rc(hetb_hgt_enorm_to_unnormalized_bwd(
               i64(col_indices, op, "col_indices"), i64(eids, op, "eids"), nullptr, i64(rel_ptrs, op, "rel_ptrs"),
               rel_ptrs.numel() - 1, sum_incoming_edges_product_softmax_score.size(0), eids.numel(),
               f32(unnormalized_attn_score, op, "unnormalized_attn_score"),
               f32(normalized_attn_score, op, "normalized_attn_score"),
               f32(grad_normalized_attn_score, op, "grad_normalized_attn_score"), f32(mu, op, "mu"),
               f32w(sum_incoming_edges_product_softmax_score, op, "sum_incoming_edges_product_softmax_score"),
               f32w(grad_unnormalized_attn_score, op, "grad_unnormalized_attn_score"), f32w(grad_mu, op, "grad_mu"), H,
               stream_of(grad_mu)),
           op);
}

// OpExport/RGNNOps.inc.h:565-607: out[eid] = in[compact row of (relation, col node) of the edge]
void rgnn_gather_compact_to_non_compact(TensorDict args, int64_t kind, Tensor &rel_ptrs, Tensor &eids,
                                        Tensor &row_indices, Tensor &col_indices, Tensor &in_feat, Tensor &out_feat) {
  const char *op = "rgnn_gather_compact_to_non_compact";
  (void)row_indices;
  c10::cuda::CUDAGuard guard(out_feat.device());
  Tensor idx;  // position-aligned source row (undefined: the edge id itself)
  if (kind == 2) {  // EnabledWithDirectIndexing: args[edata_idx_to_inverse_idx] is indexed by edge id
    idx = args.at("edata_idx_to_inverse_idx").index({eids}).contiguous();
  } else if (kind == 1) {  // Enabled: binary search of the col node in the relation's unique list
    idx = compact_index_by_position(op, col_indices, nullptr, i64(rel_ptrs, op, "rel_ptrs"), rel_ptrs.numel() - 1,
                                    args.at("unique_srcs_and_dests_rel_ptrs"),
                                    args.at("unique_srcs_and_dests_node_indices"));
  } else if (kind != 0) {
    check_rc(HETB_EUNSUPPORTED, op);
  }
  check_rc(hetb_gather_rows(i64(eids, op, "eids"), idx.defined() ? idx.data_ptr<int64_t>() : nullptr, eids.numel(),
                            f32(in_feat, op, "in_feat"), f32w(out_feat, op, "out_feat"), xlen(in_feat),
                            stream_of(out_feat)),
           op);
}

// ------------------------------------------------- P2: registered, outside the hot path ---
// Same arities as the reference's registrations (tensor arguments, trailing double where the
// reference has one) so that a call reaches the message below instead of a schema error.
void seastar_rgcn_layer0_csr(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6) { not_implemented("seastar_rgcn_layer0_csr"); }
void seastar_backward_rgcn_layer0_csr(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7) { not_implemented("seastar_backward_rgcn_layer0_csr"); }
void seastar_rgcn_layer1_csr(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7) { not_implemented("seastar_rgcn_layer1_csr"); }
void seastar_backward_rgcn_layer1_csr(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7, Tensor &a8, Tensor &a9, Tensor &a10) { not_implemented("seastar_backward_rgcn_layer1_csr"); }
void seastar_rgcn_layer1_coo(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7) { not_implemented("seastar_rgcn_layer1_coo"); }
void seastar_backward_rgcn_layer1_coo(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7, Tensor &a8, Tensor &a9, Tensor &a10) { not_implemented("seastar_backward_rgcn_layer1_coo"); }
void fused_gat_csr(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7, Tensor &a8, double slope) { not_implemented("fused_gat_csr"); }
void backward_fused_gat_csr(Tensor &a0, Tensor &a1, Tensor &a2, Tensor &a3, Tensor &a4, Tensor &a5, Tensor &a6, Tensor &a7, Tensor &a8, Tensor &a9, Tensor &a10, Tensor &a11, Tensor &a12, double slope) { not_implemented("backward_fused_gat_csr"); }

// ------------------------------------------------------------ halo exchange pack / unpack ---
void hetb200_gather_rows(Tensor &idx, Tensor &in, Tensor &out) {
  const char *op = "hetb200_gather_rows";
  TORCH_CHECK(out.size(0) == idx.numel() && xlen(out) == xlen(in), op, ": shapes do not match");
  c10::cuda::CUDAGuard guard(out.device());
  check_rc(hetb_rows_gather(i64(idx, op, "idx"), idx.numel(), f32(in, op, "in"), f32w(out, op, "out"), xlen(in),
                            stream_of(out)),
           op);
}

void hetb200_scatter_add_rows(Tensor &idx, Tensor &src, Tensor &dst) {
  const char *op = "hetb200_scatter_add_rows";
  TORCH_CHECK(src.size(0) == idx.numel() && xlen(src) == xlen(dst), op, ": shapes do not match");
  c10::cuda::CUDAGuard guard(dst.device());
  check_rc(hetb_rows_scatter_add(i64(idx, op, "idx"), idx.numel(), f32(src, op, "src"), f32w(dst, op, "dst"), xlen(src),
                                 stream_of(dst)),
           op);
}

// ------------------------------------------------------------ optimizer / weight composition ---
void hetb200_adam_bump_step(Tensor &step) {
  const char *op = "hetb200_adam_bump_step";
  c10::cuda::CUDAGuard guard(step.device());
  check_rc(hetb_adam_bump_step(f32w(step, op, "step"), stream_of(step)), op);
}

void hetb200_adam_step(Tensor &param, Tensor &grad, Tensor &exp_avg, Tensor &exp_avg_sq, Tensor &step, double lr,
                       double beta1, double beta2, double eps) {
  const char *op = "hetb200_adam_step";
  TORCH_CHECK(grad.numel() == param.numel() && exp_avg.numel() == param.numel() && exp_avg_sq.numel() == param.numel(),
              op, ": tensor sizes differ");
  c10::cuda::CUDAGuard guard(param.device());
  check_rc(hetb_adam_step(f32w(param, op, "param"), f32(grad, op, "grad"), f32w(exp_avg, op, "exp_avg"),
                          f32w(exp_avg_sq, op, "exp_avg_sq"), param.numel(), f32(step, op, "step"), (float)lr,
                          (float)beta1, (float)beta2, (float)eps, stream_of(param)),
           op);
}

// lin [T, 1, in, H*dk] (a typed-linear weight), rel [R, H, dk, dk] -> out [R, 1, in, H*dk]
void hetb200_compose_weights(Tensor &lin, Tensor &type_of_rel, Tensor &rel, Tensor &out) {
  const char *op = "hetb200_compose_weights";
  const int H = (int)rel.size(1), dk = (int)rel.size(2), in_dim = (int)lin.size(2);
  TORCH_CHECK(lin.size(3) == (int64_t)H * dk && rel.size(3) == dk && out.numel() == rel.size(0) * in_dim * H * dk, op,
              ": shapes do not match");
  c10::cuda::CUDAGuard guard(out.device());
  check_rc(hetb_compose_weights_fwd(f32(lin, op, "lin"), i64(type_of_rel, op, "type_of_rel"), f32(rel, op, "rel"),
                                    f32w(out, op, "out"), rel.size(0), in_dim, H, dk, stream_of(out)),
           op);
}

void hetb200_backward_compose_weights(Tensor &gradout, Tensor &lin, Tensor &type_of_rel, Tensor &rel, Tensor &grad_lin,
                                      Tensor &grad_rel) {
  const char *op = "hetb200_backward_compose_weights";
  const int H = (int)rel.size(1), dk = (int)rel.size(2), in_dim = (int)lin.size(2);
  c10::cuda::CUDAGuard guard(gradout.device());
  check_rc(hetb_compose_weights_bwd(f32(gradout, op, "gradout"), f32(lin, op, "lin"),
                                    i64(type_of_rel, op, "type_of_rel"), f32(rel, op, "rel"),
                                    f32w(grad_lin, op, "grad_lin"), f32w(grad_rel, op, "grad_rel"), rel.size(0), in_dim,
                                    H, dk, stream_of(gradout)),
           op);
}

// ---------------------------------------------------------------- loss ---
// hetb200 extension: fused log_softmax + nll_loss (sum / denom) of the training step.
void hetb200_softmax_cross_entropy(Tensor &logits, Tensor &labels, double denom, Tensor &lse, Tensor &loss) {
  const char *op = "hetb200_softmax_cross_entropy";
  TORCH_CHECK(logits.dim() == 2, op, ": logits must be [N,C]");
  c10::cuda::CUDAGuard guard(logits.device());
  check_rc(hetb_softmax_xent_fwd(f32(logits, op, "logits"), i64(labels, op, "labels"), logits.size(0),
                                 (int)logits.size(1), (float)denom, f32w(lse, op, "lse"), f32w(loss, op, "loss"),
                                 stream_of(logits)),
           op);
}
void hetb200_backward_softmax_cross_entropy(Tensor &logits, Tensor &labels, Tensor &lse, Tensor &grad_loss,
                                            double denom, Tensor &grad_logits) {
  const char *op = "hetb200_backward_softmax_cross_entropy";
  c10::cuda::CUDAGuard guard(logits.device());
  check_rc(hetb_softmax_xent_bwd(f32(logits, op, "logits"), i64(labels, op, "labels"), f32(lse, op, "lse"),
                                 f32(grad_loss, op, "grad_loss"), logits.size(0), (int)logits.size(1),
                                 (float)denom, f32w(grad_logits, op, "grad_logits"), stream_of(logits)),
           op);
}

// ---------------------------------------------------------- converters ---
// The reference requires CPU tensors (DataConverters.inc.h asserts
// device().is_cpu()); here CPU inputs are staged to the current CUDA device,
// converted there, and returned on the input's device.
struct Staged {
  std::vector<Tensor> dev;
  at::Device home;
};
Staged stage(std::initializer_list<Tensor> ts, const char *op) {
  Staged s{{}, ts.begin()->device()};
  for (const Tensor &t : ts) {
    TORCH_CHECK(t.scalar_type() == at::kLong, "torch_hrt::", op,
                ": index tensors must be int64");
    TORCH_CHECK(t.device() == s.home, "torch_hrt::", op,
                ": all index tensors must be on one device");
    s.dev.push_back(t.is_cuda() ? t.contiguous() : t.contiguous().cuda());
  }
  return s;
}
std::vector<Tensor> unstage(const Staged &s, std::vector<Tensor> out) {
  if (!s.home.is_cuda())
    for (auto &t : out) t = t.to(s.home);
  return out;
}

// out[j] = in[j] permuted by a stable sort on key; returns ptrs as out.back()
std::vector<Tensor> sort_by_key(const Tensor &key, int64_t num_keys,
                                const std::vector<Tensor> &payload,
                                const char *op) {
  const int64_t E = key.numel();
  c10::cuda::CUDAGuard guard(key.device());
  auto opts = key.options();
  std::vector<Tensor> out;
  const int64_t *in_ptrs[4] = {nullptr, nullptr, nullptr, nullptr};
  int64_t *out_ptrs[4] = {nullptr, nullptr, nullptr, nullptr};
  TORCH_CHECK(payload.size() <= 4, op, ": too many payload arrays");
  for (size_t j = 0; j < payload.size(); ++j) {
    out.push_back(at::empty({E}, opts));
    in_ptrs[j] = payload[j].data_ptr<int64_t>();
    out_ptrs[j] = out.back().data_ptr<int64_t>();
  }
  Tensor ptrs = at::empty({num_keys + 1}, opts);
  const int64_t bytes = hetb_sort_scratch_bytes(E, num_keys);
  Tensor scratch = at::empty({bytes}, opts.dtype(at::kByte));
  check_rc(hetb_stable_sort_by_key(key.data_ptr<int64_t>(), E, num_keys,
                                   (int)payload.size(), in_ptrs, out_ptrs,
                                   ptrs.data_ptr<int64_t>(), scratch.data_ptr(),
                                   bytes, stream_of(key)),
           op);
  out.push_back(ptrs);
  return out;
}

Tensor expand_rows(const Tensor &row_ptrs, int64_t E, const char *op) {
  c10::cuda::CUDAGuard guard(row_ptrs.device());
  Tensor rows = at::empty({E}, row_ptrs.options());
  check_rc(hetb_expand_row_ptr(row_ptrs.data_ptr<int64_t>(),
                               row_ptrs.numel() - 1, E,
                               rows.data_ptr<int64_t>(), stream_of(row_ptrs)),
           op);
  return rows;
}

// DataConverters.inc.h:283-344 → MyHyb.h:31-75 (stable by original position)
std::vector<Tensor> transpose_csr(Tensor &csr_row_ptrs, Tensor &csr_col_indices,
                                  Tensor &csr_reltypes, Tensor &csr_eids) {
  const char *op = "transpose_csr";
  Staged s = stage({csr_row_ptrs, csr_col_indices, csr_reltypes, csr_eids}, op);
  const int64_t E = s.dev[1].numel();
  const int64_t num_rows = s.dev[0].numel() - 1;
  Tensor rows = expand_rows(s.dev[0], E, op);
  auto r = sort_by_key(s.dev[1], num_rows, {rows, s.dev[2], s.dev[3]}, op);
  return unstage(s, {r[3], r[0], r[1], r[2]});
}

// DataConverters.inc.h:10-77 → MyHyb.h:1047-1097: order (rel, row, original).
std::vector<Tensor> convert_integrated_coo_to_separate_coo(
    Tensor &row_indices, Tensor &col_indices, Tensor &reltypes, Tensor &eids,
    int64_t num_rows, int64_t num_rels) {
  const char *op = "convert_integrated_coo_to_separate_coo";
  Staged s = stage({row_indices, col_indices, reltypes, eids}, op);
  auto a = sort_by_key(s.dev[0], num_rows,
                       {s.dev[0], s.dev[1], s.dev[2], s.dev[3]}, op);
  auto b = sort_by_key(a[2], num_rels, {a[0], a[1], a[3]}, op);
  return unstage(s, {b[3], b[0], b[1], b[2]});
}

// DataConverters.inc.h:216-281 → MyHyb.h:1099-1150
std::vector<Tensor> convert_integrated_csr_to_separate_coo(
    Tensor &row_ptrs, Tensor &col_indices, Tensor &reltypes, Tensor &eids) {
  const char *op = "convert_integrated_csr_to_separate_coo";
  Staged s = stage({row_ptrs, col_indices, reltypes, eids}, op);
  const int64_t E = s.dev[1].numel();
  // num_rels = max(rel)+1, as the reference infers it (MyHyb.h:347)
  const int64_t num_rels = E ? s.dev[2].max().item<int64_t>() + 1 : 0;
  Tensor rows = expand_rows(s.dev[0], E, op);
  auto b = sort_by_key(s.dev[2], num_rels, {rows, s.dev[1], s.dev[3]}, op);
  return unstage(s, {b[3], b[0], b[1], b[2]});
}

std::vector<Tensor> separate_csr_from_coo(const Staged &s, const Tensor &rows,
                                          const Tensor &cols, const Tensor &rels,
                                          const Tensor &eids, int64_t num_rows,
                                          int64_t num_rels, const char *op) {
  Tensor key = rels * num_rows + rows;
  auto b = sort_by_key(key, num_rows * num_rels, {cols, eids}, op);
  Tensor row_ptrs = b[2];
  Tensor rel_ptrs =
      row_ptrs.index({at::indexing::Slice(0, at::indexing::None, num_rows)})
          .contiguous();
  return unstage(s, {rel_ptrs, row_ptrs, b[0], b[1]});
}

// DataConverters.inc.h:79-146 → MyHyb.h:996-1045
std::vector<Tensor> convert_integrated_coo_to_separate_csr(
    Tensor &row_indices, Tensor &col_indices, Tensor &reltypes, Tensor &eids,
    int64_t num_rows, int64_t num_rels) {
  const char *op = "convert_integrated_coo_to_separate_csr";
  Staged s = stage({row_indices, col_indices, reltypes, eids}, op);
  return separate_csr_from_coo(s, s.dev[0], s.dev[1], s.dev[2], s.dev[3],
                               num_rows, num_rels, op);
}

// DataConverters.inc.h:148-214
std::vector<Tensor> convert_integrated_csr_to_separate_csr(
    Tensor &row_ptrs, Tensor &col_indices, Tensor &reltypes, Tensor &eids) {
  const char *op = "convert_integrated_csr_to_separate_csr";
  Staged s = stage({row_ptrs, col_indices, reltypes, eids}, op);
  const int64_t E = s.dev[1].numel();
  const int64_t num_rows = s.dev[0].numel() - 1;
  const int64_t num_rels = E ? s.dev[2].max().item<int64_t>() + 1 : 0;
  Tensor rows = expand_rows(s.dev[0], E, op);
  return separate_csr_from_coo(s, rows, s.dev[1], s.dev[2], s.dev[3], num_rows,
                               num_rels, op);
}

// hetb200 extension (not in the reference): stable sort of up to four int64
// payload arrays by an int64 key in [0, num_keys); returns the permuted
// payloads followed by the segment offsets ptrs[num_keys+1]. COO->CSR is
// sort_by_key(src, N, [dst, rel, eid]); relation sort is sort_by_key(rel, R, ...).
std::vector<Tensor> hetb200_stable_sort_by_key(Tensor key, int64_t num_keys,
                                               std::vector<Tensor> payload) {
  const char *op = "hetb200_stable_sort_by_key";
  std::vector<Tensor> all = {key};
  for (auto &t : payload) all.push_back(t);
  Staged s{{}, key.device()};
  for (const Tensor &t : all) {
    TORCH_CHECK(t.scalar_type() == at::kLong, op, ": tensors must be int64");
    TORCH_CHECK(t.numel() == key.numel(), op, ": payload length mismatch");
    s.dev.push_back(t.is_cuda() ? t.contiguous() : t.contiguous().cuda());
  }
  std::vector<Tensor> pl(s.dev.begin() + 1, s.dev.end());
  return unstage(s, sort_by_key(s.dev[0], num_keys, pl, op));
}

void hetb200_set_segment_mm_impl(int64_t impl) {
  check_rc(hetb_set_segment_mm_impl((int)impl), "hetb200_set_segment_mm_impl");
}

void hetb200_set_segment_mm_tma(int64_t on) {
  check_rc(hetb_set_segment_mm_tma((int)on), "hetb200_set_segment_mm_tma");
}

}  // namespace

TORCH_LIBRARY_FRAGMENT(torch_hrt, m) {
  m.def("hetb200_set_segment_mm_impl", hetb200_set_segment_mm_impl);
  m.def("hetb200_set_segment_mm_tma", hetb200_set_segment_mm_tma);
  m.def("build_debug_info", build_debug_info);
  m.def("hetb200_launch_count", hetb200_launch_count);
  m.def("hetb200_stable_sort_by_key", hetb200_stable_sort_by_key);
  // RGNN (OpExport/RGNNOps.inc.h:1189-1206)
  m.def("rgnn_relational_matmul", rgnn_relational_matmul);
  m.def("backward_rgnn_relational_matmul", backward_rgnn_relational_matmul);
  m.def("rgnn_relational_matmul_no_scatter_gather_list",
        rgnn_relational_matmul_no_scatter_gather_list);
  m.def("backward_rgnn_relational_matmul_no_scatter_gather_list",
        backward_rgnn_relational_matmul_no_scatter_gather_list);
  m.def("rgnn_inner_product_right_node_separatecoo",
        rgnn_inner_product_right_node_separatecoo);
  m.def("backward_inner_product_right_node_separatecoo",
        backward_inner_product_right_node_separatecoo);
  // RGAT (OpExport/RGATOps.inc.h:560-571). The reference's Python calls the
  // CSR ops without the trailing bool (backend/rgat_layers_and_funcs.py:49-63),
  // so it is given a default here.
  m.def(
      "relational_fused_gat_csr(Tensor(a!) incsr_row_ptr, Tensor(b!) "
      "incsr_col_indices, Tensor(c!) incsr_eids, Tensor(d!) incsr_reltypes, "
      "Tensor(e!) unique_srcs_and_dests_rel_ptrs, Tensor(f!) "
      "unique_srcs_and_dests_node_indices, Tensor(g!) feat_src, Tensor(h!) el, "
      "Tensor(i!) er, Tensor(j!) sum, Tensor(k!) exp, Tensor(l!) ret, float "
      "slope, bool CompactAsOfNodeFlag=False) -> ()",
      relational_fused_gat_csr);
  m.def(
      "backward_relational_fused_gat_csr(Tensor(a!) outcsr_row_ptr, Tensor(b!) "
      "outcsr_col_indices, Tensor(c!) outcsr_eids, Tensor(d!) outcsr_reltypes, "
      "Tensor(e!) unique_srcs_and_dests_rel_ptrs, Tensor(f!) "
      "unique_srcs_and_dests_node_indices, Tensor(g!) feat_src, Tensor(h!) el, "
      "Tensor(i!) er, Tensor(j!) sum, Tensor(k!) exp, Tensor(l!) ret, "
      "Tensor(m!) gradout, Tensor(n!) grad_feat_src, Tensor(o!) grad_el, "
      "Tensor(p!) grad_er, float slope, bool CompactAsOfNodeFlag=False) -> ()",
      backward_relational_fused_gat_csr);
  m.def("relational_fused_gat_separate_coo", relational_fused_gat_separate_coo);
  m.def("backward_relational_fused_gat_separate_coo",
        backward_relational_fused_gat_separate_coo);
  // HGT (OpExport/HGTOps.inc.h:656-674)
  m.def("hetb200_hgt_fused_attention_aggregation_csr", hetb200_hgt_fused_attention_aggregation_csr);
  m.def("hetb200_backward_hgt_fused_softmax_aggregation_csr", hetb200_backward_hgt_fused_softmax_aggregation_csr);
  m.def("hetb200_backward_hgt_fused_attention_aggregation_csr",
        hetb200_backward_hgt_fused_attention_aggregation_csr);
  // P2 names: registered for interface compatibility, outside the hot path; calling one
  // raises with the op's name (OpExport/RGCNOps.inc.h:681-699, GATOps.inc.h:130-135,
  // UtilityAndPlayground.inc.h:101-108)
  m.def("seastar_rgcn_layer1_coo", seastar_rgcn_layer1_coo);
  m.def("seastar_backward_rgcn_layer1_coo", seastar_backward_rgcn_layer1_coo);
  m.def("seastar_rgcn_layer1_csr", seastar_rgcn_layer1_csr);
  m.def("seastar_backward_rgcn_layer1_csr", seastar_backward_rgcn_layer1_csr);
  m.def("seastar_rgcn_layer0_csr", seastar_rgcn_layer0_csr);
  m.def("seastar_backward_rgcn_layer0_csr", seastar_backward_rgcn_layer0_csr);
  m.def("fused_gat_csr", fused_gat_csr);
  m.def("backward_fused_gat_csr", backward_fused_gat_csr);
  // P1 ops (OpExport/HGTOpsEdgeParallel.inc.h:380-400, HGTOps.inc.h:666-673, RGNNOps.inc.h:1202)
  m.def("hgt_full_graph_hetero_attention_ops_coo", hgt_full_graph_hetero_attention_ops_coo);
  m.def("backward_hgt_full_graph_hetero_attention_ops_coo", backward_hgt_full_graph_hetero_attention_ops_coo);
  m.def("hgt_full_graph_edge_softmax_ops_separate_coo", hgt_full_graph_edge_softmax_ops_separate_coo);
  m.def("hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo",
        hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo);
  m.def("backward_hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo",
        backward_hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo);
  m.def("backward_hgt_full_graph_enorm_to_unnormalized_attn_score_csr",
        backward_hgt_full_graph_enorm_to_unnormalized_attn_score_csr);
  m.def("backward_hgt_full_graph_enorm_to_unnormalized_attn_score_separate_coo",
        backward_hgt_full_graph_enorm_to_unnormalized_attn_score_separate_coo);
  m.def("rgnn_gather_compact_to_non_compact", rgnn_gather_compact_to_non_compact);
  m.def("hetb200_gather_rows", hetb200_gather_rows);
  m.def("hetb200_scatter_add_rows", hetb200_scatter_add_rows);
  m.def("hetb200_adam_bump_step", hetb200_adam_bump_step);
  m.def("hetb200_adam_step", hetb200_adam_step);
  m.def("hetb200_compose_weights", hetb200_compose_weights);
  m.def("hetb200_backward_compose_weights", hetb200_backward_compose_weights);
  m.def("hetb200_typed_linear_bf16out", hetb200_typed_linear_bf16out);
  m.def("hetb200_hgt_fused_fwd_plan", hetb200_hgt_fused_fwd_plan);
  m.def("hetb200_hgt_fused_bwd_plan", hetb200_hgt_fused_bwd_plan);
  m.def("hetb200_hgt_fused_bwd_plan2", hetb200_hgt_fused_bwd_plan2);
  m.def("hetb200_rgat_fused_fwd_plan", hetb200_rgat_fused_fwd_plan);
  m.def("hetb200_rgat_fused_bwd_plan", hetb200_rgat_fused_bwd_plan);
  m.def("hetb200_column_sum", hetb200_column_sum);
  m.def("hetb200_backward_rgnn_relational_matmul_no_scatter_gather_list_accumulate",
        hetb200_backward_rgnn_relational_matmul_no_scatter_gather_list_accumulate);
  m.def("hetb200_softmax_cross_entropy", hetb200_softmax_cross_entropy);
  m.def("hetb200_backward_softmax_cross_entropy", hetb200_backward_softmax_cross_entropy);
  m.def("hetb200_relational_fused_gat_csr_compact", hetb200_relational_fused_gat_csr_compact);
  m.def("hetb200_backward_relational_fused_gat_csr_compact", hetb200_backward_relational_fused_gat_csr_compact);
  m.def("hetb200_backward_relational_fused_gat_compact_rows", hetb200_backward_relational_fused_gat_compact_rows);
  m.def("hetb200_hgt_message_aggregation_csr_compact", hetb200_hgt_message_aggregation_csr_compact);
  // RGCN (OpExport/RGCNOps.inc.h:666-678)
  m.def("rgcn_layer1_separate_coo", rgcn_layer1_separate_coo);
  m.def("backward_rgcn_layer1_separate_coo", backward_rgcn_layer1_separate_coo);
  m.def("rgcn_node_mean_aggregation_compact_as_of_node_separate_coo",
        rgcn_node_mean_aggregation_compact_as_of_node_separate_coo);
  m.def("backward_rgcn_node_mean_aggregation_compact_as_of_node_separate_coo",
        backward_rgcn_node_mean_aggregation_compact_as_of_node_separate_coo);
  m.def("hetb200_rgcn_node_aggregation_csr_compact", hetb200_rgcn_node_aggregation_csr_compact);
  m.def("hetb200_backward_hgt_message_aggregation_csr_compact",
        hetb200_backward_hgt_message_aggregation_csr_compact);
  m.def("hetb200_backward_hgt_edge_softmax_ops_csr_compact", hetb200_backward_hgt_edge_softmax_ops_csr_compact);
  m.def("hgt_full_graph_edge_softmax_ops_csr",
        hgt_full_graph_edge_softmax_ops_csr);
  m.def("hgt_full_graph_message_mean_aggregation_csr",
        hgt_full_graph_message_mean_aggregation_csr);
  m.def("backward_hgt_full_graph_message_mean_aggregation_csr",
        backward_hgt_full_graph_message_mean_aggregation_csr);
  m.def("backward_hgt_full_graph_edge_softmax_ops_csr",
        backward_hgt_full_graph_edge_softmax_ops_csr);
  // Data converters (OpExport/DataConverters.inc.h:346-357)
  m.def("transpose_csr", transpose_csr);
  m.def("convert_integrated_csr_to_separate_csr",
        convert_integrated_csr_to_separate_csr);
  m.def("convert_integrated_csr_to_separate_coo",
        convert_integrated_csr_to_separate_coo);
  m.def("convert_integrated_coo_to_separate_csr",
        convert_integrated_coo_to_separate_csr);
  m.def("convert_integrated_coo_to_separate_coo",
        convert_integrated_coo_to_separate_coo);
}
