// TEST INFRASTRUCTURE ONLY — C launchers around the REFERENCE's own CUDA kernel templates.
//
// Nothing of the reference is copied: this file #includes the kernel headers where they lie under
// /root/reference/hrt/include (see oracle/build_ref.py: build_kernels) and instantiates them with
// exactly the template arguments, argument order and launch schedules the reference's torch op
// wrappers use (file:line cited per launcher). The result, oracle/_ref/libhet_ref_kernels.so, is
// compiled for sm_100a in the build container and travels to the GPU box, where the `-m gpu`
// tests run the reference kernels, the fp64 oracle (oracle/het_ops_ref.py) and the product
// (libhetb200.so) on the same inputs. It is never linked into or loaded by het_b200/.
//
// All pointers are device pointers; index tensors int64, data fp32 (the reference's only types).
// Every launcher returns the cudaError_t of its launches (0 = ok) after a stream synchronise.
#include <cuda_runtime.h>
#include <stdint.h>

#include <thrust/device_vector.h>

#include <algorithm>
#include <tuple>
#include <vector>

#include "utils.cu.h"
#include "kernel_enums.h"
#include "macros.h"
#include "DGLHackKernel/DGLHackUtils.h"
#include "DGLHackKernel/GAT/FusedGAT.cu.h"
#include "DGLHackKernel/GAT/FusedGATBackward.cu.h"
#include "DGLHackKernel/HGT/HGTForwardKernels.cu.h"
#include "DGLHackKernel/HGT/HGTBackwardKernels.cu.h"
#include "DGLHackKernel/RGNN/InnerProduct.cu.h"
#include "DGLHackKernel/RGNN/InnerProductEdgeParallel.cu.h"
#include "DGLHackKernel/RGNN/my_shmem_sgemm_func.cu.h"
#include "DGLHackKernel/RGNN/mysgemm_KernelsBlockConfigurations.h"
#include "DGLHackKernel/OpExport/ThreadingGridsBlocksSchedules.h"

typedef int64_t Idx;

static int finish(cudaStream_t st) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  e = cudaStreamSynchronize(st);
  return (int)e;
}

// ---------------------------------------------------------------------------
// RGAT forward, integrated in-CSR: OpExport/RGATOps.inc.h:50-101 (kind Disabled, RelationalFlag
// true), kernels GAT/FusedGAT.cu.h:215 and :109. The reference's Python hands `sum` over
// uninitialised although the kernel atomicAdds into it (backend/rgat_layers_and_funcs.py:658-660);
// the caller of this launcher zero-fills it.
// ---------------------------------------------------------------------------
extern "C" int ref_rgat_fwd_csr(Idx *row_ptr, Idx *col, Idx *eids, Idx *rel, int64_t num_rows, float *feat,
                                float *el, float *er, float *sum, float *exp_out, float *ret, int64_t H,
                                int64_t D, float slope, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GatFusedData<Idx, float> gdata{};
  gdata.feat_src_xlen = D;
  gdata.num_heads = H;
  gdata.eids = eids;
  gdata.leaky_relu_slope = slope;
  gdata.feat_src = feat;
  gdata.el = el;
  gdata.er = er;
  gdata.sum = sum;
  gdata.exp = exp_out;
  gdata.ret = ret;
  ETypeMapperData<Idx, CompactAsOfNodeKind::Disabled> mapper;
  ETypeData<Idx, false> etype_data{};
  etype_data.etypes = rel;
  auto [nblks, nthrs] = get_type1_schedule(H, num_rows);
  HET_gatExpLeakyReluSumKernel<Idx, float, CompactAsOfNodeKind::Disabled, true>
      <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, mapper);
  auto [nblks2, nthrs2] = get_type2_schedule(H, D, num_rows);
  HET_gatSumProdZipDivKernel<Idx, float, CompactAsOfNodeKind::Disabled, true>
      <<<nblks2, nthrs2, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, mapper);
  return finish(st);
}

// RGAT backward, integrated out-CSR: OpExport/RGATOps.inc.h:317-361 (FLAG_KERNEL_FUSED true),
// kernel GAT/FusedGATBackward.cu.h:140. grad_* must arrive zero-filled (the kernel atomicAdds).
extern "C" int ref_rgat_bwd_csr(Idx *row_ptr, Idx *col, Idx *eids, Idx *rel, int64_t num_rows, float *feat,
                                float *el, float *er, float *sum, float *exp_in, float *ret, float *gout,
                                float *grad_feat, float *grad_el, float *grad_er, int64_t H, int64_t D,
                                float slope, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BackwardGatFusedData<Idx, float> gdata{};
  gdata.feat_src_xlen = D;
  gdata.num_heads = H;
  gdata.eids = eids;
  gdata.leaky_relu_slope = slope;
  gdata.feat_src = feat;
  gdata.el = el;
  gdata.er = er;
  gdata.sum = sum;
  gdata.exp = exp_in;
  gdata.ret = ret;
  gdata.grad_out = gout;
  gdata.grad_feat_src = grad_feat;
  gdata.grad_el = grad_el;
  gdata.grad_er = grad_er;
  ETypeMapperData<Idx, CompactAsOfNodeKind::Disabled> mapper;
  ETypeData<Idx, false> etype_data{};
  etype_data.etypes = rel;
  auto [nblks, nthrs] = get_type2_schedule(H, D, num_rows);
  HET_fusedGatBackwardGradElErFeatSrcFused<Idx, float, CompactAsOfNodeKind::Disabled, true>
      <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, mapper);
  return finish(st);
}

// ---------------------------------------------------------------------------
// HGT edge softmax forward, in-CSR: OpExport/HGTOps.inc.h:23-106 with
// OutputMuAppliedAttnScoreSwitch = 3 (full_graph_edge_softmax_ops, :190-203), kernel
// HGT/HGTForwardKernels.cu.h:428. `sum` is accumulated atomically: caller zero-fills.
// ---------------------------------------------------------------------------
extern "C" int ref_hgt_edge_softmax_fwd_csr(Idx *row_ptr, Idx *col, Idx *eids, Idx *rel, int64_t num_rows,
                                            float *attn, float *mu, float *sum, float *mu_applied,
                                            float *norm, int64_t H, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  HgtEdgeSoftmaxAccumData<Idx, float, 3> gdata{};
  gdata.num_heads = H;
  gdata.eids = eids;
  gdata.mu = mu;
  gdata.unnormalized_attn_score = attn;
  gdata.edgesoftmax_sum_per_node = sum;
  gdata.mu_softmax_applied_unnormalized_attn_score = mu_applied;
  gdata.normalized_attn_score = norm;
  auto [nblks, nthrs] = get_type1_schedule(H, num_rows);
  ETypeData<Idx, false> etype_data{};
  etype_data.etypes = rel;
  HET_HGTEdgeSoftmaxAccumStageOnlyKernel<Idx, float, CompactAsOfNodeKind::Disabled, true, false, false, 3>
      <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, {});
  return finish(st);
}

// HGT message aggregation forward, in-CSR: OpExport/HGTOps.inc.h:109-188, kernel
// HGT/HGTForwardKernels.cu.h:247. use_switch 1 = what the exported op instantiates (divides its
// score argument by sum[dst]); 2 = the already-normalised score is used as is.
extern "C" int ref_hgt_aggregate_fwd_csr(Idx *row_ptr, Idx *col, Idx *rel, Idx *eids, int64_t num_rows,
                                         float *msg, float *score, float *sum, float *ret, int64_t H,
                                         int64_t D, int use_switch, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  auto [nblks, nthrs] = get_type2_schedule(H, D, num_rows);
  ETypeData<Idx, false> etype_data{};
  etype_data.etypes = rel;
  if (use_switch == 1) {
    HgtDstOutData<Idx, float, 1> gdata{};
    gdata.num_heads = H;
    gdata.message_out_dim = D;
    gdata.eids = eids;
    gdata.edgesoftmax_sum_per_node = sum;
    gdata.message = msg;
    gdata.ret = ret;
    gdata.mu_softmax_applied_unnormalized_attn_score = score;
    HET_HGTMessageAccumBasedOnOriAttnScoreAndEdgeSoftmaxSum<Idx, float, CompactAsOfNodeKind::Disabled, true,
                                                            false, false, 1>
        <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, {});
  } else {
    HgtDstOutData<Idx, float, 2> gdata{};
    gdata.num_heads = H;
    gdata.message_out_dim = D;
    gdata.eids = eids;
    gdata.message = msg;
    gdata.ret = ret;
    gdata.normalized_attn_score = score;
    HET_HGTMessageAccumBasedOnOriAttnScoreAndEdgeSoftmaxSum<Idx, float, CompactAsOfNodeKind::Disabled, true,
                                                            false, false, 2>
        <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, {});
  }
  return finish(st);
}

// HGT message aggregation backward, out-CSR: OpExport/HGTOps.inc.h:410-487 (switch 2), kernel
// HGT/HGTBackwardKernels.cu.h:344. grad_msg arrives zero-filled.
extern "C" int ref_hgt_aggregate_bwd_csr(Idx *row_ptr, Idx *col, Idx *rel, Idx *eids, int64_t num_rows,
                                         float *norm, float *gout, float *grad_msg, int64_t H, int64_t D,
                                         void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BackwardHGTMessageData<Idx, float, 2> gdata{};
  gdata.num_heads = H;
  gdata.message_src_xlen = D;
  gdata.eids = eids;
  gdata.grad_message_src = grad_msg;
  gdata.grad_out = gout;
  gdata.normalized_attn_score = norm;
  auto [nblks, nthrs] = get_type2_schedule(H, D, num_rows);
  ETypeData<Idx, false> etype_data{};
  etype_data.etypes = rel;
  HET_HGTMessageAccumBasedOnOriAttnScoreAndEdgeSoftmaxSumBackwardKernel<Idx, float, CompactAsOfNodeKind::Disabled,
                                                                        true, false, 2>
      <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, {});
  return finish(st);
}

// HGT edge softmax backward, out-CSR: OpExport/HGTOps.inc.h:489-566 (switch 2), kernel
// HGT/HGTBackwardKernels.cu.h:499. grad_attn / grad_mu arrive zero-filled. NB the kernel never
// reads the relation of an edge in the non-compact path (etype stays 0, :533), so mu[0] scales
// every edge and all of grad_mu lands in relation 0: the tests compare with that quirk applied.
extern "C" int ref_hgt_edge_softmax_bwd_csr(Idx *row_ptr, Idx *col, Idx *eids, Idx *rel, int64_t num_rows,
                                            float *msg, float *attn, float *norm, float *out, float *gout,
                                            float *mu, float *grad_attn, float *grad_mu, int64_t H, int64_t D,
                                            void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BackwardHGTAttnScoreData<Idx, float, 2> gdata{};
  gdata.num_heads = H;
  gdata.message_src_xlen = D;
  gdata.eids = eids;
  gdata.grad_attn_score = grad_attn;
  gdata.message_src = msg;
  gdata.unnormalized_attn_score = attn;
  gdata.out = out;
  gdata.grad_out = gout;
  gdata.grad_mu = grad_mu;
  gdata.mu = mu;
  gdata.normalized_attn_score = norm;
  auto [nblks, nthrs] = get_type2_schedule(H, D, num_rows);
  ETypeData<Idx, false> etype_data{};
  etype_data.etypes = rel;
  HET_HGTEdgeSoftmaxAccumStageOnlyBackwardKernel<Idx, float, CompactAsOfNodeKind::Disabled, true, false, 2>
      <<<nblks, nthrs, 0, st>>>(gdata, row_ptr, col, etype_data, num_rows, {});
  return finish(st);
}

// ---------------------------------------------------------------------------
// Inner product (edge data x node vectors), separate COO: OpExport/RGNNOps.inc.h:307-444 (kind
// Disabled), kernel RGNN/InnerProductEdgeParallel.cu.h:13. The exported op memsets the output
// first (:617-619); so does this launcher. NB the F < 32 branch of the kernel reduces with
// __shfl_down starting at hidden_xlen and mixes heads (:100-106): only F >= 32 is comparable.
// ---------------------------------------------------------------------------
extern "C" int ref_inner_product_fwd_coo(Idx *rel_ptrs, Idx *eids, Idx *row_idx, Idx *col_idx,
                                         int64_t num_edges, int64_t num_rels, float *left, float *right,
                                         float *out, int64_t H, int64_t D, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(out, 0, sizeof(float) * num_edges * H, st);
  InnerProductData<Idx, float> gdata{};
  gdata.feat_src_xlen = D;
  gdata.num_heads = H;
  gdata.eids = eids;
  gdata.feat_src = left;
  gdata.feat_dst = right;
  gdata.edge_inner_product = out;
  auto [nblks, nthrs] = get_type2_schedule(H, D, num_edges);
  ETypeMapperData<Idx, CompactAsOfNodeKind::Disabled> mapper;
  ETypeData<Idx, true> etype_data{};
  etype_data.etypes = rel_ptrs;
  etype_data.num_relations = num_rels;
  if (D / H >= 32) {
    HET_inner_product_fw_kernel_edge_parallel<Idx, float, CompactAsOfNodeKind::Disabled, true, true, false, true>
        <<<nblks, nthrs, 0, st>>>(gdata, row_idx, col_idx, etype_data, num_edges, mapper);
  } else {
    HET_inner_product_fw_kernel_edge_parallel<Idx, float, CompactAsOfNodeKind::Disabled, true, true, false, false>
        <<<nblks, nthrs, 0, st>>>(gdata, row_idx, col_idx, etype_data, num_edges, mapper);
  }
  return finish(st);
}

// Inner product backward, separate COO: OpExport/RGNNOps.inc.h:1020-1130, kernel
// RGNN/InnerProductEdgeParallel.cu.h:119. NB the kernel indexes the node side by col_indices
// although the forward used row_indices, and STORES grad_feat_dst (:180-183): comparable only on
// inputs where row == col and every node has one edge.
extern "C" int ref_inner_product_bwd_coo(Idx *rel_ptrs, Idx *eids, Idx *row_idx, Idx *col_idx,
                                         int64_t num_edges, int64_t num_rels, float *left, float *right,
                                         float *gout, float *grad_left, float *grad_right, int64_t H,
                                         int64_t D, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  BackwardInnerProductData<Idx, float> gdata{};
  gdata.feat_src_xlen = D;
  gdata.num_heads = H;
  gdata.eids = eids;
  gdata.feat_src = left;
  gdata.feat_dst = right;
  gdata.grad_inner_product = gout;
  gdata.grad_feat_dst = grad_right;
  gdata.grad_feat_src = grad_left;
  auto [nblks, nthrs] = get_type2_schedule(H, D, num_edges);
  ETypeMapperData<Idx, CompactAsOfNodeKind::Disabled> mapper;
  ETypeData<Idx, true> etype_data{};
  etype_data.etypes = rel_ptrs;
  etype_data.num_relations = num_rels;
  HET_inner_product_bck_kernel_edge_parallel<Idx, float, CompactAsOfNodeKind::Disabled, true, true>
      <<<nblks, nthrs, 0, st>>>(gdata, row_idx, col_idx, etype_data, num_edges, mapper);
  return finish(st);
}

// ---------------------------------------------------------------------------
// Segment MM. The block schedule is the reference's own host function
// (RGNN/mysgemm_KernelsBlockConfigurations.h:42-144) fed with the segment offsets copied to the
// host, as the op wrappers do with `rel_ptrs.cpu()` (OpExport/RGNNOps.inc.h:150-169).
// ---------------------------------------------------------------------------
struct RefSchedule {
  thrust::device_vector<int> all_prev;
  int grid_dim_y;
};

static RefSchedule make_schedule(const Idx *seg_ptrs_dev, int64_t num_segs, int64_t num_rows, int block_size,
                                 int64_t cap) {
  std::vector<int64_t> host(num_segs + 1);
  cudaMemcpy(host.data(), seg_ptrs_dev, sizeof(int64_t) * (num_segs + 1), cudaMemcpyDeviceToHost);
  int grid_dim_y = (int)std::min(ceil_div<int64_t>(num_rows, (int64_t)block_size), cap);
  std::vector<int> same, all_prev;
  std::tie(same, all_prev) = get_schedule_by_relation_kernel_launch_metadata<false, false, int64_t *>(
      grid_dim_y, (int)num_segs, block_size, host.data(), host.data() + num_segs + 1);
  RefSchedule s;
  s.grid_dim_y = all_prev.back();
  s.all_prev = thrust::device_vector<int>(all_prev.begin(), all_prev.end());
  return s;
}

// rgnn_relational_matmul, kind Disabled, gather list != scatter list:
// OpExport/RGNNOps.inc.h:95-236, kernel RGNN/my_shmem_sgemm_func.cu.h:504. ret arrives zero-filled
// (backend/rgnn_layers_and_funcs.py allocates it with th.zeros).
extern "C" int ref_relational_matmul_fwd(Idx *rel_ptrs, Idx *gather, Idx *eids, int64_t num_rels,
                                         int64_t num_edges, float *weights, float *node_feat, float *ret,
                                         int64_t H, int64_t K, int64_t N, int input_num_head_one, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  constexpr bool REG_TILING_FLAG = true;
  MY_SGEMM_GRID_CONFIG()
  RefSchedule s = make_schedule(rel_ptrs, num_rels, num_edges, WORK_BLOCK_SIZE_Y, 32768);
  const dim3 nblks(ceil_div<int64_t>(N, (int64_t)WORK_BLOCK_SIZE_X), s.grid_dim_y, H);
  const dim3 nthrs(THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y);
  int *sched = thrust::raw_pointer_cast(s.all_prev.data());
  if (input_num_head_one) {
    HET_RGNNFeatPerEdgeFwProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_X,
                              WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t, int64_t *, true>
        <<<nblks, nthrs, 0, st>>>(node_feat, weights, ret, gather, rel_ptrs, eids, K, N, (int)H, sched, num_rels);
  } else {
    HET_RGNNFeatPerEdgeFwProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_X,
                              WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t, int64_t *, false>
        <<<nblks, nthrs, 0, st>>>(node_feat, weights, ret, gather, rel_ptrs, eids, K, N, (int)H, sched, num_rels);
  }
  return finish(st);
}

// backward_rgnn_relational_matmul, kind Disabled: OpExport/RGNNOps.inc.h:756-944
// (WORK_BLOCK_SIZE 32, grid cap 4096), kernels RGNN/my_shmem_sgemm_func.cu.h:711 and :746.
// weights_t is W.transpose(2,3).contiguous(); grad_* arrive zero-filled.
extern "C" int ref_relational_matmul_bwd(Idx *rel_ptrs, Idx *gather, Idx *eids, int64_t num_rels,
                                         int64_t num_edges, float *weights_t, float *node_feat, float *gradout,
                                         float *grad_node_feat, float *grad_weights, int64_t H, int64_t K,
                                         int64_t N, int input_num_head_one, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  constexpr bool REG_TILING_FLAG = true;
  MY_SGEMM_GRID_CONFIG()
  RefSchedule s = make_schedule(rel_ptrs, num_rels, num_edges, 32, 4096);
  const dim3 nblks(ceil_div<int64_t>(K, (int64_t)WORK_BLOCK_SIZE_X), s.grid_dim_y, H);
  const dim3 nblks_outer(ceil_div<int64_t>(N, (int64_t)WORK_BLOCK_SIZE_X),
                         ceil_div<int64_t>(K, (int64_t)WORK_BLOCK_SIZE_Y), H * s.grid_dim_y);
  const dim3 nthrs(THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y);
  int *sched = thrust::raw_pointer_cast(s.all_prev.data());
  if (input_num_head_one) {
    HET_RGNNDeltaNodeFeatInputBckProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y,
                                      WORK_BLOCK_SIZE_X, WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t, int64_t *,
                                      true>
        <<<nblks, nthrs, 0, st>>>(gradout, weights_t, grad_node_feat, eids, rel_ptrs, gather, N, K, (int)H, sched,
                                  num_rels);
    HET_RGNNDeltaWeightBckProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_X,
                               WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t, int64_t *, true>
        <<<nblks_outer, nthrs, 0, st>>>(node_feat, gradout, grad_weights, gather, rel_ptrs, eids, K, N, (int)H,
                                        sched, num_rels);
  } else {
    HET_RGNNDeltaNodeFeatInputBckProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y,
                                      WORK_BLOCK_SIZE_X, WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t, int64_t *,
                                      false>
        <<<nblks, nthrs, 0, st>>>(gradout, weights_t, grad_node_feat, eids, rel_ptrs, gather, N, K, (int)H, sched,
                                  num_rels);
    HET_RGNNDeltaWeightBckProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_X,
                               WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t, int64_t *, false>
        <<<nblks_outer, nthrs, 0, st>>>(node_feat, gradout, grad_weights, gather, rel_ptrs, eids, K, N, (int)H,
                                        sched, num_rels);
  }
  return finish(st);
}

// rgnn_relational_matmul_no_scatter_gather_list: OpExport/RGNNOps.inc.h:21-82, kernel
// RGNN/my_shmem_sgemm_func.cu.h:538. ret arrives zero-filled.
extern "C" int ref_matmul_no_scatter_gather_fwd(Idx *offsets, int64_t num_types, int64_t num_rows, float *weights,
                                                float *inputs, float *ret, int64_t H, int64_t K, int64_t N,
                                                void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  constexpr bool REG_TILING_FLAG = true;
  MY_SGEMM_GRID_CONFIG()
  RefSchedule s = make_schedule(offsets, num_types, num_rows, WORK_BLOCK_SIZE_Y, 32768);
  const dim3 nblks(ceil_div<int64_t>(N, (int64_t)WORK_BLOCK_SIZE_X), s.grid_dim_y, H);
  const dim3 nthrs(THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y);
  HET_RGNNMatmulNoScatterGatherListFwProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y,
                                          WORK_BLOCK_SIZE_X, WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t,
                                          int64_t *>
      <<<nblks, nthrs, 0, st>>>(inputs, weights, ret, offsets, thrust::raw_pointer_cast(s.all_prev.data()),
                                num_types, (int)H, K, N);
  return finish(st);
}

// backward_rgnn_relational_matmul_no_scatter_gather_list: OpExport/RGNNOps.inc.h:664-743, kernels
// RGNN/my_shmem_sgemm_func.cu.h:566 and :598. grad_* arrive zero-filled.
extern "C" int ref_matmul_no_scatter_gather_bwd(Idx *offsets, int64_t num_types, int64_t num_rows,
                                                float *weights_t, float *inputs, float *gradout,
                                                float *grad_inputs, float *grad_weights, int64_t H, int64_t K,
                                                int64_t N, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  constexpr bool REG_TILING_FLAG = true;
  MY_SGEMM_GRID_CONFIG()
  RefSchedule s = make_schedule(offsets, num_types, num_rows, WORK_BLOCK_SIZE_Y, 32768);
  int *sched = thrust::raw_pointer_cast(s.all_prev.data());
  const dim3 nblks(ceil_div<int64_t>(K, (int64_t)WORK_BLOCK_SIZE_X), s.grid_dim_y, H);
  const dim3 nthrs(THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y);
  HET_RGNNMatmulNoScatterGatherDeltaFeatBckProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y,
                                                WORK_BLOCK_SIZE_X, WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t,
                                                int64_t *>
      <<<nblks, nthrs, 0, st>>>(gradout, weights_t, grad_inputs, offsets, sched, num_types, (int)H, K, N);
  const dim3 nblks_outer(ceil_div<int64_t>(N, (int64_t)WORK_BLOCK_SIZE_X),
                         ceil_div<int64_t>(K, (int64_t)WORK_BLOCK_SIZE_Y), s.grid_dim_y);
  HET_RGNNDeltaWeightNoScatterGatherListBckProp<REG_TILING_FLAG, THREADING_BLOCK_SIZE_X, THREADING_BLOCK_SIZE_Y,
                                                WORK_BLOCK_SIZE_X, WORK_BLOCK_SIZE_Y, WORK_BLOCK_SIZE_K, int64_t,
                                                int64_t *>
      <<<nblks_outer, nthrs, 0, st>>>(inputs, gradout, grad_weights, offsets, K, N, (int)H, sched, num_types);
  return finish(st);
}
