/*
 * hetb200.h — C ABI of libhetb200.so: B200-native (sm_100a) kernels for the
 * relational-GNN hot path of K-Wu/HET (Hector).
 *
 * This is the drop-in boundary. Every entry point takes plain device pointers
 * and sizes (no torch types), runs on the CUDA stream passed as `stream`
 * (a cudaStream_t cast to void*), enqueues work and returns without
 * synchronising. Return value: 0 on success, a negative HETB_E* code on bad
 * arguments, a positive cudaError_t if a launch failed.
 *
 * The torch shim (het_b200/csrc/torch_bindings.cpp) registers the SAME op
 * names the reference registers under `torch.ops.torch_hrt.*` and forwards
 * to these functions. Each declaration cites the reference interface it
 * replaces (paths relative to the reference repo root).
 *
 * Conventions (same as the reference):
 *   - all index arrays are int64, contiguous, device resident;
 *   - all float arrays are fp32, contiguous, device resident;
 *   - "row" = source node, "col" = destination node in the original
 *     orientation; in-CSR rows are destinations, out-CSR rows are sources;
 *   - `eids[p]` maps position p of a format to the row of a per-edge tensor;
 *   - H = heads, F = features per head, D = H*F.
 */
#ifndef HETB200_H_
#define HETB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HETB_OK 0
#define HETB_EINVAL (-1)      /* null pointer / negative size */
#define HETB_EUNSUPPORTED (-2) /* shape or kind outside what is implemented */
#define HETB_ENOMEM (-3)      /* scratch allocation failed */
#define HETB_ETOOLARGE (-4)   /* size beyond the fixed capacity of an allocation-free entry point */

/* ABI version and a printable build string (replaces
 * hrt/buildutils/genutils/gen_headers.py:17-40 `build_debug_info`). */
int hetb_abi_version(void);
const char *hetb_build_info(void);
/* Number of kernel launches issued by this library since process start
 * (bench.py reports it as `gpu_launches`). */
int64_t hetb_launch_count(void);

/* ------------------------------------------------------------------------
 * Relation-typed linear layer (segment MM, gather-MM-scatter)
 * ---------------------------------------------------------------------- */

/* Y[c_idx?[i], h, :] (=|+=) A[a_idx?[i], (a_heads==1 ? 0 : h), :] · W[s, h, :, :]
 * for every segment s and every row i in [seg_ptrs[s], seg_ptrs[s+1]).
 *   A  : [*, a_heads, K]      W : [S, H, K, N]      Y : [*, c_heads, N]
 *   a_idx / c_idx may be NULL (identity).  c_heads is H, or 1 with
 *   accumulate!=0, in which case the H per-head results are summed.
 *   accumulate!=0 adds into Y with atomics (rows may repeat in c_idx).
 * Replaces: rgnn_relational_matmul (OpExport/RGNNOps.inc.h:238-295 →
 * RGNN/my_shmem_sgemm_func.cu.h:504-531), its ACGatherScatterListIdentical
 * twin (:629-664), rgnn_relational_matmul_no_scatter_gather_list
 * (RGNNOps.inc.h:21-88 → my_shmem_sgemm_func.cu.h:538-564) and the delta-
 * input halves of their backward ops (RGNNOps.inc.h:946-1010, 664-753 →
 * my_shmem_sgemm_func.cu.h:566-595, 711-745), which are the same contraction
 * with the pre-transposed weight and the index roles swapped. */
int hetb_segment_mm(const float *A, const int64_t *a_idx, int a_heads,
                    const float *W, float *Y, const int64_t *c_idx,
                    int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                    int64_t num_rows, int H, int K, int N, int accumulate,
                    void *stream);

/* hetb_segment_mm with Y stored as bf16 ([*, H, N] __nv_bfloat16, plain store): the tables the
 * plan-based edge passes gather (north star: "2e-2 in bf16"). fp32 inputs, 3xTF32 products and fp32
 * accumulation as above; only the final store rounds. Tensor-core shapes only (H*K' = H*N = 64);
 * HETB_EUNSUPPORTED otherwise (no fallback). */
int hetb_segment_mm_bf16out(const float *A, const int64_t *a_idx, int a_heads, const float *W, void *Y_bf16,
                            const int64_t *c_idx, const int64_t *seg_ptrs, int64_t num_segs, int64_t num_rows,
                            int H, int K, int N, void *stream);

/* Implementation switch for hetb_segment_mm (testing / A-B measurement): 0 = automatic
 * (tcgen05 tensor-core kernel when the shape allows, else SIMT), 1 = SIMT kernels only,
 * 2 = tensor-core kernel only (HETB_EUNSUPPORTED if the shape is outside it). */
int hetb_set_segment_mm_impl(int impl);

/* A-operand path of the 64-wide tensor-core kernel (testing / A-B measurement): 1 (default) = the
 * 128 x 64 tile of A reaches shared memory by TMA (cp.async.bulk.tensor tile copies for contiguous rows,
 * tile::gather4 for gathered rows; plain whole-tile outputs leave by TMA too), 0 = producer warps load the
 * rows through registers. Same products bit for bit. Replaces the shared-memory staging loops of
 * RGNN/my_shmem_sgemm_func.cu.h:60-140 (no TMA in the reference). */
int hetb_set_segment_mm_tma(int on);

/* dW[s, h, :, :] += Σ_{i in segment s} A[a_idx?[i], (a_heads==1?0:h), :]^T
 *                                     · G[g_idx?[i], (g_heads==1?0:h), :]
 *   A : [*, a_heads, K]   G : [*, g_heads, N]   dW : [S, H, K, N] (caller
 *   zero-fills, as the reference's Python does: backend/rgnn_layers_and_funcs.py:52-55).
 * Replaces the delta-weight halves of the backward ops
 * (RGNNOps.inc.h:756-944 → my_shmem_sgemm_func.cu.h:597-628, 747-776). */
int hetb_segment_mm_grad_w(const float *A, const int64_t *a_idx, int a_heads,
                           const float *G, const int64_t *g_idx, int g_heads,
                           float *dW, const int64_t *seg_ptrs,
                           int64_t num_segs, int64_t num_rows, int H, int K,
                           int N, void *stream);

/* ------------------------------------------------------------------------
 * Edge-wise inner product (HGT attention score)
 * ---------------------------------------------------------------------- */

/* out[eids[i], h] = Σ_f left[L(i), h, f] · right[row_idx[i], h, f]
 * with L(i) = eids[i] (kind 0: left is per-edge data) or left_idx[eids[i]] when
 * left_idx != NULL (CompactAsOfNodeKind::EnabledWithDirectIndexing, kind 2: left
 * holds one row per unique (relation, node) pair and left_idx is the reference's
 * "edata_idx_to_inverse_idx", kernel_enums.h:6-119).
 * Replaces rgnn_inner_product_right_node_separatecoo, kinds 0 and 2
 * (RGNNOps.inc.h:609-658 → RGNN/InnerProductEdgeParallel.cu.h:13-113). */
int hetb_inner_product_fwd(const int64_t *eids, const int64_t *row_idx,
                           const int64_t *left_idx,
                           int64_t num_edges, const float *left,
                           const float *right, float *out, int H, int F,
                           void *stream);

/* grad_left[L(i),h,:]  (=|+=) g[eids[i],h] · right[row_idx[i],h,:]   (plain store for
 *   per-edge left data; summed into the compact row, which must arrive zero-filled,
 *   when left_idx != NULL)
 * grad_right[row_idx[i],h,:] += g[eids[i],h] · left[L(i),h,:]   (summed —
 * the reference stores instead of accumulating, InnerProductEdgeParallel.cu.h:180-183,
 * a known bug we do not reproduce).  grad_right must arrive zero-filled.
 * Replaces backward_inner_product_right_node_separatecoo
 * (RGNNOps.inc.h:1131-1181 → InnerProductEdgeParallel.cu.h:119-202). */
int hetb_inner_product_bwd(const int64_t *eids, const int64_t *row_idx,
                           const int64_t *left_idx,
                           int64_t num_edges, const float *left,
                           const float *right, const float *grad_out,
                           float *grad_left, float *grad_right, int H, int F,
                           void *stream);

/* ------------------------------------------------------------------------
 * RGAT edge attention pipeline (logits → edge softmax → aggregation)
 * ---------------------------------------------------------------------- */

/* Compact-as-of-node with direct indexing (CompactAsOfNodeKind 2 / 4,
 * hrt/include/kernel_enums.h:6-48; "edata_idx_to_inverse_idx[_row|_col]"): when a
 * pointer is non-NULL the corresponding tensor holds one row per unique
 * (relation, node) pair instead of one per edge. The index arrays of the CSR / edge-list
 * ops are ALIGNED WITH THE eids ARGUMENT: X_idx[p] = edata_idx_to_inverse_idx[eids[p]],
 * gathered once when the graph is built, so the edge pass has one dependent load level
 * less. Gradients of a compact tensor are SUMMED into it (it must arrive zero-filled). */
typedef struct hetb_compact_idx {
  const int64_t *feat_idx; /* feat / grad_feat rows  (by source:      ..._row) */
  const int64_t *el_idx;   /* el / grad_el rows      (by source:      ..._row) */
  const int64_t *er_idx;   /* er / grad_er rows      (by destination: ..._col) */
} hetb_compact_idx;

/* Over in-CSR rows d (destinations), positions p in [row_ptr[d], row_ptr[d+1]):
 *   e = eids[p];  x = el[e,h] + er[e,h];
 *   exp_out[e,h] = exp(x > 0 ? x : slope*x);   sum[d,h] = Σ_p exp_out[e,h];
 *   ret[d,h,:] = Σ_p exp_out[e,h] / sum[d,h] · feat[e,h,:]
 * sum/exp_out/ret may arrive uninitialised (backend/rgat_layers_and_funcs.py:656-664);
 * every element is written. Rows without in-edges get sum = 0, ret = 0.
 * Replaces relational_fused_gat_csr with CompactAsOfNodeFlag=false
 * (OpExport/RGATOps.inc.h:251-277 → GAT/FusedGAT.cu.h:33-222). */
int hetb_rgat_fwd_csr(const int64_t *row_ptr, const int64_t *eids,
                      int64_t num_rows, int64_t num_edges, const float *feat,
                      const float *el, const float *er, float *sum,
                      float *exp_out, float *ret, int H, int F, float slope,
                      const hetb_compact_idx *cidx /* NULL: per-edge inputs */,
                      void *stream);

/* Edge-parallel twin over a COO (dst = col_idx[i], e = eids[i]); same outputs.
 * Replaces relational_fused_gat_separate_coo, kind 0
 * (RGATOps.inc.h:170-245 → RGAT/RGATKernelsSeparateCOO.cu.h:17-215). */
int hetb_rgat_fwd_coo(const int64_t *col_idx, const int64_t *eids,
                      int64_t num_nodes, int64_t num_edges, const float *feat,
                      const float *el, const float *er, float *sum,
                      float *exp_out, float *ret, int H, int F, float slope,
                      const hetb_compact_idx *cidx /* NULL: per-edge inputs; else rows per COO position */,
                      void *stream);

/* For every edge position p (dst d = dst_idx[p], e = eids[p]):
 *   a = exp_in[e,h] / sum[d,h];
 *   grad_feat[e,h,:] = a · gout[d,h,:]
 *   t = Σ_f gout[d,h,f]·(feat[e,h,f] − ret[d,h,f]) · a · (el+er > 0 ? 1 : slope)
 *   grad_el[e,h] = grad_er[e,h] = t
 * (closed form: GAT/FusedGATBackward.cu.h:245-266). Serves both
 * backward_relational_fused_gat_csr (RGATOps.inc.h:430-461 →
 * FusedGATBackward.cu.h:31-148; dst_idx = out-CSR col_indices) and
 * backward_relational_fused_gat_separate_coo (RGATOps.inc.h:465-551;
 * dst_idx = COO col_indices). Each edge id is written exactly once, so the
 * reference's atomics are not needed. */
int hetb_rgat_bwd(const int64_t *dst_idx, const int64_t *eids,
                  int64_t num_edges, const float *feat, const float *el,
                  const float *er, const float *sum, const float *exp_in,
                  const float *ret, const float *gout, float *grad_feat,
                  float *grad_el, float *grad_er, int H, int F, float slope,
                  const hetb_compact_idx *cidx /* NULL: per-edge tensors */,
                  void *stream);

/* The same backward for compact feat / el (one row per unique (relation, source) pair) and compact
 * er (one per unique (relation, destination) pair), walking the edges GROUPED BY COMPACT SOURCE
 * ROW: urow_ptr[num_urows+1] is a CSR over the compact source rows, and per position p of row u:
 * u_dst[p] the destination, u_eids[p] the edge id, u_eridx[p] the compact er row. grad_feat[u]
 * and grad_el[u] are accumulated in registers and stored once per row (the reference's compact
 * backward, RGATBackwardKernels.cu.h:14-250, adds every edge's product atomically).
 * gdotr_scratch: num_dst*H floats of workspace. grad_* must arrive zero-filled. */
int hetb_rgat_bwd_compact_rows(const int64_t *urow_ptr, const int64_t *u_dst,
                               const int64_t *u_eids, const int64_t *u_eridx,
                               int64_t num_urows, int64_t num_edges, int64_t num_dst,
                               const float *feat, const float *el, const float *er,
                               const float *sum, const float *exp_in, const float *ret,
                               const float *gout, float *gdotr_scratch, float *grad_feat,
                               float *grad_el, float *grad_er, int H, int F, float slope,
                               void *stream);

/* ------------------------------------------------------------------------
 * HGT edge attention pipeline
 * ---------------------------------------------------------------------- */

/* Over in-CSR rows d: e = eids[p], r = rel[p];
 *   t[e,h] = expf(attn[e,h] · mu[r,h]);  sum[d,h] = Σ_p t;  norm[e,h] = t/sum[d,h]
 * t_out or norm_out may be NULL. sum is fully written (0 for empty rows).
 * Replaces hgt_full_graph_edge_softmax_ops_csr
 * (OpExport/HGTOps.inc.h:190-203 → HGT/HGTForwardKernels.cu.h:425-587). */
int hetb_hgt_edge_softmax_fwd_csr(const int64_t *row_ptr, const int64_t *eids,
                                  const int64_t *rel, int64_t num_rows,
                                  int64_t num_edges, const float *attn,
                                  const float *mu, float *sum, float *t_out,
                                  float *norm_out, int H, void *stream);

/* ret[d,h,:] = Σ_p w[e,h] · msg[e,h,:], with w = score[e,h] / sum[d,h] when
 * divide_by_sum != 0 (what the reference kernel does,
 * HGTForwardKernels.cu.h:304-309, switch 1) and w = score[e,h] otherwise.
 * Replaces hgt_full_graph_message_mean_aggregation_csr
 * (HGTOps.inc.h:180-188 → HGTForwardKernels.cu.h:244-369). */
int hetb_hgt_aggregate_fwd_csr(const int64_t *row_ptr, const int64_t *eids,
                               int64_t num_rows, int64_t num_edges,
                               const float *msg, const float *score,
                               const float *sum, float *ret, int H, int F,
                               int divide_by_sum,
                               const int64_t *msg_idx /* NULL, or compact msg row per position (aligned with eids) */,
                               void *stream);

/* grad_msg[e,h,:] = norm[e,h] · gout[dst_idx[p],h,:]
 * Replaces backward_hgt_full_graph_message_mean_aggregation_csr
 * (HGTOps.inc.h:477-487 → HGT/HGTBackwardKernels.cu.h:343-440). */
int hetb_hgt_aggregate_bwd(const int64_t *dst_idx, const int64_t *eids,
                           int64_t num_edges, const float *norm,
                           const float *gout, float *grad_msg, int H, int F,
                           const int64_t *msg_idx /* non-NULL: grad_msg is compact, summed */,
                           const float *sum_div /* non-NULL: norm holds exp(mu·score), divide by sum[d,h] */,
                           void *stream);

/* s = Σ_f gout[d,h,f]·(msg[e,h,f] − out[d,h,f]) · norm[e,h]
 * grad_attn[e,h] = s · mu[r,h];   grad_mu[r,h] += s · attn[e,h]
 * with r = rel[p] (the reference reads relation 0 for every edge in the
 * non-compact path, HGTBackwardKernels.cu.h:533,617-618 "FIXME"; we use the
 * edge's relation). grad_mu must arrive zero-filled.
 * Replaces backward_hgt_full_graph_edge_softmax_ops_csr
 * (HGTOps.inc.h:553-566 → HGTBackwardKernels.cu.h:499-627). */
int hetb_hgt_edge_softmax_bwd(const int64_t *dst_idx, const int64_t *eids,
                              const int64_t *rel, int64_t num_edges,
                              int64_t num_rels, const float *msg,
                              const float *attn, const float *norm,
                              const float *out, const float *gout,
                              const float *mu, float *grad_attn,
                              float *grad_mu, int H, int F,
                              const int64_t *msg_idx /* NULL, or compact msg row per position (aligned with eids) */,
                              const float *sum_div /* non-NULL: norm holds exp(mu·score), divide by sum[d,h] */,
                              void *stream);

/* Attention score + edge softmax + aggregation in one pass over the in-CSR (hetb200 fusion of
 * rgnn_inner_product_right_node + hgt_full_graph_edge_softmax_ops_csr +
 * hgt_full_graph_message_mean_aggregation_csr; same maths, reference files as those three):
 *   attn[e,h] = Σ_f left[L(e),h,f]·kv[src[p],h,f];  t[e,h] = exp(attn·mu[rel[p],h]);
 *   sum[d,h] = Σ_p t;  ret[d,h,:] = Σ_p t[e,h]·msg[M(e),h,:] / sum[d,h]
 * with L(e) = left_idx ? left_idx[p] : e and M(e) = msg_idx ? msg_idx[p] : e (aligned with eids). */
int hetb_hgt_fused_fwd_csr(const int64_t *row_ptr, const int64_t *eids, const int64_t *rel,
                           const int64_t *src, int64_t num_rows, int64_t num_edges,
                           const float *left, const int64_t *left_idx, const float *kv,
                           const float *msg, const int64_t *msg_idx, const float *mu, float *attn,
                           float *t_out, float *sum, float *ret, int H, int F, void *stream);

/* Backward of hetb_hgt_fused_fwd_csr in one pass over the OUT-CSR (rows = sources), replacing
 * hetb_hgt_aggregate_bwd + hetb_hgt_edge_softmax_bwd + hetb_inner_product_bwd: with
 * norm = t[e,h]/sum[d,h], sg = Σ_f gout[d,h,f]·(msg[M(e),h,f] − out[d,h,f])·norm, ga = sg·mu[r,h]:
 *   grad_msg[M(e)] += norm·gout[d];  grad_left[L(e)] += ga·kv[s];  grad_kv[s] += ga·left[L(e)];
 *   grad_mu[r,h] += sg·attn[e,h].   grad_* must arrive zero-filled.
 * gdoto_scratch: num_dst*H floats of workspace (<gout[d,h,:], out[d,h,:]> per destination, so the
 * edge pass gathers one destination row instead of two). */
int hetb_hgt_fused_bwd_csr(const int64_t *row_ptr, const int64_t *dst, const int64_t *eids,
                           const int64_t *rel, int64_t num_rows, int64_t num_edges, int64_t num_rels,
                           const float *left, const int64_t *left_idx, const float *kv,
                           const float *msg, const int64_t *msg_idx, const float *mu,
                           const float *attn, const float *t, const float *sum, const float *out,
                           const float *gout, int64_t num_dst, float *gdoto_scratch,
                           float *grad_left, float *grad_kv, float *grad_msg,
                           float *grad_mu, int H, int F, void *stream);

/* ------------------------------------------------------------------------
 * RGCN
 * ---------------------------------------------------------------------- */

/* ret[dst_idx[i], :] = Σ_i norm[e] · feat[F(e), :],  e = eids[i],  F(e) = feat_idx ? feat_idx[e] : e.
 * Edge-parallel aggregation over the separate COO: the second half of rgcn_layer1_separate_coo
 * (OpExport/RGCNOps.inc.h:84-138; feat = per-edge X[src]·W_r rows from hetb_segment_mm) and
 * rgcn_node_mean_aggregation_compact_as_of_node_separate_coo with direct indexing
 * (RGCNOps.inc.h:23-82; feat compact, feat_idx = inverse_indices_row). ret is zero-filled here
 * (the reference accumulates into a th.empty tensor). With a CSR at hand,
 * hetb_hgt_aggregate_fwd_csr(divide_by_sum = 0) computes the same sum without atomics; the
 * backward is hetb_hgt_aggregate_bwd (grad_feat) and hetb_inner_product_fwd (grad_norm). */
int hetb_rgcn_aggregate_coo(const int64_t *eids, const int64_t *dst_idx, const int64_t *feat_idx,
                            int64_t num_edges, int64_t num_dst, const float *feat,
                            const float *norm, float *ret, int F, void *stream);

/* ------------------------------------------------------------------------
 * Loss of the full-graph training step
 * ---------------------------------------------------------------------- */

/* *loss = (1/denom) Σ_i (logsumexp(logits[i,:]) − logits[i, labels[i]]);  lse[i] is saved for the
 * backward. Equals F.nll_loss(F.log_softmax(logits), labels, reduction="sum") / denom, the loss of
 * HET_RGNN_train (hrt/python/RGNNUtils/RGNNUtils.py:291-345). num_classes % 4 == 0. */
int hetb_softmax_xent_fwd(const float *logits, const int64_t *labels, int64_t num_rows,
                          int num_classes, float denom, float *lse, float *loss, void *stream);

/* grad_logits[i,c] = (*grad_loss / denom) · (exp(logits[i,c] − lse[i]) − [c == labels[i]]) */
int hetb_softmax_xent_bwd(const float *logits, const int64_t *labels, const float *lse,
                          const float *grad_loss, int64_t num_rows, int num_classes, float denom,
                          float *grad_logits, void *stream);

/* ------------------------------------------------------------------------
 * Relation sort / CSR construction (bit-exact integer work, on the GPU)
 * ---------------------------------------------------------------------- */

/* Bytes of device scratch the converters below need for `num_edges` edges and
 * keys below `max_key`. The caller allocates it (so the library never calls
 * cudaMalloc on the hot path). */
int64_t hetb_sort_scratch_bytes(int64_t num_edges, int64_t max_key);

/* Stable sort of edge tuples by `key` (0 <= key < num_keys): writes
 * ptrs[num_keys+1] (exclusive scan of key counts) and, for each payload
 * array in[j] (j < num_payload <= 4), out[j] = in[j] permuted into key order.
 * Order inside one key = original order. This one primitive builds:
 *   COO -> CSR          (key = src, payloads dst/rel/eid;
 *                        python/utils_lite/sparse_matrix_converters.py:7-39, stable)
 *   CSR transpose       (key = col, payloads expanded-row/rel/eid;
 *                        DataConverters.inc.h:283-344 → MyHyb/MyHyb.h:31-75)
 *   relation sort       (key = rel after an eid sort; MyHyb.h:1099-1150 +
 *                        python/utils/coo_sorters.py:156-170)              */
int hetb_stable_sort_by_key(const int64_t *key, int64_t num_edges,
                            int64_t num_keys, int num_payload,
                            const int64_t *const *in, int64_t *const *out,
                            int64_t *ptrs, void *scratch,
                            int64_t scratch_bytes, void *stream);

/* row_idx[p] = r for p in [row_ptr[r], row_ptr[r+1]) — CSR → COO row expansion. */
int hetb_expand_row_ptr(const int64_t *row_ptr, int64_t num_rows,
                        int64_t num_edges, int64_t *row_idx, void *stream);

/* ------------------------------------------------------------------------
 * Walk plans and the plan-based fused edge passes (SURVEY.md §8(b) "hetb_graph_plan": the
 * per-graph work decomposition and int32-narrowed index copies; §8(d) "i = 4")
 * ---------------------------------------------------------------------- */

#define HETB_DTYPE_F32 0
#define HETB_DTYPE_BF16 1 /* storage type of the gathered tables; arithmetic and gradients stay fp32 */

/* One walk over E positions. Built once per graph by the host (het_b200/graph.py:
 * HetGraph.get_edge_plan, from the relation-sorted COO the reference's converters produce:
 * MyHyb/MyHyb.h:1099-1150 + python/utils/coo_sorters.py:156-170, and the compact-as-of-node lists
 * of python/utils_lite/mydgl_graph_methods.py:11-157). All arrays int32, device resident.
 *   Forward plans walk positions sorted by (destination, relation): rows = destinations.
 *     idx_a = source node, idx_b = compact (relation, source) row, idx_c = compact (relation,
 *     destination) row of the position; row_rel[c] = relation of compact (relation, destination) row c.
 *   Backward plans walk positions sorted by (source, relation): rows = sources, runs = compact
 *     (relation, source) rows. idx_a = destination node, idx_b = compact (relation, source) row,
 *     idx_c = compact (relation, destination) row, idx_d = forward position of the same edge;
 *     row_rel[u] = relation of compact (relation, source) row u, bit 31 set when the run is split
 *     across items (listed in fix_runs).
 * Work items are [item_ptr[g], item_ptr[g+1]); item_row[g] indexes nz_rows / nz_ptr (the rows that
 * have positions: nz_rows[j] = row id, bit 31 set when the row is split across items;
 * nz_ptr[j] = first position). fix_rows: rows the pass accumulates into by reduction (pre-zeroed;
 * a forward pass also merges and normalises them); a backward plan lists empty rows there too.
 * fix_runs: backward plan: compact (relation, source) rows split across items; forward plan: compact
 * (relation, destination) rows split across items (hetb_hgt_fused_bwd_plan2's second walk).
 * empty_rows (forward): rows without positions (zero-filled). pslot / pslot_row (forward): the
 * valid partial slots (2*item + {0: head row, 1: tail row}) of split rows and the row of each. */
typedef struct hetb_walk_plan {
  int64_t num_positions;
  int32_t num_items, num_nz_rows, num_fix_rows, num_empty_rows, num_pslots, num_fix_runs;
  const int32_t *item_ptr, *item_row, *nz_rows, *nz_ptr;
  const int32_t *fix_rows, *empty_rows, *pslot, *pslot_row, *fix_runs;
  const int32_t *idx_a, *idx_b, *idx_c, *idx_d, *row_rel;
} hetb_walk_plan;

/* HGT attention score + edge softmax (max-subtracted, online) + message aggregation in one walk:
 * a5 + a10 + a11 of SURVEY.md §8(a) (rgnn_inner_product_right_node_separatecoo:
 * OpExport/RGNNOps.inc.h:609-658; hgt_full_graph_edge_softmax_ops_csr: OpExport/HGTOps.inc.h:190-203;
 * hgt_full_graph_message_mean_aggregation_csr: :180-188) on compact inputs (kind 2 direct indexing,
 * kernel_enums.h:6-119).  left [Uc,H,F] compact (relation, destination) rows (W_att·q), kv [Nsrc,H,F]
 * (k), msg [Ur,H,F] compact (relation, source) rows, mu [R,H].  Writes lse [N,H] (log-sum-exp of
 * mu·attn per destination and head; replaces the reference's per-edge exp and per-node sum), ret [N,H,F]
 * and, when attn is not NULL, attn [E,H] (the raw score, in FORWARD-PLAN POSITION order; the default
 * backward recomputes it and the default forward passes NULL: nothing is saved per edge).  Scratch:
 * part_m / part_s [2*num_items,H], part_acc [2*num_items,H*F], ssum [N,H]. */
int hetb_hgt_fused_fwd_plan(const hetb_walk_plan *plan, const void *left, const void *kv, const void *msg,
                            int table_dtype, const float *mu, float *attn, float *lse, float *ret, float *part_m,
                            float *part_s, float *part_acc, float *ssum, int H, int F, void *stream);

/* Backward of the above in one walk (a6 + a12: backward_inner_product_right_node_separatecoo
 * OpExport/RGNNOps.inc.h:1131-1181; backward_hgt_full_graph_{message_mean_aggregation,edge_softmax_ops}_csr
 * OpExport/HGTOps.inc.h:477-487, 553-566). grad_left [Uc,H,F] and grad_mu [R,H] must arrive
 * zero-filled; grad_kv [Nsrc,H,F] and grad_msg [Ur,H,F] are fully written. gdoto_scratch [N,2H]
 * ((lse, <gout,out>) pairs per destination). attn = NULL (the default): the raw score is recomputed from
 * left and kv, in the forward's operation order (same bits); attn given: read at the edge's forward
 * position (one more random access per edge). */
int hetb_hgt_fused_bwd_plan(const hetb_walk_plan *plan, int64_t num_rels, int64_t num_dst, const void *left,
                            const void *kv, const void *msg, int table_dtype, const float *mu, const float *attn,
                            const float *lse, const float *out, const float *gout, float *gdoto_scratch,
                            float *grad_left, float *grad_kv, float *grad_msg, float *grad_mu, int H, int F,
                            void *stream);

/* The same backward in two walks (the default of het_b200/backend.py): the (source, relation)-ordered walk
 * of `plan` owns grad_kv / grad_msg as above but does not reduce into grad_left; it leaves the gradient of
 * the raw score in attn_inout (forward-plan position order, in place of the score it read there), and a
 * second walk over `fwd_plan` (the forward pass's (destination, relation) order) owns grad_left [Uc,H,F] as a
 * register accumulator per compact (relation, destination) row: no atomics on the destination side
 * (the reference: one atomicAdd per edge and element, RGNN/InnerProductEdgeParallel.cu.h:119-202), no
 * zero-fill of grad_left, run-to-run identical results. Only grad_mu must arrive zero-filled. fwd_plan
 * must carry fix_runs (the compact (relation, destination) rows cut by an item boundary). */
int hetb_hgt_fused_bwd_plan2(const hetb_walk_plan *plan, const hetb_walk_plan *fwd_plan, int64_t num_rels,
                             int64_t num_dst, const void *left, const void *kv, const void *msg, int table_dtype,
                             const float *mu, float *attn_inout, const float *lse, const float *out,
                             const float *gout, float *gdoto_scratch, float *grad_left, float *grad_kv,
                             float *grad_msg, float *grad_mu, int H, int F, void *stream);

/* RGAT logits + edge softmax (max-subtracted, online) + aggregation in one walk: a7
 * (relational_fused_gat_csr, OpExport/RGATOps.inc.h:251-277) on compact inputs: feat [Ur,H,F], el [Ur,H]
 * compact (relation, source) rows, er [Uc,H] compact (relation, destination) rows. Writes lse [N,H],
 * ret [N,H,F]; nothing per edge (the reference's exp [E,H] is recomputed by the backward). */
int hetb_rgat_fused_fwd_plan(const hetb_walk_plan *plan, const void *feat, int table_dtype, const float *el,
                             const float *er, float slope, float *lse, float *ret, float *part_m, float *part_s,
                             float *part_acc, float *ssum, int H, int F, void *stream);

/* Backward of the above (a9: backward_relational_fused_gat_csr, OpExport/RGATOps.inc.h:430-461).
 * grad_er [Uc,H] must arrive zero-filled; grad_feat / grad_el are fully written. gdotr_scratch [N,2H]. */
int hetb_rgat_fused_bwd_plan(const hetb_walk_plan *plan, int64_t num_dst, const void *feat, int table_dtype,
                             const float *el, const float *er, const float *lse, const float *ret,
                             const float *gout, float *gdotr_scratch, float *grad_feat, float *grad_el,
                             float *grad_er, int H, int F, float slope, void *stream);

/* ------------------------------------------------------------------------
 * Separate-COO (edge-parallel) forms of the HGT ops and compact-index helpers (SURVEY.md §8a a13/a14,
 * §8f N2/N4). The relation of COO position i is rel_of_pos[i] when given, else the segment of
 * rel_ptrs[num_rels+1] that holds i (the relation-sorted separate COO).
 * ---------------------------------------------------------------------- */

/* hgt_full_graph_edge_softmax_ops_separate_coo (OpExport/HGTOpsEdgeParallel.inc.h:18-32):
 * t = exp(attn*mu[rel]), sum[dst] = sum t, norm = t / sum[dst]. Normalises over the in-edges of the
 * DESTINATION (dst = the COO's col_indices); the reference's edge-parallel kernel keys the sum by
 * row_indices (HGT/HGTForwardKernels.cu.h:619), documented quirk, not copied. t_out / norm_out may be NULL. */
int hetb_hgt_edge_softmax_fwd_coo(const int64_t *dst, const int64_t *eids, const int64_t *rel_of_pos,
                                  const int64_t *rel_ptrs, int64_t num_rels, int64_t num_nodes, int64_t num_edges,
                                  const float *attn, const float *mu, float *sum, float *t_out, float *norm_out,
                                  int H, void *stream);

/* ret[dst[i],h,:] = sum_i w[eids[i],h] * feat[eids[i],h,:] (ret is zero-filled here): the aggregation
 * half of hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo
 * (OpExport/HGTOpsEdgeParallel.inc.h:33-92; the typed linear half is hetb_segment_mm). */
int hetb_weighted_scatter_coo(const int64_t *dst, const int64_t *eids, int64_t num_edges, int64_t num_dst,
                              const float *feat, const float *w, float *ret, int H, int F, void *stream);

/* backward_hgt_full_graph_enorm_to_unnormalized_attn_score_{csr,separate_coo}
 * (OpExport/HGTOps.inc.h:283-340, 598-650): softmax Jacobian from grad of the normalised score to
 * grad of the raw score and of mu. dst[i] = destination of position i. grad_mu must arrive
 * zero-filled; sumprod_scratch [num_nodes, H] is zeroed here. */
int hetb_hgt_enorm_to_unnormalized_bwd(const int64_t *dst, const int64_t *eids, const int64_t *rel_of_pos,
                                       const int64_t *rel_ptrs, int64_t num_rels, int64_t num_nodes,
                                       int64_t num_edges, const float *attn, const float *norm,
                                       const float *grad_norm, const float *mu, float *sumprod_scratch,
                                       float *grad_attn, float *grad_mu, int H, void *stream);

/* Allocation-free variant of hetb_hgt_enorm_to_unnormalized_bwd. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture,
 * tight serving loops, harnesses that forbid it). Same inputs, outputs and maths; the [num_nodes, H]
 * workspace is a static device array inside the library (HETB_STATIC_SCRATCH_FLOATS floats = 4 MB, held
 * by every GPU that loads the library until exit) instead of a caller buffer. num_nodes * H beyond that
 * returns HETB_ETOOLARGE (and says so on stderr); nothing is launched. Calls no cudaMalloc / cudaFree.
 * All calls on a device share the array: it is zeroed on `stream` at the start of each call, and calls
 * that overlap on the GPU (two streams, two threads) race. grad_mu must arrive zero-filled. */
#define HETB_STATIC_SCRATCH_FLOATS (1 << 20)
int hetb_hgt_enorm_to_unnormalized_bwd_static(const int64_t *dst, const int64_t *eids, const int64_t *rel_of_pos,
                                              const int64_t *rel_ptrs, int64_t num_rels, int64_t num_nodes,
                                              int64_t num_edges, const float *attn, const float *norm,
                                              const float *grad_norm, const float *mu, float *grad_attn,
                                              float *grad_mu, int H, void *stream);

/* rgnn_gather_compact_to_non_compact (OpExport/RGNNOps.inc.h:449-600):
 * out[eids[i], :] = in[idx ? idx[i] : eids[i], :], rows of xlen floats (xlen % 4 == 0). */
int hetb_gather_rows(const int64_t *eids, const int64_t *idx, int64_t num_edges, const float *in, float *out,
                     int64_t xlen, void *stream);

/* out[i, :] = in[idx[i], :] and dst[idx[i], :] += src[i, :] (rows of xlen floats, xlen % 4 == 0): pack /
 * unpack of the multi-GPU halo exchange (het_b200/parallel.py; the reference has no distributed path). */
int hetb_rows_gather(const int64_t *idx, int64_t n, const float *in, float *out, int64_t xlen, void *stream);
int hetb_rows_scatter_add(const int64_t *idx, int64_t n, const float *src, float *dst, int64_t xlen, void *stream);

/* find_relational_compact_as_of_node_index with binary search (kernel_enums.h:96-119) for every
 * position: out_idx[i] = row of (relation of i, node[i]) in the per-relation sorted unique lists, -1
 * (and *missing_counter += 1) when the node is not listed. Turns CompactAsOfNodeKind::Enabled /
 * EnabledWithDualList into the direct-indexing form the edge kernels take. */
int hetb_compact_index(const int64_t *node, const int64_t *rel_of_pos, const int64_t *rel_ptrs, int64_t num_rels,
                       int64_t num_edges, const int64_t *unique_rel_ptrs, const int64_t *unique_nodes,
                       int64_t *out_idx, int *missing_counter, void *stream);

/* ------------------------------------------------------------------------
 * Optimizer step and weight composition of the full-graph trainer
 * ---------------------------------------------------------------------- */

/* Adam (torch.optim.Adam's update, no weight decay / amsgrad) over one fp32 tensor; the optimizer of
 * HET_RGNN_train (hrt/python/RGNNUtils/RGNNUtils.py:291-345). `step` is a device float holding the
 * step count (hetb_adam_bump_step adds 1), so the update is capturable in a CUDA graph. Pointers
 * 16-byte aligned. */
int hetb_adam_bump_step(float *step, void *stream);
int hetb_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, int64_t n, const float *step,
                   float lr, float beta1, float beta2, float eps, void *stream);

/* out[r, i, h*dk+e] = sum_d lin[type_of_rel[r], i, h*dk+d] * rel[r, h, d, e]: the per-relation composed
 * weights of HGT's linear operator reordering (hrt/python/HGT/models.py:124-155). lin [T, in_dim, H*dk],
 * rel [R, H, dk, dk], out [R, in_dim, H*dk]. The backward fills grad_rel and ADDS into grad_lin
 * (zero-filled by the caller). */
int hetb_compose_weights_fwd(const float *lin, const int64_t *type_of_rel, const float *rel, float *out, int64_t R,
                             int in_dim, int H, int dk, void *stream);
int hetb_compose_weights_bwd(const float *gout, const float *lin, const int64_t *type_of_rel, const float *rel,
                             float *grad_lin, float *grad_rel, int64_t R, int in_dim, int H, int dk, void *stream);

/* out[c] = sum_i x[i, c] for a row-major [num_rows, num_cols] matrix: the gradient of the layer
 * bias (h + b; hrt/python/RGAT/models.py:264-278, RGCN/RGCN.py:343-344). num_cols % 4 == 0. */
int hetb_column_sum(const float *x, int64_t num_rows, int num_cols, float *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* HETB200_H_ */
