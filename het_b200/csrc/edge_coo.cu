// Separate-COO (edge-parallel) forms of the HGT ops and the compact-index helpers
// (SURVEY.md §8a a13 / a14, §8f N2 / N4). These are the reference's "P1" slicings of maths the
// plan-based walkers already do in one pass; they exist so that every flag of the reference's
// models runs, and are written as plain edge-parallel kernels over the relation-sorted COO:
// one lane (or lane group) per position, vector reductions into destination rows.
//
//   coo_softmax_*        hgt_full_graph_edge_softmax_ops_separate_coo
//                        (OpExport/HGTOpsEdgeParallel.inc.h:18-32 -> HGT/HGTForwardKernels.cu.h:591-761)
//   coo_weighted_scatter hgt_full_graph_fused_message_calc_and_mean_aggregation_separate_coo, after
//                        the typed linear (OpExport/HGTOpsEdgeParallel.inc.h:33-92 ->
//                        RGNN/my_shmem_sgemm_func_rgcn_hgt.cu.h:709-816)
//   enorm_bwd_*          backward_hgt_full_graph_enorm_to_unnormalized_attn_score_{csr,separate_coo}
//                        (OpExport/HGTOps.inc.h:283-340, 598-650 -> HGT/HGTBackwardKernels.cu.h:172-338)
//   gather_rows_kernel   rgnn_gather_compact_to_non_compact (OpExport/RGNNOps.inc.h:449-600 ->
//                        RGNN/GatherCompactToNonCompact.cu.h:8-90)
//   compact_index_kernel find_relational_compact_as_of_node_index with binary search
//                        (kernel_enums.h:96-119): turns the binary-search compact kinds into the
//                        direct-indexing form every edge kernel here takes.
#include <cstdio>

#include "common.cuh"

namespace hetb {

constexpr int kCThreads = 256;
constexpr int kCMaxHeads = 8;

// relation of COO position i: rel_of_pos[i] when given, else the segment of rel_ptrs holding i
__device__ __forceinline__ int rel_at(const int64_t *__restrict__ rel_of_pos, const int64_t *__restrict__ rel_ptrs,
                                      int num_rels, int64_t i) {
  return rel_of_pos ? (int)__ldg(rel_of_pos + i) : segment_of(rel_ptrs, num_rels, i);
}

// t[e,h] = exp(a[e,h] * mu[r,h]);  sum[dst,h] += t   (sum zero-filled by the host entry)
__global__ void __launch_bounds__(kCThreads)
    coo_softmax_exp_sum_kernel(const int64_t *__restrict__ dst, const int64_t *__restrict__ eids,
                               const int64_t *__restrict__ rel_of_pos, const int64_t *__restrict__ rel_ptrs,
                               int num_rels, int64_t E, const float *__restrict__ attn,
                               const float *__restrict__ mu, float *__restrict__ sum, float *__restrict__ t_out,
                               int H) {
  const int64_t total = E * H;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / H;
    const int h = (int)(t % H);
    const int64_t e = __ldg(eids + i), d = __ldg(dst + i);
    const int r = rel_at(rel_of_pos, rel_ptrs, num_rels, i);
    const float w = __expf(__ldg(attn + e * H + h) * __ldg(mu + (int64_t)r * H + h));
    if (t_out) t_out[e * H + h] = w;
    atomicAdd(sum + d * H + h, w);
  }
}

// norm[e,h] = exp(a*mu) / sum[dst,h]
__global__ void __launch_bounds__(kCThreads)
    coo_softmax_norm_kernel(const int64_t *__restrict__ dst, const int64_t *__restrict__ eids,
                            const int64_t *__restrict__ rel_of_pos, const int64_t *__restrict__ rel_ptrs,
                            int num_rels, int64_t E, const float *__restrict__ attn, const float *__restrict__ mu,
                            const float *__restrict__ sum, float *__restrict__ norm, int H) {
  const int64_t total = E * H;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / H;
    const int h = (int)(t % H);
    const int64_t e = __ldg(eids + i), d = __ldg(dst + i);
    const int r = rel_at(rel_of_pos, rel_ptrs, num_rels, i);
    const float s = __ldg(sum + d * H + h);
    norm[e * H + h] = s > 0.f ? __expf(__ldg(attn + e * H + h) * __ldg(mu + (int64_t)r * H + h)) / s : 0.f;
  }
}

// ret[dst[i],h,:] += w[e,h] * feat[e,h,:]   (ret zero-filled by the host entry)
template <int LPR>
__global__ void __launch_bounds__(kCThreads)
    coo_weighted_scatter_kernel(const int64_t *__restrict__ dst, const int64_t *__restrict__ eids, int64_t E,
                                const float *__restrict__ feat, const float *__restrict__ w,
                                float *__restrict__ ret, int H, int hw) {
  constexpr int GROUPS = kCThreads / LPR;
  constexpr int D = LPR * 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int h = l / hw;
  const int64_t stride = (int64_t)gridDim.x * GROUPS;
  for (int64_t i = (int64_t)blockIdx.x * GROUPS + g; i < E; i += stride) {
    const int64_t e = __ldg(eids + i), d = __ldg(dst + i);
    const float a = __ldg(w + e * H + h);
    const float4 v = ldcs4(feat + e * D + 4 * l);
    red_add4(ret + d * D + 4 * l, make_float4(a * v.x, a * v.y, a * v.z, a * v.w));
  }
}

// Softmax Jacobian from the gradient of the NORMALISED score to the raw score and mu
// (HGT/HGTBackwardKernels.cu.h:160-170): with S = norm, g = grad_norm, over the in-edges of a node
//   sumprod[d,h] = sum_e S_e g_e          (stage 0; sumprod zero-filled by the host entry)
//   grad_a[e,h]  = (g_e - sumprod[d,h]) * S_e * mu[r,h]
//   grad_mu[r,h] += (g_e - sumprod[d,h]) * S_e * a[e,h]     (the reference leaves the `* a` out: :322-323 TODO)
__global__ void __launch_bounds__(kCThreads)
    enorm_bwd_sumprod_kernel(const int64_t *__restrict__ dst, const int64_t *__restrict__ eids, int64_t E,
                             const float *__restrict__ norm, const float *__restrict__ gnorm,
                             float *__restrict__ sumprod, int H) {
  const int64_t total = E * H;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / H;
    const int h = (int)(t % H);
    const int64_t e = __ldg(eids + i), d = __ldg(dst + i);
    atomicAdd(sumprod + d * H + h, __ldg(norm + e * H + h) * __ldg(gnorm + e * H + h));
  }
}

__global__ void __launch_bounds__(kCThreads)
    enorm_bwd_apply_kernel(const int64_t *__restrict__ dst, const int64_t *__restrict__ eids,
                           const int64_t *__restrict__ rel_of_pos, const int64_t *__restrict__ rel_ptrs,
                           int num_rels, int64_t E, const float *__restrict__ attn, const float *__restrict__ norm,
                           const float *__restrict__ gnorm, const float *__restrict__ mu,
                           const float *__restrict__ sumprod, float *__restrict__ grad_attn,
                           float *__restrict__ grad_mu, int H, int use_smem) {
  extern __shared__ float mu_acc[];
  const int RH = num_rels * H;
  if (use_smem) {
    for (int j = threadIdx.x; j < RH; j += kCThreads) mu_acc[j] = 0.f;
    __syncthreads();
  }
  const int64_t total = E * H;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / H;
    const int h = (int)(t % H);
    const int64_t e = __ldg(eids + i), d = __ldg(dst + i);
    const int r = rel_at(rel_of_pos, rel_ptrs, num_rels, i);
    const float v = (__ldg(gnorm + e * H + h) - __ldg(sumprod + d * H + h)) * __ldg(norm + e * H + h);
    grad_attn[e * H + h] = v * __ldg(mu + (int64_t)r * H + h);
    const float gm = v * __ldg(attn + e * H + h);
    if (use_smem) atomicAdd(&mu_acc[r * H + h], gm);
    else atomicAdd(grad_mu + (int64_t)r * H + h, gm);
  }
  if (use_smem) {
    __syncthreads();
    for (int j = threadIdx.x; j < RH; j += kCThreads) {
      const float v = mu_acc[j];
      if (v != 0.f) atomicAdd(grad_mu + j, v);
    }
  }
}

// out[eids[i], :] = in[idx ? idx[i] : eids[i], :]   (row length `xlen` floats, a multiple of 4)
__global__ void __launch_bounds__(kCThreads)
    gather_rows_kernel(const int64_t *__restrict__ eids, const int64_t *__restrict__ idx, int64_t E,
                       const float *__restrict__ in, float *__restrict__ out, int xlen4) {
  const int64_t total = E * xlen4;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / xlen4;
    const int c = (int)(t % xlen4);
    const int64_t e = __ldg(eids + i);
    const int64_t s = idx ? __ldg(idx + i) : e;
    stcs4(out + (e * xlen4 + c) * 4, ldg4(in + (s * xlen4 + c) * 4));
  }
}

// out_idx[i] = unique_rel_ptrs[r] + (rank of node[i] among unique_nodes[unique_rel_ptrs[r] .. [r+1]))
// with r = relation of position i: the binary-search compact kinds (CompactAsOfNodeKind::Enabled /
// EnabledWithDualList) made position-aligned direct indices. A node that is not in its relation's
// list yields -1 (the host entry reports it; the reference's search would return a neighbour's row).
__global__ void __launch_bounds__(kCThreads)
    compact_index_kernel(const int64_t *__restrict__ node, const int64_t *__restrict__ rel_of_pos,
                         const int64_t *__restrict__ rel_ptrs, int num_rels, int64_t E,
                         const int64_t *__restrict__ unique_rel_ptrs, const int64_t *__restrict__ unique_nodes,
                         int64_t *__restrict__ out_idx, int *__restrict__ missing) {
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t i = (int64_t)blockIdx.x * kCThreads + threadIdx.x; i < E; i += stride) {
    const int r = rel_at(rel_of_pos, rel_ptrs, num_rels, i);
    const int64_t v = __ldg(node + i);
    int64_t lo = __ldg(unique_rel_ptrs + r), hi = __ldg(unique_rel_ptrs + r + 1);
    const int64_t end = hi;
    while (lo < hi) {  // lower bound
      const int64_t mid = (lo + hi) >> 1;
      if (__ldg(unique_nodes + mid) < v) lo = mid + 1;
      else hi = mid;
    }
    const bool found = lo < end && __ldg(unique_nodes + lo) == v;
    out_idx[i] = found ? lo : -1;
    if (!found) atomicAdd(missing, 1);
  }
}

// out[i, :] = in[idx[i], :]  /  dst[idx[i], :] += src[i, :]: pack and unpack of the multi-GPU halo exchange
__global__ void __launch_bounds__(kCThreads)
    rows_gather_kernel(const int64_t *__restrict__ idx, int64_t n, const float *__restrict__ in,
                       float *__restrict__ out, int xlen4) {
  const int64_t total = n * xlen4;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / xlen4;
    const int c = (int)(t % xlen4);
    stcs4(out + t * 4, ldg4(in + (__ldg(idx + i) * xlen4 + c) * 4));
  }
}

__global__ void __launch_bounds__(kCThreads)
    rows_scatter_add_kernel(const int64_t *__restrict__ idx, int64_t n, const float *__restrict__ src,
                            float *__restrict__ dst, int xlen4) {
  const int64_t total = n * xlen4;
  const int64_t stride = (int64_t)gridDim.x * kCThreads;
  for (int64_t t = (int64_t)blockIdx.x * kCThreads + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / xlen4;
    const int c = (int)(t % xlen4);
    red_add4(dst + (__ldg(idx + i) * xlen4 + c) * 4, ldcs4(src + t * 4));
  }
}

static int coo_grid(int64_t n) {
  const int64_t need = ceil_div64(n, kCThreads);
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(need < cap ? (need > 0 ? need : 1) : cap);
}

static int coo_lpr(int H, int F) {
  if (H <= 0 || F <= 0 || (F % 4) != 0 || !is_pow2(H) || !is_pow2(F)) return 0;
  const int v = H * F / 4;
  return (v < 4 || v > 32) ? 0 : v;
}

}  // namespace hetb

using namespace hetb;

extern "C" int hetb_hgt_edge_softmax_fwd_coo(const int64_t *dst, const int64_t *eids, const int64_t *rel_of_pos,
                                             const int64_t *rel_ptrs, int64_t num_rels, int64_t num_nodes,
                                             int64_t num_edges, const float *attn, const float *mu, float *sum,
                                             float *t_out, float *norm_out, int H, void *stream) {
  HETB_CHECK_ARG(num_nodes >= 0 && num_edges >= 0 && num_rels >= 0 && H > 0);
  if (num_nodes == 0) return HETB_OK;
  HETB_CHECK_ARG(sum != nullptr);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(sum, 0, sizeof(float) * num_nodes * H, st);
  if (ce != cudaSuccess) return (int)ce;
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(dst && eids && attn && mu && (rel_of_pos || rel_ptrs) && num_rels <= INT32_MAX);
  const int grid = coo_grid(num_edges * H);
  coo_softmax_exp_sum_kernel<<<grid, kCThreads, 0, st>>>(dst, eids, rel_of_pos, rel_ptrs, (int)num_rels, num_edges,
                                                         attn, mu, sum, t_out, H);
  HETB_LAUNCH_OK();
  if (norm_out) {
    coo_softmax_norm_kernel<<<grid, kCThreads, 0, st>>>(dst, eids, rel_of_pos, rel_ptrs, (int)num_rels, num_edges, attn,
                                                        mu, sum, norm_out, H);
    HETB_LAUNCH_OK();
  }
  return HETB_OK;
}

extern "C" int hetb_weighted_scatter_coo(const int64_t *dst, const int64_t *eids, int64_t num_edges,
                                         int64_t num_dst, const float *feat, const float *w, float *ret, int H,
                                         int F, void *stream) {
  HETB_CHECK_ARG(num_edges >= 0 && num_dst >= 0);
  if (num_dst == 0) return HETB_OK;
  HETB_CHECK_ARG(ret != nullptr);
  const int lpr = coo_lpr(H, F);
  if (!lpr) return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(ret, 0, sizeof(float) * num_dst * H * F, st);
  if (ce != cudaSuccess) return (int)ce;
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(dst && eids && feat && w);
  const int groups = kCThreads / lpr;
  const int64_t need = ceil_div64(num_edges, groups);
  const int grid = (int)(need < (int64_t)kNumSMs * 16 ? need : (int64_t)kNumSMs * 16);
  switch (lpr) {
#define WS(L) case L: coo_weighted_scatter_kernel<L><<<grid, kCThreads, 0, st>>>(dst, eids, num_edges, feat, w, ret, H, F / 4); break;
    WS(4) WS(8) WS(16) WS(32)
#undef WS
  }
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_hgt_enorm_to_unnormalized_bwd(const int64_t *dst, const int64_t *eids, const int64_t *rel_of_pos,
                                                  const int64_t *rel_ptrs, int64_t num_rels, int64_t num_nodes,
                                                  int64_t num_edges, const float *attn, const float *norm,
                                                  const float *grad_norm, const float *mu, float *sumprod_scratch,
                                                  float *grad_attn, float *grad_mu, int H, void *stream) {
  HETB_CHECK_ARG(num_nodes >= 0 && num_edges >= 0 && num_rels >= 0 && H > 0);
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(dst && eids && attn && norm && grad_norm && mu && sumprod_scratch && grad_attn && grad_mu &&
                 (rel_of_pos || rel_ptrs) && num_rels <= INT32_MAX);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(sumprod_scratch, 0, sizeof(float) * num_nodes * H, st);
  if (ce != cudaSuccess) return (int)ce;
  int grid = coo_grid(num_edges * H);
  enorm_bwd_sumprod_kernel<<<grid, kCThreads, 0, st>>>(dst, eids, num_edges, norm, grad_norm, sumprod_scratch, H);
  HETB_LAUNCH_OK();
  const int use_smem = (num_rels * H <= 4096) ? 1 : 0;
  const size_t smem = use_smem ? (size_t)num_rels * H * sizeof(float) : 0;
  if (grid > kNumSMs * 8) grid = kNumSMs * 8;
  enorm_bwd_apply_kernel<<<grid, kCThreads, smem, st>>>(dst, eids, rel_of_pos, rel_ptrs, (int)num_rels, num_edges, attn,
                                                        norm, grad_norm, mu, sumprod_scratch, grad_attn, grad_mu, H,
                                                        use_smem);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_gather_rows(const int64_t *eids, const int64_t *idx, int64_t num_edges, const float *in,
                                float *out, int64_t xlen, void *stream) {
  HETB_CHECK_ARG(num_edges >= 0 && xlen > 0);
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(eids && in && out);
  if (xlen % 4) return HETB_EUNSUPPORTED;
  gather_rows_kernel<<<coo_grid(num_edges * (xlen / 4)), kCThreads, 0, (cudaStream_t)stream>>>(eids, idx, num_edges, in,
                                                                                              out, (int)(xlen / 4));
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_compact_index(const int64_t *node, const int64_t *rel_of_pos, const int64_t *rel_ptrs,
                                  int64_t num_rels, int64_t num_edges, const int64_t *unique_rel_ptrs,
                                  const int64_t *unique_nodes, int64_t *out_idx, int *missing_counter,
                                  void *stream) {
  HETB_CHECK_ARG(num_edges >= 0 && num_rels >= 0);
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(node && (rel_of_pos || rel_ptrs) && unique_rel_ptrs && unique_nodes && out_idx && missing_counter &&
                 num_rels <= INT32_MAX);
  compact_index_kernel<<<coo_grid(num_edges), kCThreads, 0, (cudaStream_t)stream>>>(
      node, rel_of_pos, rel_ptrs, (int)num_rels, num_edges, unique_rel_ptrs, unique_nodes, out_idx, missing_counter);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_rows_gather(const int64_t *idx, int64_t n, const float *in, float *out, int64_t xlen, void *stream) {
  HETB_CHECK_ARG(n >= 0 && xlen > 0);
  if (n == 0) return HETB_OK;
  HETB_CHECK_ARG(idx && in && out);
  if (xlen % 4) return HETB_EUNSUPPORTED;
  rows_gather_kernel<<<coo_grid(n * (xlen / 4)), kCThreads, 0, (cudaStream_t)stream>>>(idx, n, in, out, (int)(xlen / 4));
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_rows_scatter_add(const int64_t *idx, int64_t n, const float *src, float *dst, int64_t xlen,
                                     void *stream) {
  HETB_CHECK_ARG(n >= 0 && xlen > 0);
  if (n == 0) return HETB_OK;
  HETB_CHECK_ARG(idx && src && dst);
  if (xlen % 4) return HETB_EUNSUPPORTED;
  rows_scatter_add_kernel<<<coo_grid(n * (xlen / 4)), kCThreads, 0, (cudaStream_t)stream>>>(idx, n, src, dst,
                                                                                           (int)(xlen / 4));
  HETB_LAUNCH_OK();
  return HETB_OK;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of hetb_hgt_enorm_to_unnormalized_bwd (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph
// capture, tight serving loops, harnesses that forbid it). The [num_nodes, H] sum-of-products workspace
// the caller of the existing entry point has to provide lives in a file-scope device array instead:
// HETB_STATIC_SCRATCH_FLOATS floats (4 MB), held by every GPU that loads the library until exit.
//
// In host code the array's NAME is a host-side stand-in, not a device address; the address a kernel may
// use is resolved with cudaGetSymbolAddress once per device, with that device current, and cached.
// Every call shares the array and it keeps the last call's contents: the part a call uses is zeroed
// on the call's stream first. Calls that overlap on the GPU (different streams, or the same device from
// two threads) race on it. Calls nothing that allocates: no cudaMalloc / cudaFree, no torch.
// ---------------------------------------------------------------------------
static __device__ __align__(256) float g_static_scratch[HETB_STATIC_SCRATCH_FLOATS];

static int static_scratch_ptr(float **out) {
  constexpr int kMaxDevices = 64;
  static float *ptrs[kMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) return (int)ce;
  if (dev < 0 || dev >= kMaxDevices) return HETB_EUNSUPPORTED;
  if (!ptrs[dev]) {
    void *p = nullptr;
    ce = cudaGetSymbolAddress(&p, g_static_scratch);  // device `dev` is current: its copy of the array
    if (ce != cudaSuccess) return (int)ce;
    ptrs[dev] = static_cast<float *>(p);
  }
  *out = ptrs[dev];
  return HETB_OK;
}

extern "C" int hetb_hgt_enorm_to_unnormalized_bwd_static(const int64_t *dst, const int64_t *eids,
                                                         const int64_t *rel_of_pos, const int64_t *rel_ptrs,
                                                         int64_t num_rels, int64_t num_nodes, int64_t num_edges,
                                                         const float *attn, const float *norm,
                                                         const float *grad_norm, const float *mu, float *grad_attn,
                                                         float *grad_mu, int H, void *stream) {
  HETB_CHECK_ARG(num_nodes >= 0 && num_edges >= 0 && num_rels >= 0 && H > 0);
  if (num_nodes > (int64_t)HETB_STATIC_SCRATCH_FLOATS / H) {
    std::fprintf(stderr,
                 "hetb_hgt_enorm_to_unnormalized_bwd_static: num_nodes * H = %lld exceeds the static scratch of %d "
                 "floats; use hetb_hgt_enorm_to_unnormalized_bwd with a caller-provided workspace\n",
                 (long long)num_nodes * H, (int)HETB_STATIC_SCRATCH_FLOATS);
    return HETB_ETOOLARGE;
  }
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(dst && eids && attn && norm && grad_norm && mu && grad_attn && grad_mu && (rel_of_pos || rel_ptrs) &&
                 num_rels <= INT32_MAX);
  float *sumprod = nullptr;
  const int rc = static_scratch_ptr(&sumprod);
  if (rc != HETB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(sumprod, 0, sizeof(float) * num_nodes * H, st);  // last call's contents
  if (ce != cudaSuccess) return (int)ce;
  int grid = coo_grid(num_edges * H);
  enorm_bwd_sumprod_kernel<<<grid, kCThreads, 0, st>>>(dst, eids, num_edges, norm, grad_norm, sumprod, H);
  HETB_LAUNCH_OK();
  const int use_smem = (num_rels * H <= 4096) ? 1 : 0;
  const size_t smem = use_smem ? (size_t)num_rels * H * sizeof(float) : 0;
  if (grid > kNumSMs * 8) grid = kNumSMs * 8;
  enorm_bwd_apply_kernel<<<grid, kCThreads, smem, st>>>(dst, eids, rel_of_pos, rel_ptrs, (int)num_rels, num_edges, attn,
                                                        norm, grad_norm, mu, sumprod, grad_attn, grad_mu, H, use_smem);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
