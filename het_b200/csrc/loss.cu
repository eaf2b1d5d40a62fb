// Classification loss of the full-graph training step, fused:
//   loss = (1/denom) * sum_i ( logsumexp(logits[i,:]) - logits[i, label[i]] )
// i.e. F.nll_loss(F.log_softmax(logits), labels, reduction="sum") / denom, the loss
// HET_RGNN_train computes every epoch (hrt/python/RGNNUtils/RGNNUtils.py:291-345). torch runs
// it as log_softmax + a nll_loss reduction that takes 2.1 ms at N = 1.94 M rows (one pass over
// 0.5 GB should take 0.08 ms); here one lane group reads a row once, and the backward re-reads
// the logits and the saved logsumexp and writes the gradient in place of two more passes.
#include "common.cuh"

namespace hetb {

constexpr int kLossThreads = 256;
constexpr int kLossLPR = 16;  // lanes per row: a 64-class row is one float4 per lane

__global__ void __launch_bounds__(kLossThreads)
    xent_fwd_kernel(const float *__restrict__ logits, const int64_t *__restrict__ labels, int64_t n, int C,
                    float inv_denom, float *__restrict__ lse, float *__restrict__ loss) {
  constexpr int GROUPS = kLossThreads / kLossLPR;
  constexpr int U = 4;  // rows in flight per lane group (one row at a time ran at 1.8 TB/s)
  const int g = threadIdx.x / kLossLPR, l = threadIdx.x % kLossLPR;
  const int64_t stride = (int64_t)gridDim.x * GROUPS;
  const int64_t iters = ceil_div64(n, stride * U);
  float local = 0.f;
  int64_t i0 = (int64_t)blockIdx.x * GROUPS + g;
  if (C == 4 * kLossLPR) {
    // a row is one float4 per lane: read once, kept in registers for both the max and the sum
    for (int64_t it = 0; it < iters; ++it, i0 += stride * U) {
      float4 v[U];
      int lab[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        lab[u] = 0;
        if (i < n) {
          v[u] = ldcs4(logits + i * C + 4 * l);
          lab[u] = (int)__ldg(labels + i);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        float m = fmaxf(fmaxf(v[u].x, v[u].y), fmaxf(v[u].z, v[u].w));
        for (int o = kLossLPR >> 1; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float s = __expf(v[u].x - m) + __expf(v[u].y - m) + __expf(v[u].z - m) + __expf(v[u].w - m);
        // the label's logit sits in lane lab / 4, component lab % 4
        const int c = lab[u] & 3;
        float pick = (lab[u] >> 2) == l ? (c == 0 ? v[u].x : c == 1 ? v[u].y : c == 2 ? v[u].z : v[u].w) : 0.f;
        s = group_sum(s, kLossLPR);
        pick = group_sum(pick, kLossLPR);
        if (i < n && l == 0) {
          const float z = m + __logf(s);
          lse[i] = z;
          local += z - pick;
        }
      }
    }
  } else {
    for (int64_t it = 0; it < iters * U; ++it, i0 += stride) {
      const int64_t i = i0;
      const bool ok = i < n;
      const float *row = logits + (ok ? i : 0) * C;
      float m = -INFINITY;
      for (int c = 4 * l; c < C; c += 4 * kLossLPR) {
        const float4 v = ldcs4(row + c);
        m = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
      }
      for (int o = kLossLPR >> 1; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      float s = 0.f;
      for (int c = 4 * l; c < C; c += 4 * kLossLPR) {
        const float4 v = ldg4(row + c);  // second pass hits L1/L2
        s += __expf(v.x - m) + __expf(v.y - m) + __expf(v.z - m) + __expf(v.w - m);
      }
      s = group_sum(s, kLossLPR);
      if (ok && l == 0) {
        const float z = m + __logf(s);
        lse[i] = z;
        local += z - __ldg(row + __ldg(labels + i));
      }
    }
  }
  // one atomic per CTA
  __shared__ float part[kLossThreads / 32];
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kLossThreads / 32; ++w) t += part[w];
    atomicAdd(loss, t * inv_denom);
  }
}

// grad[i,c] = gscale * (exp(logits[i,c] - lse[i]) - [c == label[i]]),  gscale = grad_loss / denom
__global__ void __launch_bounds__(kLossThreads)
    xent_bwd_kernel(const float *__restrict__ logits, const int64_t *__restrict__ labels,
                    const float *__restrict__ lse, const float *__restrict__ grad_loss, int64_t n, int C,
                    float inv_denom, float *__restrict__ grad) {
  constexpr int GROUPS = kLossThreads / kLossLPR;
  const int g = threadIdx.x / kLossLPR, l = threadIdx.x % kLossLPR;
  const int64_t stride = (int64_t)gridDim.x * GROUPS;
  const float gs = __ldg(grad_loss) * inv_denom;
  for (int64_t i = (int64_t)blockIdx.x * GROUPS + g; i < n; i += stride) {
    const float z = __ldg(lse + i);
    const int lab = (int)__ldg(labels + i);
    for (int c = 4 * l; c < C; c += 4 * kLossLPR) {
      const float4 v = ldcs4(logits + i * C + c);
      float4 o = make_float4(__expf(v.x - z), __expf(v.y - z), __expf(v.z - z), __expf(v.w - z));
      if (lab == c) o.x -= 1.f;
      if (lab == c + 1) o.y -= 1.f;
      if (lab == c + 2) o.z -= 1.f;
      if (lab == c + 3) o.w -= 1.f;
      stcs4(grad + i * C + c, make_float4(gs * o.x, gs * o.y, gs * o.z, gs * o.w));
    }
  }
}

// Column sum of a row-major [rows, cols] matrix (cols a multiple of 4, <= 256): the gradient of the
// layer bias, d/db (h + b) = sum over rows of the incoming gradient (HET_RGATLayer._epilogue,
// hrt/python/RGAT/models.py:264-278). One lane per 4 columns, CTAs stride the rows, partial sums
// meet in shared memory, one vector reduction per CTA into the zeroed output.
__global__ void __launch_bounds__(kLossThreads)
    column_sum_kernel(const float *__restrict__ x, int64_t rows, int cols, float *__restrict__ out) {
  __shared__ float4 part[kLossThreads];
  const int lpr = cols / 4;                       // lanes per row
  const int groups = kLossThreads / lpr;          // rows per CTA step
  const int g = threadIdx.x / lpr, l = threadIdx.x % lpr;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (g < groups) {
    const int64_t stride = (int64_t)gridDim.x * groups;
    int64_t r = (int64_t)blockIdx.x * groups + g;
    for (; r + 3 * stride < rows; r += 4 * stride) {  // 4 rows in flight per lane
      const float4 a = ldg4(x + r * cols + 4 * l), b = ldg4(x + (r + stride) * cols + 4 * l);
      const float4 c = ldg4(x + (r + 2 * stride) * cols + 4 * l), d = ldg4(x + (r + 3 * stride) * cols + 4 * l);
      acc.x += (a.x + b.x) + (c.x + d.x); acc.y += (a.y + b.y) + (c.y + d.y);
      acc.z += (a.z + b.z) + (c.z + d.z); acc.w += (a.w + b.w) + (c.w + d.w);
    }
    for (; r < rows; r += stride) {
      const float4 a = ldg4(x + r * cols + 4 * l);
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    }
  }
  part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x < lpr) {
    float4 s = part[threadIdx.x];
    for (int k = 1; k < groups; ++k) {
      const float4 v = part[k * lpr + threadIdx.x];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    red_add4(out + 4 * threadIdx.x, s);
  }
}

}  // namespace hetb

using namespace hetb;

extern "C" int hetb_softmax_xent_fwd(const float *logits, const int64_t *labels, int64_t num_rows,
                                     int num_classes, float denom, float *lse, float *loss, void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_classes > 0 && denom > 0.f && loss);
  if (num_classes % 4) return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(loss, 0, sizeof(float), st);
  if (ce != cudaSuccess) return (int)ce;
  if (num_rows == 0) return HETB_OK;
  HETB_CHECK_ARG(logits && labels && lse);
  const int64_t need = ceil_div64(num_rows, kLossThreads / kLossLPR);
  const int grid = (int)(need < kNumSMs * 8 ? need : kNumSMs * 8);
  xent_fwd_kernel<<<grid, kLossThreads, 0, st>>>(logits, labels, num_rows, num_classes, 1.f / denom, lse, loss);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_softmax_xent_bwd(const float *logits, const int64_t *labels, const float *lse,
                                     const float *grad_loss, int64_t num_rows, int num_classes, float denom,
                                     float *grad_logits, void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_classes > 0 && denom > 0.f);
  if (num_classes % 4) return HETB_EUNSUPPORTED;
  if (num_rows == 0) return HETB_OK;
  HETB_CHECK_ARG(logits && labels && lse && grad_loss && grad_logits);
  const int64_t need = ceil_div64(num_rows, kLossThreads / kLossLPR);
  const int grid = (int)(need < kNumSMs * 16 ? need : kNumSMs * 16);
  xent_bwd_kernel<<<grid, kLossThreads, 0, (cudaStream_t)stream>>>(logits, labels, lse, grad_loss, num_rows,
                                                                   num_classes, 1.f / denom, grad_logits);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_column_sum(const float *x, int64_t num_rows, int num_cols, float *out, void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_cols > 0 && out);
  if (num_cols % 4 || num_cols > 4 * kLossThreads || (kLossThreads % (num_cols / 4)) != 0) return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(out, 0, sizeof(float) * (size_t)num_cols, st);
  if (ce != cudaSuccess) return (int)ce;
  if (num_rows == 0) return HETB_OK;
  HETB_CHECK_ARG(x != nullptr);
  const int groups = kLossThreads / (num_cols / 4);
  const int64_t need = ceil_div64(num_rows, (int64_t)groups * 8);
  const int grid = (int)(need < kNumSMs * 4 ? (need < 1 ? 1 : need) : kNumSMs * 4);
  column_sum_kernel<<<grid, kLossThreads, 0, st>>>(x, num_rows, num_cols, out);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
