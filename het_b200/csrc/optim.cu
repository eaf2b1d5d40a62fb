// Optimizer step and per-step weight composition of the full-graph trainer, so that nothing on the
// timed step is a library kernel (round-1 launch lists: torch's multi-tensor Adam 7 % of the RGAT
// step, cuBLAS GEMMs from th.einsum for the composed weights).
//
// * adam_kernel: Adam (Kingma & Ba; the update torch.optim.Adam applies, bias-corrected, no weight
//   decay, no amsgrad) over one tensor: 4 streams in, 3 out, one pass. The step count lives on the
//   device (a float the kernel's first thread block does not touch: a 1-thread kernel bumps it once
//   per optimizer step), so the whole step stays capturable in a CUDA graph. This is what
//   HET_RGNN_train's `optimizer.step()` does for the model and the node embedding
//   (hrt/python/RGNNUtils/RGNNUtils.py:291-345).
// * compose kernels: W[r, i, h*dk+e] = sum_d lin[type[r], i, h*dk+d] * rel[r, h, d, e] — Hector's
//   linear operator reordering for HGT (W_q.W_att, W_v.W_msg per relation; hrt/python/HGT/models.py:124-155
//   `multiply_among_weights_first_flag`), forward and both gradients.
#include "common.cuh"

namespace hetb {

constexpr int kOptThreads = 256;

__global__ void adam_bump_step_kernel(float *step) { *step += 1.f; }

__global__ void __launch_bounds__(kOptThreads)
    adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
                int64_t n, const float *__restrict__ step, float lr, float beta1, float beta2, float eps) {
  const float t = *step;
  const float bc1 = 1.f - powf(beta1, t), bc2 = 1.f - powf(beta2, t);
  const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * kOptThreads;
  for (int64_t i = (int64_t)blockIdx.x * kOptThreads + threadIdx.x; i < n4; i += stride) {
    float4 pp = *reinterpret_cast<float4 *>(p + 4 * i);
    const float4 gg = ldcs4(g + 4 * i);
    float4 mm = *reinterpret_cast<float4 *>(m + 4 * i), vv = *reinterpret_cast<float4 *>(v + 4 * i);
#define ADAM1(c)                                                        \
  mm.c = beta1 * mm.c + (1.f - beta1) * gg.c;                           \
  vv.c = beta2 * vv.c + (1.f - beta2) * gg.c * gg.c;                    \
  pp.c -= step_size * mm.c / (sqrtf(vv.c) * inv_sqrt_bc2 + eps);
    ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
    st4(p + 4 * i, pp);
    st4(m + 4 * i, mm);
    st4(v + 4 * i, vv);
  }
  // tail (n % 4 elements)
  const int64_t base = n4 << 2;
  const int64_t j = base + (int64_t)blockIdx.x * kOptThreads + threadIdx.x;
  if (j < n) {
    const float gj = g[j];
    const float mj = beta1 * m[j] + (1.f - beta1) * gj;
    const float vj = beta2 * v[j] + (1.f - beta2) * gj * gj;
    m[j] = mj;
    v[j] = vj;
    p[j] -= step_size * mj / (sqrtf(vj) * inv_sqrt_bc2 + eps);
  }
#undef ADAM1
}

// out[r, i, h*dk + e] = sum_d lin[type[r], i, h*dk + d] * rel[r, h, d, e]; one thread per output element
__global__ void __launch_bounds__(kOptThreads)
    compose_fwd_kernel(const float *__restrict__ lin, const int64_t *__restrict__ type_of_rel,
                       const float *__restrict__ rel, float *__restrict__ out, int R, int in_dim, int H, int dk) {
  const int D = H * dk;
  const int64_t total = (int64_t)R * in_dim * D;
  for (int64_t t = (int64_t)blockIdx.x * kOptThreads + threadIdx.x; t < total; t += (int64_t)gridDim.x * kOptThreads) {
    const int c = (int)(t % D), i = (int)((t / D) % in_dim), r = (int)(t / ((int64_t)D * in_dim));
    const int h = c / dk, e = c % dk;
    const float *lrow = lin + ((int64_t)__ldg(type_of_rel + r) * in_dim + i) * D + h * dk;
    const float *w = rel + (((int64_t)r * H + h) * dk) * dk + e;
    float acc = 0.f;
    for (int d = 0; d < dk; ++d) acc = fmaf(__ldg(lrow + d), __ldg(w + (int64_t)d * dk), acc);
    out[t] = acc;
  }
}

// grad_lin[type[r], i, h*dk + d] += sum_e gout[r, i, h*dk + e] * rel[r, h, d, e]   (atomic: relations share a type)
__global__ void __launch_bounds__(kOptThreads)
    compose_bwd_lin_kernel(const float *__restrict__ gout, const int64_t *__restrict__ type_of_rel,
                           const float *__restrict__ rel, float *__restrict__ grad_lin, int R, int in_dim, int H,
                           int dk) {
  const int D = H * dk;
  const int64_t total = (int64_t)R * in_dim * D;
  for (int64_t t = (int64_t)blockIdx.x * kOptThreads + threadIdx.x; t < total; t += (int64_t)gridDim.x * kOptThreads) {
    const int c = (int)(t % D), i = (int)((t / D) % in_dim), r = (int)(t / ((int64_t)D * in_dim));
    const int h = c / dk, d = c % dk;
    const float *grow = gout + ((int64_t)r * in_dim + i) * D + h * dk;
    const float *w = rel + (((int64_t)r * H + h) * dk + d) * dk;
    float acc = 0.f;
    for (int e = 0; e < dk; ++e) acc = fmaf(__ldg(grow + e), __ldg(w + e), acc);
    atomicAdd(grad_lin + ((int64_t)__ldg(type_of_rel + r) * in_dim + i) * D + c, acc);
  }
}

// grad_rel[r, h, d, e] = sum_i lin[type[r], i, h*dk + d] * gout[r, i, h*dk + e]; one thread per element
__global__ void __launch_bounds__(kOptThreads)
    compose_bwd_rel_kernel(const float *__restrict__ gout, const int64_t *__restrict__ type_of_rel,
                           const float *__restrict__ lin, float *__restrict__ grad_rel, int R, int in_dim, int H,
                           int dk) {
  const int D = H * dk;
  const int64_t total = (int64_t)R * H * dk * dk;
  for (int64_t t = (int64_t)blockIdx.x * kOptThreads + threadIdx.x; t < total; t += (int64_t)gridDim.x * kOptThreads) {
    const int e = (int)(t % dk), d = (int)((t / dk) % dk), h = (int)((t / ((int64_t)dk * dk)) % H);
    const int r = (int)(t / ((int64_t)dk * dk * H));
    const float *l0 = lin + (int64_t)__ldg(type_of_rel + r) * in_dim * D + h * dk + d;
    const float *g0 = gout + (int64_t)r * in_dim * D + h * dk + e;
    float acc = 0.f;
    for (int i = 0; i < in_dim; ++i) acc = fmaf(__ldg(l0 + (int64_t)i * D), __ldg(g0 + (int64_t)i * D), acc);
    grad_rel[t] = acc;
  }
}

static int opt_grid(int64_t n) {
  const int64_t need = ceil_div64(n, kOptThreads);
  const int64_t cap = (int64_t)kNumSMs * 8;
  return (int)(need < cap ? (need > 0 ? need : 1) : cap);
}

}  // namespace hetb

using namespace hetb;

extern "C" int hetb_adam_bump_step(float *step, void *stream) {
  HETB_CHECK_ARG(step != nullptr);
  adam_bump_step_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, int64_t n,
                              const float *step, float lr, float beta1, float beta2, float eps, void *stream) {
  HETB_CHECK_ARG(n >= 0);
  if (n == 0) return HETB_OK;
  HETB_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && step);
  HETB_CHECK_ARG(((uintptr_t)param % 16) == 0 && ((uintptr_t)grad % 16) == 0 && ((uintptr_t)exp_avg % 16) == 0 &&
                 ((uintptr_t)exp_avg_sq % 16) == 0);
  const int64_t work = (n >> 2) > kOptThreads ? (n >> 2) : kOptThreads;
  adam_kernel<<<opt_grid(work), kOptThreads, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, step, lr,
                                                                         beta1, beta2, eps);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

extern "C" int hetb_compose_weights_fwd(const float *lin, const int64_t *type_of_rel, const float *rel, float *out,
                                        int64_t R, int in_dim, int H, int dk, void *stream) {
  HETB_CHECK_ARG(R >= 0 && in_dim > 0 && H > 0 && dk > 0);
  if (R == 0) return HETB_OK;
  HETB_CHECK_ARG(lin && type_of_rel && rel && out);
  compose_fwd_kernel<<<opt_grid(R * in_dim * H * dk), kOptThreads, 0, (cudaStream_t)stream>>>(lin, type_of_rel, rel, out,
                                                                                             (int)R, in_dim, H, dk);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

// grad_lin must arrive zero-filled (relations that share a node type add into the same rows).
extern "C" int hetb_compose_weights_bwd(const float *gout, const float *lin, const int64_t *type_of_rel,
                                        const float *rel, float *grad_lin, float *grad_rel, int64_t R, int in_dim,
                                        int H, int dk, void *stream) {
  HETB_CHECK_ARG(R >= 0 && in_dim > 0 && H > 0 && dk > 0);
  if (R == 0) return HETB_OK;
  HETB_CHECK_ARG(gout && lin && type_of_rel && rel && grad_lin && grad_rel);
  cudaStream_t st = (cudaStream_t)stream;
  compose_bwd_lin_kernel<<<opt_grid(R * in_dim * H * dk), kOptThreads, 0, st>>>(gout, type_of_rel, rel, grad_lin, (int)R,
                                                                               in_dim, H, dk);
  HETB_LAUNCH_OK();
  compose_bwd_rel_kernel<<<opt_grid(R * H * dk * dk), kOptThreads, 0, st>>>(gout, type_of_rel, lin, grad_rel, (int)R,
                                                                           in_dim, H, dk);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
