// Plan-based fused edge passes (the benchmark's default path).
//
// A *walk plan* (include/hetb200.h: hetb_walk_plan) is built once per graph (het_b200/graph.py:
// HetGraph.get_edge_plan) and holds int32-narrowed, position-aligned index arrays in the order the
// kernel walks them, plus the work decomposition:
//   * positions are sorted by (destination, relation) for the forward passes and by (source,
//     relation) for the backward passes, so the compact (relation, node) row on the walked side is
//     constant over a run of consecutive positions: it is fetched from DRAM once per run (the
//     repeated per-position loads hit L1) and its gradient is a register accumulator;
//   * work items are runs of ~32 consecutive positions whose boundaries are snapped to row / run
//     starts, so only rows longer than an item (power-law hubs) are ever split across lane groups;
//   * rows without positions are listed, not skipped over at run time; the row of an item's first
//     position is stored, not searched.
// Softmax is the max-subtracted online form (north star; the reference kernels overflow at
// logit > 88: GAT/FusedGAT.cu.h:193-203, HGT/HGTForwardKernels.cu.h:499-508): a lane group carries
// (running max m, sum s, accumulator acc) per row and head, rescaling on a new max. A row split
// across items leaves one partial (m, s, acc) per item in a scratch slot and an atomic max of m in
// the row's slot of `lse`; a second kernel, parallel over the partial slots, adds
// exp(m - M)·(s, acc) into the row, a third normalises the split rows and leaves
// lse = M + log(sum). The backward recomputes the normalised score as exp(x - lse[dst]), so the
// per-edge exp() of the reference is never stored.
//
// T = float or __nv_bfloat16 is the STORAGE type of the gathered tables (the compact rows and the
// node rows read per edge); all arithmetic and every accumulator / gradient is fp32.
#include <cuda_bf16.h>

#include "common.cuh"

namespace hetb {

constexpr int kPThreads = 256;
constexpr int kMaxSmemMuPlan = 4096;
// tuning knobs (tools/sweep_edge_plan.py rebuilds with -D...): positions in flight per lane group in the
// forward / backward walks, resident CTAs per SM the register allocation is capped for
#ifndef HETB_PLAN_FWD_U
#define HETB_PLAN_FWD_U 1
#endif
#ifndef HETB_PLAN_BWD_U
#define HETB_PLAN_BWD_U 1
#endif
#ifndef HETB_PLAN_FWD_MINB
#define HETB_PLAN_FWD_MINB 6
#endif
#ifndef HETB_PLAN_BWD_MINB
#define HETB_PLAN_BWD_MINB 4
#endif
#ifndef HETB_PLAN_PREFETCH
#define HETB_PLAN_PREFETCH 1  // fetch the next batch's indices while this batch's rows are in flight
#endif

template <typename T>
__device__ __forceinline__ float4 ld_row4(const T *base, int64_t row, int D, int l);

template <>
__device__ __forceinline__ float4 ld_row4<float>(const float *base, int64_t row, int D, int l) {
  return ldg4(base + row * D + 4 * l);
}
template <>
__device__ __forceinline__ float4 ld_row4<__nv_bfloat16>(const __nv_bfloat16 *base, int64_t row, int D, int l) {
  const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(base + row * D + 4 * l));
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162 *>(&raw.x);
  const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162 *>(&raw.y);
  const float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}

// Sum over the `width` lanes of one head inside a lane group; only the lanes of `mask` take part,
// so the lane groups of one warp may be at different trip counts.
__device__ __forceinline__ float group_sum_m(float v, int width, unsigned mask) {
  for (int o = width >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}

// HW > 0: the head width is a compile-time constant (the shapes of the BASELINE configs), the reduction is
// HW/2 .. 1 shuffles with no loop; HW == 0: the run-time loop above. The walks are bound by issue slots as
// much as by memory (58 % of the slots busy at mag size), and the loop form costs ~11 instructions per step.
template <int HW>
__device__ __forceinline__ float group_sum_t(float v, int width, unsigned mask) {
  if (HW == 0) return group_sum_m(v, width, mask);
#pragma unroll
  for (int o = HW >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}

// The raw attention score <left, kv> with a FIXED contraction order: the forward walk and the backward walk
// (which recomputes the score instead of reading a saved copy) must produce the same bits, and the compiler
// is free to contract a*b + c*d either way in different contexts.
__device__ __forceinline__ float score_dot4(float4 a, float4 b) {
  return __fmaf_rn(a.w, b.w, __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, __fmul_rn(a.x, b.x))));
}

// 2^x, one MUFU: inputs below -126 give 0 (softmax weights that small do not matter), no denormal fix-up.
// The walks keep their logits in the log2 domain (x * log2(e)) so that this is the whole exponential.
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;

__device__ __forceinline__ void atomic_max_float(float *addr, float v) {
  if (v >= 0.f) atomicMax(reinterpret_cast<int *>(addr), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned *>(addr), __float_as_uint(v));
}

template <int LPR>
__device__ __forceinline__ unsigned group_mask() {
  if (LPR == 32) return 0xffffffffu;
  const unsigned base = (threadIdx.x & 31) / LPR * LPR;
  return ((1u << LPR) - 1u) << base;
}

constexpr float kNegInf = -3.0e38f;  // finite stand-in for -inf: exp(kNegInf - x) == 0, no inf - inf

// ---------------------------------------------------------------------------
// Pre-pass of a forward walk: rows listed in fix_rows (split across items) get ret = 0,
// lse = -"inf" (atomic-max target), ssum = 0; rows in empty_rows get ret = 0, lse = 0.
// ---------------------------------------------------------------------------
template <int LPR>
__global__ void __launch_bounds__(kPThreads)
    plan_fwd_prepare_kernel(const int *__restrict__ fix_rows, int num_fix, const int *__restrict__ empty_rows,
                            int num_empty, float *__restrict__ ret, float *__restrict__ lse,
                            float *__restrict__ ssum, int H, int hw) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int i = blockIdx.x * GROUPS + g;
  if (i >= num_fix + num_empty) return;
  const bool fix = i < num_fix;
  const int64_t d = fix ? __ldg(fix_rows + i) : __ldg(empty_rows + (i - num_fix));
  st4(ret + d * D + 4 * l, make_float4(0.f, 0.f, 0.f, 0.f));
  if ((l % hw) == 0) {
    lse[d * H + l / hw] = fix ? kNegInf : 0.f;
    if (fix) ssum[d * H + l / hw] = 0.f;
  }
}

// Partial slots of split rows -> the row: w = exp(m_slot - M_row); ret += w·acc, ssum += w·s.
template <int LPR>
__global__ void __launch_bounds__(kPThreads)
    plan_fwd_merge_kernel(const int *__restrict__ pslot, const int *__restrict__ pslot_row, int num_pslots,
                          const float *__restrict__ part_m, const float *__restrict__ part_s,
                          const float *__restrict__ part_acc, const float *__restrict__ lse /* holds M */,
                          float *__restrict__ ssum, float *__restrict__ ret, int H, int hw) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int i = blockIdx.x * GROUPS + g;
  if (i >= num_pslots) return;
  const int64_t x = __ldg(pslot + i), d = __ldg(pslot_row + i);
  const int h = l / hw;
  const float w = ex2_ftz(__ldg(part_m + x * H + h) - __ldg(lse + d * H + h));  // both in the log2 domain
  const float4 a = ldcs4(part_acc + x * D + 4 * l);
  red_add4(ret + d * D + 4 * l, make_float4(w * a.x, w * a.y, w * a.z, w * a.w));
  if ((l % hw) == 0) atomicAdd(ssum + d * H + h, w * __ldg(part_s + x * H + h));
}

// Split rows: ret /= ssum; lse = M + log(ssum) (natural units on return).
template <int LPR>
__global__ void __launch_bounds__(kPThreads)
    plan_fwd_normalize_kernel(const int *__restrict__ fix_rows, int num_fix, const float *__restrict__ ssum,
                              float *__restrict__ lse, float *__restrict__ ret, int H, int hw) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int i = blockIdx.x * GROUPS + g;
  if (i >= num_fix) return;
  const int64_t d = __ldg(fix_rows + i);
  const int h = l / hw;
  const float s = ssum[d * H + h];
  const float inv = s > 0.f ? 1.f / s : 0.f;
  float4 v = *reinterpret_cast<const float4 *>(ret + d * D + 4 * l);
  st4(ret + d * D + 4 * l, make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv));
  if ((l % hw) == 0) lse[d * H + h] = lse[d * H + h] * kLn2 + __logf(s);  // M was kept in the log2 domain
}

// ---------------------------------------------------------------------------
// HGT forward: attention score + edge softmax + weighted aggregation in one walk over the
// (destination, relation)-sorted positions. For position p of destination d:
//   a[p,h] = <left[L(p),h,:], kv[src(p),h,:]>            (RGNN/InnerProductEdgeParallel.cu.h:13-113)
//   x      = a[p,h] · mu[rel(L(p)),h]                     (HGT/HGTForwardKernels.cu.h:425-587)
//   ret[d,h,:] = Σ_p softmax_p(x) · msg[M(p),h,:]         (HGT/HGTForwardKernels.cu.h:244-369)
// Saved for the backward: a (position order, 4·H bytes per edge) and lse[d,h].
// MODE_RGAT (same walk, RGAT's logits): x = leaky_relu(el[M(p),h] + er[L(p),h]), msg = feat
// (GAT/FusedGAT.cu.h:123-222,33-116); nothing per edge is saved (the backward recomputes x).
// ---------------------------------------------------------------------------
template <typename T, int LPR, bool MODE_RGAT, int HW>
__global__ void __launch_bounds__(kPThreads, HETB_PLAN_FWD_MINB)
    plan_fwd_walk_kernel(const int *__restrict__ item_ptr, const int *__restrict__ item_row, int num_items,
                         const int *__restrict__ nz_rows, const int *__restrict__ nz_ptr,
                         const int *__restrict__ p_src, const int *__restrict__ p_msg,
                         const int *__restrict__ p_left, const int *__restrict__ ucol_rel,
                         const T *__restrict__ left, const T *__restrict__ kv, const T *__restrict__ msg,
                         const float *__restrict__ mu,            // HGT: [R,H]
                         const float *__restrict__ el, const float *__restrict__ er, float slope,  // RGAT
                         float *__restrict__ attn, float *__restrict__ lse, float *__restrict__ ret,
                         float *__restrict__ part_m, float *__restrict__ part_s, float *__restrict__ part_acc,
                         int H, int hw_rt) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  constexpr int U = HETB_PLAN_FWD_U;  // positions in flight per lane group
  const int hw = HW ? HW : hw_rt;
  const int grp = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int g = blockIdx.x * GROUPS + grp;
  if (g >= num_items) return;  // whole lane group leaves: its lanes are in no other group's mask
  // Sub-warp shuffle masks on purpose: the lane groups of a warp run their items INDEPENDENTLY (they diverge at
  // the first row change and are scheduled as separate instruction streams from then on), which is worth more
  // than the ~8 instructions per shuffle the compiler spends on the convergence check of a partial mask: a
  // lockstep version with full-warp shuffles and 30 % fewer instructions measured 2.55 ms against 2.07
  // (profiles/r02u_walk_variants.log).
  const unsigned gm = group_mask<LPR>();
  const int h = l / hw;
  const bool leader = (l % hw) == 0;
  const int p0 = __ldg(item_ptr + g), p1 = __ldg(item_ptr + g + 1);
  int j = __ldg(item_row + g);
  int64_t d = __ldg(nz_rows + j) & 0x7fffffff;
  int row_start = __ldg(nz_ptr + j), row_end = __ldg(nz_ptr + j + 1);
  float m = kNegInf, s = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);

  auto flush = [&]() {
    if (row_start >= p0 && row_end <= p1) {  // the whole row was walked here
      const float inv = s > 0.f ? 1.f / s : 0.f;
      st4(ret + d * D + 4 * l, make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv));
      if (leader) lse[d * H + h] = m * kLn2 + __logf(s);  // m is kept in the log2 domain
    } else {  // a piece of a split row: partial (m, s, acc) into this item's head / tail slot
      const int64_t x = 2 * (int64_t)g + (row_start < p0 ? 0 : 1);
      stcs4(part_acc + x * D + 4 * l, acc);
      if (leader) {
        part_m[x * H + h] = m;
        part_s[x * H + h] = s;
        atomic_max_float(lse + d * H + h, m);
      }
    }
  };

  // Software pipeline: the index triple of batch b+1 is fetched while the rows of batch b are in
  // flight, so a batch costs one memory round trip (the rows), not two (indices, then rows).
  int sn[U], mi[U], li[U];
  auto load_idx = [&](int p, int *s_, int *m_, int *l_) {
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int pp = (p + q < p1) ? p + q : p1 - 1;  // clamp: the tail repeats the last position (masked below)
      m_[q] = __ldg(p_msg + pp);
      l_[q] = __ldg(p_left + pp);
      if (!MODE_RGAT) s_[q] = __ldg(p_src + pp);
    }
  };
  if (HETB_PLAN_PREFETCH) load_idx(p0, sn, mi, li);
  for (int p = p0; p < p1; p += U) {
    float xs[U];
    float4 lv[U], kvv[U], mv[U];
    if (!HETB_PLAN_PREFETCH) load_idx(p, sn, mi, li);
#pragma unroll
    for (int q = 0; q < U; ++q) {
      mv[q] = ld_row4<T>(msg, mi[q], D, l);
      if (!MODE_RGAT) {
        lv[q] = ld_row4<T>(left, li[q], D, l);
        kvv[q] = ld_row4<T>(kv, sn[q], D, l);
        xs[q] = __ldg(mu + __ldg(ucol_rel + li[q]) * H + h) * kLog2e;
      } else {
        xs[q] = __ldg(el + (int64_t)mi[q] * H + h) + __ldg(er + (int64_t)li[q] * H + h);
      }
    }
    if (HETB_PLAN_PREFETCH && p + U < p1) load_idx(p + U, sn, mi, li);  // next batch (the loads above already hold their addresses)
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int pp = p + q;
      if (pp < p1) {  // uniform inside the lane group
        if (pp == row_end) {
          flush();
          ++j;
          d = __ldg(nz_rows + j) & 0x7fffffff;
          row_start = pp;
          row_end = __ldg(nz_ptr + j + 1);
          m = kNegInf; s = 0.f;
          acc = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float x;
        if (!MODE_RGAT) {
          const float a = group_sum_t<HW>(score_dot4(lv[q], kvv[q]), hw, gm);
          if (leader && attn) attn[(int64_t)pp * H + h] = a;
          x = a * xs[q];
        } else {
          x = (xs[q] > 0.f ? xs[q] : slope * xs[q]) * kLog2e;
        }
        const float mn = fmaxf(m, x);  // log2 domain: exp(v) = 2^(v·log2 e)
        const float sc = ex2_ftz(m - mn), w = ex2_ftz(x - mn);
        acc.x = fmaf(acc.x, sc, w * mv[q].x);
        acc.y = fmaf(acc.y, sc, w * mv[q].y);
        acc.z = fmaf(acc.z, sc, w * mv[q].z);
        acc.w = fmaf(acc.w, sc, w * mv[q].w);
        s = fmaf(s, sc, w);
        m = mn;
      }
    }
  }
  flush();
}

// ---------------------------------------------------------------------------
// plan_fwd_walk_kernel for 64-float rows with EIGHT lanes per lane group (two float4 per lane and row): four
// independent instruction streams per warp instead of two, and the per-position scalar work (indices, logits,
// exponentials, row bookkeeping) is executed by half as many lanes. Same arithmetic per element and the same
// reduction tree over a head's 16 floats at 4 heads (the 4 per-float4 partial sums are combined in the same
// order), so the raw score there is the 16-lane kernel's — and the backward's recomputed one — bit for bit.
// HETB_FWD_LANES8 selects it (default on for 64-float rows with 4 heads of 16 or 1 head of 64).
// ---------------------------------------------------------------------------
#ifndef HETB_FWD_LANES8
#define HETB_FWD_LANES8 1  // measured at mag size: 1.91 ms against 2.06 ms for the 16-lane kernel (4 CTAs/SM; 2.17 at 3 or 5)
#endif
#ifndef HETB_PLAN_FWD8_MINB
#define HETB_PLAN_FWD8_MINB 4
#endif

template <typename T>
__device__ __forceinline__ void ld_row8(const T *base, int64_t row, int l, float4 &a, float4 &b) {
  a = ld_row4<T>(base, row, 64, 2 * l);
  b = ld_row4<T>(base, row, 64, 2 * l + 1);
}

template <typename T, bool MODE_RGAT, int HW /* lanes per head: 2 (H = 4) or 8 (H = 1) */>
__global__ void __launch_bounds__(kPThreads, HETB_PLAN_FWD8_MINB)
    plan_fwd_walk8_kernel(const int *__restrict__ item_ptr, const int *__restrict__ item_row, int num_items,
                          const int *__restrict__ nz_rows, const int *__restrict__ nz_ptr,
                          const int *__restrict__ p_src, const int *__restrict__ p_msg,
                          const int *__restrict__ p_left, const int *__restrict__ ucol_rel,
                          const T *__restrict__ left, const T *__restrict__ kv, const T *__restrict__ msg,
                          const float *__restrict__ mu, const float *__restrict__ el,
                          const float *__restrict__ er, float slope, float *__restrict__ attn,
                          float *__restrict__ lse, float *__restrict__ ret, float *__restrict__ part_m,
                          float *__restrict__ part_s, float *__restrict__ part_acc, int H) {
  constexpr int LPR = 8, D = 64, GROUPS = kPThreads / LPR;
  const int grp = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int g = blockIdx.x * GROUPS + grp;
  if (g >= num_items) return;
  const unsigned gm = group_mask<LPR>();
  const int h = l / HW;
  const bool leader = (l % HW) == 0;
  const int p0 = __ldg(item_ptr + g), p1 = __ldg(item_ptr + g + 1);
  int j = __ldg(item_row + g);
  int64_t d = __ldg(nz_rows + j) & 0x7fffffff;
  int row_start = __ldg(nz_ptr + j), row_end = __ldg(nz_ptr + j + 1);
  float m = kNegInf, s = 0.f;
  float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0;

  auto flush = [&]() {
    if (row_start >= p0 && row_end <= p1) {
      const float inv = s > 0.f ? 1.f / s : 0.f;
      st4(ret + d * D + 8 * l, make_float4(acc0.x * inv, acc0.y * inv, acc0.z * inv, acc0.w * inv));
      st4(ret + d * D + 8 * l + 4, make_float4(acc1.x * inv, acc1.y * inv, acc1.z * inv, acc1.w * inv));
      if (leader) lse[d * H + h] = m * kLn2 + __logf(s);
    } else {
      const int64_t x = 2 * (int64_t)g + (row_start < p0 ? 0 : 1);
      stcs4(part_acc + x * D + 8 * l, acc0);
      stcs4(part_acc + x * D + 8 * l + 4, acc1);
      if (leader) {
        part_m[x * H + h] = m;
        part_s[x * H + h] = s;
        atomic_max_float(lse + d * H + h, m);
      }
    }
  };

  int sn = 0, mi, li;
  auto load_idx = [&](int p) {
    const int pp = p < p1 ? p : p1 - 1;
    mi = __ldg(p_msg + pp);
    li = __ldg(p_left + pp);
    if (!MODE_RGAT) sn = __ldg(p_src + pp);
  };
  load_idx(p0);
  for (int p = p0; p < p1; ++p) {
    float4 mv0, mv1, lv0, lv1, kv0, kv1;
    float xs;
    ld_row8<T>(msg, mi, l, mv0, mv1);
    if (!MODE_RGAT) {
      ld_row8<T>(left, li, l, lv0, lv1);
      ld_row8<T>(kv, sn, l, kv0, kv1);
      xs = __ldg(mu + __ldg(ucol_rel + li) * H + h) * kLog2e;
    } else {
      xs = __ldg(el + (int64_t)mi * H + h) + __ldg(er + (int64_t)li * H + h);
    }
    if (p + 1 < p1) load_idx(p + 1);
    if (p == row_end) {
      flush();
      ++j;
      d = __ldg(nz_rows + j) & 0x7fffffff;
      row_start = p;
      row_end = __ldg(nz_ptr + j + 1);
      m = kNegInf; s = 0.f;
      acc0 = acc1 = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float x;
    if (!MODE_RGAT) {
      // the 16-lane kernel sums a head's four float4 dots as (d0 + d2) + (d1 + d3); lane l here holds d_{2l}, d_{2l+1}
      // of the row, i.e. (d0, d1) / (d2, d3) of its head: xor-1 first would give (d0 + d1) + (d2 + d3). Reproduce
      // the 16-lane order: exchange the odd partial first.
      const float da = score_dot4(lv0, kv0), db = score_dot4(lv1, kv1);
      float a;
      if (HW == 2) {
        const float oa = __shfl_xor_sync(gm, da, 1), ob = __shfl_xor_sync(gm, db, 1);
        a = (da + oa) + (db + ob);
      } else {
        a = group_sum_t<HW>(da + db, HW, gm);
      }
      if (leader && attn) attn[(int64_t)p * H + h] = a;
      x = a * xs;
    } else {
      x = (xs > 0.f ? xs : slope * xs) * kLog2e;
    }
    const float mn = fmaxf(m, x);
    const float sc = ex2_ftz(m - mn), w = ex2_ftz(x - mn);
    acc0.x = fmaf(acc0.x, sc, w * mv0.x);
    acc0.y = fmaf(acc0.y, sc, w * mv0.y);
    acc0.z = fmaf(acc0.z, sc, w * mv0.z);
    acc0.w = fmaf(acc0.w, sc, w * mv0.w);
    acc1.x = fmaf(acc1.x, sc, w * mv1.x);
    acc1.y = fmaf(acc1.y, sc, w * mv1.y);
    acc1.z = fmaf(acc1.z, sc, w * mv1.z);
    acc1.w = fmaf(acc1.w, sc, w * mv1.w);
    s = fmaf(s, sc, w);
    m = mn;
  }
  flush();
}

// ---------------------------------------------------------------------------
// Pre-pass of a backward walk: zero the rows of `buf` ([*, D]) listed in rows[0..n).
// ---------------------------------------------------------------------------
template <int LPR>
__global__ void __launch_bounds__(kPThreads)
    plan_zero_rows_kernel(const int *__restrict__ rows, int n, float *__restrict__ buf, float *__restrict__ aux,
                          int H, int hw) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int i = blockIdx.x * GROUPS + g;
  if (i >= n) return;
  const int64_t r = __ldg(rows + i) & 0x7fffffff;
  st4(buf + r * D + 4 * l, make_float4(0.f, 0.f, 0.f, 0.f));
  if (aux && (l % hw) == 0) aux[r * H + l / hw] = 0.f;
}

// <a[d,h,:], b[d,h,:]> per (row, head): the softmax Jacobian's per-destination term <gout, out>, stored
// NEXT TO lse[d,h] as out[d] = (lse[d,0..H), dot[d,0..H)): the backward walks read both per edge, and what a
// random access costs the memory system is the access, not its 16 bytes (37 G rows/s for anything up to
// 128 B, profiles/r02_random_rows.log) — one 32-byte read per edge instead of two 16-byte reads.
template <int LPR>
__global__ void __launch_bounds__(kPThreads)
    plan_rowdot_kernel(const float *__restrict__ a, const float *__restrict__ b, const float *__restrict__ lse,
                       int64_t num_rows, float *__restrict__ out, int H, int hw) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int64_t d = (int64_t)blockIdx.x * GROUPS + g;
  if (d >= num_rows) return;
  const unsigned gm = group_mask<LPR>();
  const float4 x = ldg4(a + d * D + 4 * l), y = ldg4(b + d * D + 4 * l);
  const float v = group_sum_m(dot4(x, y), hw, gm);
  if ((l % hw) == 0) {
    out[d * 2 * H + l / hw] = __ldg(lse + d * H + l / hw) * kLog2e;  // the walks exponentiate in base 2
    out[d * 2 * H + H + l / hw] = v;
  }
}

// ---------------------------------------------------------------------------
// HGT backward of plan_fwd_walk_kernel in ONE walk over the (source, relation)-sorted positions
// (replaces HGTBackwardKernels.cu.h:343-440, 499-627 and InnerProductEdgeParallel.cu.h:119-202).
// For position p of source s, run u = M(p) (compact (relation, source) row), destination d:
//   nrm = exp(a[q]·mu − lse[d,h])  (q = the forward position of the edge)
//   sg  = (<gout[d,h,:], msg[u,h,:]> − <gout[d,h,:], out[d,h,:]>) · nrm
//   ga  = sg · mu[r,h];   grad_mu[r,h] += sg · a[q]
//   grad_msg[u,h,:]    += nrm · gout[d,h,:]      register accumulator of the run, one store
//   grad_kv[s,h,:]     += ga  · left[L(p),h,:]   register accumulator of the row, one store
//   grad_left[L(p),h,:] += ga · kv[s,h,:]         vector reduction (destination side: the only
//                                                 gradient the source-major order cannot own)
// Rows / runs split across items (bit 31 of nz_rows[j] / urow_rel[u]) are pre-zeroed and reduced.
// TWO_PASS: the destination-side gradient is not reduced here. The walk leaves ga — the gradient of the raw
// score — in `gattn` at the edge's forward position (in place of the score it has just read there: the
// sector is in L2), and plan_hgt_grad_left_kernel, a second walk in the FORWARD (destination, relation)
// order, owns grad_left[L] as a register accumulator: 256 B of kv[src] per edge instead of the 512 B
// read-modify-write of a vector reduction, no zero-fill of grad_left, and a result that does not depend
// on the order atomics land in.
// ---------------------------------------------------------------------------
// SCORE: where the raw score a comes from. kRecompute (the default): a = <left[L], kv[s]> again — both rows
// are in registers (kv[s] is the row constant, left[L] is gathered for grad_kv anyway), same operation order
// as the forward, so the same bits — and the forward saves NOTHING per edge; kSaved: read from attn at the
// edge's forward position (one more random access per edge: 3.99 vs 3.0x ms at mag size).
enum ScoreSource : int { kSaved = 0, kRecompute = 1 };

template <typename T, int LPR, bool TWO_PASS, int SCORE, int HW>
__global__ void __launch_bounds__(kPThreads, HETB_PLAN_BWD_MINB)
    plan_hgt_bwd_walk_kernel(const int *__restrict__ item_ptr, const int *__restrict__ item_row, int num_items,
                             const int *__restrict__ nz_rows, const int *__restrict__ nz_ptr,
                             const int *__restrict__ p_dst, const int *__restrict__ p_left,
                             const int *__restrict__ p_msg, const int *__restrict__ p_perm,
                             const int *__restrict__ urow_rel, int num_rels, const T *__restrict__ left,
                             const T *__restrict__ kv, const T *__restrict__ msg, const float *__restrict__ mu,
                             const float *__restrict__ attn,
                             const float *__restrict__ lse_gdoto /* [N, 2H]: plan_rowdot_kernel */,
                             const float *__restrict__ gout,
                             float *__restrict__ grad_left, float *__restrict__ grad_kv,
                             float *__restrict__ grad_msg, float *__restrict__ grad_mu, int H, int hw_rt,
                             int use_smem, float *gattn /* TWO_PASS: score in, ga out; not restrict, never __ldg */) {
  const int hw = HW ? HW : hw_rt;
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  constexpr int U = HETB_PLAN_BWD_U;
  extern __shared__ float mu_acc[];
  const int RH = num_rels * H;
  if (use_smem) {
    for (int i = threadIdx.x; i < RH; i += kPThreads) mu_acc[i] = 0.f;
    __syncthreads();
  }
  const int grp = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int g = blockIdx.x * GROUPS + grp;
  const unsigned gm = group_mask<LPR>();  // sub-warp masks on purpose: see plan_fwd_walk_kernel
  const int h = l / hw;
  const bool leader = (l % hw) == 0;
  if (g < num_items) {
    const int p0 = __ldg(item_ptr + g), p1 = __ldg(item_ptr + g + 1);
    int j = __ldg(item_row + g);
    int rowword = __ldg(nz_rows + j);
    int64_t srow = rowword & 0x7fffffff;
    int row_end = __ldg(nz_ptr + j + 1);
    constexpr bool kNeedKv = !TWO_PASS || SCORE == kRecompute;
    float4 kvr = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kNeedKv) kvr = ld_row4<T>(kv, srow, D, l);
    float4 acc_k = make_float4(0.f, 0.f, 0.f, 0.f);
    // run state
    int cur_u = -1, relword = 0;
    float4 msgr = make_float4(0.f, 0.f, 0.f, 0.f), acc_m = make_float4(0.f, 0.f, 0.f, 0.f);
    float mu_r = 0.f, mu_r2 = 0.f, acc_mu = 0.f;

    auto flush_run = [&]() {
      if (cur_u < 0) return;
      if (relword < 0) red_add4(grad_msg + (int64_t)cur_u * D + 4 * l, acc_m);  // split run
      else st4(grad_msg + (int64_t)cur_u * D + 4 * l, acc_m);
      if (leader) {
        const int r = relword & 0x7fffffff;
        if (use_smem) atomicAdd(&mu_acc[r * H + h], acc_mu);
        else atomicAdd(grad_mu + r * H + h, acc_mu);
      }
    };
    auto flush_row = [&]() {
      if (rowword < 0) red_add4(grad_kv + srow * D + 4 * l, acc_k);  // split row
      else st4(grad_kv + srow * D + 4 * l, acc_k);
    };

    int dn[U], li[U], ui[U], pq[U];
    auto load_idx = [&](int p) {
#pragma unroll
      for (int q = 0; q < U; ++q) {
        const int pp = (p + q < p1) ? p + q : p1 - 1;
        dn[q] = __ldg(p_dst + pp);
        li[q] = __ldg(p_left + pp);
        ui[q] = __ldg(p_msg + pp);
        if (SCORE == kSaved || TWO_PASS) pq[q] = __ldg(p_perm + pp);
      }
    };
    if (HETB_PLAN_PREFETCH) load_idx(p0);
    for (int p = p0; p < p1; p += U) {
      int lic[U], uic[U], pqc[U];
      if (!HETB_PLAN_PREFETCH) load_idx(p);
      float aa[U], ls[U], cc[U];
      float4 go[U], lv[U];
#pragma unroll
      for (int q = 0; q < U; ++q) {
        go[q] = ldg4(gout + (int64_t)dn[q] * D + 4 * l);
        lv[q] = ld_row4<T>(left, li[q], D, l);
        if (SCORE == kSaved) aa[q] = TWO_PASS ? gattn[(int64_t)pq[q] * H + h] : __ldg(attn + (int64_t)pq[q] * H + h);
        ls[q] = __ldg(lse_gdoto + (int64_t)dn[q] * 2 * H + h);
        cc[q] = __ldg(lse_gdoto + (int64_t)dn[q] * 2 * H + H + h);
        lic[q] = li[q];
        uic[q] = ui[q];
        pqc[q] = pq[q];
      }
      if (HETB_PLAN_PREFETCH && p + U < p1) load_idx(p + U);  // indices of the next batch while this batch's rows are in flight
#pragma unroll
      for (int q = 0; q < U; ++q) {
        const int pp = p + q;
        if (pp < p1) {  // uniform inside the lane group
          if (pp == row_end) {
            flush_run();
            cur_u = -1;
            flush_row();
            ++j;
            rowword = __ldg(nz_rows + j);
            srow = rowword & 0x7fffffff;
            row_end = __ldg(nz_ptr + j + 1);
            if (kNeedKv) kvr = ld_row4<T>(kv, srow, D, l);
            acc_k = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          if (uic[q] != cur_u) {
            flush_run();
            cur_u = uic[q];
            relword = __ldg(urow_rel + cur_u);
            msgr = ld_row4<T>(msg, cur_u, D, l);
            mu_r = __ldg(mu + (relword & 0x7fffffff) * H + h);
            mu_r2 = mu_r * kLog2e;
            acc_m = make_float4(0.f, 0.f, 0.f, 0.f);
            acc_mu = 0.f;
          }
          if (SCORE == kRecompute) aa[q] = group_sum_t<HW>(score_dot4(lv[q], kvr), hw, gm);  // the forward's a, bit for bit
          const float sdot = group_sum_t<HW>(dot4(go[q], msgr), hw, gm) - cc[q];
          const float nrm = ex2_ftz(aa[q] * mu_r2 - ls[q]);  // ls = lse·log2(e) (plan_rowdot_kernel)
          const float sg = sdot * nrm;
          const float ga = sg * mu_r;
          acc_mu = fmaf(sg, aa[q], acc_mu);
          fma4(acc_m, nrm, go[q]);
          fma4(acc_k, ga, lv[q]);
          if (TWO_PASS) {
            if (leader) gattn[(int64_t)pqc[q] * H + h] = ga;
          } else {
            red_add4(grad_left + (int64_t)lic[q] * D + 4 * l,
                     make_float4(ga * kvr.x, ga * kvr.y, ga * kvr.z, ga * kvr.w));
          }
        }
      }
    }
    flush_run();
    flush_row();
  }
  if (use_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < RH; i += kPThreads) {
      const float v = mu_acc[i];
      if (v != 0.f) atomicAdd(grad_mu + i, v);
    }
  }
}

// ---------------------------------------------------------------------------
// Second walk of the two-pass HGT backward, over the FORWARD plan ((destination, relation)-sorted):
//   grad_left[L,h,:] = sum over the positions p of run L of ga[p,h] * kv[src(p),h,:]
// (the destination half of RGNN/InnerProductEdgeParallel.cu.h:119-202, which the reference accumulates
// with one atomicAdd per edge and element). A (destination, relation) run is a register accumulator
// stored once; a run cut by an item boundary (hub rows only: listed in the plan's fix_runs and
// pre-zeroed) reduces its pieces. ga is read in position order (coalesced).
// ---------------------------------------------------------------------------
template <typename T, int LPR>
__global__ void __launch_bounds__(kPThreads, HETB_PLAN_FWD_MINB)
    plan_hgt_grad_left_kernel(const int *__restrict__ item_ptr, int num_items, int num_positions,
                              const int *__restrict__ p_src, const int *__restrict__ p_left,
                              const T *__restrict__ kv, const float *__restrict__ gattn,
                              float *__restrict__ grad_left, int H, int hw) {
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  constexpr int U = HETB_PLAN_FWD_U;
  const int grp = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int g = blockIdx.x * GROUPS + grp;
  if (g >= num_items) return;
  const int h = l / hw;
  const int p0 = __ldg(item_ptr + g), p1 = __ldg(item_ptr + g + 1);
  int cur = __ldg(p_left + p0);
  bool open_head = p0 > 0 && __ldg(p_left + p0 - 1) == cur;  // the first run began in an earlier item
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  auto flush = [&](bool open_tail) {
    float *dst = grad_left + (int64_t)cur * D + 4 * l;
    if (open_head || open_tail) red_add4(dst, acc);
    else st4(dst, acc);
  };
  int sn[U], li[U];
  auto load_idx = [&](int p) {
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int pp = (p + q < p1) ? p + q : p1 - 1;
      sn[q] = __ldg(p_src + pp);
      li[q] = __ldg(p_left + pp);
    }
  };
  if (HETB_PLAN_PREFETCH) load_idx(p0);
  for (int p = p0; p < p1; p += U) {
    if (!HETB_PLAN_PREFETCH) load_idx(p);
    float4 kvv[U];
    float ga[U];
    int lic[U];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      kvv[q] = ld_row4<T>(kv, sn[q], D, l);
      const int pp = (p + q < p1) ? p + q : p1 - 1;
      ga[q] = __ldcs(gattn + (int64_t)pp * H + h);
      lic[q] = li[q];
    }
    if (HETB_PLAN_PREFETCH && p + U < p1) load_idx(p + U);
#pragma unroll
    for (int q = 0; q < U; ++q) {
      if (p + q < p1) {  // uniform inside the lane group
        if (lic[q] != cur) {
          flush(false);
          open_head = false;
          cur = lic[q];
          acc = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        fma4(acc, ga[q], kvv[q]);
      }
    }
  }
  flush(p1 < num_positions && __ldg(p_left + p1) == cur);
}

// ---------------------------------------------------------------------------
// RGAT backward in plan form: one walk over the positions grouped by compact (relation, source)
// row u (GAT/FusedGATBackward.cu.h:31-148 restated): feat[u], el[u] are run constants,
// grad_feat[u] / grad_el[u] register accumulators with one store per run; the normalised score
// is recomputed as exp(lrelu(el[u] + er[L]) − lse[d]); <gout[d], ret[d]> comes from the pre-pass.
// grad_er takes one scalar atomic per edge and head (destination side).
// ---------------------------------------------------------------------------
template <typename T, int LPR, int HW>
__global__ void __launch_bounds__(kPThreads)
    plan_rgat_bwd_walk_kernel(const int *__restrict__ item_ptr, int num_items, const int *__restrict__ p_dst,
                              const int *__restrict__ p_left, const int *__restrict__ p_msg,
                              const int *__restrict__ urow_rel, const T *__restrict__ feat,
                              const float *__restrict__ el, const float *__restrict__ er,
                              const float *__restrict__ lse, const float *__restrict__ gdotr,
                              const float *__restrict__ gout, float *__restrict__ grad_feat,
                              float *__restrict__ grad_el, float *__restrict__ grad_er, int H, int hw_rt,
                              float slope) {
  const int hw = HW ? HW : hw_rt;
  constexpr int GROUPS = kPThreads / LPR;
  constexpr int D = LPR * 4;
  constexpr int U = 4;
  const int grp = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int g = blockIdx.x * GROUPS + grp;
  if (g >= num_items) return;
  const unsigned gm = group_mask<LPR>();
  const int h = l / hw;
  const bool leader = (l % hw) == 0;
  const int p0 = __ldg(item_ptr + g), p1 = __ldg(item_ptr + g + 1);
  int cur_u = -1, relword = 0;
  float4 featr = make_float4(0.f, 0.f, 0.f, 0.f), acc_f = make_float4(0.f, 0.f, 0.f, 0.f);
  float elr = 0.f, acc_el = 0.f;

  auto flush_run = [&]() {
    if (cur_u < 0) return;
    if (relword < 0) {
      red_add4(grad_feat + (int64_t)cur_u * D + 4 * l, acc_f);
      if (leader) atomicAdd(grad_el + (int64_t)cur_u * H + h, acc_el);
    } else {
      st4(grad_feat + (int64_t)cur_u * D + 4 * l, acc_f);
      if (leader) grad_el[(int64_t)cur_u * H + h] = acc_el;
    }
  };

  for (int p = p0; p < p1; p += U) {
    int dn[U], li[U], ui[U];
    float ls[U], cc[U], erv[U];
    float4 go[U];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int pp = (p + q < p1) ? p + q : p1 - 1;
      dn[q] = __ldg(p_dst + pp);
      li[q] = __ldg(p_left + pp);
      ui[q] = __ldg(p_msg + pp);
    }
#pragma unroll
    for (int q = 0; q < U; ++q) {
      go[q] = ldg4(gout + (int64_t)dn[q] * D + 4 * l);
      ls[q] = __ldg(gdotr + (int64_t)dn[q] * 2 * H + h);  // (lse, <gout, ret>) pairs: plan_rowdot_kernel
      cc[q] = __ldg(gdotr + (int64_t)dn[q] * 2 * H + H + h);
      erv[q] = __ldg(er + (int64_t)li[q] * H + h);
    }
#pragma unroll
    for (int q = 0; q < U; ++q) {
      if (p + q < p1) {  // uniform inside the lane group
        if (ui[q] != cur_u) {
          flush_run();
          cur_u = ui[q];
          relword = __ldg(urow_rel + cur_u);
          featr = ld_row4<T>(feat, cur_u, D, l);
          elr = __ldg(el + (int64_t)cur_u * H + h);
          acc_f = make_float4(0.f, 0.f, 0.f, 0.f);
          acc_el = 0.f;
        }
        const float sdot = group_sum_t<HW>(dot4(go[q], featr), hw, gm) - cc[q];
        const float x = elr + erv[q];
        const float a = ex2_ftz((x > 0.f ? x : slope * x) * kLog2e - ls[q]);  // ls = lse·log2(e)
        fma4(acc_f, a, go[q]);
        if (leader) {
          const float v = sdot * a * (x > 0.f ? 1.f : slope);
          acc_el += v;
          atomicAdd(grad_er + (int64_t)li[q] * H + h, v);
        }
      }
    }
  }
  flush_run();
}

static int plan_lpr(int H, int F) {
  if (H <= 0 || F <= 0 || (F % 4) != 0) return 0;
  if (!is_pow2(H) || !is_pow2(F)) return 0;
  const int v = H * F / 4;
  if (v < 4 || v > 32) return 0;
  return v;
}

#define PLAN_LPR(lpr, CALL)                         \
  switch (lpr) {                                    \
    case 4: { constexpr int L = 4; CALL; } break;   \
    case 8: { constexpr int L = 8; CALL; } break;   \
    case 16: { constexpr int L = 16; CALL; } break; \
    case 32: { constexpr int L = 32; CALL; } break; \
    default: return HETB_EUNSUPPORTED;              \
  }

// The walk kernels also take the head width as a constant for the two shapes of the BASELINE configs
// (64-float rows: 4 heads of 16, 1 head of 64); every other shape runs the generic HWC = 0 instantiation.
#define PLAN_LPR_HW(lpr, hw, CALL)                                         \
  if ((lpr) == 16 && (hw) == 4) { constexpr int L = 16, HWC = 4; CALL; }   \
  else if ((lpr) == 16 && (hw) == 16) { constexpr int L = 16, HWC = 16; CALL; } \
  else { constexpr int HWC = 0; PLAN_LPR(lpr, CALL) }

static inline unsigned grid_for(int64_t n, int groups) { return (unsigned)(n > 0 ? ceil_div64(n, groups) : 1); }

template <typename T, bool RGAT>
static int fwd_plan_launch(const hetb_walk_plan *pl, const T *left, const T *kv, const T *msg, const float *mu,
                           const float *el, const float *er, float slope, float *attn, float *lse, float *ret,
                           float *part_m, float *part_s, float *part_acc, float *ssum, int H, int F,
                           cudaStream_t st) {
  const int lpr = plan_lpr(H, F);
  if (!lpr) return HETB_EUNSUPPORTED;
  const int groups = kPThreads / lpr;
  const int hw = F / 4;
  if (pl->num_fix_rows + pl->num_empty_rows > 0) {
    PLAN_LPR(lpr, (plan_fwd_prepare_kernel<L><<<grid_for(pl->num_fix_rows + pl->num_empty_rows, groups), kPThreads,
                                              0, st>>>(pl->fix_rows, pl->num_fix_rows, pl->empty_rows,
                                                       pl->num_empty_rows, ret, lse, ssum, H, hw)));
    HETB_LAUNCH_OK();
  }
  if (pl->num_items == 0) return HETB_OK;
  if (HETB_FWD_LANES8 && lpr == 16 && (hw == 4 || hw == 16)) {
    const unsigned grid8 = grid_for(pl->num_items, kPThreads / 8);
    if (hw == 4)
      plan_fwd_walk8_kernel<T, RGAT, 2><<<grid8, kPThreads, 0, st>>>(
          pl->item_ptr, pl->item_row, pl->num_items, pl->nz_rows, pl->nz_ptr, pl->idx_a, pl->idx_b, pl->idx_c,
          pl->row_rel, left, kv, msg, mu, el, er, slope, attn, lse, ret, part_m, part_s, part_acc, H);
    else
      plan_fwd_walk8_kernel<T, RGAT, 8><<<grid8, kPThreads, 0, st>>>(
          pl->item_ptr, pl->item_row, pl->num_items, pl->nz_rows, pl->nz_ptr, pl->idx_a, pl->idx_b, pl->idx_c,
          pl->row_rel, left, kv, msg, mu, el, er, slope, attn, lse, ret, part_m, part_s, part_acc, H);
  } else
  PLAN_LPR_HW(lpr, hw, (plan_fwd_walk_kernel<T, L, RGAT, HWC><<<grid_for(pl->num_items, groups), kPThreads, 0, st>>>(
                    pl->item_ptr, pl->item_row, pl->num_items, pl->nz_rows, pl->nz_ptr, pl->idx_a, pl->idx_b,
                    pl->idx_c, pl->row_rel, left, kv, msg, mu, el, er, slope, attn, lse, ret, part_m, part_s,
                    part_acc, H, hw)));
  HETB_LAUNCH_OK();
  if (pl->num_pslots > 0) {
    PLAN_LPR(lpr, (plan_fwd_merge_kernel<L><<<grid_for(pl->num_pslots, groups), kPThreads, 0, st>>>(
                      pl->pslot, pl->pslot_row, pl->num_pslots, part_m, part_s, part_acc, lse, ssum, ret, H, hw)));
    HETB_LAUNCH_OK();
  }
  if (pl->num_fix_rows > 0) {
    PLAN_LPR(lpr, (plan_fwd_normalize_kernel<L><<<grid_for(pl->num_fix_rows, groups), kPThreads, 0, st>>>(
                      pl->fix_rows, pl->num_fix_rows, ssum, lse, ret, H, hw)));
    HETB_LAUNCH_OK();
  }
  return HETB_OK;
}

static bool plan_ok_fwd(const hetb_walk_plan *pl) {
  return pl && pl->num_items >= 0 &&
         (pl->num_items == 0 || (pl->item_ptr && pl->item_row && pl->nz_rows && pl->nz_ptr && pl->idx_b && pl->idx_c)) &&
         (pl->num_fix_rows == 0 || pl->fix_rows) && (pl->num_empty_rows == 0 || pl->empty_rows) &&
         (pl->num_pslots == 0 || (pl->pslot && pl->pslot_row));
}

}  // namespace hetb

using namespace hetb;

extern "C" int hetb_hgt_fused_fwd_plan(const hetb_walk_plan *plan, const void *left, const void *kv,
                                       const void *msg, int table_dtype, const float *mu, float *attn, float *lse,
                                       float *ret, float *part_m, float *part_s, float *part_acc, float *ssum,
                                       int H, int F, void *stream) {
  HETB_CHECK_ARG(plan_ok_fwd(plan) && lse && ret);
  HETB_CHECK_ARG(plan->num_items == 0 || (left && kv && msg && mu && plan->idx_a && plan->row_rel));  // attn optional
  HETB_CHECK_ARG(plan->num_pslots == 0 || (part_m && part_s && part_acc && ssum));
  cudaStream_t st = (cudaStream_t)stream;
  if (table_dtype == HETB_DTYPE_F32)
    return fwd_plan_launch<float, false>(plan, (const float *)left, (const float *)kv, (const float *)msg, mu,
                                         nullptr, nullptr, 0.f, attn, lse, ret, part_m, part_s, part_acc, ssum, H, F,
                                         st);
  if (table_dtype == HETB_DTYPE_BF16)
    return fwd_plan_launch<__nv_bfloat16, false>(plan, (const __nv_bfloat16 *)left, (const __nv_bfloat16 *)kv,
                                                 (const __nv_bfloat16 *)msg, mu, nullptr, nullptr, 0.f, attn, lse,
                                                 ret, part_m, part_s, part_acc, ssum, H, F, st);
  return HETB_EUNSUPPORTED;
}

extern "C" int hetb_rgat_fused_fwd_plan(const hetb_walk_plan *plan, const void *feat, int table_dtype,
                                        const float *el, const float *er, float slope, float *lse, float *ret,
                                        float *part_m, float *part_s, float *part_acc, float *ssum, int H, int F,
                                        void *stream) {
  HETB_CHECK_ARG(plan_ok_fwd(plan) && lse && ret);
  HETB_CHECK_ARG(plan->num_items == 0 || (feat && el && er));
  HETB_CHECK_ARG(plan->num_pslots == 0 || (part_m && part_s && part_acc && ssum));
  cudaStream_t st = (cudaStream_t)stream;
  if (table_dtype == HETB_DTYPE_F32)
    return fwd_plan_launch<float, true>(plan, nullptr, nullptr, (const float *)feat, nullptr, el, er, slope, nullptr,
                                        lse, ret, part_m, part_s, part_acc, ssum, H, F, st);
  if (table_dtype == HETB_DTYPE_BF16)
    return fwd_plan_launch<__nv_bfloat16, true>(plan, nullptr, nullptr, (const __nv_bfloat16 *)feat, nullptr, el, er,
                                                slope, nullptr, lse, ret, part_m, part_s, part_acc, ssum, H, F, st);
  return HETB_EUNSUPPORTED;
}

template <typename T>
static int hgt_bwd_plan_launch(const hetb_walk_plan *pl, const hetb_walk_plan *fwd, int64_t num_rels, int64_t num_dst,
                               const T *left, const T *kv, const T *msg, const float *mu, const float *attn,
                               float *gattn, const float *lse, const float *out, const float *gout, float *gdoto,
                               float *grad_left, float *grad_kv, float *grad_msg, float *grad_mu, int H, int F,
                               cudaStream_t st) {
  const int lpr = plan_lpr(H, F);
  if (!lpr) return HETB_EUNSUPPORTED;
  const int groups = kPThreads / lpr;
  const int hw = F / 4;
  if (pl->num_fix_rows > 0) {
    PLAN_LPR(lpr, (plan_zero_rows_kernel<L><<<grid_for(pl->num_fix_rows, groups), kPThreads, 0, st>>>(
                      pl->fix_rows, pl->num_fix_rows, grad_kv, nullptr, H, hw)));
    HETB_LAUNCH_OK();
  }
  if (pl->num_fix_runs > 0) {
    PLAN_LPR(lpr, (plan_zero_rows_kernel<L><<<grid_for(pl->num_fix_runs, groups), kPThreads, 0, st>>>(
                      pl->fix_runs, pl->num_fix_runs, grad_msg, nullptr, H, hw)));
    HETB_LAUNCH_OK();
  }
  if (pl->num_items == 0) return HETB_OK;
  PLAN_LPR(lpr, (plan_rowdot_kernel<L><<<grid_for(num_dst, groups), kPThreads, 0, st>>>(gout, out, lse, num_dst, gdoto,
                                                                                     H, hw)));
  HETB_LAUNCH_OK();
  const int use_smem = (num_rels * H <= kMaxSmemMuPlan) ? 1 : 0;
  const size_t smem = use_smem ? (size_t)num_rels * H * sizeof(float) : 0;
  if (!fwd) {
    if (attn) {
      PLAN_LPR_HW(lpr, hw, (plan_hgt_bwd_walk_kernel<T, L, false, kSaved, HWC><<<grid_for(pl->num_items, groups), kPThreads, smem, st>>>(
                        pl->item_ptr, pl->item_row, pl->num_items, pl->nz_rows, pl->nz_ptr, pl->idx_a, pl->idx_c,
                        pl->idx_b, pl->idx_d, pl->row_rel, (int)num_rels, left, kv, msg, mu, attn, gdoto, gout,
                        grad_left, grad_kv, grad_msg, grad_mu, H, hw, use_smem, nullptr)));
    } else {
      PLAN_LPR_HW(lpr, hw, (plan_hgt_bwd_walk_kernel<T, L, false, kRecompute, HWC><<<grid_for(pl->num_items, groups), kPThreads, smem, st>>>(
                        pl->item_ptr, pl->item_row, pl->num_items, pl->nz_rows, pl->nz_ptr, pl->idx_a, pl->idx_c,
                        pl->idx_b, pl->idx_d, pl->row_rel, (int)num_rels, left, kv, msg, mu, nullptr, gdoto, gout,
                        grad_left, grad_kv, grad_msg, grad_mu, H, hw, use_smem, nullptr)));
    }
    HETB_LAUNCH_OK();
    return HETB_OK;
  }
  // two passes: the source-major walk leaves ga in gattn, the destination-major walk owns grad_left
  PLAN_LPR_HW(lpr, hw, (plan_hgt_bwd_walk_kernel<T, L, true, kSaved, HWC><<<grid_for(pl->num_items, groups), kPThreads, smem, st>>>(
                    pl->item_ptr, pl->item_row, pl->num_items, pl->nz_rows, pl->nz_ptr, pl->idx_a, pl->idx_c,
                    pl->idx_b, pl->idx_d, pl->row_rel, (int)num_rels, left, kv, msg, mu, nullptr, gdoto, gout,
                    nullptr, grad_kv, grad_msg, grad_mu, H, hw, use_smem, gattn)));
  HETB_LAUNCH_OK();
  if (fwd->num_fix_runs > 0) {
    PLAN_LPR(lpr, (plan_zero_rows_kernel<L><<<grid_for(fwd->num_fix_runs, groups), kPThreads, 0, st>>>(
                      fwd->fix_runs, fwd->num_fix_runs, grad_left, nullptr, H, hw)));
    HETB_LAUNCH_OK();
  }
  PLAN_LPR(lpr, (plan_hgt_grad_left_kernel<T, L><<<grid_for(fwd->num_items, groups), kPThreads, 0, st>>>(
                    fwd->item_ptr, fwd->num_items, (int)fwd->num_positions, fwd->idx_a, fwd->idx_c, kv, gattn,
                    grad_left, H, hw)));
  HETB_LAUNCH_OK();
  return HETB_OK;
}

// grad_left and grad_mu must arrive zero-filled (every edge reduces into them); grad_kv and
// grad_msg need no initialisation (every row is stored, split / empty rows are zeroed here).
extern "C" int hetb_hgt_fused_bwd_plan(const hetb_walk_plan *plan, int64_t num_rels, int64_t num_dst,
                                       const void *left, const void *kv, const void *msg, int table_dtype,
                                       const float *mu, const float *attn, const float *lse, const float *out,
                                       const float *gout, float *gdoto_scratch, float *grad_left, float *grad_kv,
                                       float *grad_msg, float *grad_mu, int H, int F, void *stream) {
  HETB_CHECK_ARG(plan && plan->num_items >= 0 && num_rels >= 0 && num_dst >= 0);
  HETB_CHECK_ARG((plan->num_fix_rows == 0 || (plan->fix_rows && grad_kv)) &&
                 (plan->num_fix_runs == 0 || (plan->fix_runs && grad_msg)));
  HETB_CHECK_ARG(plan->num_items == 0 ||
                 (plan->item_ptr && plan->item_row && plan->nz_rows && plan->nz_ptr && plan->idx_a && plan->idx_b &&
                  plan->idx_c && plan->row_rel && (plan->idx_d || !attn) && left && kv && msg && mu && lse && out &&
                  gout && gdoto_scratch && grad_left && grad_kv && grad_msg && grad_mu && num_dst > 0));
  cudaStream_t st = (cudaStream_t)stream;
  if (table_dtype == HETB_DTYPE_F32)
    return hgt_bwd_plan_launch<float>(plan, nullptr, num_rels, num_dst, (const float *)left, (const float *)kv,
                                      (const float *)msg, mu, attn, nullptr, lse, out, gout, gdoto_scratch,
                                      grad_left, grad_kv, grad_msg, grad_mu, H, F, st);
  if (table_dtype == HETB_DTYPE_BF16)
    return hgt_bwd_plan_launch<__nv_bfloat16>(plan, nullptr, num_rels, num_dst, (const __nv_bfloat16 *)left,
                                              (const __nv_bfloat16 *)kv, (const __nv_bfloat16 *)msg, mu, attn,
                                              nullptr, lse, out, gout, gdoto_scratch, grad_left, grad_kv, grad_msg,
                                              grad_mu, H, F, st);
  return HETB_EUNSUPPORTED;
}

// Two-pass form (see plan_hgt_bwd_walk_kernel<.., TWO_PASS> and plan_hgt_grad_left_kernel): only grad_mu
// must arrive zero-filled; grad_left, grad_kv and grad_msg are fully written. attn_inout holds the raw scores
// of the forward on entry and the gradient of the raw scores (forward-plan position order) on return.
extern "C" int hetb_hgt_fused_bwd_plan2(const hetb_walk_plan *plan, const hetb_walk_plan *fwd_plan, int64_t num_rels,
                                        int64_t num_dst, const void *left, const void *kv, const void *msg,
                                        int table_dtype, const float *mu, float *attn_inout, const float *lse,
                                        const float *out, const float *gout, float *gdoto_scratch, float *grad_left,
                                        float *grad_kv, float *grad_msg, float *grad_mu, int H, int F, void *stream) {
  HETB_CHECK_ARG(plan && fwd_plan && plan->num_items >= 0 && num_rels >= 0 && num_dst >= 0);
  HETB_CHECK_ARG(plan->num_positions == fwd_plan->num_positions && plan->num_positions <= INT32_MAX);
  HETB_CHECK_ARG((plan->num_fix_rows == 0 || (plan->fix_rows && grad_kv)) &&
                 (plan->num_fix_runs == 0 || (plan->fix_runs && grad_msg)) &&
                 (fwd_plan->num_fix_runs == 0 || (fwd_plan->fix_runs && grad_left)));
  HETB_CHECK_ARG(plan->num_items == 0 ||
                 (plan->item_ptr && plan->item_row && plan->nz_rows && plan->nz_ptr && plan->idx_a && plan->idx_b &&
                  plan->idx_c && plan->idx_d && plan->row_rel && fwd_plan->item_ptr && fwd_plan->idx_a &&
                  fwd_plan->idx_c && left && kv && msg && mu && attn_inout && lse && out && gout && gdoto_scratch &&
                  grad_left && grad_kv && grad_msg && grad_mu && num_dst > 0));
  cudaStream_t st = (cudaStream_t)stream;
  if (table_dtype == HETB_DTYPE_F32)
    return hgt_bwd_plan_launch<float>(plan, fwd_plan, num_rels, num_dst, (const float *)left, (const float *)kv,
                                      (const float *)msg, mu, nullptr, attn_inout, lse, out, gout, gdoto_scratch,
                                      grad_left, grad_kv, grad_msg, grad_mu, H, F, st);
  if (table_dtype == HETB_DTYPE_BF16)
    return hgt_bwd_plan_launch<__nv_bfloat16>(plan, fwd_plan, num_rels, num_dst, (const __nv_bfloat16 *)left,
                                              (const __nv_bfloat16 *)kv, (const __nv_bfloat16 *)msg, mu, nullptr,
                                              attn_inout, lse, out, gout, gdoto_scratch, grad_left, grad_kv,
                                              grad_msg, grad_mu, H, F, st);
  return HETB_EUNSUPPORTED;
}

template <typename T>
static int rgat_bwd_plan_launch(const hetb_walk_plan *pl, int64_t num_dst, const T *feat, const float *el,
                                const float *er, const float *lse, const float *ret, const float *gout,
                                float *gdotr, float *grad_feat, float *grad_el, float *grad_er, int H, int F,
                                float slope, cudaStream_t st) {
  const int lpr = plan_lpr(H, F);
  if (!lpr) return HETB_EUNSUPPORTED;
  const int groups = kPThreads / lpr;
  const int hw = F / 4;
  if (pl->num_fix_runs > 0) {
    PLAN_LPR(lpr, (plan_zero_rows_kernel<L><<<grid_for(pl->num_fix_runs, groups), kPThreads, 0, st>>>(
                      pl->fix_runs, pl->num_fix_runs, grad_feat, grad_el, H, hw)));
    HETB_LAUNCH_OK();
  }
  if (pl->num_items == 0) return HETB_OK;
  PLAN_LPR(lpr, (plan_rowdot_kernel<L><<<grid_for(num_dst, groups), kPThreads, 0, st>>>(gout, ret, lse, num_dst, gdotr,
                                                                                     H, hw)));
  HETB_LAUNCH_OK();
  PLAN_LPR_HW(lpr, hw, (plan_rgat_bwd_walk_kernel<T, L, HWC><<<grid_for(pl->num_items, groups), kPThreads, 0, st>>>(
                    pl->item_ptr, pl->num_items, pl->idx_a, pl->idx_c, pl->idx_b, pl->row_rel, feat, el, er, lse,
                    gdotr, gout, grad_feat, grad_el, grad_er, H, hw, slope)));
  HETB_LAUNCH_OK();
  return HETB_OK;
}

// grad_er must arrive zero-filled; grad_feat / grad_el need no initialisation.
extern "C" int hetb_rgat_fused_bwd_plan(const hetb_walk_plan *plan, int64_t num_dst, const void *feat,
                                        int table_dtype, const float *el, const float *er, const float *lse,
                                        const float *ret, const float *gout, float *gdotr_scratch,
                                        float *grad_feat, float *grad_el, float *grad_er, int H, int F, float slope,
                                        void *stream) {
  HETB_CHECK_ARG(plan && plan->num_items >= 0 && num_dst >= 0);
  HETB_CHECK_ARG(plan->num_fix_runs == 0 || (plan->fix_runs && grad_feat && grad_el));
  HETB_CHECK_ARG(plan->num_items == 0 ||
                 (plan->item_ptr && plan->idx_a && plan->idx_b && plan->idx_c && plan->row_rel && feat && el && er &&
                  lse && ret && gout && gdotr_scratch && grad_feat && grad_el && grad_er && num_dst > 0));
  cudaStream_t st = (cudaStream_t)stream;
  if (table_dtype == HETB_DTYPE_F32)
    return rgat_bwd_plan_launch<float>(plan, num_dst, (const float *)feat, el, er, lse, ret, gout, gdotr_scratch,
                                       grad_feat, grad_el, grad_er, H, F, slope, st);
  if (table_dtype == HETB_DTYPE_BF16)
    return rgat_bwd_plan_launch<__nv_bfloat16>(plan, num_dst, (const __nv_bfloat16 *)feat, el, er, lse, ret, gout,
                                               gdotr_scratch, grad_feat, grad_el, grad_er, H, F, slope, st);
  return HETB_EUNSUPPORTED;
}
