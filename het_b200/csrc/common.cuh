// Shared device/host helpers for libhetb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hetb200.h"

namespace hetb {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

extern int64_t g_launch_count;

#define HETB_CHECK_ARG(cond) \
  do {                       \
    if (!(cond)) return HETB_EINVAL; \
  } while (0)

// Called after every <<<>>>: counts the launch and surfaces launch errors.
#define HETB_LAUNCH_OK()                         \
  do {                                           \
    ++::hetb::g_launch_count;                    \
    cudaError_t e__ = cudaPeekAtLastError();     \
    if (e__ != cudaSuccess) return (int)e__;     \
  } while (0)

__host__ __device__ __forceinline__ int64_t ceil_div64(int64_t a, int64_t b) {
  return (a + b - 1) / b;
}

// Largest s with ptrs[s] <= i, for a non-decreasing ptrs[0..n] with
// ptrs[0] <= i < ptrs[n]. Empty segments are skipped naturally.
__device__ __forceinline__ int segment_of(const int64_t *__restrict__ ptrs,
                                          int n, int64_t i) {
  int lo = 0, hi = n;  // invariant: ptrs[lo] <= i < ptrs[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (__ldg(ptrs + mid) <= i)
      lo = mid;
    else
      hi = mid;
  }
  return lo;
}

__device__ __forceinline__ float4 ldg4(const float *p) {
  return __ldg(reinterpret_cast<const float4 *>(p));
}
// Streaming (evict-first) 128-bit load for per-edge data touched exactly once.
__device__ __forceinline__ float4 ldcs4(const float *p) {
  return __ldcs(reinterpret_cast<const float4 *>(p));
}
__device__ __forceinline__ void st4(float *p, float4 v) {
  *reinterpret_cast<float4 *>(p) = v;
}
__device__ __forceinline__ void stcs4(float *p, float4 v) {
  __stcs(reinterpret_cast<float4 *>(p), v);
}
// 128-bit vector reduction into global memory (sm_90+: red.global.add.v4.f32).
__device__ __forceinline__ void red_add4(float *p, float4 v) {
  atomicAdd(reinterpret_cast<float4 *>(p), v);
}

__device__ __forceinline__ float dot4(float4 a, float4 b) {
  return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
}
__device__ __forceinline__ void fma4(float4 &acc, float s, float4 v) {
  acc.x = fmaf(s, v.x, acc.x);
  acc.y = fmaf(s, v.y, acc.y);
  acc.z = fmaf(s, v.z, acc.z);
  acc.w = fmaf(s, v.w, acc.w);
}

// Sum over the `width` consecutive lanes (power of two, aligned) that hold one
// head; every lane of the group ends with the total.
__device__ __forceinline__ float group_sum(float v, int width) {
  for (int o = width >> 1; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

inline bool is_pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

}  // namespace hetb
