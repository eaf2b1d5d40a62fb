// Microbenchmark: how many random rows per second does HBM3e deliver as a function of the row size?
// Each group of ROW_BYTES/16 lanes reads one table row per position with 128-bit loads; positions are
// consumed `unroll` at a time so that every group keeps `unroll` rows in flight. Built by
// tools/microbench/build.sh into tools/microbench/librandom_rows.so, driven by tools/bench_random_rows.py.
#include <cstdint>
#include <cuda_runtime.h>

template <int LPR, int U>
__global__ void __launch_bounds__(256) random_rows_kernel(const float4 *__restrict__ table, const int *__restrict__ idx,
                                                          int64_t num_pos, float *__restrict__ sink) {
  const int lane = threadIdx.x % LPR;
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LPR;
  const int64_t num_groups = (int64_t)gridDim.x * blockDim.x / LPR;
  float acc = 0.f;
  for (int64_t p = group * U; p < num_pos; p += num_groups * U) {
    int r[U];
#pragma unroll
    for (int u = 0; u < U; ++u) r[u] = (p + u < num_pos) ? __ldg(idx + p + u) : 0;
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldg(table + (int64_t)r[u] * LPR + lane);
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u].x + v[u].y + v[u].z + v[u].w;
  }
  if (acc == 123456.789f) sink[0] = acc;
}

template <int LPR>
static int launch(const void *table, const int *idx, int64_t num_pos, float *sink, int unroll, int blocks, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const float4 *t = (const float4 *)table;
  switch (unroll) {
    case 1: random_rows_kernel<LPR, 1><<<blocks, 256, 0, st>>>(t, idx, num_pos, sink); break;
    case 2: random_rows_kernel<LPR, 2><<<blocks, 256, 0, st>>>(t, idx, num_pos, sink); break;
    case 4: random_rows_kernel<LPR, 4><<<blocks, 256, 0, st>>>(t, idx, num_pos, sink); break;
    case 8: random_rows_kernel<LPR, 8><<<blocks, 256, 0, st>>>(t, idx, num_pos, sink); break;
    default: return -1;
  }
  return (int)cudaGetLastError();
}

extern "C" int random_rows(const void *table, const int *idx, int64_t num_pos, int row_bytes, float *sink, int unroll,
                           int blocks, void *stream) {
  switch (row_bytes) {
    case 64: return launch<4>(table, idx, num_pos, sink, unroll, blocks, stream);
    case 128: return launch<8>(table, idx, num_pos, sink, unroll, blocks, stream);
    case 256: return launch<16>(table, idx, num_pos, sink, unroll, blocks, stream);
    case 512: return launch<32>(table, idx, num_pos, sink, unroll, blocks, stream);
    default: return -2;
  }
}
