// Relation sort / CSR construction on the GPU: a hand-written stable LSD radix
// sort (8-bit digits) over (key, original position) pairs, followed by a
// payload gather and a boundary kernel that writes the segment offsets.
// Integer work only; results are bit-exact against the stable CPU oracle.
//
// Replaces the CPU converters of the reference (MyHyb/MyHyb.h:31-75,
// 996-1150: std::map / vector<vector<vector<pair>>> bucketing of size
// R x N) and the Python argsort/bincount/cumsum path
// (python/utils_lite/sparse_matrix_converters.py:7-39).
#include "common.cuh"

namespace hetb {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 keys per CTA
constexpr int kRadix = 256;

__global__ void __launch_bounds__(kSortThreads)
    sort_init_kernel(const int64_t *__restrict__ key, int64_t E,
                     uint32_t *__restrict__ k32, uint32_t *__restrict__ idx) {
  const int64_t stride = (int64_t)gridDim.x * kSortThreads;
  for (int64_t i = (int64_t)blockIdx.x * kSortThreads + threadIdx.x; i < E;
       i += stride) {
    k32[i] = (uint32_t)__ldg(key + i);
    idx[i] = (uint32_t)i;
  }
}

// hist[digit * num_ctas + cta] = number of keys of this CTA's tile with digit.
__global__ void __launch_bounds__(kSortThreads)
    sort_hist_kernel(const uint32_t *__restrict__ keys, int64_t E, int shift,
                     uint32_t *__restrict__ hist, int num_ctas) {
  __shared__ uint32_t cnt[kRadix];
  cnt[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const int64_t i = base + r * kSortThreads + threadIdx.x;
    if (i < E) atomicAdd(&cnt[(keys[i] >> shift) & 0xff], 1u);
  }
  __syncthreads();
  hist[(size_t)threadIdx.x * num_ctas + blockIdx.x] = cnt[threadIdx.x];
}

// In-place exclusive scan of n uint32 values by one CTA of 1024 threads.
__global__ void __launch_bounds__(1024)
    scan_u32_kernel(uint32_t *__restrict__ data, int64_t n) {
  __shared__ uint32_t part[1024];
  const int t = threadIdx.x;
  const int64_t chunk = ceil_div64(n, 1024);
  const int64_t b = (int64_t)t * chunk;
  const int64_t e = (b + chunk < n) ? b + chunk : n;
  uint32_t s = 0;
  for (int64_t i = b; i < e; ++i) s += data[i];
  part[t] = s;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    uint32_t v = (t >= off) ? part[t - off] : 0u;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  uint32_t run = part[t] - s;  // exclusive prefix of this thread's chunk
  for (int64_t i = b; i < e; ++i) {
    const uint32_t v = data[i];
    data[i] = run;
    run += v;
  }
}

// Three-phase scan for arrays longer than one CTA's worth (the digit histogram of a 21 M-edge sort is
// 2.6 M counters: one CTA took 1.35 ms per radix pass). Phase 1: every CTA scans its 4096 values in place
// (exclusive, relative to the CTA's first value) and leaves its total in sums[]; phase 2: scan_u32_kernel
// over sums[]; phase 3: add each CTA's offset.
constexpr int kScanBlock = 4096;  // 1024 threads x 4 values

__global__ void __launch_bounds__(1024)
    scan_blocks_kernel(uint32_t *__restrict__ data, int64_t n, uint32_t *__restrict__ sums) {
  __shared__ uint32_t part[1024];
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kScanBlock + (int64_t)t * 4;
  uint32_t v[4], s = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    v[k] = (base + k < n) ? data[base + k] : 0u;
    s += v[k];
  }
  part[t] = s;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    const uint32_t u = (t >= off) ? part[t - off] : 0u;
    __syncthreads();
    part[t] += u;
    __syncthreads();
  }
  uint32_t run = part[t] - s;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (base + k < n) data[base + k] = run;
    run += v[k];
  }
  if (t == 1023) sums[blockIdx.x] = part[1023];
}

__global__ void __launch_bounds__(1024)
    scan_add_offsets_kernel(uint32_t *__restrict__ data, int64_t n, const uint32_t *__restrict__ sums) {
  const int64_t base = (int64_t)blockIdx.x * kScanBlock + (int64_t)threadIdx.x * 4;
  const uint32_t off = sums[blockIdx.x];
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (base + k < n) data[base + k] += off;
}

// Stable scatter: keys of a tile are ranked in index order. Inside a round of
// 256 keys the eight warps take turns, each ranking its lanes with match_any.
__global__ void __launch_bounds__(kSortThreads)
    sort_scatter_kernel(const uint32_t *__restrict__ keys_in,
                        const uint32_t *__restrict__ vals_in, int64_t E,
                        int shift, const uint32_t *__restrict__ hist,
                        int num_ctas, uint32_t *__restrict__ keys_out,
                        uint32_t *__restrict__ vals_out) {
  __shared__ uint32_t offset[kRadix];
  offset[threadIdx.x] = hist[(size_t)threadIdx.x * num_ctas + blockIdx.x];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t base = (int64_t)blockIdx.x * kSortTile;
  for (int r = 0; r < kSortItems; ++r) {
    const int64_t i = base + r * kSortThreads + threadIdx.x;
    const bool ok = i < E;
    uint32_t k = 0, v = 0;
    if (ok) {
      k = keys_in[i];
      v = vals_in[i];
    }
    // digits of out-of-range lanes are set to a value no valid lane shares
    const uint32_t digit = ok ? ((k >> shift) & 0xff) : (0x100u + lane);
    for (int w = 0; w < kSortThreads / 32; ++w) {
      if (warp == w) {
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        if (ok) {
          const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
          const uint32_t pos = offset[digit] + rank;
          keys_out[pos] = k;
          vals_out[pos] = v;
        }
        __syncwarp();
        if (ok && (peers & ((1u << lane) - 1u)) == 0)
          offset[digit] += __popc(peers);
      }
      __syncthreads();
    }
  }
}

// ptrs[k] = first position p with sorted_key[p] >= k  (ptrs[num_keys] = E).
__global__ void __launch_bounds__(kSortThreads)
    boundaries_kernel(const uint32_t *__restrict__ sorted, int64_t E,
                      int64_t num_keys, int64_t *__restrict__ ptrs) {
  const int64_t stride = (int64_t)gridDim.x * kSortThreads;
  for (int64_t p = (int64_t)blockIdx.x * kSortThreads + threadIdx.x; p <= E;
       p += stride) {
    const int64_t lo = (p == 0) ? 0 : (int64_t)sorted[p - 1] + 1;
    const int64_t hi = (p == E) ? num_keys : (int64_t)sorted[p];
    for (int64_t k = lo; k <= hi; ++k) ptrs[k] = p;
  }
}

__global__ void __launch_bounds__(kSortThreads)
    gather_i64_kernel(const int64_t *__restrict__ in,
                      const uint32_t *__restrict__ perm, int64_t E,
                      int64_t *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * kSortThreads;
  for (int64_t p = (int64_t)blockIdx.x * kSortThreads + threadIdx.x; p < E;
       p += stride)
    out[p] = __ldg(in + perm[p]);
}

__global__ void __launch_bounds__(kSortThreads)
    expand_row_ptr_kernel(const int64_t *__restrict__ row_ptr, int64_t num_rows,
                          int64_t E, int64_t *__restrict__ row_idx) {
  const int64_t stride = (int64_t)gridDim.x * kSortThreads;
  for (int64_t p = (int64_t)blockIdx.x * kSortThreads + threadIdx.x; p < E;
       p += stride) {
    // largest r with row_ptr[r] <= p
    int64_t lo = 0, hi = num_rows;
    while (hi - lo > 1) {
      const int64_t mid = (lo + hi) >> 1;
      if (__ldg(row_ptr + mid) <= p)
        lo = mid;
      else
        hi = mid;
    }
    row_idx[p] = lo;
  }
}

__global__ void fill_i64_kernel(int64_t *p, int64_t n, int64_t v) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += stride)
    p[i] = v;
}

static int stream_grid(int64_t n, int threads) {
  int64_t need = ceil_div64(n, threads);
  int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}

static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace hetb

using namespace hetb;

extern "C" int64_t hetb_sort_scratch_bytes(int64_t num_edges, int64_t max_key) {
  (void)max_key;
  if (num_edges < 0) return HETB_EINVAL;
  const int64_t ctas = ceil_div64(num_edges > 0 ? num_edges : 1, kSortTile);
  const int64_t scan_blocks = ceil_div64((int64_t)kRadix * ctas, kScanBlock);
  return (int64_t)(4 * align_up((size_t)num_edges * 4) +
                   align_up((size_t)kRadix * ctas * 4) + align_up((size_t)scan_blocks * 4) + 256);
}

extern "C" int hetb_stable_sort_by_key(const int64_t *key, int64_t num_edges,
                                       int64_t num_keys, int num_payload,
                                       const int64_t *const *in,
                                       int64_t *const *out, int64_t *ptrs,
                                       void *scratch, int64_t scratch_bytes,
                                       void *stream) {
  HETB_CHECK_ARG(num_edges >= 0 && num_keys >= 0 && num_payload >= 0 &&
                 num_payload <= 4);
  HETB_CHECK_ARG(ptrs != nullptr);
  if (num_edges >= (int64_t)1 << 32 || num_keys > (int64_t)1 << 32)
    return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (num_edges == 0) {
    fill_i64_kernel<<<stream_grid(num_keys + 1, 256), 256, 0, st>>>(
        ptrs, num_keys + 1, 0);
    HETB_LAUNCH_OK();
    return HETB_OK;
  }
  HETB_CHECK_ARG(key && scratch);
  HETB_CHECK_ARG(scratch_bytes >= hetb_sort_scratch_bytes(num_edges, num_keys));
  for (int j = 0; j < num_payload; ++j) HETB_CHECK_ARG(in[j] && out[j]);

  const int64_t E = num_edges;
  const int num_ctas = (int)ceil_div64(E, kSortTile);
  char *base = reinterpret_cast<char *>(
      ((uintptr_t)scratch + 255) & ~(uintptr_t)255);
  const size_t arr = align_up((size_t)E * 4);
  uint32_t *keys_a = reinterpret_cast<uint32_t *>(base);
  uint32_t *keys_b = reinterpret_cast<uint32_t *>(base + arr);
  uint32_t *vals_a = reinterpret_cast<uint32_t *>(base + 2 * arr);
  uint32_t *vals_b = reinterpret_cast<uint32_t *>(base + 3 * arr);
  uint32_t *hist = reinterpret_cast<uint32_t *>(base + 4 * arr);
  const int64_t hist_n = (int64_t)kRadix * num_ctas;
  const int64_t scan_blocks = ceil_div64(hist_n, kScanBlock);
  uint32_t *scan_sums = reinterpret_cast<uint32_t *>(base + 4 * arr + align_up((size_t)hist_n * 4));

  sort_init_kernel<<<stream_grid(E, kSortThreads), kSortThreads, 0, st>>>(
      key, E, keys_a, vals_a);
  HETB_LAUNCH_OK();

  int bits = 0;
  while (bits < 32 && ((int64_t)1 << bits) < num_keys) ++bits;
  int passes = (bits + 7) / 8;
  if (passes < 1) passes = 1;
  for (int p = 0; p < passes; ++p) {
    const int shift = 8 * p;
    sort_hist_kernel<<<num_ctas, kSortThreads, 0, st>>>(keys_a, E, shift, hist,
                                                        num_ctas);
    HETB_LAUNCH_OK();
    if (scan_blocks <= 1) {
      scan_u32_kernel<<<1, 1024, 0, st>>>(hist, hist_n);
      HETB_LAUNCH_OK();
    } else {
      scan_blocks_kernel<<<(unsigned)scan_blocks, 1024, 0, st>>>(hist, hist_n, scan_sums);
      HETB_LAUNCH_OK();
      scan_u32_kernel<<<1, 1024, 0, st>>>(scan_sums, scan_blocks);
      HETB_LAUNCH_OK();
      scan_add_offsets_kernel<<<(unsigned)scan_blocks, 1024, 0, st>>>(hist, hist_n, scan_sums);
      HETB_LAUNCH_OK();
    }
    sort_scatter_kernel<<<num_ctas, kSortThreads, 0, st>>>(
        keys_a, vals_a, E, shift, hist, num_ctas, keys_b, vals_b);
    HETB_LAUNCH_OK();
    uint32_t *t = keys_a; keys_a = keys_b; keys_b = t;
    t = vals_a; vals_a = vals_b; vals_b = t;
  }
  boundaries_kernel<<<stream_grid(E + 1, kSortThreads), kSortThreads, 0, st>>>(
      keys_a, E, num_keys, ptrs);
  HETB_LAUNCH_OK();
  for (int j = 0; j < num_payload; ++j) {
    gather_i64_kernel<<<stream_grid(E, kSortThreads), kSortThreads, 0, st>>>(
        in[j], vals_a, E, out[j]);
    HETB_LAUNCH_OK();
  }
  return HETB_OK;
}

extern "C" int hetb_expand_row_ptr(const int64_t *row_ptr, int64_t num_rows,
                                   int64_t num_edges, int64_t *row_idx,
                                   void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_edges >= 0);
  if (num_edges == 0) return HETB_OK;
  HETB_CHECK_ARG(row_ptr && row_idx && num_rows > 0);
  cudaStream_t st = (cudaStream_t)stream;
  expand_row_ptr_kernel<<<stream_grid(num_edges, kSortThreads), kSortThreads, 0,
                          st>>>(row_ptr, num_rows, num_edges, row_idx);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
