// Weight gradient of the relation-typed linear layer on the 5th-generation tensor cores
// (tcgen05 + TMEM), sm_100a:
//
//   dW[s] += X_s^T · G_s        X_s = A[a_idx?[i], :], G_s = G[g_idx?[i], :]  for rows i of segment s
//
// (reference: HET_RGNNDeltaWeightBckProp, hrt/include/DGLHackKernel/RGNN/my_shmem_sgemm_func.cu.h:711-776,
// an outer-product SGEMM that re-reads X and dY per 32x32 tile and atomically adds every product
// block). Both rows are 64 floats wide (hidden 64, every BASELINE.json config), so the job is a
// 64 x 64 contraction over the ROWS of a segment: 512 B of HBM reads per row against 8 KFLOP,
// i.e. bandwidth bound once the contraction leaves the fp32 SIMT pipe (the SIMT kernel it
// replaces ran at the FMA peak, 8 % of the HBM roofline).
//
// Layout trick: a row-major [rows][32 floats] panel whose 32-byte chunks are XOR-swizzled with
// the row is exactly the canonical MN-major tf32 UMMA operand (SWIZZLE_128B_BASE32B) with
// MN = feature and K = row (tc_ptx.cuh). So rows are copied as they lie in global memory (one
// 16-byte cp.async per chunk, swizzled destination) and the tensor core contracts over the row
// index — no transpose anywhere.
//
// Precision: 3xTF32 (tc has no fp32 MMA). Each value is split into hi (tf32) + lo (exact
// remainder). The hi and lo copies of X are stacked as ONE M = 128 operand
// [X_hi f0-31 | X_hi f32-63 | X_lo f0-31 | X_lo f32-63], so two MMAs per 8 rows
//   D[0:64]   = X_hi^T G_hi + X_hi^T G_lo
//   D[64:128] = X_lo^T G_hi + X_lo^T G_lo
// give all four partial products (one more than 3xTF32 needs) and the two halves of D are summed
// by the reductions into dW.
//
// Structure (persistent, one CTA per SM, 544 threads, warp specialised, mbarrier pipeline):
//   warps 0-3   flush    : at the end of a segment run, tcgen05.ld the 128 x 64 accumulator and
//                          red.add.v4 it into dW[s] (only the live blocks for per-head weights)
//   warp  4     MMA      : one lane issues 2 tcgen05.mma.kind::tf32 (M128 N64 K8, A and B MN-major)
//                          per 8 rows; the accumulator stays in TMEM across the CTA's whole run of
//                          a segment (two accumulators, so a flush overlaps the next segment)
//   warps 5-8   loaders  : each takes every 4th stage: 32 rows x (X row + G row) as 16-byte
//                          cp.async (zero fill past the end of a piece), completion signalled
//                          through cp.async.mbarrier.arrive; gather indices are fetched one own
//                          stage (= 4 stages) ahead. Nothing waits on a global load.
//   warps 9-16  split    : LDS the raw rows, write hi in place and lo beside it, proxy fence
// A CTA owns a contiguous row range; its stages are cut at segment boundaries, so an accumulator
// only ever holds rows of one segment. Six 32 KB stages = 96 KB of raw rows in flight per SM.
#include "common.cuh"
#include "tc_ptx.cuh"

namespace hetb {
namespace tcgw {

using namespace tc;

constexpr int SR = 32;       // rows per stage = 4 MMA k-steps
constexpr int kStages = 6;
constexpr int kEpiWarps = 4;
constexpr int kLoadWarps = 4;
constexpr int kConvWarps = 8;
constexpr int kConvThreads = kConvWarps * 32;
constexpr int kThreads = (kEpiWarps + 1 + kLoadWarps + kConvWarps) * 32;  // 544

constexpr uint32_t kAtom = SR * 128;  // SR rows of one 32-feature panel: 4 KB
constexpr uint32_t kOffXhi = 0, kOffXlo = 2 * kAtom, kOffGhi = 4 * kAtom, kOffGlo = 6 * kAtom;
constexpr uint32_t kStageBytes = 8 * kAtom;  // 32 KB
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;

// D = f32, A = B = tf32, A and B MN-major (bits 15, 16), N = 64, M = 128
constexpr uint32_t kIdescMN =
    (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((128u >> 4) << 24);

// The CTA's rows cut into stages of at most SR rows that never cross a segment boundary.
// Every role walks the same sequence.
struct RowItems {
  const int64_t *__restrict__ seg_ptrs;
  int num_segs;
  int64_t M, row, row_stop, seg_end;
  int s;

  __device__ RowItems(const int64_t *sp, int ns, int64_t m, int64_t r0, int64_t r1)
      : seg_ptrs(sp), num_segs(ns), M(m), row(r0), row_stop(r1), seg_end(0), s(0) {
    if (row < row_stop) {
      s = segment_of(seg_ptrs, num_segs, row);
      skip_empty();
    }
  }
  __device__ void skip_empty() {
    while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= row) ++s;
    seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;
  }
  // rows [r0, r0 + nrows) of segment seg; last = the accumulator of this segment run is complete
  __device__ bool next(int64_t &r0, int &nrows, int &seg, bool &last) {
    if (row >= row_stop) return false;
    int64_t stop = row + SR;
    if (seg_end < stop) stop = seg_end;
    if (row_stop < stop) stop = row_stop;
    r0 = row;
    nrows = (int)(stop - row);
    seg = s;
    row = stop;
    last = (row == seg_end) || (row >= row_stop);
    if (row == seg_end && row < row_stop) {
      ++s;
      skip_empty();
    }
    return true;
  }
};

__global__ void __launch_bounds__(kThreads, 1)
    gradw_tc_kernel(const float *__restrict__ A, const int64_t *__restrict__ a_idx, int a_heads,
                    const float *__restrict__ G, const int64_t *__restrict__ g_idx, float *__restrict__ dW,
                    const int64_t *__restrict__ seg_ptrs, int num_segs, int64_t M, int H, int K, int N,
                    int64_t rows_per_cta, int probe) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SW128 atoms need 1024 B alignment
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_base = base + kStages * kStageBytes;
  auto full_bar = [&](int i) { return bar_base + 8u * i; };                   // cp.async landed
  auto conv_bar = [&](int i) { return bar_base + 8u * (kStages + i); };       // hi/lo written
  auto empty_bar = [&](int i) { return bar_base + 8u * (2 * kStages + i); };  // MMAs retired
  auto tfull_bar = [&](int i) { return bar_base + 8u * (3 * kStages + i); };
  auto tempty_bar = [&](int i) { return bar_base + 8u * (3 * kStages + 2 + i); };
  const uint32_t tmem_slot = bar_base + 8u * (3 * kStages + 4);
  volatile uint32_t *tmem_slot_g = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kEpiWarps) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        mbar_init(full_bar(i), 32);  // the 32 lanes of the loader warp that owns the stage
        mbar_init(conv_bar(i), kConvThreads);
        mbar_init(empty_bar(i), 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(tfull_bar(i), 1);
        mbar_init(tempty_bar(i), kEpiWarps * 32);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 128);  // two 64-column accumulators
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_g;

  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  RowItems it(seg_ptrs, num_segs, M, row0, row1);
  int64_t r0;
  int nrows, seg;
  bool last;

  if (warp > kEpiWarps + kLoadWarps) {
    // ===================== split hi / lo =====================
    const int c = threadIdx.x - (kEpiWarps + 1 + kLoadWarps) * 32;
    uint32_t n = 0;
    while (it.next(r0, nrows, seg, last)) {
      const uint32_t stage = n % kStages, phase = (n / kStages) & 1;
      mbar_wait(full_bar(stage), phase);
      uint8_t *sb = gbase + stage * kStageBytes;
#pragma unroll
      for (int j = 0; j < SR * 32 / kConvThreads; ++j) {
        const int id = c + kConvThreads * j;
        const int row = id >> 5, q = id & 31, mat = q >> 4, ch = q & 15;
        const uint32_t off = (uint32_t)(mat * (4 * kAtom) + (ch >> 3) * kAtom) + mn32b_off(row, ch & 7);
        const float4 x = *reinterpret_cast<const float4 *>(sb + off);
        const float4 hi = make_float4(tf32_hi(x.x), tf32_hi(x.y), tf32_hi(x.z), tf32_hi(x.w));
        const float4 lo = make_float4(x.x - hi.x, x.y - hi.y, x.z - hi.z, x.w - hi.w);
        if (!(probe & 1)) *reinterpret_cast<float4 *>(sb + off) = hi;
        *reinterpret_cast<float4 *>(sb + off + 2 * kAtom) = lo;
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      mbar_arrive(conv_bar(stage));
      ++n;
    }
  } else if (warp > kEpiWarps) {
    // ===================== loaders =====================
    const int lw = warp - kEpiWarps - 1;
    RowItems itp = it;  // walks ahead: (pr0, pn) is this warp's next stage
    int64_t pr0 = 0;
    int pn = 0, ps;
    bool pl, pvalid = true;
    for (int i = 0; i <= lw && pvalid; ++i) pvalid = itp.next(pr0, pn, ps, pl);
    auto fetch_idx = [&](const int64_t *idx, int64_t rbase, int cnt) -> int {
      // a row index fits 31 bits: 2^31 rows of 256 B would be 550 GB
      if (lane >= cnt) return 0;
      return idx ? (int)__ldg(idx + rbase + lane) : (int)(rbase + lane);
    };
    int ia = 0, ig = 0;
    if (pvalid) {
      ia = fetch_idx(a_idx, pr0, pn);
      ig = fetch_idx(g_idx, pr0, pn);
    }
    const int cl = lane & 15;
    const float *srcbase = (lane < 16) ? A : G;
    const uint32_t dst_lane = (lane < 16 ? kOffXhi : kOffGhi) + (uint32_t)(cl >> 3) * kAtom;
    uint32_t n = (uint32_t)lw;
    while (pvalid) {
      const int cur_n = pn, cia = ia, cig = ig;
      for (int i = 0; i < kLoadWarps && pvalid; ++i) pvalid = itp.next(pr0, pn, ps, pl);
      if (pvalid) {  // indices of this warp's next stage: four stages of time to land
        ia = fetch_idx(a_idx, pr0, pn);
        ig = fetch_idx(g_idx, pr0, pn);
      }
      const uint32_t stage = n % kStages, phase = (n / kStages) & 1;
      mbar_wait(empty_bar(stage), phase ^ 1);
      const uint32_t sb = base + stage * kStageBytes + dst_lane;
#pragma unroll 8
      for (int i = 0; i < SR; ++i) {
        const int ra = __shfl_sync(0xffffffffu, cia, i), rg = __shfl_sync(0xffffffffu, cig, i);
        const bool ok = i < cur_n;
        const int r = (lane < 16) ? ra : rg;
        const float *src = ok ? srcbase + (int64_t)r * 64 + cl * 4 : srcbase;
        cp_async16(sb + mn32b_off(i, cl & 7), src, ok ? 16u : 0u);
      }
      cp_async_arrive_noinc(full_bar(stage));
      n += kLoadWarps;
    }
  } else if (warp == kEpiWarps) {
    // ===================== MMA issuer =====================
    uint32_t n = 0, f = 0;
    bool fresh = true;
    while (it.next(r0, nrows, seg, last)) {
      const uint32_t stage = n % kStages, phase = (n / kStages) & 1;
      if (fresh) mbar_wait(tempty_bar(f & 1), ((f >> 1) & 1) ^ 1);  // accumulator drained by the flush warps
      mbar_wait(conv_bar(stage), phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sb = base + stage * kStageBytes;
        const uint32_t d = tmem_base + (f & 1) * 64u;
        const int nk = (nrows + 7) >> 3;
        for (int j = 0; j < nk; ++j) {
          const uint64_t a = make_desc_mn_sw128_32b(sb + kOffXhi + j * 1024, kAtom);
          const uint64_t bh = make_desc_mn_sw128_32b(sb + kOffGhi + j * 1024, kAtom);
          const uint64_t bl = make_desc_mn_sw128_32b(sb + kOffGlo + j * 1024, kAtom);
          umma_tf32(d, a, bh, kIdescMN, (fresh && j == 0) ? 0u : 1u);
          umma_tf32(d, a, bl, kIdescMN, 1u);
        }
        umma_commit(empty_bar(stage));
        if (last) umma_commit(tfull_bar(f & 1));
      }
      __syncwarp();
      fresh = false;
      if (last) {
        ++f;
        fresh = true;
      }
      ++n;
    }
  } else {
    // ===================== flush =====================
    uint32_t f = 0;
    const int m = warp * 32 + lane;  // accumulator row: feature (m & 63) of X_hi (m < 64) or X_lo
    const int kf = m & 63;
    const int kk = kf % K;
    while (it.next(r0, nrows, seg, last)) {
      if (!last) continue;
      const uint32_t acc = f & 1;
      mbar_wait(tfull_bar(acc), (f >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * 64u;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t v[32];
        tmem_ld32(taddr + 32 * half, v);
        tmem_ld_wait();
        if (half == 1) {
          tc_fence_before();
          mbar_arrive(tempty_bar(acc));  // accumulator free: the next segment run may start
        }
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          const int col = half * 32 + c4 * 4;
          const int h = col / N, nn = col % N;
          if (a_heads == 1 || kf / K == h) {
            float *dst = dW + (((int64_t)seg * H + h) * K + kk) * N + nn;
            red_add4(dst, make_float4(__uint_as_float(v[4 * c4]), __uint_as_float(v[4 * c4 + 1]),
                                      __uint_as_float(v[4 * c4 + 2]), __uint_as_float(v[4 * c4 + 3])));
          }
        }
      }
      ++f;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

}  // namespace tcgw
}  // namespace hetb

using namespace hetb;

// Returns HETB_EUNSUPPORTED when the shape is outside the tensor-core path; the caller
// (hetb_segment_mm_grad_w) then uses the SIMT kernels.
int hetb_segment_mm_grad_w_tc(const float *A, const int64_t *a_idx, int a_heads, const float *G,
                              const int64_t *g_idx, int g_heads, float *dW, const int64_t *seg_ptrs,
                              int64_t num_segs, int64_t num_rows, int H, int K, int N, int probe, void *stream) {
  if (a_heads * K != 64 || g_heads * N != 64 || g_heads != H || (N % 4) != 0) return HETB_EUNSUPPORTED;
  if (a_heads != 1 && a_heads != H) return HETB_EUNSUPPORTED;
  if (num_segs > INT32_MAX) return HETB_EUNSUPPORTED;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(tcgw::gradw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)tcgw::kSmemBytes);
    if (ce != cudaSuccess) return (int)ce;
    attr_set = true;
  }
  int64_t ctas = ceil_div64(num_rows, 8 * tcgw::SR);
  if (ctas > kNumSMs) ctas = kNumSMs;
  const int64_t rows_per_cta = ceil_div64(num_rows, ctas);
  ctas = ceil_div64(num_rows, rows_per_cta);
  tcgw::gradw_tc_kernel<<<(unsigned)ctas, tcgw::kThreads, tcgw::kSmemBytes, (cudaStream_t)stream>>>(
      A, a_idx, a_heads, G, g_idx, dW, seg_ptrs, (int)num_segs, num_rows, H, K, N, rows_per_cta, probe);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
