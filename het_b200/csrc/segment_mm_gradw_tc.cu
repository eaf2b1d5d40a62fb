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
// MN = feature and K = row (tc_ptx.cuh). So rows are stored as they lie in global memory (one
// 16-byte chunk per lane, swizzled destination) and the tensor core contracts over the row
// index — no transpose anywhere.
//
// Precision: split-tf32 (the tensor core has no fp32 MMA). x = hi + lo with hi = x truncated to
// tf32 and lo the exact remainder. The tensor core ignores the low 13 mantissa bits of a tf32
// operand, so the raw fp32 row IS the hi operand and only lo is computed. hi and lo are stacked
// into ONE M = 128 operand [X f0-31 | X f32-63 | X_lo f0-31 | X_lo f32-63] and ONE N = 128
// operand [G | G_lo], so a single MMA per 8 rows yields all four partial products
//   D[0:64, 0:64] = X_hi^T G_hi     D[0:64, 64:128] = X_hi^T G_lo
//   D[64:128, 0:64] = X_lo^T G_hi   D[64:128, 64:128] = X_lo^T G_lo
// which the flush adds into the same dW block (error vs the fp64 oracle ~1e-6 relative).
//
// Structure (persistent, one CTA per SM, 416 threads, warp specialised, mbarrier pipeline):
//   warps 0-3   flush    : at the end of a segment run, tcgen05.ld the 128 x 128 accumulator and
//                          red.add.v4 it into dW[s] (only the live blocks for per-head weights);
//                          they wait with a back-off so their polling does not steal issue slots
//   warp  4     MMA      : one lane issues tcgen05.mma.kind::tf32 (M128 N128 K8, A and B MN-major),
//                          one per 8 rows; the accumulator stays in TMEM across the CTA's whole run
//                          of a segment (two accumulators, so a flush overlaps the next segment).
//                          MN-major tf32 MMAs run at about half rate and the issue queue is
//                          shallow, so this warp does nothing else: what to issue comes from a
//                          one-word stage descriptor the producers leave in shared memory.
//   warps 5-12  producers: each warp owns SR/8 rows of every stage. Predicated 128-bit loads of the
//                          X and G rows (gather indices loaded straight into the address register,
//                          two register sets so the next stage's rows are in flight while this
//                          one is stored), then raw row + exact lo remainder into the swizzled
//                          stage, proxy fence, ONE mbarrier arrival per warp. (An earlier version
//                          used cp.async loaders + separate split warps: LDGSTS issue cost ~64 clk
//                          per 512 B made it slower than the register path)
// A CTA owns a contiguous row range; its stages are cut at segment boundaries, so an accumulator
// only ever holds rows of one segment.
#include "common.cuh"
#include "tc_ptx.cuh"

namespace hetb {
namespace tcgw {

using namespace tc;

#ifndef HETB_GW_SR
#define HETB_GW_SR 64
#endif
constexpr int SR = HETB_GW_SR;     // rows per stage (8 per MMA k-step)
constexpr int kStages = 192 / SR;  // 192 KB of stages: 96 KB of raw rows in flight per SM
constexpr int kEpiWarps = 4;
#ifndef HETB_GW_PW
#define HETB_GW_PW 8
#endif
constexpr int kProdWarps = HETB_GW_PW;
constexpr int kThreads = (kEpiWarps + 1 + kProdWarps) * 32;  // 416 with 8 producer warps
static_assert(SR == 32 || SR == 64, "stage rows");

constexpr uint32_t kAtom = SR * 128;  // SR rows of one 32-feature panel
constexpr uint32_t kOffX = 0, kOffXlo = 2 * kAtom, kOffG = 4 * kAtom, kOffGlo = 6 * kAtom;
constexpr uint32_t kStageBytes = 8 * kAtom;  // 1 KB per row
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 512 /*barriers, stage words*/;

// D = f32, A = B = tf32, A and B MN-major (bits 15, 16), N = 128, M = 128
constexpr uint32_t kIdescMN =
    (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t kAccCols = 128;  // TMEM columns per accumulator

// The CTA's rows cut into stages of at most SR rows that never cross a segment boundary.
// Positions are 32-bit offsets from the CTA's first row (64-bit compare/select chains cost
// ~200 cycles per call), and the segment table is only touched when a segment ends.
struct RowItems {
  const int64_t *__restrict__ seg_ptrs;
  int64_t row_base;  // first row of the CTA
  int num_segs, n, rel, seg_end, s;  // rows of the CTA, position, end of segment s (clamped to n)

  __device__ RowItems(const int64_t *sp, int ns, int64_t r0, int64_t r1)
      : seg_ptrs(sp), row_base(r0), num_segs(ns), n(r1 > r0 ? (int)(r1 - r0) : 0), rel(0), seg_end(0), s(0) {
    if (n > 0) {
      s = segment_of(seg_ptrs, num_segs, r0);
      skip_empty();
    }
  }
  __device__ void skip_empty() {
    const int64_t row = row_base + rel;
    while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= row) ++s;
    int64_t e = n;
    if (s + 1 < num_segs) {
      const int64_t se = __ldg(seg_ptrs + s + 1) - row_base;
      if (se < e) e = se;
    }
    seg_end = (int)e;
  }
  // rows [rel0, rel0 + nrows) (relative to row_base) of segment seg; last = the accumulator of
  // this segment run is complete
  __device__ __forceinline__ bool next(int &rel0, int &nrows, int &seg, bool &last) {
    if (rel >= n) return false;
    const int stop = min(rel + SR, seg_end);
    rel0 = rel;
    nrows = stop - rel;
    seg = s;
    rel = stop;
    last = (stop == seg_end);
    if (last && stop < n) {
      ++s;
      skip_empty();
    }
    return true;
  }
  // number of stages the whole range yields (walks segments, not stages)
  __device__ int count() const {
    RowItems c = *this;
    int items = 0;
    while (c.rel < c.n) {
      items += (c.seg_end - c.rel + SR - 1) / SR;
      c.rel = c.seg_end;
      if (c.rel < c.n) {
        ++c.s;
        c.skip_empty();
      }
    }
    return items;
  }
};

#define HETB_WAIT(acc, bar, par)       \
  do {                                 \
    if (dbg) {                         \
      const long long t__ = clock64(); \
      mbar_wait(bar, par);             \
      acc += clock64() - t__;          \
    } else {                           \
      mbar_wait(bar, par);             \
    }                                  \
  } while (0)

// GATHER = false: neither operand has a gather list; row numbers are computed, never loaded (a
// predicated-off index load still makes its register a load destination, and the scoreboard it shares
// with the row loads stalls the consumer: see segment_mm_tc.cu).
template <bool GATHER>
__global__ void __launch_bounds__(kThreads, 1)
    gradw_tc_kernel(const float *__restrict__ A, const int64_t *__restrict__ a_idx, int a_heads,
                    const float *__restrict__ G, const int64_t *__restrict__ g_idx, float *__restrict__ dW,
                    const int64_t *__restrict__ seg_ptrs, int num_segs, int64_t M, int H, int K, int N,
                    int64_t rows_per_cta, int probe, long long *__restrict__ dbg) {
  const long long t_start = dbg ? clock64() : 0;
  long long w0 = 0, w1 = 0, w2 = 0, w3 = 0;  // cycles this warp's lane 0 spent in barrier waits etc. (dbg only)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // swizzle atoms: 1024 B alignment
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_base = base + kStages * kStageBytes;
  auto full_bar = [&](int i) { return bar_base + 8u * i; };                   // rows and lo parts stored
  auto empty_bar = [&](int i) { return bar_base + 8u * (2 * kStages + i); };  // MMAs retired
  auto tfull_bar = [&](int i) { return bar_base + 8u * (3 * kStages + i); };
  auto tempty_bar = [&](int i) { return bar_base + 8u * (3 * kStages + 2 + i); };
  const uint32_t tmem_slot = bar_base + 8u * (3 * kStages + 4);
  volatile uint32_t *tmem_slot_g = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));
  // one word per stage, written by the split warps before they arrive: k-steps | last << 8
  volatile uint32_t *stage_word = tmem_slot_g + 2;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kEpiWarps) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        mbar_init(full_bar(i), kProdWarps);
        mbar_init(empty_bar(i), 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(tfull_bar(i), 1);
        mbar_init(tempty_bar(i), kEpiWarps * 32);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 2 * kAccCols);  // two accumulators
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_g;

  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  RowItems it(seg_ptrs, num_segs, row0, row1);
  int r0, nrows, seg;
  bool last;

  if (warp > kEpiWarps) {
    // ===================== producers: gather rows, split, store =====================
    // The gather is latency bound, so what matters is bytes in flight: every producer thread
    // keeps two stages of rows in registers (16 x 16 B). Right after stage n has been split and
    // stored, its registers are re-used for the row loads of stage n+2 (whose indices arrived two
    // steps ago) and its index registers for the index loads of stage n+4. The loop is unrolled
    // two ways so the register sets rotate by name.
    const int pw = warp - kEpiWarps - 1;          // rows pw, pw + 8, ...
    const int mat = lane >> 4, ch = lane & 15;    // lanes 0-15: X chunks, 16-31: G chunks
    constexpr int RP = SR / kProdWarps;           // rows per thread and stage
    const float *srcbase = (mat ? G : A) + ch * 4;
    const int64_t *idxbase = mat ? g_idx : a_idx;
    const uint32_t dst_lane = (uint32_t)(mat * (4 * kAtom) + (ch >> 3) * kAtom);
    RowItems it_ix = it;  // runs 2..4 stages ahead of `it`
    uint32_t n = 0;

    // Predicated, not branched (a branch per slot costs more than the load), and the load writes
    // the index register directly: anything computed from the loaded value here would stall this
    // step for a full memory latency.
    const bool ident = (idxbase == nullptr);
    auto issue_idx = [&](int *ix) {
      int q0, qn, qs;
      bool ql;
      const bool valid = it_ix.next(q0, qn, qs, ql);
      const int lim = (valid && !(probe & 4)) ? qn : 0;  // probe 4: no global traffic
      const int64_t rb = row0 + q0 + pw;
#pragma unroll
      for (int i = 0; i < RP; ++i) {
        const bool ok = pw + kProdWarps * i < lim;
        const int64_t r = rb + kProdWarps * i;
        // a row index fits 31 bits: 2^31 rows of 256 B would be 550 GB
        int v = ok ? (int)r : -1;
        if (GATHER) {
          if (ok && !ident) v = (int)__ldg(idxbase + r);
        }
        ix[i] = v;
      }
    };
    auto issue_rows = [&](const int *ix, float4 *v) {
#pragma unroll
      for (int i = 0; i < RP; ++i) {
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ix[i] >= 0) v[i] = ldg4(srcbase + (int64_t)ix[i] * 64);
      }
    };
    auto step = [&](float4 *rS, int *ixS) -> bool {
      if (!it.next(r0, nrows, seg, last)) return false;
      const uint32_t stage = n % kStages, phase = (n / kStages) & 1;
      HETB_WAIT(w0, empty_bar(stage), phase ^ 1);
      uint8_t *sb = gbase + stage * kStageBytes + dst_lane;
      long long t1 = dbg ? clock64() : 0;
#pragma unroll
      for (int i = 0; i < RP; ++i) {
        const uint32_t off = mn32b_off(pw + kProdWarps * i, ch & 7);
        const float4 x = rS[i];
        const float4 hi = make_float4(tf32_hi(x.x), tf32_hi(x.y), tf32_hi(x.z), tf32_hi(x.w));
        const float4 lo = make_float4(x.x - hi.x, x.y - hi.y, x.z - hi.z, x.w - hi.w);
        if (!(probe & 8)) {
          // the tensor core ignores the low 13 mantissa bits: the raw row is the hi operand
          *reinterpret_cast<float4 *>(sb + off) = x;
          *reinterpret_cast<float4 *>(sb + off + 2 * kAtom) = lo;
        }
      }
      long long t2 = dbg ? clock64() : 0;
      issue_rows(ixS, rS);  // stage n + 2
      issue_idx(ixS);       // stage n + 4
      long long t3 = dbg ? clock64() : 0;
      if (pw == 0 && lane == 0) stage_word[stage] = (uint32_t)((nrows + 7) >> 3) | (last ? 256u : 0u);
      if (!(probe & 32)) fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar(stage));  // one arrival per warp
      if (dbg) {
        w1 += t2 - t1;             // wait for the rows + split + store
        w2 += t3 - t2;             // issue loads
        w3 += clock64() - t3;      // fence + arrive
      }
      ++n;
      return true;
    };

    int ix0[RP], ix1[RP];
    float4 v0[RP], v1[RP];
    issue_idx(ix0);
    issue_idx(ix1);
    issue_rows(ix0, v0);
    issue_idx(ix0);
    issue_rows(ix1, v1);
    issue_idx(ix1);
    for (;;) {
      if (!step(v0, ix0)) break;
      if (!step(v1, ix1)) break;
    }
  } else if (warp == kEpiWarps) {
    // ===================== MMA issuer =====================
    const int items = it.count();
    uint32_t f = 0, stage = 0, phase = 0;
    bool fresh = true;
    const uint64_t desc0 = make_desc_mn_sw128_32b(base, kAtom);  // + byte offset >> 4 in the low word
    for (int n = 0; n < items; ++n) {
      if (fresh) HETB_WAIT(w0, tempty_bar(f & 1), ((f >> 1) & 1) ^ 1);  // accumulator drained by the flush warps
      HETB_WAIT(w0, full_bar(stage), phase);
      const uint32_t word = stage_word[stage];
      tc_fence_after();
      const bool lastf = (word & 256u) != 0;
      if (lane == 0) {
        const uint32_t d = tmem_base + (f & 1) * kAccCols;
        const int nk = (probe & 2) ? 0 : (int)(word & 255u);  // probe 2: pipeline without MMAs
        const uint64_t da = desc0 + ((stage * kStageBytes + kOffX) >> 4);
        const uint64_t db = desc0 + ((stage * kStageBytes + kOffG) >> 4);  // [G | G_lo], N = 128
        for (int j = 0; j < nk; ++j)
          umma_tf32(d, da + (uint64_t)(j * (1024 >> 4)), db + (uint64_t)(j * (1024 >> 4)), kIdescMN,
                    (fresh && j == 0) ? 0u : 1u);
        if (probe & 64) mbar_arrive(empty_bar(stage));  // probe 64 (with 2): plain arrive instead of tcgen05.commit
        else umma_commit(empty_bar(stage));
        if (lastf) umma_commit(tfull_bar(f & 1));
      }
      __syncwarp();
      fresh = lastf;
      f += lastf ? 1u : 0u;
      if (++stage == kStages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else {
    // ===================== flush =====================
    uint32_t f = 0;
    const int m = warp * 32 + lane;  // accumulator row: feature (m & 63) of X (m < 64) or X_lo
    const int kf = m & 63;
    const int kk = kf % K;
    while (it.next(r0, nrows, seg, last)) {
      if (!last) continue;
      const uint32_t acc = f & 1;
      mbar_wait_relaxed(tfull_bar(acc), (f >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * kAccCols;
#pragma unroll
      for (int part = 0; part < 4; ++part) {
        uint32_t v[32];
        tmem_ld32(taddr + 32 * part, v);
        tmem_ld_wait();
        if (part == 3) {
          tc_fence_before();
          mbar_arrive(tempty_bar(acc));  // accumulator free: the next segment run may start
        }
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          const int col = (part & 1) * 32 + c4 * 4;  // columns 64..127: the G_lo products of the same dW column
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

  if (dbg && lane == 0) {  // per warp: total cycles, cycles in barrier waits
    long long *o = dbg + ((int64_t)blockIdx.x * (kThreads / 32) + warp) * 5;
    o[0] = clock64() - t_start;
    o[1] = w0;
    o[2] = w1;
    o[3] = w2;
    o[4] = w3;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * kAccCols);
  }
}

}  // namespace tcgw
}  // namespace hetb

using namespace hetb;

static long long *g_gradw_dbg = nullptr;
// Debug: per-warp cycle counters ([CTAs][17 warps][2] int64) written by the next launches.
extern "C" int hetb_debug_set_gradw_buffer(void *p) {
  g_gradw_dbg = (long long *)p;
  return HETB_OK;
}

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
    cudaError_t ce = cudaFuncSetAttribute(tcgw::gradw_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)tcgw::kSmemBytes);
    if (ce != cudaSuccess) return (int)ce;
    ce = cudaFuncSetAttribute(tcgw::gradw_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)tcgw::kSmemBytes);
    if (ce != cudaSuccess) return (int)ce;
    attr_set = true;
  }
  int64_t ctas = ceil_div64(num_rows, 8 * tcgw::SR);
  if (ctas > kNumSMs) ctas = kNumSMs;
  const int64_t rows_per_cta = ceil_div64(num_rows, ctas);
  if (rows_per_cta > INT32_MAX) return HETB_EUNSUPPORTED;
  ctas = ceil_div64(num_rows, rows_per_cta);
  if (a_idx || g_idx)
    tcgw::gradw_tc_kernel<true><<<(unsigned)ctas, tcgw::kThreads, tcgw::kSmemBytes, (cudaStream_t)stream>>>(
        A, a_idx, a_heads, G, g_idx, dW, seg_ptrs, (int)num_segs, num_rows, H, K, N, rows_per_cta, probe, g_gradw_dbg);
  else
    tcgw::gradw_tc_kernel<false><<<(unsigned)ctas, tcgw::kThreads, tcgw::kSmemBytes, (cudaStream_t)stream>>>(
        A, a_idx, a_heads, G, g_idx, dW, seg_ptrs, (int)num_segs, num_rows, H, K, N, rows_per_cta, probe, g_gradw_dbg);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
