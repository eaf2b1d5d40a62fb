// Grouped segment MM on tcgen05 + TMEM for hidden sizes above 64 (sm_100a):
//
//   Y[c_idx?[i], :] (=|+=) A[a_idx?[i], :] · B_s        rows i of segment s,
//
// A rows Kt floats wide (Kt a multiple of 32), B_s the dense [Kt x Nt] matrix W[s] denotes (one head:
// [K x N]; a_heads = 1, c_heads = H: the H per-head [K x N] blocks side by side, Nt = H·N), Nt a
// multiple of 64. segment_mm_tc.cu keeps the 64 x 64 case (all BASELINE.json configs), where B_s
// stays resident in shared memory; here B_s is up to 1 MB, so both operands stream through a
// shared-memory ring, one 32-float K panel per stage, and the accumulator tile is 128 x NT with
// NT = 256 / 128 / 64 (the largest that divides Nt): at NT = 256 a stage carries 3 x 4 MMAs of
// M128 N256 K8 (1.6 k tensor clocks) against 16 KB of A and 32 KB of B_s coming from L2, which is what
// keeps the tensor pipe, not L2, the limiter (reference: RGNN/my_shmem_sgemm_func.cu.h:14-466, a
// 32 x 32-tile SIMT SGEMM).
//
// Precision: 3xTF32 as in segment_mm_tc.cu. The tensor core ignores the low 13 mantissa bits of a
// tf32 operand, so the raw fp32 panel IS the hi operand; only lo = x - trunc_tf32(x) is computed.
//
// Roles (persistent, one CTA per SM, 416 threads):
//   warps 0-3   epilogue : tcgen05.ld the 128 x NT accumulator 64 columns at a time, stage through
//                          swizzled shared memory, 128-bit coalesced row stores / vector reductions
//   warp  4     MMA      : one lane issues 12 tcgen05.mma.kind::tf32 (M128 N<NT> K8, both operands
//                          K-major, SWIZZLE_128B) per stage, commits to mbarriers; two TMEM accumulators
//   warps 5-12  producers: A panel rows with 128-bit loads (8 lanes per 128 B, gathered through a_idx),
//                          B_s panel read as it lies in memory ([k][n], n contiguous: 32-bit loads
//                          coalesced along n, four consecutive k per lane) and stored TRANSPOSED as the
//                          K-major operand with one 128-bit store per 4 k — no workspace, no pre-pass.
//                          A rows (HBM) are loaded two stages ahead into two register sets, the B_s
//                          panel (L2) one stage ahead into one.
#include "common.cuh"
#include "tc_items.cuh"
#include "tc_ptx.cuh"

namespace hetb {
namespace tcw {

using namespace tc;

constexpr int TM = kTileRows;
constexpr int kEpiWarps = 4;
constexpr int kProdWarps = 8;
constexpr int kThreads = (kEpiWarps + 1 + kProdWarps) * 32;  // 416
constexpr uint32_t kPanelA = TM * 128;                       // 128 rows x 32 floats: 16 KB
constexpr uint32_t kEpiBytes = TM * 64 * 4;                  // 32 KB: 64 output columns of the tile

template <int NT>
struct Cfg {
  static constexpr uint32_t kPanelB = NT * 128;
  static constexpr uint32_t kOffAlo = kPanelA, kOffB = 2 * kPanelA, kOffBlo = 2 * kPanelA + kPanelB;
  static constexpr uint32_t kStageBytes = 2 * kPanelA + 2 * kPanelB;  // 96 / 64 / 48 KB
  static constexpr int kStages = (192 * 1024) / kStageBytes;          // 2 / 3 / 4
  static constexpr uint32_t kOffEpi = kStages * kStageBytes;
  static constexpr uint32_t kSmemBytes = kOffEpi + kEpiBytes + 1024 /*align*/ + 256 /*barriers*/;
  // D = f32, A = B = tf32, both K-major, N = NT, M = 128
  static constexpr uint32_t kIdesc =
      (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NT >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
};

// (piece, N tile, K panel) in the order every role walks them
struct StageIter {
  PieceIter it;
  int t0, p0, p1, seg, nt, kp, ntiles, kpanels;
  bool valid;
  __device__ StageIter(const PieceIter &src, int ntiles_, int kpanels_)
      : it(src), t0(0), p0(0), p1(0), seg(0), nt(0), kp(0), ntiles(ntiles_), kpanels(kpanels_), valid(false) {
    valid = it.next(t0, p0, p1, seg);
  }
  __device__ __forceinline__ void advance() {
    if (++kp == kpanels) {
      kp = 0;
      if (++nt == ntiles) {
        nt = 0;
        valid = it.next(t0, p0, p1, seg);
      }
    }
  }
};

template <int NT, bool GATHER>
__global__ void __launch_bounds__(kThreads, 1)
    segmm_tc_wide_kernel(const float *__restrict__ A, const int64_t *__restrict__ a_idx, const float *__restrict__ W,
                         float *__restrict__ Y, const int64_t *__restrict__ c_idx, const int64_t *__restrict__ seg_ptrs,
                         int num_segs, int64_t M, int H, int K, int N, int accumulate) {
  using C = Cfg<NT>;
  constexpr int kStages = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SW128 atoms need 1024 B alignment
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_base = base + C::kOffEpi + kEpiBytes;
  auto full_bar = [&](int i) { return bar_base + 8u * i; };
  auto empty_bar = [&](int i) { return bar_base + 8u * (kStages + i); };
  auto tfull_bar = [&](int i) { return bar_base + 8u * (2 * kStages + i); };
  auto tempty_bar = [&](int i) { return bar_base + 8u * (2 * kStages + 2 + i); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 4);
  volatile uint32_t *tmem_slot_g = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Nt = H * N;            // output row width
  const int ntiles = Nt / NT;      // N tiles per row tile
  const int kpanels = K / 32;      // stages per (piece, N tile)

  if (warp == kEpiWarps) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        mbar_init(full_bar(i), kProdWarps);
        mbar_init(empty_bar(i), 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(tfull_bar(i), 1);
        mbar_init(tempty_bar(i), kEpiWarps);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 2 * NT);  // two accumulators of NT columns
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_g;

  const int64_t num_tiles = ceil_div64(M, TM);
  const int64_t tile_begin = num_tiles * blockIdx.x / gridDim.x;
  const int64_t tile_end = num_tiles * (blockIdx.x + 1) / gridDim.x;
  const PieceIter pieces(seg_ptrs, num_segs, M, tile_begin, tile_end);
  const int64_t row_base = tile_begin * TM;

  if (warp > kEpiWarps) {
    // ===================== producers =====================
    const int p = threadIdx.x - (kEpiWarps + 1) * 32;
    const int pw = p >> 5;
    const int chunkA = p & 7, rA = p >> 3;  // A: 16-byte chunk of the 128 B panel row, rows rA + 32 i
    constexpr int RA = TM / 32;             // 4 A rows per thread
    constexpr int RB = NT / 32;             // B: column i*32 + lane, k chunk pw, per thread
    StageIter ahead(pieces, ntiles, kpanels);  // the stage whose loads are issued next (two ahead of the stores)
    // Row numbers of the current piece (-1 = no row). With a gather list they are loaded when `ahead` enters
    // a piece (one exposed load per piece: K/32 x Nt/NT stages); a prefetch one piece ahead made the compiler
    // rotate load-written registers every stage, and with six scoreboards per warp that false dependency
    // stalled every stage on the A rows just issued (profiles/r02j_segmm_wide_ncu_summary.txt).
    int ixA[RA];
    auto load_ix = [&]() {
#pragma unroll
      for (int i = 0; i < RA; ++i) {
        const int rr = ahead.t0 + rA + 32 * i;
        const bool ok = rr >= ahead.p0 && rr < ahead.p1;
        const int64_t gr = row_base + rr;
        int x = ok ? (int)gr : -1;  // a row index fits 31 bits
        if (GATHER) {
          if (ok) x = (int)__ldg(a_idx + gr);
        }
        ixA[i] = x;
      }
    };
    // A rows come from HBM (gathered): two register sets, loads issued two stages ahead. B_s comes from L2
    // (every CTA of a segment reads the same weight): one register set, loaded one stage ahead, right after
    // the previous panel has been stored — two B sets would not fit the 128 registers a 13-warp CTA gets.
    struct BNext { int seg, nt, kp, valid; } nb = {0, 0, 0, 0};  // the stage whose B panel is loaded next
    float4 vb[RB];
    auto issue_a = [&](float4 *va) -> int {  // loads of stage `ahead`; 1 if there was one
      const int had = ahead.valid ? 1 : 0;
      if (ahead.valid) {
        if (ahead.kp == 0 && ahead.nt == 0) load_ix();
        const float *ap = A + ahead.kp * 32 + chunkA * 4;
#pragma unroll
        for (int i = 0; i < RA; ++i) {
          va[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (ixA[i] >= 0) va[i] = ldg4(ap + (int64_t)ixA[i] * K);
        }
      }
      return had;
    };
    auto issue_b = [&]() {
      if (nb.valid) {
        const int k0 = nb.kp * 32 + pw * 4;
        if (H == 1) {  // one pointer per k row, the 32-column groups at immediate offsets
          const float *w0 = W + ((int64_t)nb.seg * K + k0) * N + (nb.nt * NT + lane);
          const float *w1 = w0 + N, *w2 = w1 + N, *w3 = w2 + N;
#pragma unroll
          for (int i = 0; i < RB; ++i)
            vb[i] = make_float4(__ldg(w0 + i * 32), __ldg(w1 + i * 32), __ldg(w2 + i * 32), __ldg(w3 + i * 32));
        } else {
#pragma unroll
          for (int i = 0; i < RB; ++i) {
            const int cb = nb.nt * NT + i * 32;  // first output column of this warp's 32 (one head: N % 32 == 0)
            const int h = cb / N;                // head cb / N, column c - h N of W[seg, head]
            const float *wp = W + (((int64_t)nb.seg * H + h) * K + k0) * N + (cb - h * N + lane);
            vb[i] = make_float4(__ldg(wp), __ldg(wp + N), __ldg(wp + 2 * (int64_t)N), __ldg(wp + 3 * (int64_t)N));
          }
        }
      }
    };
    auto snapshot = [&]() { nb.seg = ahead.seg; nb.nt = ahead.nt; nb.kp = ahead.kp; nb.valid = ahead.valid ? 1 : 0; };
    uint32_t stage = 0, phase = 0;
    int inflight = 0;  // stages whose A rows are in registers, not yet stored
    auto step = [&](float4 *va) -> bool {
      if (inflight == 0) return false;
      --inflight;
      mbar_wait(empty_bar(stage), phase ^ 1);
      uint8_t *gs = gbase + stage * C::kStageBytes;
#pragma unroll
      for (int i = 0; i < RB; ++i) {
        const int r = i * 32 + lane;
        const uint32_t off = (uint32_t)(r * 128 + ((pw ^ (r & 7)) << 4));
        const float4 x = vb[i];
        const float4 lo = make_float4(x.x - tf32_hi(x.x), x.y - tf32_hi(x.y), x.z - tf32_hi(x.z), x.w - tf32_hi(x.w));
        *reinterpret_cast<float4 *>(gs + C::kOffB + off) = x;  // raw = hi: the tensor core drops the low 13 bits
        *reinterpret_cast<float4 *>(gs + C::kOffBlo + off) = lo;
      }
      issue_b();  // B panel of stage n + 1 (nb), into the registers just drained
#pragma unroll
      for (int i = 0; i < RA; ++i) {
        const int r = rA + 32 * i;
        const uint32_t off = (uint32_t)(r * 128 + ((chunkA ^ (r & 7)) << 4));
        const float4 x = va[i];
        const float4 lo = make_float4(x.x - tf32_hi(x.x), x.y - tf32_hi(x.y), x.z - tf32_hi(x.z), x.w - tf32_hi(x.w));
        *reinterpret_cast<float4 *>(gs + off) = x;
        *reinterpret_cast<float4 *>(gs + C::kOffAlo + off) = lo;
      }
      snapshot();                 // stage n + 2: its B panel is loaded by the next step
      inflight += issue_a(va);    // A rows of stage n + 2
      ahead.advance();
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar(stage));
      if (++stage == kStages) {
        stage = 0;
        phase ^= 1;
      }
      return true;
    };

    float4 a0[RA], a1[RA];
    snapshot();    // stage 0
    inflight += issue_a(a0);
    ahead.advance();
    issue_b();     // B panel of stage 0
    snapshot();    // stage 1
    inflight += issue_a(a1);
    ahead.advance();
    for (;;) {
      if (!step(a0)) break;
      if (!step(a1)) break;
    }
  } else if (warp == kEpiWarps) {
    // ===================== MMA issuer =====================
    StageIter st(pieces, ntiles, kpanels);
    uint32_t stage = 0, phase = 0, accn = 0;
    while (st.valid) {
      const uint32_t acc = accn & 1;
      if (st.kp == 0) mbar_wait(tempty_bar(acc), ((accn >> 1) & 1) ^ 1);  // accumulator drained by the epilogue
      mbar_wait(full_bar(stage), phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sbase = base + stage * C::kStageBytes;
        const uint32_t d = tmem_base + acc * NT;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint64_t a_hi = make_desc_sw128(sbase + j * 32), a_lo = make_desc_sw128(sbase + C::kOffAlo + j * 32);
          const uint64_t b_hi = make_desc_sw128(sbase + C::kOffB + j * 32);
          const uint64_t b_lo = make_desc_sw128(sbase + C::kOffBlo + j * 32);
          umma_tf32(d, a_lo, b_hi, C::kIdesc, (st.kp == 0 && j == 0) ? 0u : 1u);
          umma_tf32(d, a_hi, b_lo, C::kIdesc, 1);
          umma_tf32(d, a_hi, b_hi, C::kIdesc, 1);
        }
        umma_commit(empty_bar(stage));                        // operands may be overwritten once these retire
        if (st.kp == kpanels - 1) umma_commit(tfull_bar(acc));  // accumulator complete
      }
      __syncwarp();
      if (st.kp == kpanels - 1) ++accn;
      st.advance();
      if (++stage == kStages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue =====================
    uint8_t *stg = gbase + C::kOffEpi + warp * (32 * 64 * 4);  // this warp's 32 rows x 256 B
    PieceIter it = pieces;
    int t0, p0, p1, seg;
    uint32_t accn = 0;
    while (it.next(t0, p0, p1, seg)) {
      const int my_rr = t0 + warp * 32 + lane;  // lane j <-> tile row warp*32 + j
      const int64_t my_gr = row_base + my_rr;
      int64_t my_cr = my_gr;
      if (c_idx && my_rr >= p0 && my_rr < p1) my_cr = __ldg(c_idx + my_gr);
      for (int nt = 0; nt < ntiles; ++nt, ++accn) {
        const uint32_t acc = accn & 1;
        mbar_wait(tfull_bar(acc), (accn >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * NT;
#pragma unroll 1
        for (int ch = 0; ch < NT / 64; ++ch) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t v[32];
            tmem_ld32(taddr + ch * 64 + 32 * half, v);
            tmem_ld_wait();
            if (ch == NT / 64 - 1 && half == 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(tempty_bar(acc));  // accumulator free: the next tile's MMAs may start
            }
#pragma unroll
            for (int c8 = 0; c8 < 8; ++c8) {
              const int c = half * 8 + c8;
              const int pc = (c & 8) | ((c & 7) ^ (lane & 7));
              *reinterpret_cast<uint4 *>(stg + lane * 256 + pc * 16) =
                  make_uint4(v[4 * c8], v[4 * c8 + 1], v[4 * c8 + 2], v[4 * c8 + 3]);
            }
          }
          __syncwarp();
          const int c = lane & 15;
          const int col0 = nt * NT + ch * 64 + c * 4;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int lr = i * 2 + (lane >> 4);
            const int rr = t0 + warp * 32 + lr;
            const int64_t cr = __shfl_sync(0xffffffffu, my_cr, lr);
            if (rr >= p0 && rr < p1) {
              const int pc = (c & 8) | ((c & 7) ^ (lr & 7));
              const float4 val = *reinterpret_cast<const float4 *>(stg + lr * 256 + pc * 16);
              float *dst = Y + cr * Nt + col0;
              if (accumulate) red_add4(dst, val);
              else st4(dst, val);
            }
          }
          __syncwarp();
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * NT);
  }
}

template <int NT, bool GATHER>
static int launch_wide_g(const float *A, const int64_t *a_idx, const float *W, float *Y, const int64_t *c_idx,
                       const int64_t *seg_ptrs, int num_segs, int64_t num_rows, int H, int K, int N, int accumulate,
                       cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(segmm_tc_wide_kernel<NT, GATHER>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)Cfg<NT>::kSmemBytes);
    if (ce != cudaSuccess) return (int)ce;
    attr_set = true;
  }
  const int64_t tiles = ceil_div64(num_rows, TM);
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  segmm_tc_wide_kernel<NT, GATHER><<<grid, kThreads, Cfg<NT>::kSmemBytes, st>>>(A, a_idx, W, Y, c_idx, seg_ptrs, num_segs,
                                                                        num_rows, H, K, N, accumulate);
  HETB_LAUNCH_OK();
  return HETB_OK;
}

template <int NT>
static int launch_wide(const float *A, const int64_t *a_idx, const float *W, float *Y, const int64_t *c_idx,
                       const int64_t *seg_ptrs, int num_segs, int64_t num_rows, int H, int K, int N, int accumulate,
                       cudaStream_t st) {
  return a_idx ? launch_wide_g<NT, true>(A, a_idx, W, Y, c_idx, seg_ptrs, num_segs, num_rows, H, K, N, accumulate, st)
               : launch_wide_g<NT, false>(A, a_idx, W, Y, c_idx, seg_ptrs, num_segs, num_rows, H, K, N, accumulate, st);
}

}  // namespace tcw
}  // namespace hetb

using namespace hetb;

// Returns HETB_EUNSUPPORTED when the shape is outside this kernel; hetb_segment_mm then falls back
// to the 64-wide tensor-core kernel or the SIMT kernels.
int hetb_segment_mm_tc_wide(const float *A, const int64_t *a_idx, int a_heads, const float *W, float *Y,
                            const int64_t *c_idx, int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                            int64_t num_rows, int H, int K, int N, int accumulate, void *stream) {
  if (a_heads != 1 || c_heads != H) return HETB_EUNSUPPORTED;       // dense [K x H·N] products only
  if (H > 1 && (N % 32) != 0) return HETB_EUNSUPPORTED;              // a warp's 32 columns lie in one head
  const int Nt = H * N;
  if ((K % 32) != 0 || (Nt % 64) != 0 || (K <= 64 && Nt <= 64)) return HETB_EUNSUPPORTED;
  if (K > 4096 || Nt > 4096 || num_segs > INT32_MAX) return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (Nt % 256 == 0)
    return tcw::launch_wide<256>(A, a_idx, W, Y, c_idx, seg_ptrs, (int)num_segs, num_rows, H, K, N, accumulate, st);
  if (Nt % 128 == 0)
    return tcw::launch_wide<128>(A, a_idx, W, Y, c_idx, seg_ptrs, (int)num_segs, num_rows, H, K, N, accumulate, st);
  return tcw::launch_wide<64>(A, a_idx, W, Y, c_idx, seg_ptrs, (int)num_segs, num_rows, H, K, N, accumulate, st);
}
