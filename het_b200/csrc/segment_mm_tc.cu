// Grouped segment MM on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
//   Y[c_idx?[i], :] (=|+=) A[a_idx?[i], :] · B_s        for rows i of segment s
//
// with a 64-wide contraction and a 64-wide output (hidden 64, every BASELINE.json
// config), fp32 in / fp32 out. B_s is the [64 x 64] matrix the (heads, K, N) weight
// W[s] denotes: dense when the input has one head, block-diagonal per head when input
// and output are both per-head (HGT's 4 x (16 x 16)), or the stacked [H*K, N] matrix
// when the heads are summed into one output row (delta-input of RGAT).
//
// Precision: the tensor core has no fp32 MMA. Each operand is split in registers
// into a tf32 "hi" part (low 13 mantissa bits cleared) and the exact remainder
// "lo", and D = A_lo·B_hi + A_hi·B_lo + A_hi·B_hi is accumulated in fp32 in TMEM
// (3xTF32): relative error ~2^-21, inside the 1e-4 parity bar. The kernel is HBM
// bound (16 FLOP/B), so the 3x MMA count costs nothing.
//
// Structure (persistent, one CTA per SM, warp specialised, 672 threads):
//   warps 0-3   epilogue : tcgen05.ld the 128x64 accumulator, stage through
//                          swizzled shared memory, 128-bit coalesced row stores
//                          (scatter through c_idx, or vector reductions)
//   warp  4     MMA      : one elected lane issues 24 tcgen05.mma.kind::tf32
//                          (M128 N64 K8) per work item, commits to mbarriers
//   warps 5-20  producers: gather A rows with 128-bit loads (16 lanes per 256 B
//                          row), split hi/lo, store into the 128B-swizzled K-major
//                          layout the UMMA descriptors expect; stage B_s transposed
// Two shared-memory stages (A_hi, A_lo, B_hi, B_lo = 96 KB each) and two TMEM
// accumulator stages (64 columns each) let the gather of item i+1, the MMAs of item
// i and the stores of item i-1 overlap. Gathered rows cannot be fetched as TMA
// tiles (every row has its own address), hence the register path for A; the B
// operand is 16 KB and is re-staged only when the segment changes.
//
// Work items: the row space is cut into 128-row tiles on a GLOBAL grid; a tile
// that straddles segment boundaries yields one item per segment piece (the full
// tile is multiplied, only that piece's rows are stored). No host schedule, no
// device->host copy of the segment offsets, no allocation (the reference needs all
// three per call: OpExport/RGNNOps.inc.h:150-169).
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_ptx.cuh"
#include "tc_items.cuh"

namespace hetb {
namespace tc {

constexpr int TM = kTileRows;  // rows per tile (UMMA M)
constexpr int KK = 64;   // contraction width
constexpr int NN = 64;   // output width (UMMA N)
constexpr int kEpiWarps = 4;
constexpr int kProdWarps = 8;
constexpr int kSets = 2;        // producer register sets (work items of row data in flight)
constexpr int kProdThreads = kProdWarps * 32;
constexpr int kThreads = (kEpiWarps + 1 + kProdWarps) * 32;  // 416
constexpr int kStages = 2;

constexpr uint32_t kPanelA = TM * 128;            // one 32-float K panel of A: 16 KB
constexpr uint32_t kPanelB = NN * 128;            // 8 KB
constexpr uint32_t kBytesA = 2 * kPanelA;         // 32 KB (hi or lo)
constexpr uint32_t kBytesB = 2 * kPanelB;         // 16 KB
constexpr uint32_t kStageBytes = 2 * kBytesA;     // A_hi + A_lo: 64 KB per stage
constexpr uint32_t kOffAlo = kBytesA;
constexpr uint32_t kBBufBytes = 2 * kBytesB;      // B_hi + B_lo: 32 KB per buffer, ring of 2
constexpr uint32_t kOffBlo = kBytesB;
constexpr uint32_t kEpiBytes = TM * NN * 4;       // 32 KB
constexpr uint32_t kOffB = kStages * kStageBytes;
constexpr uint32_t kOffEpi = kOffB + 2 * kBBufBytes;
constexpr uint32_t kSmemBytes = kOffEpi + kEpiBytes + 1024 /*align*/ + 256 /*barriers*/;

// B_s index maps (see header)
enum BMode : int { kDense = 0, kBlockDiag = 1, kStacked = 2 };

// Instruction descriptor (cute::UMMA::InstrDescriptor): D=f32 [4,6)=1, A=tf32 [7,10)=2, B=tf32 [10,13)=2,
// A,B K-major (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29).
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

// --------------------------------------------------------------- kernel ----
__global__ void __launch_bounds__(kThreads, 1)
    segmm_tc_kernel(const float *__restrict__ A, const int64_t *__restrict__ a_idx,
                    const float *__restrict__ W, float *__restrict__ Y,
                    const int64_t *__restrict__ c_idx, const int64_t *__restrict__ seg_ptrs, int num_segs,
                    int64_t M, int H, int K, int N, int bmode, int accumulate, int passes, int probe, int y_bf16) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SW128 atoms need 1024 B alignment
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t epi_base = base + kOffEpi;
  const uint32_t bar_base = epi_base + kEpiBytes;
  // barriers: full[2], empty[2], tmem_full[2], tmem_empty[2], then the TMEM base address
  auto full_bar = [&](int i) { return bar_base + 8u * i; };
  auto empty_bar = [&](int i) { return bar_base + 16u + 8u * i; };
  auto tfull_bar = [&](int i) { return bar_base + 32u + 8u * i; };
  auto tempty_bar = [&](int i) { return bar_base + 48u + 8u * i; };
  const uint32_t tmem_slot = bar_base + 64u;
  volatile uint32_t *tmem_slot_g = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kEpiWarps) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        mbar_init(full_bar(i), kProdWarps);
        mbar_init(empty_bar(i), 1);
        mbar_init(tfull_bar(i), 1);
        mbar_init(tempty_bar(i), kEpiWarps);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 2 * NN);  // two 64-column accumulator stages
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_g;

  const int64_t num_tiles = ceil_div64(M, TM);
  const int64_t tile_begin = num_tiles * blockIdx.x / gridDim.x;
  const int64_t tile_end = num_tiles * (blockIdx.x + 1) / gridDim.x;
  PieceIter it(seg_ptrs, num_segs, M, tile_begin, tile_end);
  const int64_t row_base = tile_begin * TM;
  int tile_row0, p0, p1;  // relative to row_base
  int seg;

  if (warp > kEpiWarps) {
    // ===================== producers =====================
    // The gather is latency bound (random 256 B rows), so what matters is bytes in
    // flight per SM. The producer threads keep kSets work items of row data in
    // registers: right after item n has been split and stored, its register set is
    // re-used for the row loads of item n+kSets, so every load has kSets pipeline steps
    // to land; gather indices are fetched a further kSets items ahead. The loop is
    // unrolled kSets ways so the sets rotate by name (a register copy of an in-flight
    // load would stall on it).
    const int p = threadIdx.x - (kEpiWarps + 1) * 32;
    const int chunk = p & 15;                          // 16-byte chunk of the 256 B row
    const int panel = chunk >> 3, kc = chunk & 7;
    const int r_in = p >> 4;                           // rows r_in + kRowStep*i, i < RP
    constexpr int kRowStep = kProdThreads / 16;
    constexpr int RP = TM / kRowStep;                  // rows per thread
    int cur_seg = -1, bsel = 1;
    uint32_t n = 0;
    PieceIter it_ix = it;  // runs kSets..2*kSets items ahead of `it`

    // Predicated loads that write the index register directly: a branch per slot costs more
    // than the load, and anything computed from the loaded value here would stall this step for
    // a full memory latency.
    auto issue_idx = [&](int *ix) {
      int t0, a0, a1, sg;
      const bool valid = it_ix.next(t0, a0, a1, sg);
      const int lim = valid ? it_ix.n : 0;
#pragma unroll
      for (int i = 0; i < RP; ++i) {
        const int rr = t0 + r_in + kRowStep * i;
        const bool ok = valid && rr < lim;
        const int64_t gr = row_base + rr;
        // a row index fits 31 bits: 2^31 rows of 256 B would be 550 GB
        int v = ok ? (int)gr : -1;
        if (ok && a_idx) v = (int)__ldg(a_idx + gr);
        ix[i] = v;
      }
    };
    auto issue_rows = [&](const int *ix, float4 *v) {
#pragma unroll
      for (int i = 0; i < RP; ++i) {
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ix[i] >= 0 && !(probe & 2)) v[i] = ldg4(A + (int64_t)ix[i] * KK + chunk * 4);
      }
    };
    // one pipeline step: split + store item n from rS, then re-use rS / ixS for the
    // row loads of item n+kSets and the index loads of item n+2*kSets.
    auto step = [&](float4 *rS, int *ixS) -> bool {
      if (!it.next(tile_row0, p0, p1, seg)) return false;
      const int stage = n & 1;
      const uint32_t phase = (n >> 1) & 1;
      mbar_wait(empty_bar(stage), phase ^ 1);
      uint8_t *gs = gbase + stage * kStageBytes;
#pragma unroll
      for (int i = 0; i < RP; ++i) {
        const int r = r_in + kRowStep * i;
        const uint32_t off = (uint32_t)(panel * kPanelA + r * 128 + ((kc ^ (r & 7)) << 4));
        const float4 x = rS[i];
        const float4 hi = make_float4(tf32_hi(x.x), tf32_hi(x.y), tf32_hi(x.z), tf32_hi(x.w));
        const float4 lo = make_float4(x.x - hi.x, x.y - hi.y, x.z - hi.z, x.w - hi.w);
        if (!(probe & 8)) {
          *reinterpret_cast<float4 *>(gs + off) = hi;
          *reinterpret_cast<float4 *>(gs + kOffAlo + off) = lo;
        }
      }
      issue_rows(ixS, rS);  // item n + kSets (its indices arrived kSets steps ago)
      issue_idx(ixS);       // item n + 2*kSets
      // ---- B_s: transposed ([n][k], K-major) into the B ring, once per segment change.
      // The buffer being replaced was last read by MMAs two segment runs back, which
      // retired before empty[stage] completed (tcgen05.commit covers all earlier MMAs).
      if (seg != cur_seg) {
        cur_seg = seg;
        bsel ^= 1;
        uint8_t *gb = gbase + kOffB + bsel * kBBufBytes;
        const float *Ws = W + (int64_t)seg * H * K * N;
        for (int e = p; e < KK * NN / 4; e += kProdThreads) {
          const int kk = e >> 4, nn = (e & 15) << 2;
          float4 wv;
          if (bmode == kStacked) {
            wv = ldg4(Ws + (int64_t)kk * N + nn);
          } else {
            const int h = nn / N;
            const bool live = (bmode == kDense) || (kk / K == h);
            const int kr = (bmode == kDense) ? kk : kk % K;
            wv = live ? ldg4(Ws + ((int64_t)h * K + kr) * N + (nn % N)) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
          const float w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const float hi = tf32_hi(w4[t]);
            const uint32_t off = sw128_off(NN, nn + t, kk);
            *reinterpret_cast<float *>(gb + off) = hi;
            *reinterpret_cast<float *>(gb + kOffBlo + off) = w4[t] - hi;
          }
        }
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar(stage));  // one arrival per warp: 256 per stage cost more than the split
      ++n;
      return true;
    };

    static_assert(kSets == 2 || kSets == 3, "producer register sets");
    int ix0[RP], ix1[RP], ix2[RP];
    float4 r0[RP], r1[RP], r2[RP];
    issue_idx(ix0);
    issue_idx(ix1);
    if (kSets == 3) issue_idx(ix2);
    issue_rows(ix0, r0);
    issue_idx(ix0);
    issue_rows(ix1, r1);
    issue_idx(ix1);
    if (kSets == 3) {
      issue_rows(ix2, r2);
      issue_idx(ix2);
    }
    for (;;) {
      if (!step(r0, ix0)) break;
      if (!step(r1, ix1)) break;
      if (kSets == 3 && !step(r2, ix2)) break;
    }
  } else if (warp == kEpiWarps) {
    // ===================== MMA issuer =====================
    uint32_t n = 0;
    int cur_seg = -1, bsel = 1;
    while (it.next(tile_row0, p0, p1, seg)) {
      const int stage = n & 1;
      const uint32_t phase = (n >> 1) & 1;
      if (seg != cur_seg) { cur_seg = seg; bsel ^= 1; }
      mbar_wait(tempty_bar(stage), phase ^ 1);  // accumulator stage drained by the epilogue
      mbar_wait(full_bar(stage), phase);        // operands landed
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sbase = base + stage * kStageBytes;
        const uint32_t bbase = base + kOffB + bsel * kBBufBytes;
        const uint32_t d = tmem_base + (uint32_t)(stage * NN);
        uint32_t acc = 0;
#pragma unroll
        for (int j = 0; j < KK / 8; ++j) {
          const uint32_t ka = (uint32_t)((j >> 2) * kPanelA + (j & 3) * 32);
          const uint32_t kb = (uint32_t)((j >> 2) * kPanelB + (j & 3) * 32);
          const uint64_t a_hi = make_desc_sw128(sbase + ka), a_lo = make_desc_sw128(sbase + kOffAlo + ka);
          const uint64_t b_hi = make_desc_sw128(bbase + kb), b_lo = make_desc_sw128(bbase + kOffBlo + kb);
          if (passes == 3) {
            umma_tf32(d, a_lo, b_hi, kIdesc, acc);
            umma_tf32(d, a_hi, b_lo, kIdesc, 1);
            umma_tf32(d, a_hi, b_hi, kIdesc, 1);
          } else if (passes == 1) {
            umma_tf32(d, a_hi, b_hi, kIdesc, acc);
          }
          acc = 1;
        }
        umma_commit(empty_bar(stage));  // operands may be overwritten once these MMAs retire
        umma_commit(tfull_bar(stage));  // accumulator ready for the epilogue
      }
      __syncwarp();
      ++n;
    }
  } else {
    // ===================== epilogue =====================
    uint8_t *stg = gbase + kOffEpi + warp * (32 * NN * 4);  // this warp's 32 rows x 256 B
    uint32_t n = 0;
    while (it.next(tile_row0, p0, p1, seg)) {
      const int stage = n & 1;
      const uint32_t phase = (n >> 1) & 1;
      // destination row of this lane's tile row (lane j <-> row warp*32 + j), fetched early
      const int my_rr = tile_row0 + warp * 32 + lane;
      const int64_t my_gr = row_base + my_rr;
      int64_t my_cr = my_gr;
      if (c_idx && my_rr >= p0 && my_rr < p1) my_cr = __ldg(c_idx + my_gr);
      mbar_wait(tfull_bar(stage), phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(stage * NN);
      // registers (thread = row) -> swizzled staging, 32 columns at a time
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t v[32];
        tmem_ld32(taddr + 32 * half, v);
        tmem_ld_wait();
        if (half == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(stage));  // TMEM stage free: the next MMA may start
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
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int lr = i * 2 + (lane >> 4);  // row within this warp's 32
        const int rr = tile_row0 + warp * 32 + lr;
        const int64_t cr = __shfl_sync(0xffffffffu, my_cr, lr);
        if (rr >= p0 && rr < p1 && !(probe & 4)) {
          const int pc = (c & 8) | ((c & 7) ^ (lr & 7));
          const float4 val = *reinterpret_cast<const float4 *>(stg + lr * 256 + pc * 16);
          if (y_bf16) {  // bf16 storage of a table the edge passes gather (fp32 accumulate up to here)
            const __nv_bfloat162 lo2 = __floats2bfloat162_rn(val.x, val.y), hi2 = __floats2bfloat162_rn(val.z, val.w);
            uint2 packed;
            packed.x = *reinterpret_cast<const uint32_t *>(&lo2);
            packed.y = *reinterpret_cast<const uint32_t *>(&hi2);
            *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(Y) + cr * NN + c * 4) = packed;
          } else {
            float *dst = Y + cr * NN + c * 4;
            if (accumulate) red_add4(dst, val);
            else st4(dst, val);
          }
        }
      }
      __syncwarp();
      ++n;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * NN);
  }
}

}  // namespace tc
}  // namespace hetb

using namespace hetb;

// defined in segment_mm_tc_tma.cu: the same item scheme with the A tile delivered by TMA
int hetb_segment_mm_tc_tma(const float *A, const int64_t *a_idx, int a_heads, const float *W, float *Y,
                           const int64_t *c_idx, int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                           int64_t num_rows, int H, int K, int N, int accumulate, int y_bf16, void *stream);

static int g_segmm_tma = 1;  // 1: TMA-fed kernel first (falls through when unsupported), 0: register path only

extern "C" int hetb_set_segment_mm_tma(int on) {
  if (on != 0 && on != 1) return HETB_EINVAL;
  g_segmm_tma = on;
  return HETB_OK;
}

// Returns HETB_EUNSUPPORTED when the shape is outside the tensor-core path; the caller
// (hetb_segment_mm) then uses the SIMT kernels.
int hetb_segment_mm_tc(const float *A, const int64_t *a_idx, int a_heads, const float *W, float *Y,
                       const int64_t *c_idx, int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                       int64_t num_rows, int H, int K, int N, int accumulate, int passes, void *stream) {
  int bmode;
  const int y_bf16 = (passes >> 12) & 1;  // bit 12: Y is __nv_bfloat16 (hetb_segment_mm_bf16out)
  passes &= 0xfff;
  if (y_bf16 && accumulate) return HETB_EUNSUPPORTED;
  const int probe = passes >> 4;
  passes &= 15;
  // (a plain `Y += A.B` on contiguous rows is the one shape the register path measures faster: 0.246 vs 0.256 ms
  //  at 1.94 M rows, both at 6 TB/s — profiles/r02r_segmm_tma_ab.log)
  if (g_segmm_tma && passes == 3 && probe == 0 && !(accumulate && !c_idx && !a_idx)) {
    const int rc = hetb_segment_mm_tc_tma(A, a_idx, a_heads, W, Y, c_idx, c_heads, seg_ptrs, num_segs, num_rows, H, K,
                                          N, accumulate, y_bf16, stream);
    if (rc != HETB_EUNSUPPORTED) return rc;
  }
  if (a_heads == H && c_heads == H && H * K == tc::KK && H * N == tc::NN)
    bmode = (H == 1) ? tc::kDense : tc::kBlockDiag;
  else if (a_heads == 1 && c_heads == H && K == tc::KK && H * N == tc::NN)
    bmode = tc::kDense;
  else if (a_heads == H && c_heads == 1 && accumulate && H * K == tc::KK && N == tc::NN)
    bmode = tc::kStacked;
  else
    return HETB_EUNSUPPORTED;
  if (num_segs > INT32_MAX || (N % 4) != 0) return HETB_EUNSUPPORTED;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(tc::segmm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)tc::kSmemBytes);
    if (ce != cudaSuccess) return (int)ce;
    attr_set = true;
  }
  const int64_t tiles = ceil_div64(num_rows, tc::TM);
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  tc::segmm_tc_kernel<<<grid, tc::kThreads, tc::kSmemBytes, (cudaStream_t)stream>>>(
      A, a_idx, W, Y, c_idx, seg_ptrs, (int)num_segs, num_rows, H, K, N, bmode, accumulate, passes, probe, y_bf16);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
