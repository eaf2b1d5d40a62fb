// Grouped segment MM on tcgen05 + TMEM with the A operand delivered by TMA (sm_100a).
//
//   Y[c_idx?[i], :] (=|+=) A[a_idx?[i], :] · B_s        rows i of segment s,  64-wide contraction and output
//
// Same maths, work items and B_s staging as segment_mm_tc.cu (3xTF32, fp32 accumulate in TMEM); what
// changes is how the 128 x 64 fp32 tile of A reaches shared memory. There, eight producer warps load the
// rows into registers, split them into tf32 hi / lo and store both; the bytes in flight per SM are bounded
// by the registers that hold them. Here four warps (one per 128-byte K panel and 64-row half of the tile)
// issue TMA copies and no register ever holds a row:
//
//   * contiguous rows (a_idx == NULL: the `..._no_scatter_gather_list` ops, type-sorted node tables):
//     one `cp.async.bulk.tensor.2d` tile copy per item and panel (box 32 floats x 128 rows), rows past
//     the end of the table zero-filled by the tensor map;
//   * gathered rows (a_idx given: `rgnn_relational_matmul` and the compact-as-of-node products):
//     `cp.async.bulk.tensor.2d ... tile::gather4` — four row indices per instruction, each lane of the
//     warp owns four consecutive tile rows, 16 instructions per item and warp. The row indices of the
//     next item are fetched while the copies of this one are issued.
//
// Both land in the 128B-swizzled K-major layout the UMMA descriptors of segment_mm_tc.cu describe
// (CU_TENSOR_MAP_SWIZZLE_128B == chunk ^ (row & 7) on 1024-byte aligned panels), and the raw fp32 tile IS
// the tf32 "hi" operand: the tensor core ignores the low 13 mantissa bits (segment_mm_tc_wide.cu relies on
// the same fact). Seven warps read the landed tile from shared memory and write only lo = x - trunc(x).
// Three raw stages (96 KB of row data in flight per SM, against 64 KB of registers before) + two lo
// stages + ONE B_s buffer (re-staged at a segment change after the MMAs that read it retired) + the
// epilogue staging = 224 KB.
//
// Epilogue as in segment_mm_tc.cu; a plain store of whole 32-row groups (c_idx == NULL, no accumulation,
// fp32) leaves through TMA as well (`cp.async.bulk.tensor.2d.global.shared::cta`, one 32 x 32-float box per
// K panel and warp, clipped at the end of the table by the tensor map).
//
// Reference for the op: RGNN/my_shmem_sgemm_func.cu.h:504-531 (HET_RGNNFeatPerEdgeFwProp), launched by
// OpExport/RGNNOps.inc.h:150-169 with a host block schedule; no counterpart of the data path.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_items.cuh"
#include "tc_ptx.cuh"

namespace hetb {
namespace tct {

using namespace tc;

constexpr int TM = kTileRows;
constexpr int KK = 64;
constexpr int NN = 64;
constexpr int kEpiWarps = 4;
constexpr int kMmaWarp = kEpiWarps;       // warp 4
constexpr int kTmaWarp0 = kEpiWarps + 1;  // TMA warps: (K panel, row group) each
#ifndef HETB_TMA_WARPS
#define HETB_TMA_WARPS 4
#endif
constexpr int kTmaWarps = HETB_TMA_WARPS;
constexpr int kRowGroups = kTmaWarps / 2;          // row groups of a tile, one TMA warp per (panel, group)
constexpr int kGroupRows = TM / kRowGroups;        // 64 rows: 16 lanes x 4 rows of a gather4
constexpr int kProdWarp0 = kTmaWarp0 + kTmaWarps;
constexpr int kProdWarps = 16 - kProdWarp0;        // 16 warps in all (4 per scheduler): 128 registers per thread
constexpr int kProdThreads = kProdWarps * 32;
constexpr int kThreads = (kEpiWarps + 1 + kTmaWarps + kProdWarps) * 32;  // 512: 4 + 1 + 4 + 7 warps
static_assert(kTmaWarps >= 2 && (kTmaWarps & 1) == 0 && TM % kRowGroups == 0 && kGroupRows % 4 == 0, "TMA warp layout");
constexpr int kRawStages = 3;
constexpr int kLoStages = 2;

constexpr uint32_t kPanelA = TM * 128;     // one 32-float K panel of A: 16 KB
constexpr uint32_t kPanelB = NN * 128;     // 8 KB
constexpr uint32_t kBytesA = 2 * kPanelA;  // 32 KB
constexpr uint32_t kBytesB = 2 * kPanelB;  // 16 KB
constexpr uint32_t kOffLo = kRawStages * kBytesA;            // 96 KB
constexpr uint32_t kOffB = kOffLo + kLoStages * kBytesA;     // 160 KB: B_hi, then B_lo
constexpr uint32_t kOffBlo = kBytesB;
constexpr uint32_t kOffEpi = kOffB + 2 * kBytesB;            // 192 KB
constexpr uint32_t kEpiBytes = TM * NN * 4;                  // 32 KB: per warp 2 panels x 32 rows x 128 B
constexpr uint32_t kSmemBytes = kOffEpi + kEpiBytes + 1024 /*align*/ + 256 /*barriers*/;
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

enum BMode : int { kDense = 0, kBlockDiag = 1, kStacked = 2 };

// ------------------------------------------------------------------ PTX ----
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_tile(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_gather4(uint32_t dst, const CUtensorMap *tm, int c0, int r0, int r1, int r2,
                                                 int r3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}
__device__ __forceinline__ void tma_store_tile(const CUtensorMap *tm, int c0, int c1, uint32_t src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap *tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}

// ring position of item n in a ring of S stages
template <int S>
struct Ring {
  uint32_t stage = 0, phase = 0;
  __device__ __forceinline__ void advance() {
    if (++stage == S) {
      stage = 0;
      phase ^= 1;
    }
  }
};

// --------------------------------------------------------------- kernel ----
template <bool GATHER>
__global__ void __launch_bounds__(kThreads, 1)
    segmm_tc_tma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmY,
                        const int64_t *__restrict__ a_idx, const float *__restrict__ W, float *__restrict__ Y,
                        const int64_t *__restrict__ c_idx, const int64_t *__restrict__ seg_ptrs, int num_segs,
                        int64_t M, int H, int K, int N, int bmode, int accumulate, int y_bf16, int tma_store) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SW128 atoms need 1024 B alignment
  uint8_t *gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_base = base + kOffEpi + kEpiBytes;
  auto raw_full = [&](int i) { return bar_base + 8u * i; };
  auto raw_empty = [&](int i) { return bar_base + 8u * (kRawStages + i); };
  auto lo_full = [&](int i) { return bar_base + 8u * (2 * kRawStages + i); };
  auto lo_empty = [&](int i) { return bar_base + 8u * (2 * kRawStages + kLoStages + i); };
  auto tfull_bar = [&](int i) { return bar_base + 8u * (2 * kRawStages + 2 * kLoStages + i); };
  auto tempty_bar = [&](int i) { return bar_base + 8u * (2 * kRawStages + 2 * kLoStages + 2 + i); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kRawStages + 2 * kLoStages + 4);
  volatile uint32_t *tmem_slot_g = reinterpret_cast<volatile uint32_t *>(gbase + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kMmaWarp) {
    if (lane == 0) {
      for (int i = 0; i < kRawStages; ++i) {
        mbar_init(raw_full(i), kTmaWarps);  // each TMA warp's arrive.expect_tx; the copies complete the bytes
        mbar_init(raw_empty(i), 1);  // tcgen05.commit
      }
      for (int i = 0; i < kLoStages; ++i) {
        mbar_init(lo_full(i), kProdWarps);
        mbar_init(lo_empty(i), 1);
        mbar_init(tfull_bar(i), 1);
        mbar_init(tempty_bar(i), kEpiWarps);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 2 * NN);  // two 64-column accumulator stages
  } else if (warp == kTmaWarp0 && lane == 0) {
    prefetch_tensormap(&tmA);
    if (tma_store) prefetch_tensormap(&tmY);
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
  int tile_row0, p0, p1, seg;  // relative to row_base

  if (warp >= kTmaWarp0 && warp < kProdWarp0) {
    // ===================== TMA issuers =====================
    // One warp per (K panel, group of kGroupRows tile rows). A gather4 has per-lane operands, which the
    // hardware issues lane by lane (an ELECT loop of R2UR + UTMALDG in SASS, measured ~130 clocks per
    // instruction): with two warps issuing 32 each, an item took 1.65 us to issue against the 2.2 us it had,
    // and every other role waited on the copies (profiles/r02r_segmm_gather_ncu_summary.txt). Four warps
    // issuing 16 each and eight issuing 8 each measure the same (0.39 ms for 4.46 M gathered rows against
    // 0.51 ms); four leave seven warps for the lo tile, which the contiguous case needs (0.170 vs 0.180 ms
    // for 1.94 M rows with three).
    // Rows of a tile that belong to another segment's piece are loaded too (their products are not
    // stored), rows past M repeat row M-1 of the gather list / are zero-filled by the tensor map: every
    // item moves exactly kGroupRows x 128 bytes per warp.
    const uint32_t panel = (uint32_t)(warp - kTmaWarp0) & 1u;
    const int group_row0 = ((warp - kTmaWarp0) >> 1) * kGroupRows;
    const bool issuer = lane < kGroupRows / 4;
    Ring<kRawStages> raw;
    PieceIter it_ix = it;
    int nx[4] = {0, 0, 0, 0};
    auto load_idx = [&]() {
      int t0, a0, a1, sg;
      if (GATHER && it_ix.next(t0, a0, a1, sg) && issuer) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          int64_t gr = row_base + t0 + group_row0 + 4 * lane + j;
          gr = gr < M ? gr : M - 1;
          nx[j] = (int)__ldg(a_idx + gr);  // a row index fits 31 bits: 2^31 rows of 256 B would be 550 GB
        }
      }
    };
    load_idx();
    while (it.next(tile_row0, p0, p1, seg)) {
      const int r0 = nx[0], r1 = nx[1], r2 = nx[2], r3 = nx[3];
      load_idx();  // next item's indices travel while this item's copies are issued
      mbar_wait(raw_empty(raw.stage), raw.phase ^ 1);
      const uint32_t bar = raw_full(raw.stage);
      const uint32_t dst = base + raw.stage * kBytesA + panel * kPanelA + (uint32_t)group_row0 * 128u;
      if (GATHER) {
        if (lane == 0) mbar_arrive_expect_tx(bar, kGroupRows * 128u);
        __syncwarp();
        if (issuer) tma_load_gather4(dst + lane * 512u, &tmA, (int)panel * 32, r0, r1, r2, r3, bar);
      } else if (lane == 0) {  // contiguous rows: one whole-panel copy from the first row group's warp
        if (group_row0 == 0) {
          mbar_arrive_expect_tx(bar, kPanelA);
          tma_load_tile(dst, &tmA, (int)panel * 32, (int)(row_base + tile_row0), bar);
        } else {
          mbar_arrive(bar);
        }
      }
      raw.advance();
    }
  } else if (warp >= kProdWarp0) {
    // ===================== lo / B_s producers =====================
    const int p = threadIdx.x - kProdWarp0 * 32;
    Ring<kRawStages> raw;
    uint32_t n = 0;
    int cur_seg = -1;
    while (it.next(tile_row0, p0, p1, seg)) {
      const uint32_t ls = n & 1, lph = (n >> 1) & 1;
      if (seg != cur_seg) {
        // One B_s buffer: the MMAs of item n-1 (whose commit covers all earlier ones) must have retired.
        if (n > 0) mbar_wait(lo_empty((n - 1) & 1), ((n - 1) >> 1) & 1);
        cur_seg = seg;
        uint8_t *gb = gbase + kOffB;
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
      mbar_wait(lo_empty(ls), lph ^ 1);
      mbar_wait(raw_full(raw.stage), raw.phase);
      // lo[off] = x[off] - trunc_tf32(x[off]) at the same (swizzled) offset: no address arithmetic
      const uint8_t *src = gbase + raw.stage * kBytesA;
      uint8_t *dst = gbase + kOffLo + ls * kBytesA;
#pragma unroll 4
      for (uint32_t off = (uint32_t)p * 16u; off < kBytesA; off += kProdThreads * 16u) {
        const float4 x = *reinterpret_cast<const float4 *>(src + off);
        *reinterpret_cast<float4 *>(dst + off) =
            make_float4(x.x - tf32_hi(x.x), x.y - tf32_hi(x.y), x.z - tf32_hi(x.z), x.w - tf32_hi(x.w));
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) mbar_arrive(lo_full(ls));
      raw.advance();
      ++n;
    }
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer =====================
    Ring<kRawStages> raw;
    uint32_t n = 0;
    while (it.next(tile_row0, p0, p1, seg)) {
      const uint32_t stage = n & 1, phase = (n >> 1) & 1;
      mbar_wait(tempty_bar(stage), phase ^ 1);     // accumulator stage drained by the epilogue
      mbar_wait(raw_full(raw.stage), raw.phase);   // hi (raw) tile landed
      mbar_wait(lo_full(stage), phase);            // lo tile and B_s written
      tc_fence_after();
      if (lane == 0) {
        const uint32_t hbase = base + raw.stage * kBytesA;
        const uint32_t lbase = base + kOffLo + stage * kBytesA;
        const uint32_t bbase = base + kOffB;
        const uint32_t d = tmem_base + (uint32_t)(stage * NN);
        uint32_t acc = 0;
#pragma unroll
        for (int j = 0; j < KK / 8; ++j) {
          const uint32_t ka = (uint32_t)((j >> 2) * kPanelA + (j & 3) * 32);
          const uint32_t kb = (uint32_t)((j >> 2) * kPanelB + (j & 3) * 32);
          const uint64_t a_hi = make_desc_sw128(hbase + ka), a_lo = make_desc_sw128(lbase + ka);
          const uint64_t b_hi = make_desc_sw128(bbase + kb), b_lo = make_desc_sw128(bbase + kOffBlo + kb);
          umma_tf32(d, a_lo, b_hi, kIdesc, acc);
          umma_tf32(d, a_hi, b_lo, kIdesc, 1);
          umma_tf32(d, a_hi, b_hi, kIdesc, 1);
          acc = 1;
        }
        umma_commit(raw_empty(raw.stage));  // the raw stage may be overwritten by the next copy
        umma_commit(lo_empty(stage));       // the lo stage (and, at a segment change, B_s) may be rewritten
        umma_commit(tfull_bar(stage));      // accumulator ready for the epilogue
      }
      __syncwarp();
      raw.advance();
      ++n;
    }
  } else {
    // ===================== epilogue =====================
    // staging of this warp: 2 K panels x 32 rows x 128 B, 16-byte chunks at chunk ^ (row & 7): the layout a
    // SWIZZLE_128B tensor map with a 32-float x 32-row box reads, and conflict-free for both passes below
    uint8_t *stg = gbase + kOffEpi + warp * 8192;
    const uint32_t stg_s = base + kOffEpi + warp * 8192;
    const int n_rows = it.n;
    uint32_t n = 0;
    bool store_pending = false;
    while (it.next(tile_row0, p0, p1, seg)) {
      const int stage = n & 1;
      const uint32_t phase = (n >> 1) & 1;
      const int w_r0 = tile_row0 + warp * 32;          // this warp's 32 tile rows
      const int w_r1 = min(w_r0 + 32, n_rows);
      const bool bulk = tma_store && w_r0 >= p0 && w_r1 <= p1 && w_r0 < w_r1;
      const int my_rr = w_r0 + lane;
      const int64_t my_gr = row_base + my_rr;
      int64_t my_cr = my_gr;
      if (c_idx && my_rr >= p0 && my_rr < p1) my_cr = __ldg(c_idx + my_gr);
      mbar_wait(tfull_bar(stage), phase);
      tc_fence_after();
      if (store_pending) {  // the previous bulk store must have read the staging before it is rewritten
        if (lane == 0) tma_store_wait_read();
        __syncwarp();
        store_pending = false;
      }
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(stage * NN);
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
        for (int c8 = 0; c8 < 8; ++c8)
          *reinterpret_cast<uint4 *>(stg + half * 4096 + lane * 128 + ((c8 ^ (lane & 7)) << 4)) =
              make_uint4(v[4 * c8], v[4 * c8 + 1], v[4 * c8 + 2], v[4 * c8 + 3]);
      }
      if (bulk) {
        fence_proxy_async();  // staging stores -> visible to the TMA engine
        __syncwarp();
        if (lane == 0) {
          const int row = (int)(row_base + w_r0);
          tma_store_tile(&tmY, 0, row, stg_s);
          tma_store_tile(&tmY, 32, row, stg_s + 4096);
          tma_store_commit();
        }
        store_pending = true;
      } else {
        __syncwarp();
        const int c = lane & 15;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int lr = i * 2 + (lane >> 4);  // row within this warp's 32
          const int rr = w_r0 + lr;
          const int64_t cr = __shfl_sync(0xffffffffu, my_cr, lr);
          if (rr >= p0 && rr < p1) {
            const float4 val = *reinterpret_cast<const float4 *>(stg + (c >> 3) * 4096 + lr * 128 + (((c & 7) ^ (lr & 7)) << 4));
            if (y_bf16) {  // bf16 storage of a table the edge passes gather (fp32 accumulate up to here)
              const __nv_bfloat162 lo2 = __floats2bfloat162_rn(val.x, val.y), hi2 = __floats2bfloat162_rn(val.z, val.w);
              uint2 packed;
              packed.x = *reinterpret_cast<const uint32_t *>(&lo2);
              packed.y = *reinterpret_cast<const uint32_t *>(&hi2);
              *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(Y) + cr * NN + c * 4) = packed;
            } else {
              float *dstp = Y + cr * NN + c * 4;
              if (accumulate) red_add4(dstp, val);
              else st4(dstp, val);
            }
          }
        }
        __syncwarp();
      }
      ++n;
    }
    if (store_pending && lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * NN);
  }
}

// ------------------------------------------------------------ host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point query: libhetb200.so links cudart only.
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    else
      (void)cudaGetLastError();
  }
  return fn;
}

// [rows, 64] fp32 row-major table, box = 32 floats x box_rows, 128-byte swizzle, zero fill out of bounds
static bool encode_rows64(CUtensorMap *tm, const void *ptr, uint64_t rows, uint32_t box_rows) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)KK, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)KK * 4};
  const cuuint32_t box[2] = {32u, box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(ptr), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace tct
}  // namespace hetb

using namespace hetb;

// Returns HETB_EUNSUPPORTED when the shape is outside this kernel or the driver cannot encode tensor maps;
// hetb_segment_mm_tc then runs the register-path kernel of segment_mm_tc.cu.
int hetb_segment_mm_tc_tma(const float *A, const int64_t *a_idx, int a_heads, const float *W, float *Y,
                           const int64_t *c_idx, int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                           int64_t num_rows, int H, int K, int N, int accumulate, int y_bf16, void *stream) {
  using namespace tct;
  int bmode;
  if (a_heads == H && c_heads == H && H * K == KK && H * N == NN)
    bmode = (H == 1) ? kDense : kBlockDiag;
  else if (a_heads == 1 && c_heads == H && K == KK && H * N == NN)
    bmode = kDense;
  else if (a_heads == H && c_heads == 1 && accumulate && H * K == KK && N == NN)
    bmode = kStacked;
  else
    return HETB_EUNSUPPORTED;
  if (num_segs > INT32_MAX || (N % 4) != 0 || num_rows >= INT32_MAX) return HETB_EUNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(A) & 15) != 0 || (reinterpret_cast<uintptr_t>(Y) & 15) != 0) return HETB_EUNSUPPORTED;
  CUtensorMap tmA, tmY;
  // gathered: one row per box, four boxes per gather4; the table's row count is not part of the op's
  // arguments, so the map spans every 31-bit row index (the gather list holds only valid rows)
  if (!(a_idx ? encode_rows64(&tmA, A, (uint64_t)INT32_MAX, 1) : encode_rows64(&tmA, A, (uint64_t)num_rows, TM)))
    return HETB_EUNSUPPORTED;
  const int tma_store = (!c_idx && !accumulate && !y_bf16) ? 1 : 0;
  if (!encode_rows64(&tmY, tma_store ? (const void *)Y : (const void *)A, (uint64_t)num_rows, 32)) return HETB_EUNSUPPORTED;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(segmm_tc_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)kSmemBytes);
    if (ce == cudaSuccess)
      ce = cudaFuncSetAttribute(segmm_tc_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (ce != cudaSuccess) return (int)ce;
    attr_set = true;
  }
  const int64_t tiles = ceil_div64(num_rows, TM);
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  cudaStream_t st = (cudaStream_t)stream;
  if (a_idx)
    segmm_tc_tma_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(tmA, tmY, a_idx, W, Y, c_idx, seg_ptrs, (int)num_segs,
                                                                   num_rows, H, K, N, bmode, accumulate, y_bf16, tma_store);
  else
    segmm_tc_tma_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(tmA, tmY, a_idx, W, Y, c_idx, seg_ptrs, (int)num_segs,
                                                                    num_rows, H, K, N, bmode, accumulate, y_bf16, tma_store);
  HETB_LAUNCH_OK();
  return HETB_OK;
}
