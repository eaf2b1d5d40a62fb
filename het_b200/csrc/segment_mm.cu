// Relation-typed linear layer: grouped segment MM with optional row gather on
// the input and row scatter on the output, and its weight-gradient contraction.
// fp32 in, fp32 accumulate (parity mode: matches the reference's fp32 SIMT
// SGEMM to rounding). Segments are walked on a GLOBAL row grid: CTA t owns rows
// [t*TM, (t+1)*TM) and handles every segment piece that intersects them, so
// there is no per-call host schedule, no device->host copy of rel_ptrs and no
// per-call allocation (the reference does all three,
// OpExport/RGNNOps.inc.h:150-169), and ragged / tiny segments cost nothing
// beyond one extra weight-tile load.
#include "common.cuh"

namespace hetb {

constexpr int kMMThreads = 256;

// ---------------------------------------------------------------------------
// Fully general fallback: one thread per output element. Correct for every
// (H, K, N, a_heads, c_heads); used only for shapes the tiled paths decline.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kMMThreads)
    segmm_generic_kernel(const float *__restrict__ A,
                         const int64_t *__restrict__ a_idx, int a_heads,
                         const float *__restrict__ W, float *__restrict__ Y,
                         const int64_t *__restrict__ c_idx, int c_heads,
                         const int64_t *__restrict__ seg_ptrs, int num_segs,
                         int64_t M, int H, int K, int N, int accumulate) {
  const int64_t C = (int64_t)H * N;
  const int64_t total = M * C;
  const int64_t stride = (int64_t)gridDim.x * kMMThreads;
  for (int64_t t = (int64_t)blockIdx.x * kMMThreads + threadIdx.x; t < total;
       t += stride) {
    const int64_t i = t / C;
    const int c = (int)(t % C);
    const int h = c / N, n = c % N;
    const int s = segment_of(seg_ptrs, num_segs, i);
    const int64_t ar = a_idx ? __ldg(a_idx + i) : i;
    const int64_t cr = c_idx ? __ldg(c_idx + i) : i;
    const float *a = A + (ar * a_heads + (a_heads == 1 ? 0 : h)) * K;
    const float *w = W + (((int64_t)s * H + h) * K) * N + n;
    float acc = 0.f;
    for (int k = 0; k < K; ++k) acc = fmaf(__ldg(a + k), __ldg(w + (int64_t)k * N), acc);
    float *y = Y + (cr * c_heads + (c_heads == 1 ? 0 : h)) * N + n;
    if (accumulate)
      atomicAdd(y, acc);
    else
      *y = acc;
  }
}

// ---------------------------------------------------------------------------
// Tiled SIMT path. CTA tile: TM rows x CW output columns (columns live in the
// flattened [H*N] space; a 4-column micro tile never straddles a head because
// N % 4 == 0). A rows are staged row-major in shared memory with 128-bit
// copies; each thread owns RPT rows x 4 columns and walks k four at a time
// with 128-bit shared loads (broadcast for A, contiguous for W).
// ---------------------------------------------------------------------------
template <int CW, int TM>
__global__ void __launch_bounds__(kMMThreads)
    segmm_tiled_kernel(const float *__restrict__ A,
                       const int64_t *__restrict__ a_idx, int a_heads,
                       const float *__restrict__ W, float *__restrict__ Y,
                       const int64_t *__restrict__ c_idx, int c_heads,
                       const int64_t *__restrict__ seg_ptrs, int num_segs,
                       int64_t M, int H, int K, int N, int accumulate) {
  constexpr int TX = CW / 4;               // threads across columns
  constexpr int TY = kMMThreads / TX;      // threads across rows
  constexpr int RPT = TM / TY;             // rows per thread
  static_assert(TM % TY == 0, "tile rows must divide");
  extern __shared__ __align__(16) float smem[];
  const int AK = a_heads * K;              // floats per staged A row
  const int AS = AK + 4;                   // padded row stride (keeps 16 B alignment)
  float *As = smem;                        // [TM][AS]
  float *Ws = smem + (size_t)TM * AS;      // [K][CW]

  const int tx = threadIdx.x % TX;
  const int ty = threadIdx.x / TX;
  const int64_t row0 = (int64_t)blockIdx.x * TM;
  const int64_t row1 = (row0 + TM < M) ? row0 + TM : M;
  const int col0 = blockIdx.y * CW;        // first flattened output column
  const int c = col0 + tx * 4;
  const int h = c / N;
  const int n0 = c % N;
  const int a_off = (a_heads == 1) ? 0 : h * K;

  // Stage the A tile once (all heads): coalesced 128-bit loads per row.
  const int vec_per_row = AK / 4;
  for (int v = threadIdx.x; v < TM * vec_per_row; v += kMMThreads) {
    const int r = v / vec_per_row, kv = v % vec_per_row;
    float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < row1) {
      const int64_t ar = a_idx ? __ldg(a_idx + row0 + r) : (row0 + r);
      val = ldg4(A + ar * AK + kv * 4);
    }
    *reinterpret_cast<float4 *>(As + (size_t)r * AS + kv * 4) = val;
  }

  int s = segment_of(seg_ptrs, num_segs, row0);
  int64_t p0 = row0;
  while (p0 < row1) {
    // advance to the segment that contains p0 (skips empty segments)
    while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= p0) ++s;
    const int64_t seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;
    const int64_t p1 = seg_end < row1 ? seg_end : row1;
    __syncthreads();  // As ready (first pass) / previous Ws consumers done
    // Stage W[s, heads of this column chunk, :, :] as Ws[k][col - col0].
    for (int v = threadIdx.x; v < K * TX; v += kMMThreads) {
      const int k = v / TX, cv = v % TX;
      const int cc = col0 + cv * 4;
      const int hh = cc / N, nn = cc % N;
      const float4 w4 = ldg4(W + (((int64_t)s * H + hh) * K + k) * N + nn);
      *reinterpret_cast<float4 *>(Ws + (size_t)k * CW + cv * 4) = w4;
    }
    __syncthreads();

    const int lr0 = (int)(p0 - row0), lr1 = (int)(p1 - row0);
    // rows of this thread: ty + TY*j ; only those inside [lr0, lr1) are stored
    float4 acc[RPT];
#pragma unroll
    for (int j = 0; j < RPT; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < K; k += 4) {
      float4 b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        b[q] = *reinterpret_cast<const float4 *>(Ws + (size_t)(k + q) * CW + tx * 4);
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        const float4 a = *reinterpret_cast<const float4 *>(
            As + (size_t)(ty + TY * j) * AS + a_off + k);
        fma4(acc[j], a.x, b[0]);
        fma4(acc[j], a.y, b[1]);
        fma4(acc[j], a.z, b[2]);
        fma4(acc[j], a.w, b[3]);
      }
    }
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int lr = ty + TY * j;
      if (lr >= lr0 && lr < lr1) {
        const int64_t gi = row0 + lr;
        const int64_t cr = c_idx ? __ldg(c_idx + gi) : gi;
        float *y = Y + (cr * c_heads + (c_heads == 1 ? 0 : h)) * N + n0;
        if (accumulate)
          red_add4(y, acc[j]);
        else
          st4(y, acc[j]);
      }
    }
    p0 = p1;
  }
}

// ---------------------------------------------------------------------------
// N == 1 (RGAT attention logits el/er = <feat[e,h,:], attn[r,h,:]>):
// lane group per row, 128-bit loads, shuffle reduction per head.
// ---------------------------------------------------------------------------
// A CTA owns a contiguous range of rows and each lane group walks it with stride GROUPS, so the
// segment of a row is found by binary search once and then advanced (535 relations on the
// wikikg2-shaped graph made the per-row search a 10-deep dependent load chain: 3x the roofline time).
struct SegCursor {
  const int64_t *__restrict__ ptrs;
  int num_segs, s;
  int64_t end;  // first row past segment s
  __device__ SegCursor(const int64_t *p, int n, int64_t row, int64_t M) : ptrs(p), num_segs(n), s(0), end(M) {
    s = segment_of(ptrs, num_segs, row);
    end = (s + 1 < num_segs) ? __ldg(ptrs + s + 1) : M;
  }
  __device__ __forceinline__ int at(int64_t row, int64_t M) {  // rows must not decrease between calls
    while (row >= end && s + 1 < num_segs) {
      ++s;
      end = (s + 1 < num_segs) ? __ldg(ptrs + s + 1) : M;
    }
    return s;
  }
};

template <int LPR>
__global__ void __launch_bounds__(kMMThreads)
    segmm_rowdot_kernel(const float *__restrict__ A,
                        const int64_t *__restrict__ a_idx,
                        const float *__restrict__ W, float *__restrict__ Y,
                        const int64_t *__restrict__ c_idx,
                        const int64_t *__restrict__ seg_ptrs, int num_segs,
                        int64_t M, int H, int hw, int accumulate, int64_t rows_per_cta) {
  constexpr int GROUPS = kMMThreads / LPR;
  constexpr int D = LPR * 4;
  constexpr int U = 4;  // rows in flight per lane group
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int h = l / hw;
  const bool leader = (l % hw) == 0;
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;  // rows_per_cta is a multiple of GROUPS * U
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  if (row0 >= row1) return;
  SegCursor cur(seg_ptrs, num_segs, row0 + g < M ? row0 + g : M - 1, M);
  for (int64_t base = row0; base < row1; base += U * GROUPS) {  // uniform trip count: the shuffles need the whole warp
    const int64_t i = base + g;
    float4 a[U], w[U];
    bool ok[U];
    int sg[U];
    // segments first, row loads after: the cursor's `end` can be written by a load, and a compare on it
    // placed between the row loads waits (shared scoreboard) for the rows just requested — that kept one
    // row per lane group in flight instead of U (ncu source view: 60 % of the samples on those compares)
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t r = i + j * GROUPS;
      ok[j] = r < row1;
      sg[j] = ok[j] ? cur.at(r, M) : 0;
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t r = i + j * GROUPS;
      a[j] = w[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ok[j]) {
        const int64_t ar = a_idx ? __ldg(a_idx + r) : r;
        a[j] = ldcs4(A + ar * D + 4 * l);
        w[j] = ldg4(W + (int64_t)sg[j] * D + 4 * l);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const float v = group_sum(dot4(a[j], w[j]), hw);
      if (ok[j] && leader) {
        const int64_t r = i + j * GROUPS;
        const int64_t cr = c_idx ? __ldg(c_idx + r) : r;
        if (accumulate)
          atomicAdd(Y + cr * H + h, v);
        else
          Y[cr * H + h] = v;
      }
    }
  }
}

// K == 1 (backward of the above): Y[c,h,:] (+)= A[a,h] * W[s,h,0,:].
template <int LPR>
__global__ void __launch_bounds__(kMMThreads)
    segmm_expand_kernel(const float *__restrict__ A,
                        const int64_t *__restrict__ a_idx,
                        const float *__restrict__ W, float *__restrict__ Y,
                        const int64_t *__restrict__ c_idx,
                        const int64_t *__restrict__ seg_ptrs, int num_segs,
                        int64_t M, int H, int hw, int accumulate, int64_t rows_per_cta) {
  constexpr int GROUPS = kMMThreads / LPR;
  constexpr int D = LPR * 4;
  constexpr int U = 4;
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int h = l / hw;
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  if (row0 + g >= row1) return;
  SegCursor cur(seg_ptrs, num_segs, row0 + g, M);
  for (int64_t i = row0 + g; i < row1; i += U * GROUPS) {
    float a[U];
    float4 w[U];
    int64_t cr[U];
    int sg[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {  // segments first, loads after (see segmm_rowdot_kernel)
      const int64_t r = i + j * GROUPS;
      sg[j] = r < row1 ? cur.at(r, M) : 0;
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t r = i + j * GROUPS;
      cr[j] = -1;
      a[j] = 0.f;
      w[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < row1) {
        const int64_t ar = a_idx ? __ldg(a_idx + r) : r;
        cr[j] = c_idx ? __ldg(c_idx + r) : r;
        a[j] = __ldg(A + ar * H + h);
        w[j] = ldg4(W + (int64_t)sg[j] * D + 4 * l);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      if (cr[j] >= 0) {
        const float4 y = make_float4(a[j] * w[j].x, a[j] * w[j].y, a[j] * w[j].z, a[j] * w[j].w);
        if (accumulate)
          red_add4(Y + cr[j] * D + 4 * l, y);
        else
          stcs4(Y + cr[j] * D + 4 * l, y);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Weight gradient: dW[s,h] += A_s^T · G_s. Persistent CTAs own contiguous row
// ranges; accumulators stay in registers across row tiles and are flushed with
// atomics only when the segment changes or the range ends, so the number of
// global atomics is (#CTAs + #segments) * H*K*N instead of one per product.
// ---------------------------------------------------------------------------
constexpr int kGWRows = 32;  // rows staged per shared-memory tile

template <int TPT>  // 4x4 register tiles per thread
__global__ void __launch_bounds__(kMMThreads)
    gradw_tiled_kernel(const float *__restrict__ A,
                       const int64_t *__restrict__ a_idx, int a_heads,
                       const float *__restrict__ G,
                       const int64_t *__restrict__ g_idx, int g_heads,
                       float *__restrict__ dW,
                       const int64_t *__restrict__ seg_ptrs, int num_segs,
                       int64_t M, int H, int K, int N, int64_t rows_per_cta) {
  extern __shared__ __align__(16) float smem[];
  const int AK = a_heads * K, GN = g_heads * N;
  float *As = smem;                         // [kGWRows][AK]
  float *Gs = smem + (size_t)kGWRows * AK;  // [kGWRows][GN]

  const int ntn = N / 4, ntk = K / 4;
  const int tiles = H * ntk * ntn;
  // If there are fewer tiles than threads, RS threads share one tile and split
  // the rows of every staged tile between them.
  const int RS = tiles >= kMMThreads ? 1 : kMMThreads / tiles;
  const int rs = tiles >= kMMThreads ? 0 : threadIdx.x / tiles;
  const bool active = (rs < RS);

  int t_h[TPT], t_k[TPT], t_n[TPT];
  bool t_ok[TPT];
  float acc[TPT][16];
#pragma unroll
  for (int j = 0; j < TPT; ++j) {
    const int id = (tiles >= kMMThreads) ? threadIdx.x + j * kMMThreads
                                         : threadIdx.x % tiles;
    t_ok[j] = active && id < tiles && (tiles >= kMMThreads || j == 0);
    const int idc = id < tiles ? id : 0;
    t_n[j] = (idc % ntn) * 4;
    t_k[j] = ((idc / ntn) % ntk) * 4;
    t_h[j] = idc / (ntn * ntk);
#pragma unroll
    for (int q = 0; q < 16; ++q) acc[j][q] = 0.f;
  }

  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  if (row0 >= row1) return;
  int s = segment_of(seg_ptrs, num_segs, row0);
  while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= row0) ++s;
  int64_t seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;

  auto flush = [&](int seg) {
#pragma unroll
    for (int j = 0; j < TPT; ++j) {
      if (!t_ok[j]) continue;
      float *dst = dW + (((int64_t)seg * H + t_h[j]) * K + t_k[j]) * N + t_n[j];
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        red_add4(dst + (int64_t)kk * N,
                 make_float4(acc[j][kk * 4 + 0], acc[j][kk * 4 + 1],
                             acc[j][kk * 4 + 2], acc[j][kk * 4 + 3]));
        acc[j][kk * 4 + 0] = acc[j][kk * 4 + 1] = acc[j][kk * 4 + 2] =
            acc[j][kk * 4 + 3] = 0.f;
      }
    }
  };

  const int avec = AK / 4, gvec = GN / 4;
  for (int64_t base = row0; base < row1; base += kGWRows) {
    const int nrows = (int)((row1 - base < kGWRows) ? row1 - base : kGWRows);
    __syncthreads();
    for (int v = threadIdx.x; v < nrows * avec; v += kMMThreads) {
      const int r = v / avec, kv = v % avec;
      const int64_t ar = a_idx ? __ldg(a_idx + base + r) : (base + r);
      *reinterpret_cast<float4 *>(As + (size_t)r * AK + kv * 4) =
          ldg4(A + ar * AK + kv * 4);
    }
    for (int v = threadIdx.x; v < nrows * gvec; v += kMMThreads) {
      const int r = v / gvec, nv = v % gvec;
      const int64_t gr = g_idx ? __ldg(g_idx + base + r) : (base + r);
      *reinterpret_cast<float4 *>(Gs + (size_t)r * GN + nv * 4) =
          ldg4(G + gr * GN + nv * 4);
    }
    __syncthreads();
    int r = 0;
    while (r < nrows) {
      // rows [r, rend) of this tile belong to segment s
      int rend = nrows;
      if (seg_end - base < rend) rend = (int)(seg_end - base);
      for (int rr = r + rs; rr < rend; rr += RS) {
#pragma unroll
        for (int j = 0; j < TPT; ++j) {
          if (!t_ok[j]) continue;
          const float4 a = *reinterpret_cast<const float4 *>(
              As + (size_t)rr * AK + (a_heads == 1 ? 0 : t_h[j] * K) + t_k[j]);
          const float4 g = *reinterpret_cast<const float4 *>(
              Gs + (size_t)rr * GN + (g_heads == 1 ? 0 : t_h[j] * N) + t_n[j]);
          const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            acc[j][kk * 4 + 0] = fmaf(av[kk], g.x, acc[j][kk * 4 + 0]);
            acc[j][kk * 4 + 1] = fmaf(av[kk], g.y, acc[j][kk * 4 + 1]);
            acc[j][kk * 4 + 2] = fmaf(av[kk], g.z, acc[j][kk * 4 + 2]);
            acc[j][kk * 4 + 3] = fmaf(av[kk], g.w, acc[j][kk * 4 + 3]);
          }
        }
      }
      r = rend;
      if (base + r == seg_end && base + r < row1) {
        flush(s);
        ++s;
        while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= base + r) ++s;
        seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;
      }
    }
  }
  flush(s);
}

// General fallback for the weight gradient (any K, N; e.g. N == 1 or K == 1):
// each thread owns up to ACC accumulators of the flattened [H*K*N] space.
__global__ void __launch_bounds__(kMMThreads)
    gradw_generic_kernel(const float *__restrict__ A,
                         const int64_t *__restrict__ a_idx, int a_heads,
                         const float *__restrict__ G,
                         const int64_t *__restrict__ g_idx, int g_heads,
                         float *__restrict__ dW,
                         const int64_t *__restrict__ seg_ptrs, int num_segs,
                         int64_t M, int H, int K, int N, int64_t rows_per_cta) {
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  if (row0 >= row1) return;
  const int total = H * K * N;
  for (int e0 = blockIdx.y * kMMThreads; e0 < total;
       e0 += gridDim.y * kMMThreads) {
    const int e = e0 + threadIdx.x;
    const bool ok = e < total;
    const int ec = ok ? e : 0;
    const int n = ec % N, k = (ec / N) % K, h = ec / (N * K);
    int s = segment_of(seg_ptrs, num_segs, row0);
    while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= row0) ++s;
    int64_t seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;
    float acc = 0.f;
    for (int64_t i = row0; i < row1; ++i) {
      if (i == seg_end) {
        if (ok && acc != 0.f) atomicAdd(dW + (int64_t)s * total + e, acc);
        acc = 0.f;
        ++s;
        while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= i) ++s;
        seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;
      }
      const int64_t ar = a_idx ? __ldg(a_idx + i) : i;
      const int64_t gr = g_idx ? __ldg(g_idx + i) : i;
      const float a = __ldg(A + (ar * a_heads + (a_heads == 1 ? 0 : h)) * K + k);
      const float g = __ldg(G + (gr * g_heads + (g_heads == 1 ? 0 : h)) * N + n);
      acc = fmaf(a, g, acc);
    }
    if (ok && acc != 0.f) atomicAdd(dW + (int64_t)s * total + e, acc);
  }
}

// ---------------------------------------------------------------------------
// Weight gradient of the attention-vector products (N == 1 or K == 1):
//   dW[s, c] += sum_{i in segment s} S[s_idx[i], c / hw4] * V[v_idx[i], c]     c < D = H*F
// i.e. a segment sum of head-scaled rows. A CTA owns a contiguous row range and walks
// it segment piece by piece: its 256/LPR lane groups stride the piece's rows with
// 128-bit loads, partial sums are combined in shared memory, and ONE group issues the
// vector reductions, so global atomics are (#CTAs + #segments) * D.
// ---------------------------------------------------------------------------
template <int LPR>
__global__ void __launch_bounds__(kMMThreads)
    gradw_rowscale_kernel(const float *__restrict__ V, const int64_t *__restrict__ v_idx,
                          const float *__restrict__ S, const int64_t *__restrict__ s_idx,
                          float *__restrict__ dW, const int64_t *__restrict__ seg_ptrs, int num_segs,
                          int64_t M, int H, int hw, int64_t rows_per_cta) {
  constexpr int GROUPS = kMMThreads / LPR;
  constexpr int D = LPR * 4;
  __shared__ float4 part[GROUPS][LPR];
  const int g = threadIdx.x / LPR, l = threadIdx.x % LPR;
  const int h = l / hw;
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row1 = (row0 + rows_per_cta < M) ? row0 + rows_per_cta : M;
  if (row0 >= row1) return;
  int s = segment_of(seg_ptrs, num_segs, row0);
  while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= row0) ++s;
  int64_t a = row0;
  while (a < row1) {
    const int64_t seg_end = (s + 1 < num_segs) ? __ldg(seg_ptrs + s + 1) : M;
    const int64_t b = seg_end < row1 ? seg_end : row1;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int64_t i = a + g;
    for (; i + 3 * GROUPS < b; i += 4 * GROUPS) {  // 4 rows in flight per group
      float sc[4];
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t r = i + j * GROUPS;
        const int64_t vr = v_idx ? __ldg(v_idx + r) : r;
        const int64_t sr = s_idx ? __ldg(s_idx + r) : r;
        sc[j] = __ldg(S + sr * H + h);
        v[j] = ldcs4(V + vr * D + 4 * l);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) fma4(acc, sc[j], v[j]);
    }
    for (; i < b; i += GROUPS) {
      const int64_t vr = v_idx ? __ldg(v_idx + i) : i;
      const int64_t sr = s_idx ? __ldg(s_idx + i) : i;
      fma4(acc, __ldg(S + sr * H + h), ldcs4(V + vr * D + 4 * l));
    }
    part[g][l] = acc;
    __syncthreads();
    for (int stride = GROUPS / 2; stride > 0; stride >>= 1) {
      if (g < stride) {
        const float4 x = part[g][l], y = part[g + stride][l];
        part[g][l] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
      }
      __syncthreads();
    }
    if (g == 0) red_add4(dW + (int64_t)s * D + 4 * l, part[0][l]);
    __syncthreads();
    a = b;
    if (a == seg_end && a < row1) {
      ++s;
      while (s + 1 < num_segs && __ldg(seg_ptrs + s + 1) <= a) ++s;
    }
  }
}

static int pick_lpr(int d) {
  if (d % 4) return 0;
  const int v = d / 4;
  return (v == 4 || v == 8 || v == 16 || v == 32) ? v : 0;
}

}  // namespace hetb

using namespace hetb;

// defined in segment_mm_tc.cu (tcgen05 path)
int hetb_segment_mm_tc(const float *A, const int64_t *a_idx, int a_heads, const float *W, float *Y,
                       const int64_t *c_idx, int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                       int64_t num_rows, int H, int K, int N, int accumulate, int passes, void *stream);

// defined in segment_mm_tc_wide.cu (tcgen05 path for contraction / output widths above 64)
int hetb_segment_mm_tc_wide(const float *A, const int64_t *a_idx, int a_heads, const float *W, float *Y,
                            const int64_t *c_idx, int c_heads, const int64_t *seg_ptrs, int64_t num_segs,
                            int64_t num_rows, int H, int K, int N, int accumulate, void *stream);

int hetb_segment_mm_grad_w_tc(const float *A, const int64_t *a_idx, int a_heads, const float *G,
                              const int64_t *g_idx, int g_heads, float *dW, const int64_t *seg_ptrs,
                              int64_t num_segs, int64_t num_rows, int H, int K, int N, int probe, void *stream);

static int g_segmm_impl = 0;  // 0 auto, 1 SIMT only, 2 tensor core only (3xTF32), 3 TC 1xTF32, 4 TC pipeline without MMAs (timing probe)

extern "C" int hetb_set_segment_mm_impl(int impl) {
  if (impl < 0 || impl > 65535) return HETB_EINVAL;
  g_segmm_impl = impl;
  return HETB_OK;
}

extern "C" int hetb_segment_mm(const float *A, const int64_t *a_idx,
                               int a_heads, const float *W, float *Y,
                               const int64_t *c_idx, int c_heads,
                               const int64_t *seg_ptrs, int64_t num_segs,
                               int64_t num_rows, int H, int K, int N,
                               int accumulate, void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_segs >= 0 && H > 0 && K > 0 && N > 0);
  HETB_CHECK_ARG(a_heads == 1 || a_heads == H);
  HETB_CHECK_ARG(c_heads == H || (c_heads == 1 && (accumulate || H == 1)));
  if (num_rows == 0) return HETB_OK;
  HETB_CHECK_ARG(A && W && Y && seg_ptrs && num_segs > 0);
  if (num_segs > INT32_MAX) return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  const int S = (int)num_segs;
  const int C = H * N;

  // Tensor-core path (tcgen05/TMEM, 3xTF32): 64-wide contraction and output, segments of
  // useful size (a tile pays one pass per segment piece it straddles).
  if (g_segmm_impl != 1) {
    const bool worth = num_rows >= 4096 && num_rows / num_segs >= 16;
    if (g_segmm_impl >= 2 || worth) {
      // impl >= 16: timing probes, (impl >> 4) = probe mask << 4 | passes
      const int passes = g_segmm_impl >= 16 ? (g_segmm_impl >> 4) : (g_segmm_impl == 3 ? 1 : (g_segmm_impl == 4 ? 0 : 3));
      const int rc = hetb_segment_mm_tc(A, a_idx, a_heads, W, Y, c_idx, c_heads, seg_ptrs, num_segs,
                                        num_rows, H, K, N, accumulate, passes, stream);
      if (rc != HETB_EUNSUPPORTED || g_segmm_impl >= 16) return rc;
      if (passes == 3) {  // hidden sizes above 64: both operands streamed, 128 x 256 accumulator tiles
        const int rcw = hetb_segment_mm_tc_wide(A, a_idx, a_heads, W, Y, c_idx, c_heads, seg_ptrs, num_segs, num_rows,
                                                H, K, N, accumulate, stream);
        if (rcw != HETB_EUNSUPPORTED) return rcw;
      }
      if (g_segmm_impl >= 2) return rc;
    }
  }

  // N == 1 with per-head input: row-dot fast path.
  if (N == 1 && a_heads == H && c_heads == H && is_pow2(H) && is_pow2(K) &&
      K >= 4 && pick_lpr(H * K)) {
    const int lpr = pick_lpr(H * K);
    const int64_t chunk = (int64_t)(kMMThreads / lpr) * 4;  // rows one CTA pass covers (GROUPS x U)
    int64_t rows_per_cta = ceil_div64(ceil_div64(num_rows, (int64_t)kNumSMs * 16), chunk) * chunk;
    const int grid = (int)ceil_div64(num_rows, rows_per_cta);
    switch (lpr) {
#define ROWDOT(L)                                                              \
  case L:                                                                      \
    segmm_rowdot_kernel<L><<<grid, kMMThreads, 0, st>>>(                       \
        A, a_idx, W, Y, c_idx, seg_ptrs, S, num_rows, H, K / 4, accumulate,    \
        rows_per_cta);                                                         \
    break;
      ROWDOT(4) ROWDOT(8) ROWDOT(16) ROWDOT(32)
#undef ROWDOT
    }
    HETB_LAUNCH_OK();
    return HETB_OK;
  }
  // K == 1 with per-head output: expansion fast path.
  if (K == 1 && a_heads == H && c_heads == H && is_pow2(H) && is_pow2(N) &&
      N >= 4 && pick_lpr(H * N)) {
    const int lpr = pick_lpr(H * N);
    const int64_t chunk = (int64_t)(kMMThreads / lpr) * 4;
    int64_t rows_per_cta = ceil_div64(ceil_div64(num_rows, (int64_t)kNumSMs * 16), chunk) * chunk;
    const int grid = (int)ceil_div64(num_rows, rows_per_cta);
    switch (lpr) {
#define EXPAND(L)                                                              \
  case L:                                                                      \
    segmm_expand_kernel<L><<<grid, kMMThreads, 0, st>>>(                       \
        A, a_idx, W, Y, c_idx, seg_ptrs, S, num_rows, H, N / 4, accumulate,    \
        rows_per_cta);                                                         \
    break;
      EXPAND(4) EXPAND(8) EXPAND(16) EXPAND(32)
#undef EXPAND
    }
    HETB_LAUNCH_OK();
    return HETB_OK;
  }
  // Tiled path.
  const int AK = a_heads * K;
  if ((N % 4) == 0 && (K % 4) == 0 && AK <= 256 && K <= 256 &&
      (C % 16) == 0) {
    constexpr int TM = 128;
    const int CW = (C % 64 == 0) ? 64 : (C % 32 == 0 ? 32 : 16);
    const size_t smem = ((size_t)TM * (AK + 4) + (size_t)K * CW) * sizeof(float);
    if (smem <= 200 * 1024) {
      dim3 grid((unsigned)ceil_div64(num_rows, TM), C / CW);
      cudaError_t ce = cudaSuccess;
#define TILED(CWV)                                                             \
  do {                                                                         \
    static bool attr_set = false;                                              \
    if (!attr_set) {                                                           \
      ce = cudaFuncSetAttribute(segmm_tiled_kernel<CWV, TM>,                   \
                                cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                200 * 1024);                                   \
      attr_set = true;                                                         \
    }                                                                          \
    if (ce != cudaSuccess) return (int)ce;                                     \
    segmm_tiled_kernel<CWV, TM><<<grid, kMMThreads, smem, st>>>(               \
        A, a_idx, a_heads, W, Y, c_idx, c_heads, seg_ptrs, S, num_rows, H, K,  \
        N, accumulate);                                                        \
  } while (0)
      if (CW == 64)
        TILED(64);
      else if (CW == 32)
        TILED(32);
      else
        TILED(16);
#undef TILED
      HETB_LAUNCH_OK();
      return HETB_OK;
    }
  }
  // Fallback.
  {
    const int64_t total = num_rows * C;
    const int64_t need = ceil_div64(total, kMMThreads);
    const int grid = (int)(need < kNumSMs * 32 ? need : kNumSMs * 32);
    segmm_generic_kernel<<<grid, kMMThreads, 0, st>>>(
        A, a_idx, a_heads, W, Y, c_idx, c_heads, seg_ptrs, S, num_rows, H, K, N,
        accumulate);
    HETB_LAUNCH_OK();
  }
  return HETB_OK;
}

// Same contraction with the output stored as bf16 (plain store, no accumulation): the tables the
// plan-based edge passes gather (k, W_att·q, W_msg·v per compact row). Tensor-core path only
// (64-wide contraction and output); anything else is HETB_EUNSUPPORTED, there is no fallback.
extern "C" int hetb_segment_mm_bf16out(const float *A, const int64_t *a_idx, int a_heads, const float *W,
                                       void *Y_bf16, const int64_t *c_idx, const int64_t *seg_ptrs,
                                       int64_t num_segs, int64_t num_rows, int H, int K, int N, void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_segs >= 0 && H > 0 && K > 0 && N > 0);
  HETB_CHECK_ARG(a_heads == 1 || a_heads == H);
  if (num_rows == 0) return HETB_OK;
  HETB_CHECK_ARG(A && W && Y_bf16 && seg_ptrs && num_segs > 0);
  return hetb_segment_mm_tc(A, a_idx, a_heads, W, (float *)Y_bf16, c_idx, H, seg_ptrs, num_segs, num_rows, H, K, N, 0,
                            3 | (1 << 12), stream);
}

extern "C" int hetb_segment_mm_grad_w(const float *A, const int64_t *a_idx,
                                      int a_heads, const float *G,
                                      const int64_t *g_idx, int g_heads,
                                      float *dW, const int64_t *seg_ptrs,
                                      int64_t num_segs, int64_t num_rows, int H,
                                      int K, int N, void *stream) {
  HETB_CHECK_ARG(num_rows >= 0 && num_segs >= 0 && H > 0 && K > 0 && N > 0);
  HETB_CHECK_ARG(a_heads == 1 || a_heads == H);
  HETB_CHECK_ARG(g_heads == 1 || g_heads == H);
  if (num_rows == 0) return HETB_OK;
  HETB_CHECK_ARG(A && G && dW && seg_ptrs && num_segs > 0);
  if (num_segs > INT32_MAX) return HETB_EUNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  const int S = (int)num_segs;
  const int AK = a_heads * K, GN = g_heads * N;
  // Tensor-core path (tcgen05/TMEM, 3xTF32, contraction over the rows): 64-wide rows on both sides.
  if (g_segmm_impl != 1 && (g_segmm_impl >= 2 || num_rows >= 2048)) {
    const int probe = g_segmm_impl >= 16 ? (g_segmm_impl >> 8) : 0;
    const int rc = hetb_segment_mm_grad_w_tc(A, a_idx, a_heads, G, g_idx, g_heads, dW, seg_ptrs, num_segs,
                                             num_rows, H, K, N, probe, stream);
    if (rc != HETB_EUNSUPPORTED) return rc;
  }
  // Attention-vector gradients: N == 1 (dW = sum G[i,h] * A[i,h,:]) or K == 1 (dW = sum A[i,h] * G[i,h,:]).
  if ((N == 1 || K == 1) && !(N == 1 && K == 1) && a_heads == H && g_heads == H) {
    const int F = (N == 1) ? K : N;
    const int lpr = (is_pow2(H) && is_pow2(F) && F >= 4) ? pick_lpr(H * F) : 0;
    if (lpr) {
      const float *V = (N == 1) ? A : G;
      const int64_t *vi = (N == 1) ? a_idx : g_idx;
      const float *Sc = (N == 1) ? G : A;
      const int64_t *si = (N == 1) ? g_idx : a_idx;
      int64_t ctas = kNumSMs * 4;
      int64_t rows_per_cta = ceil_div64(num_rows, ctas);
      if (rows_per_cta < 256) rows_per_cta = 256;
      ctas = ceil_div64(num_rows, rows_per_cta);
      switch (lpr) {
#define ROWSCALE(L)                                                                        \
  case L:                                                                                  \
    gradw_rowscale_kernel<L><<<(unsigned)ctas, kMMThreads, 0, st>>>(                       \
        V, vi, Sc, si, dW, seg_ptrs, S, num_rows, H, F / 4, rows_per_cta);                 \
    break;
        ROWSCALE(4) ROWSCALE(8) ROWSCALE(16) ROWSCALE(32)
#undef ROWSCALE
      }
      HETB_LAUNCH_OK();
      return HETB_OK;
    }
  }
  const int tiles = (K % 4 == 0 && N % 4 == 0) ? H * (K / 4) * (N / 4) : 0;
  const size_t smem = (size_t)kGWRows * (AK + GN) * sizeof(float);
  if (tiles > 0 && tiles <= 4 * kMMThreads && smem <= 96 * 1024 &&
      (tiles >= kMMThreads ? tiles % kMMThreads == 0 : kMMThreads % tiles == 0)) {
    const int tpt = tiles >= kMMThreads ? tiles / kMMThreads : 1;
    int64_t ctas = kNumSMs * 4;
    int64_t rows_per_cta = ceil_div64(num_rows, ctas);
    rows_per_cta = ceil_div64(rows_per_cta, kGWRows) * kGWRows;
    ctas = ceil_div64(num_rows, rows_per_cta);
    cudaError_t ce = cudaSuccess;
#define GRADW(T)                                                               \
  do {                                                                         \
    static bool attr_set = false;                                              \
    if (!attr_set) {                                                           \
      ce = cudaFuncSetAttribute(gradw_tiled_kernel<T>,                         \
                                cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                96 * 1024);                                    \
      attr_set = true;                                                         \
    }                                                                          \
    if (ce != cudaSuccess) return (int)ce;                                     \
    gradw_tiled_kernel<T><<<(unsigned)ctas, kMMThreads, smem, st>>>(           \
        A, a_idx, a_heads, G, g_idx, g_heads, dW, seg_ptrs, S, num_rows, H, K, \
        N, rows_per_cta);                                                      \
  } while (0)
    if (tpt == 1)
      GRADW(1);
    else if (tpt == 2)
      GRADW(2);
    else if (tpt == 4)
      GRADW(4);
    else
      goto generic;
#undef GRADW
    HETB_LAUNCH_OK();
    return HETB_OK;
  }
generic : {
  int64_t ctas = kNumSMs * 8;
  int64_t rows_per_cta = ceil_div64(num_rows, ctas);
  if (rows_per_cta < 64) rows_per_cta = 64;
  ctas = ceil_div64(num_rows, rows_per_cta);
  const int total = H * K * N;
  int gy = (int)ceil_div64(total, kMMThreads);
  if (gy > 64) gy = 64;
  dim3 grid((unsigned)ctas, gy);
  gradw_generic_kernel<<<grid, kMMThreads, 0, st>>>(
      A, a_idx, a_heads, G, g_idx, g_heads, dW, seg_ptrs, S, num_rows, H, K, N,
      rows_per_cta);
  HETB_LAUNCH_OK();
}
  return HETB_OK;
}
