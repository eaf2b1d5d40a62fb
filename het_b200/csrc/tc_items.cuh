// Work items of the tensor-core segment MM kernels: 128-row tiles on a global grid, cut at segment
// boundaries (segment_mm_tc.cu, segment_mm_tc_wide.cu).
#pragma once
#include "common.cuh"

namespace hetb {
namespace tc {

constexpr int kTileRows = 128;  // rows per tile (UMMA M)

// ------------------------------------------------------------ work items ----
// Every role walks the same deterministic sequence of (tile, segment piece) items.
// Positions are 32-bit offsets from the CTA's first row (64-bit compare/select chains and the
// 64-bit division that maps a row to its tile cost a few hundred cycles per call, and next() sits
// on the critical path of every role); the segment table is only touched when a segment ends.
struct PieceIter {
  const int64_t *__restrict__ seg_ptrs;
  int64_t row_base;                  // first row of the CTA (a multiple of kTileRows)
  int num_segs, n, rel, seg_end, s;  // rows of the CTA, position, end of segment s (clamped to n)

  __device__ PieceIter(const int64_t *sp, int ns, int64_t m, int64_t tile_begin, int64_t tile_end)
      : seg_ptrs(sp), row_base(tile_begin * kTileRows), num_segs(ns), n(0), rel(0), seg_end(0), s(0) {
    int64_t stop = tile_end * kTileRows < m ? tile_end * kTileRows : m;
    n = stop > row_base ? (int)(stop - row_base) : 0;
    if (n > 0) {
      s = segment_of(seg_ptrs, num_segs, row_base);
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
  // next piece: rows [p0, p1) (relative to row_base) of the tile starting at t0, segment seg
  __device__ __forceinline__ bool next(int &t0, int &p0, int &p1, int &seg) {
    if (rel >= n) return false;
    t0 = rel & ~(kTileRows - 1);
    p0 = rel;
    p1 = min(min(t0 + kTileRows, n), seg_end);
    seg = s;
    rel = p1;
    if (p1 == seg_end && p1 < n) {
      ++s;
      skip_empty();
    }
    return true;
  }
};


}  // namespace tc
}  // namespace hetb
