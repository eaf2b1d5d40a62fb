// TEST INFRASTRUCTURE ONLY — thin extern "C" shim over the REFERENCE's own CPU
// converters, compiled from the reference sources where they lie
// (/root/reference/hrt/include/MyHyb/MyHyb.h) by oracle/build_ref.py into
// oracle/_ref/libhet_ref_converters.so. No reference source is copied here:
// this file only #includes the header and forwards plain arrays to
//   TransposeCSR                          MyHyb.h:31-75
//   ConvertIntegratedCOOToSeparateCOO     MyHyb.h:1047-1097
//   ConvertIntegratedCOOToSeparateCSR     MyHyb.h:996-1045
// (the functions behind the torch ops in OpExport/DataConverters.inc.h:10-344).
// Used by tests/ to pin oracle/converters_ref.py against the reference itself.
#include <cstdint>
#include <cstring>
#include <vector>

// headers the real <cusp/csr_matrix.h> used to pull in transitively
#include <algorithm>
#include <cassert>
#include <thrust/adjacent_difference.h>
#include <thrust/host_vector.h>
#include <thrust/sequence.h>
#include <thrust/sort.h>

#include "MyHyb/MyHyb.h"

using HV = thrust::host_vector<int64_t>;

static HV to_hv(const int64_t *p, int64_t n) { return HV(p, p + n); }
static void from_hv(const HV &v, int64_t *out) {
  for (size_t i = 0; i < v.size(); ++i) out[i] = v[i];
}

extern "C" {

// in/out: row_ptrs[n+1], col[e]; out: perm[e] (position in the old CSR of each
// new position), exactly what TransposeCSR returns.
void ref_transpose_csr(int64_t *row_ptrs, int64_t n, int64_t *col, int64_t e,
                       int64_t *perm) {
  HV rp = to_hv(row_ptrs, n + 1), ci = to_hv(col, e);
  HV p = TransposeCSR<int64_t>(rp, ci);
  from_hv(rp, row_ptrs);
  from_hv(ci, col);
  from_hv(p, perm);
}

void ref_coo_to_separate_coo(const int64_t *row, const int64_t *col,
                             const int64_t *rel, const int64_t *eid, int64_t e,
                             int64_t num_rows, int64_t num_rels,
                             int64_t *rel_ptrs, int64_t *out_row,
                             int64_t *out_col, int64_t *out_eid) {
  auto r = ConvertIntegratedCOOToSeparateCOO<int64_t>(
      to_hv(row, e), to_hv(col, e), to_hv(rel, e), to_hv(eid, e), num_rows,
      num_rels);
  from_hv(r[0], rel_ptrs);
  from_hv(r[1], out_row);
  from_hv(r[2], out_col);
  from_hv(r[3], out_eid);
}

void ref_coo_to_separate_csr(const int64_t *row, const int64_t *col,
                             const int64_t *rel, const int64_t *eid, int64_t e,
                             int64_t num_rows, int64_t num_rels,
                             int64_t *rel_ptrs, int64_t *row_ptrs,
                             int64_t *out_col, int64_t *out_eid) {
  auto r = ConvertIntegratedCOOToSeparateCSR<int64_t>(
      to_hv(row, e), to_hv(col, e), to_hv(rel, e), to_hv(eid, e), num_rows,
      num_rels);
  from_hv(r[0], rel_ptrs);
  from_hv(r[1], row_ptrs);
  from_hv(r[2], out_col);
  from_hv(r[3], out_eid);
}
}
