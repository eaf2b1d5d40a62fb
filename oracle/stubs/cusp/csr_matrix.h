// TEST INFRASTRUCTURE ONLY. Minimal stand-in for <cusp/csr_matrix.h>: the reference's
// MyHyb/MyHyb.h includes it on line 2 and names cusp::csr_matrix in one constructor
// signature (MyHyb.h:128) that the converter functions the oracle pins against
// (TransposeCSR, ConvertIntegratedCOOToSeparate{COO,CSR}) never call. The vendored
// cusplibrary does not compile against CUDA 12.9's thrust, so oracle/build_ref.py puts
// this declaration-only header ahead of it on the include path.
#pragma once
#include <thrust/host_vector.h>
#include <cstddef>
namespace cusp {
template <typename IndexType, typename ValueType, typename MemorySpace>
struct csr_matrix {
  std::size_t num_rows = 0, num_cols = 0, num_entries = 0;
  thrust::host_vector<IndexType> row_offsets, column_indices;
  thrust::host_vector<ValueType> values;
};
}  // namespace cusp
