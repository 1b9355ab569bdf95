// Internal declarations: segmented reductions and row gathers (not part of the C ABI).
#pragma once
#include "common.cuh"

namespace ps {

// out[s,:] = reduce over e in [ptr[s], ptr[s+1]) of src[perm ? perm[e] : e, :]
// E_hint = number of source entries covered (chooses lanes per segment).
int segment_reduce_launch(const void *src, int dtype, int64_t width, const int64_t *ptr, int64_t S, int reduce,
                          void *out, const uint32_t *perm32, int64_t E_hint, cudaStream_t s);

// out[i, :] = src[index[i], :] for rows of row_bytes bytes; IndexT = int64_t or uint32_t
template <typename IndexT>
int gather_rows_launch(const void *src, int64_t row_bytes, const IndexT *index, int64_t n_out, void *out,
                       cudaStream_t s);

// largest power-of-two unit (<= 16 B) that divides row_bytes and both base addresses
inline int copy_unit(int64_t row_bytes, const void *a, const void *b) {
  uintptr_t x = (uintptr_t)row_bytes | (uintptr_t)a | (uintptr_t)b;
  int u = 16;
  while (u > 1 && (x & (uintptr_t)(u - 1))) u >>= 1;
  return u;
}

}  // namespace ps
