// Entry-set operations between two sparse matrices: the sparse (.) sparse
// multiply of mul.py:35-73.  The reference concatenates both COO sets, argsorts the 64-bit keys and keeps
// adjacent equal pairs.  Both operands are required to be coalesced (mul.py:38-41), i.e. sorted and
// duplicate-free, so no sort is needed here: every entry of A looks its column up in the row of B it
// belongs to (B's rowptr bounds a sorted run of columns) and a matched pair is an output entry.
#include "common.cuh"

namespace ps {

// One thread per entry of A: lower bound of col in col_b[rowptr_b[r], rowptr_b[r+1]).
// out_pos[e] = position of the equal entry of B, or -1; out_mask[e] = 1 / 0.
__global__ void __launch_bounds__(256)
coo_match_kernel(const int64_t *__restrict__ row, const int64_t *__restrict__ col, int64_t n,
                 const int64_t *__restrict__ rowptr_b, const int64_t *__restrict__ col_b, int64_t M_b,
                 int64_t *__restrict__ out_pos, uint8_t *__restrict__ out_mask) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const int64_t r = ld_stream(row + e), c = ld_stream(col + e);
  int64_t found = -1;
  if (r >= 0 && r < M_b) {
    int64_t lo = rowptr_b[r], hi = rowptr_b[r + 1];
    while (lo < hi) {
      const int64_t mid = lo + ((hi - lo) >> 1);
      if (col_b[mid] < c) lo = mid + 1;
      else hi = mid;
    }
    if (lo < rowptr_b[r + 1] && col_b[lo] == c) found = lo;
  }
  out_pos[e] = found;
  out_mask[e] = found >= 0 ? 1 : 0;
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_coo_match(const int64_t *row, const int64_t *col, int64_t n, const int64_t *rowptr_b, const int64_t *col_b,
                 int64_t M_b, int64_t *out_pos, uint8_t *out_mask, void *stream) {
  PS_CHECK(n >= 0 && M_b >= 0, "coo_match: negative size");
  if (n == 0) return PS_OK;
  coo_match_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(row, col, n, rowptr_b, col_b, M_b,
                                                                                  out_pos, out_mask);
  PS_LAUNCH_OK();
  return PS_OK;
}

}  // extern "C"
