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

// cat.py:49-276 stitches row / rowptr / col / colptr / csr2csc / csc2csr of several matrices by
// `concat([x_k + delta_k])` (one elementwise add per piece, then a concat).  Here one launch writes the whole
// output: out[start_k + i] = src_k[i] + delta_k for up to CAT_MAX pieces passed by value.
constexpr int CAT_MAX = 32;
struct CatPieces {
  const int64_t *src[CAT_MAX];
  int64_t start[CAT_MAX + 1];  // start[k] = first output position of piece k, start[n] = total
  int64_t delta[CAT_MAX];
  int n;
};

__global__ void __launch_bounds__(256) concat_add_kernel(CatPieces p, int64_t *__restrict__ out) {
  const int64_t total = p.start[p.n];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int lo = 0, hi = p.n - 1;  // last piece with start <= i (empty pieces share a start: the last one wins)
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (p.start[mid] <= i) lo = mid;
      else hi = mid - 1;
    }
    out[i] = ld_stream(p.src[lo] + (i - p.start[lo])) + p.delta[lo];
  }
}

// storage.py:281-330 `sparse_resize`: a pointer / count array grown to `total` entries, the new tail holding `fill`
// (rowptr / colptr: nnz; rowcount / colcount: 0), or cut to `total` entries.  One launch, one read, one write.
__global__ void __launch_bounds__(256)
pad_i64_kernel(const int64_t *__restrict__ src, int64_t n, int64_t fill, int64_t total, int64_t *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
    out[i] = i < n ? ld_stream(src + i) : fill;
}

// storage.py:332-368 `sparse_reshape`: idx = old_N * row + col; row' = idx / new_N, col' = idx % new_N (row-major
// order is preserved, so the result stays sorted).  One launch instead of five elementwise framework ops.
__global__ void __launch_bounds__(256)
reshape_index_kernel(const int64_t *__restrict__ row, const int64_t *__restrict__ col, int64_t n, int64_t old_n,
                     int64_t new_n, int64_t *__restrict__ out_row, int64_t *__restrict__ out_col) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t idx = old_n * ld_stream(row + i) + ld_stream(col + i);
    const int64_t r = idx / new_n;
    out_row[i] = r;
    out_col[i] = idx - r * new_n;
  }
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

int ps_concat_add_i64(const int64_t *const *srcs, const int64_t *lens, const int64_t *deltas, int64_t K,
                      int64_t *out, void *stream) {
  PS_CHECK(K >= 0, "concat_add: negative piece count");
  cudaStream_t s = (cudaStream_t)stream;
  int64_t done = 0;
  for (int64_t k0 = 0; k0 < K; k0 += CAT_MAX) {
    CatPieces p;
    p.n = (int)((K - k0) < CAT_MAX ? (K - k0) : CAT_MAX);
    int64_t at = 0;
    for (int j = 0; j < p.n; ++j) {
      PS_CHECK(lens[k0 + j] >= 0, "concat_add: negative length");
      p.src[j] = srcs[k0 + j];
      p.start[j] = at;
      p.delta[j] = deltas[k0 + j];
      at += lens[k0 + j];
    }
    p.start[p.n] = at;
    if (at > 0) {
      int64_t blocks = ceil_div(at, 256 * 4);
      const int64_t cap = 16 * (int64_t)sm_count();
      if (blocks > cap) blocks = cap;
      concat_add_kernel<<<(unsigned)blocks, 256, 0, s>>>(p, out + done);
      PS_LAUNCH_OK();
    }
    done += at;
  }
  return PS_OK;
}

int ps_concat_bytes(const void *const *srcs, const int64_t *nbytes, int64_t K, void *out, void *stream) {
  PS_CHECK(K >= 0, "concat_bytes: negative piece count");
  cudaStream_t s = (cudaStream_t)stream;
  char *dst = reinterpret_cast<char *>(out);
  for (int64_t k = 0; k < K; ++k) {
    PS_CHECK(nbytes[k] >= 0, "concat_bytes: negative length");
    if (nbytes[k] > 0) PS_CUDA(cudaMemcpyAsync(dst, srcs[k], (size_t)nbytes[k], cudaMemcpyDeviceToDevice, s));
    dst += nbytes[k];
  }
  return PS_OK;
}

int ps_pad_i64(const int64_t *src, int64_t n, int64_t fill, int64_t total, int64_t *out, void *stream) {
  PS_CHECK(n >= 0 && total >= 0, "pad_i64: negative size");
  if (total == 0) return PS_OK;
  int64_t blocks = ceil_div(total, 256 * 4);
  if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
  pad_i64_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, n < total ? n : total, fill, total, out);
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_reshape_index(const int64_t *row, const int64_t *col, int64_t n, int64_t old_n, int64_t new_n, int64_t *out_row,
                     int64_t *out_col, void *stream) {
  PS_CHECK(n >= 0 && old_n >= 0 && new_n > 0, "reshape_index: bad sizes");
  if (n == 0) return PS_OK;
  int64_t blocks = ceil_div(n, 256 * 4);
  if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
  reshape_index_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(row, col, n, old_n, new_n, out_row, out_col);
  PS_LAUNCH_OK();
  return PS_OK;
}

}  // extern "C"
