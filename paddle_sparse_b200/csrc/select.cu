// index_select / masked_select building blocks.
//
// Reference behaviour restated (not ported):
//   index_select.py:17-51   dim 0: rowcount[idx] -> cumsum -> repeat_interleave -> gather_csr -> 2-3 gathers
//   index_select.py:53-89   dim 1: the same on the CSC arrays, then argsort back to row-major
//   masked_select.py:12-114 boolean-mask compactions (one host sync each in the reference)
// Here: one stream compaction kernel (decoupled look-back) turns a mask into ascending
// indices, and one load-balanced fill kernel writes (major, minor, perm) of the selected
// rows/cols per OUTPUT tile, so neither duplicated ids nor hub rows serialise a thread.
#include "common.cuh"
#include "convert.cuh"
#include "segment.cuh"

namespace ps {

// Column-range selection of `narrow(dim=1)` rides on the compaction kernel: mask == nullptr switches the
// predicate to lo <= index[i] < hi and the kept (row, col - lo) pairs are written next to the entry ids.
struct RangeSelect {
  int64_t lo, hi;
  const int64_t *row;
  int64_t *out_row, *out_col;
};

template <int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
compact_kernel(const uint8_t *__restrict__ mask, const int64_t *__restrict__ index, int64_t n,
               int64_t *__restrict__ out_idx, uint64_t *state, long long *d_count, RangeSelect range) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int64_t warp_sums[33];
  __shared__ int64_t s_tile, s_excl;
  if (threadIdx.x == 0) s_tile = (int64_t)atomicAdd((unsigned long long *)state, 1ull);
  __syncthreads();
  const int64_t tile = s_tile;
  const int64_t base = tile * TILE + (int64_t)threadIdx.x * ITEMS;
  unsigned flags = 0;
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int64_t i = base + k;
    bool f = false;
    if (i < n) {
      if (mask) {
        f = mask[index ? index[i] : i] != 0;
      } else {  // narrow.py:56-57: lo <= col < hi
        const int64_t c = index[i];
        f = c >= range.lo && c < range.hi;
      }
    }
    flags |= (unsigned)f << k;
    cnt += f;
  }
  int64_t total;
  int64_t excl = block_exclusive_scan<int64_t, THREADS>((int64_t)cnt, warp_sums, &total);
  if (threadIdx.x < 32) {
    const uint64_t p = warp_lookback(state + 1, tile, (uint64_t)total);
    if (threadIdx.x == 0) s_excl = (int64_t)p;
  }
  __syncthreads();
  excl += s_excl;
  if ((tile + 1) * (int64_t)TILE >= n && threadIdx.x == 0) *d_count = (long long)(s_excl + total);
  if (mask) {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k)
      if ((flags >> k) & 1u) out_idx[excl++] = base + k;
  } else {  // the kept entries' (row, col - lo) go out with their ids (narrow.py:59-60)
#pragma unroll
    for (int k = 0; k < ITEMS; ++k)
      if ((flags >> k) & 1u) {
        const int64_t i = base + k;
        if (out_idx) out_idx[excl] = i;
        range.out_row[excl] = range.row[i];
        range.out_col[excl] = index[i] - range.lo;
        ++excl;
      }
  }
}

static int compact_launch(const uint8_t *mask, const int64_t *index, int64_t n, int64_t *out_idx, int64_t *host_count,
                          cudaStream_t s, RangeSelect range = RangeSelect{0, 0, nullptr, nullptr, nullptr}) {
  *host_count = 0;
  if (n == 0) return PS_OK;
  constexpr int THREADS = 256, ITEMS = 8;
  const int64_t tiles = ceil_div(n, THREADS * ITEMS);
  Scratch scratch(s);
  PS_SCRATCH(state, scratch, uint64_t, tiles + 2);
  PS_CUDA(cudaMemsetAsync(state, 0, sizeof(uint64_t) * (size_t)(tiles + 2), s));
  long long *d_count = reinterpret_cast<long long *>(state + tiles + 1);
  compact_kernel<THREADS, ITEMS><<<(unsigned)tiles, THREADS, 0, s>>>(mask, index, n, out_idx, state, d_count, range);
  PS_LAUNCH_OK();
  long long h = 0;
  PS_TRY(read_back(&h, d_count, sizeof(h), s));
  *host_count = (int64_t)h;
  return PS_OK;
}

template <int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
index_select_fill_kernel(const int64_t *__restrict__ old_ptr, const int64_t *__restrict__ minor,
                         const int64_t *__restrict__ idx, int64_t K, const int64_t *__restrict__ out_ptr, int64_t E,
                         int64_t *__restrict__ out_major, int64_t *__restrict__ out_minor,
                         int64_t *__restrict__ out_perm) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = (int64_t)blockIdx.x * TILE;
  const int tile_n = (int)((E - e0) < TILE ? (E - e0) : TILE);
  expand_tile<THREADS, ITEMS>(out_ptr, K, e0, tile_n, seg, aux);
  for (int j = threadIdx.x; j < tile_n; j += THREADS) {
    const int64_t r = seg[j];
    const int64_t e = e0 + j;
    const int64_t src = old_ptr[idx[r]] + (e - out_ptr[r]);
    out_major[e] = r;
    out_minor[e] = minor[src];
    if (out_perm) out_perm[e] = src;
  }
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_mask_to_index(const uint8_t *mask, int64_t n, int64_t *out_idx, int64_t *host_count, void *stream) {
  PS_CHECK(n >= 0 && host_count, "mask_to_index: bad arguments");
  return compact_launch(mask, nullptr, n, out_idx, host_count, (cudaStream_t)stream);
}

int ps_mask_entries(const uint8_t *mask, const int64_t *index, int64_t nnz, int64_t *out_idx, int64_t *host_count,
                    void *stream) {
  PS_CHECK(nnz >= 0 && host_count, "mask_entries: bad arguments");
  return compact_launch(mask, index, nnz, out_idx, host_count, (cudaStream_t)stream);
}

int ps_narrow_cols(const int64_t *row, const int64_t *col, const void *value, int64_t row_bytes, int64_t nnz,
                   int64_t lo, int64_t hi, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *host_count,
                   void *stream) {
  PS_CHECK(nnz >= 0 && row_bytes >= 0 && host_count && lo <= hi, "narrow_cols: bad arguments");
  *host_count = 0;
  if (nnz == 0) return PS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  PS_SCRATCH(ids_buf, scratch, int64_t, value ? nnz : 1);
  int64_t *ids = value ? ids_buf : nullptr;
  PS_TRY(compact_launch(nullptr, col, nnz, ids, host_count, s, RangeSelect{lo, hi, row, out_row, out_col}));
  if (value && *host_count > 0) PS_TRY(gather_rows_launch<int64_t>(value, row_bytes, ids, *host_count, out_value, s));
  return PS_OK;
}

int ps_index_select_fill(const int64_t *old_ptr, const int64_t *minor, const void *value, int64_t row_bytes,
                         const int64_t *idx, int64_t K, const int64_t *out_ptr, int64_t E, int64_t *out_major,
                         int64_t *out_minor, void *out_value, int64_t *out_perm, void *stream) {
  PS_CHECK(K >= 0 && E >= 0 && row_bytes >= 0, "index_select_fill: negative size");
  PS_CHECK(K < 2147483647LL, "index_select_fill: too many selected ids");
  if (E == 0) return PS_OK;
  PS_CHECK(K > 0, "index_select_fill: E > 0 with K == 0");
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  int64_t *perm = out_perm;
  if (value && !perm) {
    PS_SCRATCH(tmp, scratch, int64_t, E);
    perm = tmp;
  }
  constexpr int THREADS = 256, ITEMS = 8;
  index_select_fill_kernel<THREADS, ITEMS><<<(unsigned)ceil_div(E, THREADS * ITEMS), THREADS, 0, s>>>(
      old_ptr, minor, idx, K, out_ptr, E, out_major, out_minor, perm);
  PS_LAUNCH_OK();
  if (value) PS_TRY(gather_rows_launch<int64_t>(value, row_bytes, perm, E, out_value, s));
  return PS_OK;
}

}  // extern "C"
