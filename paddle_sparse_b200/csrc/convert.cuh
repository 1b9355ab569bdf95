// Internal declarations shared between translation units (not part of the C ABI).
#pragma once
#include "common.cuh"

namespace ps {

// Number of elements of the ascending array a[0..n) that are <= key, found by a full warp
// with a 32-ary search (5 dependent loads for 2^25 elements instead of 25).
// Must be called by all 32 lanes; the result is uniform across the warp.
__device__ __forceinline__ int64_t warp_upper_bound(const int64_t *__restrict__ a, int64_t n, int64_t key) {
  const unsigned lane = lane_id();
  int64_t lo = 0, hi = n;  // answer in [lo, hi]
  while (lo < hi) {
    const int64_t len = hi - lo;
    const int64_t step = (len + 31) / 32;
    const int64_t idx = lo + (int64_t)lane * step;
    const bool le = (idx < hi) && (a[idx] <= key);
    const unsigned m = __ballot_sync(0xffffffffu, le);
    const int c = __popc(m);  // monotone: lanes 0..c-1 are true
    if (c == 0) {
      hi = lo;
    } else {
      const int64_t new_lo = lo + (int64_t)(c - 1) * step + 1;
      const int64_t cand_hi = lo + (int64_t)c * step;
      if (c < 32 && cand_hi < hi) hi = cand_hi;
      lo = new_lo;
    }
  }
  return lo;
}

// Tile geometry shared by ptr2ind and the CSR-expanding kernels (gather_csr, mul_row,
// index_select_fill): a block resolves, for a tile of TILE consecutive output entries,
// the segment id of every entry into shared memory (int32).
// `bounds` (optional, from tile_bounds_launch): bounds[t] = first segment touching tile t, bounds[tiles] = S - 1.
// With it a tile starts from two loads instead of two 32-ary searches of ~5 dependent loads each, which is what
// bounds these kernels when the whole grid is resident in a few waves (the searches serialise per tile).
// expand_tile_owned: the same resolution, but each thread keeps the segments of ITS ITEMS consecutive entries
// (tile entries threadIdx.x * ITEMS + k) in registers and nothing is written back: kernels whose threads consume
// exactly those entries (mul_row on 16-byte vectors) save the shared-memory round trip and one barrier.
template <int THREADS, int ITEMS>
__device__ __forceinline__ void expand_tile_owned(const int64_t *__restrict__ ptr, int64_t S, int64_t e0,
                                                  int tile_n, int32_t *seg /*[TILE]*/, int32_t *aux /*[40]*/,
                                                  const int32_t *__restrict__ bounds, int32_t (&own)[ITEMS]) {
  constexpr int TILE = THREADS * ITEMS;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  // first / last segment touching the tile: largest s in [0,S-1] with ptr[s] <= e
  if (bounds) {
    if (threadIdx.x < 2) aux[32 + threadIdx.x] = bounds[e0 / TILE + threadIdx.x];  // s_hi may overshoot: filtered below
  } else if (warp == 0) {
    const int64_t c = warp_upper_bound(ptr, S, e0);
    if (lane == 0) aux[32] = (int32_t)(c > 0 ? c - 1 : 0);
  } else if (warp == 1) {
    const int64_t c = warp_upper_bound(ptr, S, e0 + tile_n - 1);
    if (lane == 0) aux[33] = (int32_t)(c > 0 ? c - 1 : 0);
  }
  __syncthreads();
  const int32_t s_lo = aux[32], s_hi = aux[33];
  // Few segments start inside the tile (the usual case; always when rows are long): their clamped starts go into
  // `seg` as a sorted table and every thread finds its first entry's segment by a binary search of that table,
  // then walks it for the other ITEMS - 1 entries.  One barrier, no tile-wide initialisation, atomics or scan.
  // segment of entry p = s_lo + #{s in (s_lo, s_hi] : ptr[s] - e0 <= p} (ptr is monotone; empty segments share a
  // start and the last one wins, as below).
  const int nseg = s_hi - s_lo;
  if (nseg <= TILE) {  // uniform over the CTA
    for (int i = threadIdx.x; i < nseg; i += THREADS) {
      const int64_t p = ptr[(int64_t)s_lo + 1 + i] - e0;
      seg[i] = p < 0 ? 0 : (p > TILE ? TILE : (int32_t)p);
    }
    __syncthreads();
    const int p0 = (int)threadIdx.x * ITEMS;
    auto count_le = [&](int lo, int hi, int p) {  // first index in [lo, hi] whose start exceeds p
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (seg[mid] <= p) lo = mid + 1;
        else hi = mid;
      }
      return lo;
    };
    int lo = count_le(0, nseg, p0);
    const int last = count_le(lo, nseg, p0 + ITEMS - 1);
    if (last - lo <= 2 * ITEMS) {
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        while (lo < last && seg[lo] <= p0 + k) ++lo;
        own[k] = s_lo + lo;
      }
    } else {  // a run of empty segments lands inside this thread's entries: search per entry instead of walking it
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        lo = count_le(lo, last, p0 + k);
        own[k] = s_lo + lo;
      }
    }
    return;
  }
  for (int j = threadIdx.x; j < TILE; j += THREADS) seg[j] = -1;
  __syncthreads();
  if (threadIdx.x == 0) seg[0] = s_lo;
  __syncthreads();
  // segments starting strictly inside the tile: empty segments share a start, the last wins
  for (int64_t sgm = (int64_t)s_lo + 1 + threadIdx.x; sgm <= s_hi; sgm += THREADS) {
    const int64_t p = ptr[sgm] - e0;
    if (p > 0 && p < tile_n) atomicMax(&seg[p], (int32_t)sgm);
    else if (p <= 0) atomicMax(&seg[0], (int32_t)sgm);
  }
  __syncthreads();
  // inclusive max-scan over the tile (thread owns ITEMS consecutive slots)
  int32_t run = -1;
  int32_t loc[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int32_t v = seg[threadIdx.x * ITEMS + k];
    run = v > run ? v : run;
    loc[k] = run;
  }
  int32_t inc = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int32_t o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= (unsigned)d) inc = o > inc ? o : inc;
  }
  if (lane == 31) aux[warp] = inc;
  __syncthreads();
  int32_t carry = -1;
  for (unsigned w = 0; w < warp; ++w) carry = aux[w] > carry ? aux[w] : carry;
  const int32_t lane_excl = __shfl_up_sync(0xffffffffu, inc, 1);
  if (lane > 0) carry = lane_excl > carry ? lane_excl : carry;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) own[k] = loc[k] > carry ? loc[k] : carry;
}

// seg[j] = segment of tile entry j for every j, visible to the whole CTA on return
template <int THREADS, int ITEMS>
__device__ __forceinline__ void expand_tile(const int64_t *__restrict__ ptr, int64_t S, int64_t e0,
                                            int tile_n, int32_t *seg /*[TILE]*/, int32_t *aux /*[40]*/,
                                            const int32_t *__restrict__ bounds = nullptr) {
  int32_t own[ITEMS];
  expand_tile_owned<THREADS, ITEMS>(ptr, S, e0, tile_n, seg, aux, bounds, own);
  __syncthreads();  // `seg` may still be read as the table of segment starts
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) seg[threadIdx.x * ITEMS + k] = own[k];
  __syncthreads();
}

// bounds[t] = largest s in [0, S-1] with ptr[s] <= t * tile (t < tiles), bounds[tiles] = S - 1; one warp per tile
int tile_bounds_launch(const int64_t *ptr, int64_t S, int64_t E, int tile, int32_t *bounds, cudaStream_t s);

// ind2ptr on int64 or packed uint32 keys (the latter is used on the sorted col keys of csr2csc)
template <typename IndT>
int ind2ptr_launch(const IndT *ind, int64_t numel, int64_t M, int64_t *out, cudaStream_t s);

// out_ptr[0] = 0, out_ptr[i+1] = inclusive scan of count[idx ? idx[i] : i]
int count2ptr_launch(const int64_t *count, const int64_t *idx, int64_t n, int64_t *out_ptr, cudaStream_t s);

}  // namespace ps
