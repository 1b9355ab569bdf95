// sort-on-construct and coalesce.
//
// Reference behaviour restated (not ported):
//   storage.py:158-171  sort-on-construct: key = row*N + col, perm = argsort(key), 3 gathers
//   storage.py:454-486  coalesce: run-start mask over the sorted keys, two boolean
//                        compactions, nonzero(), paddle_scatter.segment_csr(value, ptr, reduce)
//   coalesce.py:8-29    free function = both in sequence
// Here: one packed-key radix sort (radix_sort.cuh) whose first pass packs
// (row << cbits | col) straight from the int64 arrays and carries the 32-bit value bits
// (or the entry index) as payload, followed by ONE single-pass kernel that detects run
// heads, compacts them with a decoupled look-back scan, unpacks row/col to int64 and
// reduces each run's values in sorted (= stable) order.  Wide / 8-byte values go
// through the permutation + segment-reduce path instead.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "radix_sort.cuh"
#include "segment.cuh"

namespace ps {

enum { DD_INDEX_ONLY = 0, DD_FAST = 1, DD_PTR = 2 };

}  // namespace ps
#include "local_sort.cuh"
#include "local_sort2.cuh"
#include "refine.cuh"
namespace ps {

// One tile = THREADS*ITEMS consecutive sorted entries, held in shared memory; thread t looks at entries
// t, t + THREADS, ... (conflict-free).  Run heads are ranked with ballots + one small scan, the tile's offset comes
// from a decoupled look-back, every head folds its run in sorted (= stable) order, the compacted (key, value)
// pairs are staged in shared memory and written out coalesced as int64 row / col + value.  With out_rowptr the
// kernel also emits rowptr (csrc/convert.cpp:13-43 semantics) of the coalesced result: the head that opens row r
// fills rowptr[prev_row + 1 .. r], the CTA of the last tile fills the tail.
// A run of equal keys that crosses tiles (heavy duplicates) is not walked by one thread: every tile reduces its own
// part - the entries BEFORE its first head continue the previous tile's last run - and leaves it here; the tile where
// the run ends is found by dedup_cross_kernel, which folds the parts in tile order into the run's output slot.
// (Round 1 let the head's thread walk the rest of the run through global memory: O(run) on one thread.)
struct DdCross {
  uint32_t cont_bits, tail_bits;  // partial reductions (accumulator bits): leading continuation / last head's run
  uint32_t cont_len, tail_len;
  long long first_out;            // output slot of the tile's first head (the continued run's slot is first_out - 1)
  int flags;                      // 1: no head in the tile; 2: the continued run ends here; 4: the last run goes on
  int pad;
};

template <typename KeyT, typename T, int RED, int MODE, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
dedup_kernel(KeySrc ksrc, PaySrc psrc, int64_t n, int cbits, int64_t M, int64_t *__restrict__ out_row,
             int64_t *__restrict__ out_col, T *__restrict__ out_value, int64_t *__restrict__ run_start,
             int64_t *__restrict__ out_rowptr, uint64_t *state /*[0]=ticket,[1..]=tiles*/, long long *d_nnz_out,
             DdCross *__restrict__ cross /*[tiles], DD_FAST only*/) {
  constexpr int TILE = THREADS * ITEMS, WARPS = THREADS / 32;
  using A = typename Acc<T>::type;
  static_assert(MODE != DD_FAST || sizeof(A) == 4, "partial reductions are kept as 32-bit words");
  __shared__ int s_fh;            // first head of the tile (tile_n: none)
  __shared__ A s_part[WARPS];
  __shared__ int64_t s_gap[8][3];  // long runs of empty rows, filled by the whole CTA: (first row, last row, value)
  __shared__ int s_ngap;
  __shared__ KeyT s_key[TILE + 2];  // s_key[0] = key preceding the tile, s_key[tile_n + 1] = key following it;
                                    // later the compacted keys
  __shared__ uint32_t s_pay[MODE == DD_FAST ? TILE : 1];
  __shared__ uint32_t s_wcnt[ITEMS * WARPS];
  __shared__ int64_t s_tile, s_excl;
  __shared__ int s_heads;
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  if (tid == 0) {
    s_tile = (int64_t)atomicAdd((unsigned long long *)state, 1ull);
    s_fh = TILE;
    s_ngap = 0;
  }
  __syncthreads();
  const int64_t tile = s_tile;
  const int64_t tile_base = tile * TILE;
  const int tile_n = (int)((n - tile_base) < TILE ? (n - tile_base) : TILE);

#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = k * THREADS + tid;
    if (j < tile_n) {
      s_key[j + 1] = load_key<KeyT>(ksrc, tile_base + j);
      if (MODE == DD_FAST) s_pay[j] = load_pay(psrc, tile_base + j);
    }
  }
  if (tid == 0) s_key[0] = tile_base > 0 ? load_key<KeyT>(ksrc, tile_base - 1) : KeyT(0);
  if (tid == 32 && tile_base + tile_n < n) s_key[tile_n + 1] = load_key<KeyT>(ksrc, tile_base + tile_n);
  __syncthreads();

  // ---- run heads, ranked in entry order (k, warp, lane)
  unsigned heads = 0, below[ITEMS];
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = k * THREADS + tid;
    const bool h = j < tile_n && ((tile_base + j == 0) || s_key[j + 1] != s_key[j]);
    const unsigned bal = __ballot_sync(0xffffffffu, h);
    heads |= (unsigned)h << k;
    below[k] = __popc(bal & lt);
    if (lane == 0) s_wcnt[k * WARPS + warp] = __popc(bal);
    if (MODE == DD_FAST && h) atomicMin(&s_fh, j);
  }
  __syncthreads();
  if (warp == 0) {  // exclusive scan of the ITEMS*WARPS counts, then the look-back for the tile's offset
    constexpr int PER = (ITEMS * WARPS + 31) / 32;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
      const int idx = lane * PER + q;
      v[q] = idx < ITEMS * WARPS ? s_wcnt[idx] : 0u;
      sum += v[q];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= (unsigned)d) inc += o;
    }
    uint32_t run = inc - sum;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
      const int idx = lane * PER + q;
      if (idx < ITEMS * WARPS) s_wcnt[idx] = run;
      run += v[q];
    }
    const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
    const uint64_t p = warp_lookback(state + 1, tile, (uint64_t)total);
    if (lane == 0) {
      s_excl = (int64_t)p;
      s_heads = (int)total;
    }
  }
  __syncthreads();
  const int64_t tile_excl = s_excl;
  const int tile_heads = s_heads;
  const bool last_tile = tile_base + tile_n == n;
  const KeyT cmask = (KeyT)((uint64_t(1) << cbits) - 1);
  if (MODE == DD_FAST) {  // the entries before the first head continue the previous tile's last run: reduce them
    const int fh = s_fh < tile_n ? s_fh : tile_n;
    A part = Red<RED, A>::init();
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const int j = k * THREADS + tid;
      if (j < fh) part = Red<RED, A>::op(part, to_acc<T>(pay_as<T>(s_pay[j])));
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part = Red<RED, A>::op(part, __shfl_xor_sync(0xffffffffu, part, d));
    if (lane == 0) s_part[warp] = part;
    __syncthreads();
    if (tid == 0) {
      A a = s_part[0];
      for (int w = 1; w < WARPS; ++w) a = Red<RED, A>::op(a, s_part[w]);
      DdCross c;
      memcpy(&c.cont_bits, &a, 4);
      c.tail_bits = 0;
      c.cont_len = (uint32_t)fh;
      c.tail_len = 0;
      c.first_out = tile_excl;
      const bool goes_on = !last_tile && s_key[tile_n + 1] == s_key[tile_n];
      c.flags = (tile_heads == 0 ? 1 : 0) |
                ((fh > 0 && tile_base > 0 && (tile_heads > 0 || !goes_on)) ? 2 : 0);
      c.pad = 0;
      cross[tile] = c;  // (a head whose run leaves the tile adds its part below)
    }
    __syncthreads();
  }

  // ---- every head folds its run and opens its row
  KeyT hkey[ITEMS];
  A hacc[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    hkey[k] = KeyT(0);
    hacc[k] = A(0);
    if (!((heads >> k) & 1u)) continue;
    const int j = k * THREADS + tid;
    const KeyT key = s_key[j + 1];
    hkey[k] = key;
    below[k] += s_wcnt[k * WARPS + warp];  // rank of this head among the tile's heads
    const int64_t o = tile_excl + below[k];
    if (MODE == DD_PTR) run_start[o] = tile_base + j;
    if (MODE == DD_FAST) {
      A acc = to_acc<T>(pay_as<T>(s_pay[j]));
      int jj = j + 1;
      while (jj < tile_n && s_key[jj + 1] == key) {
        acc = Red<RED, A>::op(acc, to_acc<T>(pay_as<T>(s_pay[jj])));
        ++jj;
      }
      const int64_t len = jj - j;
      if (jj == tile_n && tile_base + tile_n < n && s_key[tile_n + 1] == key) {  // the run continues past the tile:
        memcpy(&cross[tile].tail_bits, &acc, 4);                                  // dedup_cross_kernel finishes it
        cross[tile].tail_len = (uint32_t)len;
        cross[tile].flags |= 4;  // (the only writer of this tile's record after the barrier above)
      }
      hacc[k] = Red<RED, A>::finish(acc, len);
    }
    if (out_rowptr) {
      const int64_t r = (int64_t)((uint64_t)key >> cbits);
      const int64_t pr = (tile_base + j == 0) ? -1 : (int64_t)((uint64_t)s_key[j] >> cbits);
      if (r - pr > 64) {  // a long run of empty rows: the whole CTA fills it (below), not this thread
        const int g = atomicAdd(&s_ngap, 1);
        if (g < 8) {
          s_gap[g][0] = pr + 1;
          s_gap[g][1] = r;
          s_gap[g][2] = o;
        } else {
          for (int64_t q = pr + 1; q <= r; ++q) out_rowptr[q] = o;
        }
      } else {
        for (int64_t q = pr + 1; q <= r; ++q) out_rowptr[q] = o;
      }
    }
  }
  int64_t last_row = -1;
  if (last_tile && tid == 0) last_row = (int64_t)((uint64_t)s_key[tile_n] >> cbits);
  __syncthreads();  // every read of the tile is done: compact in place
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    if (!((heads >> k) & 1u)) continue;
    s_key[below[k]] = hkey[k];
    if (MODE == DD_FAST) {
      const T v = from_acc<T>(hacc[k]);
      uint32_t bits = 0;
      memcpy(&bits, &v, sizeof(T));
      s_pay[below[k]] = bits;
    }
  }
  if (last_tile && tid == 0) {
    *d_nnz_out = (long long)(tile_excl + tile_heads);
    if (MODE == DD_PTR) run_start[tile_excl + tile_heads] = n;
    s_tile = last_row;
  }
  __syncthreads();
  for (int i = tid; i < tile_heads; i += THREADS) {
    const KeyT key = s_key[i];
    out_row[tile_excl + i] = (int64_t)((uint64_t)key >> cbits);
    out_col[tile_excl + i] = (int64_t)(key & cmask);
    if (MODE == DD_FAST) out_value[tile_excl + i] = pay_as<T>(s_pay[i]);
  }
  if (out_rowptr) {
    const int ng = s_ngap < 8 ? s_ngap : 8;  // (visible: two barriers since the heads' loop)
    for (int g = 0; g < ng; ++g)
      for (int64_t q = s_gap[g][0] + tid; q <= s_gap[g][1]; q += THREADS) out_rowptr[q] = s_gap[g][2];
  }
  if (last_tile && out_rowptr) {  // rows after the last entry point at nnz'
    const int64_t total = tile_excl + tile_heads;
    for (int64_t q = s_tile + 1 + tid; q <= M; q += THREADS) out_rowptr[q] = total;
  }
}

// runs that crossed tiles: the tile where one ends folds the parts of the tiles before it, in tile order
template <typename T, int RED>
static __global__ void __launch_bounds__(256)
dedup_cross_kernel(const DdCross *__restrict__ cross, int64_t tiles, T *__restrict__ out_value) {
  using A = typename Acc<T>::type;
  using R = Red<RED, A>;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < tiles; t += (int64_t)gridDim.x * blockDim.x) {
    if (!(cross[t].flags & 2)) continue;
    int64_t u = t - 1;
    while (u > 0 && (cross[u].flags & 1)) --u;  // tiles without a head lie wholly inside the run
    A acc;
    memcpy(&acc, &cross[u].tail_bits, 4);
    int64_t len = cross[u].tail_len;
    for (int64_t w = u + 1; w <= t; ++w) {
      A part;
      memcpy(&part, &cross[w].cont_bits, 4);
      acc = R::op(acc, part);
      len += cross[w].cont_len;
    }
    out_value[cross[t].first_out - 1] = from_acc<T>(R::finish(acc, len));
  }
}

template <typename KeyT, typename T, int RED, int MODE>
int dedup_launch(KeySrc ksrc, PaySrc psrc, int64_t n, int cbits, int64_t M, int64_t *out_row, int64_t *out_col,
                 void *out_value, int64_t *run_start, int64_t *out_rowptr, long long *d_nnz, Scratch &scratch,
                 cudaStream_t s) {
  constexpr int THREADS = 512, ITEMS = 6;  // 3072 entries: 36 KB of static shared memory with 8-byte keys
  const int64_t tiles = ceil_div(n, THREADS * ITEMS);
  PS_SCRATCH(state, scratch, uint64_t, tiles + 1);
  PS_CUDA(cudaMemsetAsync(state, 0, sizeof(uint64_t) * (size_t)(tiles + 1), s));
  PS_SCRATCH(cross, scratch, DdCross, MODE == DD_FAST ? tiles : 1);
  {
    ProfScope prof(PROF_DEDUP, s);
    dedup_kernel<KeyT, T, RED, MODE, THREADS, ITEMS><<<(unsigned)tiles, THREADS, 0, s>>>(
        ksrc, psrc, n, cbits, M, out_row, out_col, reinterpret_cast<T *>(out_value), run_start, out_rowptr, state, d_nnz,
        cross);
  }
  PS_LAUNCH_OK();
  if (MODE == DD_FAST && tiles > 1) {
    dedup_cross_kernel<T, RED><<<(unsigned)ceil_div(tiles, 256), 256, 0, s>>>(cross, tiles,
                                                                              reinterpret_cast<T *>(out_value));
    PS_LAUNCH_OK();
  }
  return PS_OK;
}

// values that ride along as the 32-bit payload: one element of <= 4 bytes per entry
static bool fast_value(int dtype, int64_t width) {
  return width == 1 && (dtype == PS_F32 || dtype == PS_I32 || dtype == PS_BF16 || dtype == PS_F16);
}
static int pay_src_mode(int dtype) { return dtype_size(dtype) == 4 ? PSRC_U32 : PSRC_U16; }
static int pay_dst_mode(int dtype) { return dtype_size(dtype) == 4 ? PDST_U32 : PDST_U16; }

template <typename KeyT>
int dedup_dispatch(KeySrc ksrc, PaySrc psrc, int64_t n, int cbits, int64_t M, int mode, int dtype, int reduce,
                   int64_t *out_row, int64_t *out_col, void *out_value, int64_t *run_start, int64_t *out_rowptr,
                   long long *d_nnz, Scratch &scratch, cudaStream_t s) {
  if (mode == DD_INDEX_ONLY)
    return dedup_launch<KeyT, float, PS_SUM, DD_INDEX_ONLY>(ksrc, psrc, n, cbits, M, out_row, out_col, nullptr, nullptr,
                                                            out_rowptr, d_nnz, scratch, s);
  if (mode == DD_PTR)
    return dedup_launch<KeyT, float, PS_SUM, DD_PTR>(ksrc, psrc, n, cbits, M, out_row, out_col, nullptr, run_start,
                                                     out_rowptr, d_nnz, scratch, s);
#define PS_DD(T)                                                                                               \
  PS_DISPATCH_REDUCE(reduce, RED, return (dedup_launch<KeyT, T, RED, DD_FAST>(ksrc, psrc, n, cbits, M, out_row, \
                                                                             out_col, out_value, nullptr,      \
                                                                             out_rowptr, d_nnz, scratch, s)))
  switch (dtype) {
    case PS_F32: PS_DD(float); break;
    case PS_I32: PS_DD(int32_t); break;
    case PS_BF16: PS_DD(__nv_bfloat16); break;
    case PS_F16: PS_DD(__half); break;
    default: break;
  }
#undef PS_DD
  set_error("coalesce: dtype %d has no fast path", dtype);
  return PS_ERR_ARG;
}

struct KeyGeom {
  int rbits, cbits, bits;
};
static int key_geometry(int64_t M, int64_t N, KeyGeom *g) {
  g->rbits = bits_for(M);
  g->cbits = bits_for(N);
  g->bits = g->rbits + g->cbits;
  if (g->bits > 64) {
    set_error("sparse sizes (%lld, %lld) need %d key bits; at most 64 are supported", (long long)M, (long long)N, g->bits);
    return PS_ERR_ARG;
  }
  return PS_OK;
}

static int check_value_args(const void *value, const void *out_value, int dtype, int64_t width) {
  if (value) {
    PS_CHECK(out_value != nullptr, "value given but out_value is NULL");
    PS_CHECK(dtype_size(dtype) > 0 && dtype != PS_U8, "unsupported value dtype %d", dtype);
    PS_CHECK(width >= 1, "value width must be >= 1");
  }
  return PS_OK;
}

// ------------------------------------------------------------------------------------------------
// Hybrid coalesce (local_sort.cuh): top-T-bit buckets by one or two one-sweep passes, then one CTA per bucket sorts,
// merges and writes the final arrays.  *used = false (nothing written) when the shape or the bucket sizes rule it
// out; the caller then runs the plain LSD pipeline.  One extra host read (the largest bucket) decides.
int hybrid_variant();  // PS_HYBRID=0 disables (bench knob)

int ls_variant();  // PS_LS_VARIANT: 2 (default) = counting pass + in-bin compare (local_sort2.cuh), 1 = four LSD passes

static int ls_cap() { return ls_variant() == 1 ? LS_CAP : L2_CAP; }

template <typename T, int RED, bool HAS_VALUE, bool DEDUP, typename KeyT>
static int ls2_launch(const KeyT *lowkeys, const uint32_t *lowvals, const unsigned int *bucket_off, int nb, int R,
                      int cbits, int64_t M, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *out_rowptr,
                      uint64_t *state, long long *d_nnz, cudaStream_t s) {
  auto kern = ls2_sort_dedup_kernel<T, RED, HAS_VALUE, DEDUP, KeyT>;
  const size_t smem = sizeof(L2Smem<HAS_VALUE>);
  PS_FUNC_SMEM(kern, smem);
  int grid = sm_count() * 2;
  if (grid > nb) grid = nb;
  {
    ProfScope prof(PROF_LOCAL_SORT, s);
    kern<<<grid, L2_THREADS, smem, s>>>(lowkeys, lowvals, bucket_off, nb, R, cbits, M, out_row, out_col,
                                       reinterpret_cast<T *>(out_value), out_rowptr, state, d_nnz, L2Desc{});
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

// the same kernel on buckets given by descriptors (refine.cuh: skewed rows)
template <typename T, int RED, bool HAS_VALUE, bool DEDUP>
static int ls2_launch_desc(const uint32_t *lowkeys, const uint32_t *lowvals, const unsigned int *bucket_off /*of the
                           original buckets*/, L2Desc desc, int nb, int R, int cbits,
                           int64_t M, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *out_rowptr,
                           uint64_t *state, long long *d_nnz, cudaStream_t s) {
  auto kern = ls2_sort_dedup_kernel<T, RED, HAS_VALUE, DEDUP, uint32_t, true>;
  const size_t smem = sizeof(L2Smem<HAS_VALUE>);
  PS_FUNC_SMEM(kern, smem);
  int grid = sm_count() * 2;
  if (grid > nb) grid = nb;
  {
    ProfScope prof(PROF_LOCAL_SORT, s);
    kern<<<grid, L2_THREADS, smem, s>>>(lowkeys, lowvals, bucket_off, nb, R, cbits, M, out_row, out_col,
                                       reinterpret_cast<T *>(out_value), out_rowptr, state, d_nnz, desc);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

static __global__ void set_u32_kernel(uint32_t *p, uint32_t v) { *p = v; }
static __global__ void coarsen_counts_kernel(const unsigned int *__restrict__ fine, int shift, int nb2,
                                             unsigned int *__restrict__ coarse) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb2) return;
  unsigned int sum = 0;
  for (int q = 0; q < (1 << shift); ++q) sum += fine[((size_t)b << shift) + q];
  coarse[b] = sum;
}

// PS_REFINE (bench knob, read once): 0 = an oversized bucket sends the whole input to the LSD sort (round 1),
// 1 (default) = oversized buckets are refined, 2 = refined from a coarser top level (see hybrid_coalesce)
static int refine_variant() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_REFINE");
    v = e ? atoi(e) : 1;
  }
  return v;
}

// Oversized buckets re-partitioned level by level (refine.cuh).  On return `*nfinal` buckets are described by the
// sorted arrays of `desc` (start has nfinal + 1 entries); *ok = false: a key with more duplicates than a bucket holds
// (or a list overflow) - the caller falls back.  One host synchronisation.
static int refine_buckets(const uint32_t *keys0, const uint32_t *vals0, int64_t nnz, const unsigned int *bucket_off,
                          int nb, int R, unsigned cap, Scratch &scratch, cudaStream_t s, L2Desc *desc, int *nfinal,
                          bool *ok) {
  *ok = false;
  const bool has_v = vals0 != nullptr;
  RfLists L;
  L.seg_cap = (unsigned)(nnz / cap + 2);
  L.chunk_cap = (unsigned)(nnz / RF_CH + L.seg_cap + 2);
  L.final_cap = (unsigned)(nb + 4 * (nnz / cap) + 4 * (int64_t)L.seg_cap + 16);
  const int nlev = (R + 7) / 8;
  PS_CHECK(nlev >= 1 && nlev <= RF_MAX_LEVELS, "refine: %d local key bits unsupported", R);
  PS_SCRATCH(ctl, scratch, RfCtl, 1);
  PS_CUDA(cudaMemsetAsync(ctl, 0, sizeof(RfCtl), s));
  for (int l = 0; l < RF_MAX_LEVELS + 2; ++l) {
    L.seg[l] = nullptr;
    L.chunk_seg[l] = nullptr;
  }
  for (int l = 1; l <= nlev; ++l) {
    PS_SCRATCH(sg, scratch, RfSeg, L.seg_cap);
    PS_SCRATCH(cs, scratch, uint32_t, L.chunk_cap);
    L.seg[l] = sg;
    L.chunk_seg[l] = cs;
  }
  PS_SCRATCH(chunk_hist, scratch, unsigned int, (size_t)L.chunk_cap * 256);
  PS_SCRATCH(seg_dstart, scratch, unsigned int, (size_t)L.seg_cap * 256);
  PS_SCRATCH(f_start, scratch, uint32_t, L.final_cap);
  PS_SCRATCH(f_klo, scratch, uint32_t, L.final_cap);
  PS_SCRATCH(f_khi, scratch, uint32_t, L.final_cap);
  PS_SCRATCH(f_orig, scratch, uint32_t, L.final_cap);
  L.f_start = f_start; L.f_klo = f_klo; L.f_khi = f_khi; L.f_orig = f_orig;
  PS_SCRATCH(keys1, scratch, uint32_t, nnz);
  PS_SCRATCH(vals1, scratch, uint32_t, has_v ? nnz : 1);
  uint32_t *kb[2] = {const_cast<uint32_t *>(keys0), keys1};
  uint32_t *vb[2] = {const_cast<uint32_t *>(vals0), vals1};

  const int gbits = 32 - R < 3 ? (32 - R > 0 ? 32 - R : 0) : 3;  // neighbours a final bucket may merge: 2^gbits
  rf_select_kernel<<<(unsigned)ceil_div(ceil_div(nb, 1 << gbits), 256), 256, 0, s>>>(bucket_off, nb, R, gbits, cap, L, ctl);
  PS_LAUNCH_OK();
  auto scat_v = rf_scatter_kernel<true>;
  auto scat_n = rf_scatter_kernel<false>;
  PS_FUNC_SMEM(scat_v, sizeof(RfSmem));
  PS_FUNC_SMEM(scat_n, sizeof(RfSmem));
  for (int l = 1; l <= nlev; ++l) {
    const int hi = R - 8 * (l - 1);
    const int lo = hi - 8 > 0 ? hi - 8 : 0;
    const unsigned dmask = (1u << (hi - lo)) - 1u;
    const int src = (l - 1) & 1, dst = l & 1;
    rf_hist_kernel<<<sm_count() * 8, RF_THREADS, 0, s>>>(kb[src], L.seg[l], L.chunk_seg[l], ctl, l, lo, dmask,
                                                          L.chunk_cap, chunk_hist);
    PS_LAUNCH_OK();
    rf_scan_kernel<<<sm_count() * 8, 256, 0, s>>>(L.seg[l], ctl, ctl, l, lo, dmask, cap, L.seg_cap, chunk_hist,
                                                   seg_dstart, L, nlev);
    PS_LAUNCH_OK();
    if (has_v)
      scat_v<<<sm_count() * 3, RF_THREADS, sizeof(RfSmem), s>>>(kb[src], vb[src], kb[dst], vb[dst], L.seg[l],
                                                                L.chunk_seg[l], ctl, l, lo, dmask, L.chunk_cap, chunk_hist,
                                                                seg_dstart);
    else
      scat_n<<<sm_count() * 3, RF_THREADS, sizeof(RfSmem), s>>>(kb[src], nullptr, kb[dst], nullptr, L.seg[l],
                                                                L.chunk_seg[l], ctl, l, lo, dmask, L.chunk_cap, chunk_hist,
                                                                seg_dstart);
    PS_LAUNCH_OK();
  }
  RfCtl h;
  PS_TRY(read_back(&h, ctl, sizeof(h), s));
  if (getenv("PS_REFINE_DEBUG"))
    fprintf(stderr, "refine: segments %u %u %u %u | chunks %u %u %u %u | final buckets %u (of %d original) | overflow %u\n",
            h.nseg[1], h.nseg[2], h.nseg[3], h.nseg[4], h.nchunk[1], h.nchunk[2], h.nchunk[3], h.nchunk[4], h.nfinal, nb,
            h.overflow);
  if (h.overflow || h.nfinal == 0 || h.nfinal > L.final_cap) return PS_OK;
  // the final buckets in key order = sorted by their start
  const int nf = (int)h.nfinal;
  PS_SCRATCH(s_start, scratch, uint32_t, (size_t)nf + 1);
  PS_SCRATCH(s_klo, scratch, uint32_t, nf);
  PS_SCRATCH(s_khi, scratch, uint32_t, nf);
  PS_SCRATCH(s_orig, scratch, uint32_t, nf);
  PS_SCRATCH(t_k0, scratch, uint32_t, nf);
  PS_SCRATCH(t_k1, scratch, uint32_t, nf);
  PS_SCRATCH(t_p0, scratch, uint32_t, (size_t)nf * 3);
  PS_SCRATCH(t_p1, scratch, uint32_t, (size_t)nf * 3);
  uint32_t *tp0[3] = {t_p0, t_p0 + nf, t_p0 + 2 * (size_t)nf}, *tp1[3] = {t_p1, t_p1 + nf, t_p1 + 2 * (size_t)nf};
  PaySrcN<3> ps_;
  ps_.s[0] = PaySrc{f_klo, PSRC_U32};
  ps_.s[1] = PaySrc{f_khi, PSRC_U32};
  ps_.s[2] = PaySrc{f_orig, PSRC_U32};
  PayDstN<3> pd;
  pd.d[0] = PayDst{s_klo, PDST_U32};
  pd.d[1] = PayDst{s_khi, PDST_U32};
  pd.d[2] = PayDst{s_orig, PDST_U32};
  PS_TRY((radix_sort_tuples<uint32_t, 3>(KeySrc{f_start, nullptr, KSRC_KEYS, 0}, ps_, nf, bits_for(nnz), t_k0, t_k1, tp0,
                                         tp1, KeyDst{s_start, nullptr, KDST_KEYS, 0}, pd, scratch, s)));
  set_u32_kernel<<<1, 1, 0, s>>>(s_start + nf, (uint32_t)nnz);
  PS_LAUNCH_OK();
  desc->start = s_start;
  desc->klo = s_klo;
  desc->khi = s_khi;
  desc->orig = s_orig;
  desc->keys1 = keys1;
  desc->vals1 = has_v ? vals1 : nullptr;
  *nfinal = nf;
  *ok = true;
  return PS_OK;
}

template <typename T, int RED, bool HAS_VALUE, bool DEDUP>
static int ls_launch(const uint32_t *lowkeys, const uint32_t *lowvals, const unsigned int *bucket_off, int nb, int R,
                     int cbits, int64_t M, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *out_rowptr,
                     uint64_t *state, long long *d_nnz, cudaStream_t s) {
  if (ls_variant() != 1) {
    return ls2_launch<T, RED, HAS_VALUE, DEDUP, uint32_t>(lowkeys, lowvals, bucket_off, nb, R, cbits, M, out_row, out_col,
                                                        out_value, out_rowptr, state, d_nnz, s);
  }
  auto kern = ls_sort_dedup_kernel<T, RED, HAS_VALUE, DEDUP>;
  const size_t smem = sizeof(LsSmem<HAS_VALUE>);
  PS_FUNC_SMEM(kern, smem);
  int grid = sm_count() * 2;
  if (grid > nb) grid = nb;
  {
    ProfScope prof(PROF_LOCAL_SORT, s);
    kern<<<grid, LS_SORT_THREADS, smem, s>>>(lowkeys, lowvals, bucket_off, nb, R, cbits, M, out_row, out_col,
                                            reinterpret_cast<T *>(out_value), out_rowptr, state, d_nnz);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

// The hybrid plan for (nnz, M, rbits, cbits): T top key bits select a bucket (a whole range of 2^(rbits - T) rows), the
// low R = K - T bits are sorted inside it.  T is the smallest width whose AVERAGE bucket - nnz * 2^(rbits - T) / M
// entries: only M of the 2^rbits row values exist - fills at most 90 % of a CTA's capacity.  Returns false when the
// hybrid path does not apply (narrow keys, tiny inputs, R > 32, T > 15, more than two global passes).
struct HybridPlan {
  int T, R, K, nb;
  LsPlan lp;
};
static bool hybrid_plan(int64_t nnz, int64_t M, int rbits, int cbits, HybridPlan *hp) {
  const int K = rbits + cbits;
  if (!hybrid_variant() || nnz < (int64_t(1) << 16) || K <= 32) return false;
  const int cap = ls_cap();
  int T = K - 32 > 1 ? K - 32 : 1;  // the bucket-local key must fit 32 bits
  const int tmax = LS_MAX_T < rbits ? LS_MAX_T : rbits;
  while (T <= tmax && (double)nnz * (double)(int64_t(1) << (rbits - T)) / (double)(M > 0 ? M : 1) > 0.9 * cap) ++T;
  if (T > tmax) return false;
  const int R = K - T;
  const SortPlan gp = make_sort_plan(T, rs_geometry<uint64_t, 1>().rbits);  // the plan radix_sort_pairs will use
  if (gp.passes > 2) return false;
  hp->T = T;
  hp->R = R;
  hp->K = K;
  hp->nb = 1 << T;
  hp->lp.passes = gp.passes;
  for (int p = 0; p < gp.passes; ++p) { hp->lp.pbegin[p] = gp.begin[p]; hp->lp.pbits[p] = gp.bits[p]; }
  return true;
}

// bucket sizes -> offsets, largest bucket, bin bases of the global passes (and, with_max, max(row) / max(col))
static int hybrid_inspect(const int64_t *row, const int64_t *col, int64_t nnz, int cbits, const HybridPlan &hp,
                          LsInspect *st, bool with_max, cudaStream_t s) {
  PS_CUDA(cudaMemsetAsync(st->max_rc, 0xff, sizeof(st->max_rc), s));  // -1: empty
  PS_CUDA(cudaMemsetAsync(st->gcount, 0, sizeof(unsigned int) * (size_t)hp.nb, s));
  {
    PS_FUNC_SMEM(ls_bucket_hist_kernel, sizeof(unsigned int) << LS_MAX_T);
    ProfScope prof(PROF_BUCKET_HIST, s);
    ls_bucket_hist_kernel<<<sm_count() * 2, 1024, sizeof(unsigned int) * (size_t)hp.nb, s>>>(
        row, col, nnz, cbits, hp.R, hp.T, st->gcount, with_max ? st->max_rc : nullptr);
  }
  PS_LAUNCH_OK();
  ls_derive_kernel<<<1, 1024, 0, s>>>(st->gcount, hp.T, hp.lp, st->bucket_off, st->ghist, st->max_bucket, 0, LsChunkOut{});
  PS_LAUNCH_OK();
  return PS_OK;
}

// `pre` / `pre_max_bucket`: the inspection ps_coo_inspect ran on these very arrays (nullable / -1): saves the
// histogram read and one host synchronisation.
static int hybrid_coalesce(const int64_t *row, const int64_t *col, int64_t nnz, int rbits, int cbits, int64_t M,
                           const void *value, int dtype, bool dedup, int reduce, int64_t *out_row, int64_t *out_col,
                           void *out_value, int64_t *out_rowptr, int64_t *host_nnz_out, bool *used,
                           const LsInspect *pre, int64_t pre_max_bucket, cudaStream_t s) {
  *used = false;
  HybridPlan hp;
  if (!hybrid_plan(nnz, M, rbits, cbits, &hp)) return PS_OK;
  int T = hp.T, R = hp.R;
  int nb = hp.nb;
  Scratch scratch(s);
  const LsInspect *st = pre;
  unsigned int h_max = 0;
  if (pre && pre_max_bucket >= 0) {
    h_max = (unsigned int)pre_max_bucket;
  } else {
    PS_SCRATCH(own, scratch, LsInspect, 1);
    PS_TRY(hybrid_inspect(row, col, nnz, cbits, hp, own, false, s));
    PS_TRY(read_back(&h_max, own->max_bucket, sizeof(h_max), s));
    st = own;
  }
  // skewed rows: oversized buckets are re-partitioned (refine.cuh); PS_REFINE=0 / the old bucket kernel: the plain LSD
  // sort, which handles any distribution
  const bool refine = h_max > (unsigned)ls_cap();
  if (refine && (!refine_variant() || ls_variant() == 1)) return PS_OK;
  const unsigned int *bucket_off = st->bucket_off;
  const unsigned long long *ghist = st->ghist;
  if (refine && refine_variant() == 2) {
    // PS_REFINE=2 (measured, NOT the default): start from COARSER buckets (3 fewer top bits, as far as the local key
    // still fits 32 bits).  Then every bucket goes through the refinement, whose greedy packing also merges the many
    // light row ranges of a power-law graph into full buckets: 18.5 K final buckets on the R-MAT stream instead of
    // 30.8 K - but twice the entries reach level 2, and the coalesce takes 3.90 ms against 3.67.  The coarse counts
    // are sums of the inspection's fine ones.
    const int tmin = hp.K - 32 > 1 ? hp.K - 32 : 1;
    int T2 = T - 3 > tmin ? T - 3 : tmin;
    if (T2 < T) {
      PS_SCRATCH(co, scratch, LsInspect, 1);
      const int nb2 = 1 << T2;
      coarsen_counts_kernel<<<(unsigned)ceil_div(nb2, 256), 256, 0, s>>>(st->gcount, T - T2, nb2, co->gcount);
      PS_LAUNCH_OK();
      const SortPlan gp = make_sort_plan(T2, rs_geometry<uint64_t, 1>().rbits);
      LsPlan lp2;
      lp2.passes = gp.passes;
      for (int p = 0; p < 2; ++p) { lp2.pbegin[p] = p < gp.passes ? gp.begin[p] : 0; lp2.pbits[p] = p < gp.passes ? gp.bits[p] : 0; }
      ls_derive_kernel<<<1, 1024, 0, s>>>(co->gcount, T2, lp2, co->bucket_off, co->ghist, co->max_bucket, 0, LsChunkOut{});
      PS_LAUNCH_OK();
      T = T2;
      R = hp.K - T2;
      nb = nb2;
      bucket_off = co->bucket_off;
      ghist = co->ghist;
    }
  }

  // global passes on key bits [R, K): the last one keeps only the low R bits of the key
  PS_SCRATCH(k0, scratch, uint64_t, nnz);
  PS_SCRATCH(p0, scratch, uint32_t, nnz);
  PS_SCRATCH(lowkeys, scratch, uint32_t, nnz);
  PS_SCRATCH(lowvals, scratch, uint32_t, nnz);
  KeySrc ksrc{row, col, KSRC_PACK_COO, cbits, 0, (rbits <= 32 && cbits <= 32) ? 1 : 0};
  PaySrc psrc = value ? PaySrc{value, pay_src_mode(dtype)} : PaySrc{nullptr, PSRC_NONE};
  PayDst pdst = value ? PayDst{lowvals, PDST_U32} : PayDst{nullptr, PDST_NONE};
  PS_TRY(radix_sort_pairs<uint64_t>(ksrc, psrc, nnz, T, k0, nullptr, p0, nullptr, KeyDst{lowkeys, nullptr, KDST_LOW32, R},
                                    pdst, scratch, s, R, ghist));

  // buckets past the one holding the largest possible key are empty: only the used ones are processed (the last of
  // them owns the rows up to M)
  {
    const uint64_t top = M > 0 ? ((((uint64_t)(M - 1)) << cbits) | ((uint64_t(1) << cbits) - 1)) >> R : 0;
    if ((int64_t)top + 1 < nb) nb = (int)top + 1;
  }
  PS_SCRATCH(d_nnz, scratch, long long, 1);
  int rc = PS_ERR_ARG;
  if (refine) {
    L2Desc desc{};
    int nf = 0;
    bool ok = false;
    PS_TRY(refine_buckets(lowkeys, value ? lowvals : nullptr, nnz, bucket_off, nb, R, (unsigned)L2_CAP, scratch, s, &desc,
                          &nf, &ok));
    if (!ok) return PS_OK;  // one key with more duplicates than a bucket holds: the LSD sort
    PS_SCRATCH(dstate, scratch, uint64_t, (size_t)nf + 1);
    PS_CUDA(cudaMemsetAsync(dstate, 0, sizeof(uint64_t) * ((size_t)nf + 1), s));
#define PS_LD_ARGS lowkeys, lowvals, bucket_off, desc, nf, R, cbits, M, out_row, out_col, out_value, out_rowptr, dstate, d_nnz, s
#define PS_LD(T_)                                                                                \
  do {                                                                                           \
    if (dedup) { PS_DISPATCH_REDUCE(reduce, RED, rc = (ls2_launch_desc<T_, RED, true, true>(PS_LD_ARGS))); } \
    else rc = ls2_launch_desc<T_, PS_SUM, true, false>(PS_LD_ARGS);                              \
  } while (0)
    if (!value) {
      rc = dedup ? ls2_launch_desc<float, PS_SUM, false, true>(lowkeys, nullptr, bucket_off, desc, nf, R, cbits, M, out_row,
                                                               out_col, nullptr, out_rowptr, dstate, d_nnz, s)
                 : ls2_launch_desc<float, PS_SUM, false, false>(lowkeys, nullptr, bucket_off, desc, nf, R, cbits, M, out_row,
                                                                out_col, nullptr, out_rowptr, dstate, d_nnz, s);
    } else {
      switch (dtype) {
        case PS_F32: PS_LD(float); break;
        case PS_I32: PS_LD(int32_t); break;
        case PS_BF16: PS_LD(__nv_bfloat16); break;
        case PS_F16: PS_LD(__half); break;
        default: break;
      }
    }
#undef PS_LD
#undef PS_LD_ARGS
    PS_TRY(rc);
    if (dedup) {
      long long h_nnz = 0;
      PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
      *host_nnz_out = (int64_t)h_nnz;
    }
    *used = true;
    return PS_OK;
  }
  PS_SCRATCH(state, scratch, uint64_t, (size_t)nb + 1);
  PS_CUDA(cudaMemsetAsync(state, 0, sizeof(uint64_t) * ((size_t)nb + 1), s));
#define PS_LS_ARGS lowkeys, lowvals, bucket_off, nb, R, cbits, M, out_row, out_col, out_value, out_rowptr, state, d_nnz, s
#define PS_LS(T_)                                                                          \
  do {                                                                                     \
    if (dedup) { PS_DISPATCH_REDUCE(reduce, RED, rc = (ls_launch<T_, RED, true, true>(PS_LS_ARGS))); } \
    else rc = ls_launch<T_, PS_SUM, true, false>(PS_LS_ARGS);                              \
  } while (0)
  if (!value) {
    rc = dedup ? ls_launch<float, PS_SUM, false, true>(lowkeys, nullptr, bucket_off, nb, R, cbits, M, out_row, out_col,
                                                       nullptr, out_rowptr, state, d_nnz, s)
               : ls_launch<float, PS_SUM, false, false>(lowkeys, nullptr, bucket_off, nb, R, cbits, M, out_row, out_col,
                                                        nullptr, out_rowptr, state, d_nnz, s);
  } else {
    switch (dtype) {
      case PS_F32: PS_LS(float); break;
      case PS_I32: PS_LS(int32_t); break;
      case PS_BF16: PS_LS(__nv_bfloat16); break;
      case PS_F16: PS_LS(__half); break;
      default: break;
    }
  }
#undef PS_LS
#undef PS_LS_ARGS
  PS_TRY(rc);
  if (dedup) {
    long long h_nnz = 0;
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
    *host_nnz_out = (int64_t)h_nnz;
  }
  *used = true;
  return PS_OK;
}

// ------------------------------------------------------------------------------------------------
// Hybrid coalesce for keys whose bucket-local part does not fit 32 bits (row shards of very wide matrices: papers100M
// shards have 51- to 53-bit keys).  Same scheme - top T key bits by global passes, buckets finished in shared memory -
// with three differences: T may reach 18 bits (9-bit digits, two passes), the global passes keep the FULL 64-bit key
// (the bucket kernel takes the low R <= 36 bits itself), and, because 2^T counters no longer fit a shared-memory
// histogram, the bucket offsets are read off the bucket-sorted keys by boundary detection AFTER the passes.  One host
// read (the largest bucket) decides; oversized buckets (skew) send the call to the LSD pipeline.
constexpr int LSW_MAX_T = 18;
constexpr int LSW_MAX_R = 36;

__global__ void __launch_bounds__(256)
lsw_bounds_kernel(const uint64_t *__restrict__ keys, int64_t n, int R, int nb, unsigned int *__restrict__ bucket_off) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < n + 1; i0 += 4 * stride) {  // warp-uniform trip count
    // four strided elements per thread and round: their loads are in flight together (one at a time the kernel
    // ran at 2.6 TB/s)
    int64_t b[4], pb0[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride + threadIdx.x;
      // bucket of element i (n: one past the last bucket) and of its predecessor (-1 before the first)
      b[u] = i < n ? (int64_t)(ld_stream(keys + i) >> R) : (i == n ? (int64_t)nb : (int64_t)nb + 1);
      pb0[u] = 0;
      if ((threadIdx.x & 31u) == 0)
        pb0[u] = i == 0 ? -1 : (i <= n ? (int64_t)(ld_stream(keys + i - 1) >> R) : (int64_t)nb + 1);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride + threadIdx.x;
      int64_t pb = __shfl_up_sync(0xffffffffu, b[u], 1);
      if ((threadIdx.x & 31u) == 0) pb = pb0[u];
      if (i <= n && b[u] != pb)
        for (int64_t q = pb + 1; q <= b[u] && q <= nb; ++q) bucket_off[q] = (unsigned int)i;  // i opens buckets (pb, b]
    }
  }
}

__global__ void __launch_bounds__(256)
lsw_max_kernel(const unsigned int *__restrict__ bucket_off, int nb, unsigned int *__restrict__ max_bucket) {
  unsigned int mx = 0;
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += gridDim.x * blockDim.x) {
    const unsigned int c = bucket_off[b + 1] - bucket_off[b];
    mx = c > mx ? c : mx;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const unsigned int o = __shfl_xor_sync(0xffffffffu, mx, d);
    mx = o > mx ? o : mx;
  }
  if ((threadIdx.x & 31u) == 0 && mx) atomicMax(max_bucket, mx);
}

static int hybrid_coalesce_wide(KeySrc ksrc, int64_t nnz, int rbits, int cbits, int64_t M,
                                const void *value, int dtype, bool dedup, int reduce, int64_t *out_row, int64_t *out_col,
                                void *out_value, int64_t *out_rowptr, int64_t *host_nnz_out, bool *used, cudaStream_t s) {
  *used = false;
  const int K = rbits + cbits;
  if (!hybrid_variant() || ls_variant() == 1 || nnz < (int64_t(1) << 16) || K <= 32) return PS_OK;
  int T = K - LSW_MAX_R > 1 ? K - LSW_MAX_R : 1;
  const int tmax = LSW_MAX_T < rbits ? LSW_MAX_T : rbits;
  while (T <= tmax && (double)nnz * (double)(int64_t(1) << (rbits - T)) / (double)(M > 0 ? M : 1) > 0.9 * L2_CAP) ++T;
  if (T > tmax) return PS_OK;
  const int R = K - T;
  if (rs_num_passes<uint64_t>(T) > 2) return PS_OK;
  int nb = 1 << T;
  Scratch scratch(s);
  PS_SCRATCH(k0, scratch, uint64_t, nnz);
  PS_SCRATCH(p0, scratch, uint32_t, nnz);
  PS_SCRATCH(skeys, scratch, uint64_t, nnz);
  PS_SCRATCH(svals, scratch, uint32_t, nnz);
  PaySrc psrc = value ? PaySrc{value, pay_src_mode(dtype)} : PaySrc{nullptr, PSRC_NONE};
  PayDst pdst = value ? PayDst{svals, PDST_U32} : PayDst{nullptr, PDST_NONE};
  PS_TRY(radix_sort_pairs<uint64_t>(ksrc, psrc, nnz, T, k0, nullptr, p0, nullptr, KeyDst{skeys, nullptr, KDST_KEYS, 0}, pdst,
                                    scratch, s, R, nullptr));
  PS_SCRATCH(bucket_off, scratch, unsigned int, (size_t)nb + 2);
  PS_SCRATCH(d_max, scratch, unsigned int, 1);
  PS_CUDA(cudaMemsetAsync(d_max, 0, sizeof(unsigned int), s));
  lsw_bounds_kernel<<<sm_count() * 8, 256, 0, s>>>(skeys, nnz, R, nb, bucket_off);
  PS_LAUNCH_OK();
  lsw_max_kernel<<<64, 256, 0, s>>>(bucket_off, nb, d_max);
  PS_LAUNCH_OK();
  unsigned int h_max = 0;
  PS_TRY(read_back(&h_max, d_max, sizeof(h_max), s));
  if (h_max > (unsigned)L2_CAP) return PS_OK;  // skewed rows: the plain LSD sort handles any distribution
  {
    const uint64_t top = M > 0 ? ((((uint64_t)(M - 1)) << cbits) | ((uint64_t(1) << cbits) - 1)) >> R : 0;
    if ((int64_t)top + 1 < nb) nb = (int)top + 1;
  }
  PS_SCRATCH(state, scratch, uint64_t, (size_t)nb + 1);
  PS_CUDA(cudaMemsetAsync(state, 0, sizeof(uint64_t) * ((size_t)nb + 1), s));
  PS_SCRATCH(d_nnz, scratch, long long, 1);
  int rc = PS_ERR_ARG;
#define PS_LW_ARGS skeys, svals, bucket_off, nb, R, cbits, M, out_row, out_col, out_value, out_rowptr, state, d_nnz, s
#define PS_LW(T_)                                                                                             \
  do {                                                                                                        \
    if (dedup) { PS_DISPATCH_REDUCE(reduce, RED, rc = (ls2_launch<T_, RED, true, true, uint64_t>(PS_LW_ARGS))); } \
    else rc = ls2_launch<T_, PS_SUM, true, false, uint64_t>(PS_LW_ARGS);                                      \
  } while (0)
  if (!value) {
    rc = dedup ? ls2_launch<float, PS_SUM, false, true, uint64_t>(skeys, nullptr, bucket_off, nb, R, cbits, M, out_row,
                                                                  out_col, nullptr, out_rowptr, state, d_nnz, s)
               : ls2_launch<float, PS_SUM, false, false, uint64_t>(skeys, nullptr, bucket_off, nb, R, cbits, M, out_row,
                                                                   out_col, nullptr, out_rowptr, state, d_nnz, s);
  } else {
    switch (dtype) {
      case PS_F32: PS_LW(float); break;
      case PS_I32: PS_LW(int32_t); break;
      case PS_BF16: PS_LW(__nv_bfloat16); break;
      case PS_F16: PS_LW(__half); break;
      default: break;
    }
  }
#undef PS_LW
#undef PS_LW_ARGS
  PS_TRY(rc);
  if (dedup) {
    long long h_nnz = 0;
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
    *host_nnz_out = (int64_t)h_nnz;
  }
  *used = true;
  return PS_OK;
}

// ------------------------------------------------------------------------------------------------
// sort + (optionally) dedup, shared by ps_sort_coo and ps_coalesce
template <typename KeyT>
int sort_impl(const int64_t *row, const int64_t *col, int64_t nnz, KeyGeom g, const void *value, int dtype,
              int64_t width, bool dedup, int reduce, int64_t *out_row, int64_t *out_col, void *out_value,
              int64_t *out_perm, int64_t *host_nnz_out, int64_t M, int64_t *out_rowptr, cudaStream_t s,
              const LsInspect *pre = nullptr, int64_t pre_max_bucket = -1, const KeySrc *src32 = nullptr) {
  // src32: the tuples as 32-bit (global row, col) words with a row offset (what the peer exchange delivers) instead
  // of the int64 arrays; everything downstream of the first key load is the same
  KeySrc ksrc = src32 ? *src32 : KeySrc{row, col, KSRC_PACK_COO, g.cbits, 0, (g.rbits <= 32 && g.cbits <= 32) ? 1 : 0};
  const bool fast = value && fast_value(dtype, width) && !out_perm;
  if (sizeof(KeyT) == 8 && (fast || !value) && !out_perm) {  // hybrid path (sort only, or sort + dedup)
    bool used = false;
    int64_t ignored = 0;
    if (!src32) {  // (its bucket histogram reads the int64 arrays)
      PS_TRY(hybrid_coalesce(row, col, nnz, g.rbits, g.cbits, M, value, dtype, dedup, reduce, out_row, out_col, out_value,
                             out_rowptr, dedup ? host_nnz_out : &ignored, &used, pre, pre_max_bucket, s));
      if (used) return PS_OK;
    }
    PS_TRY(hybrid_coalesce_wide(ksrc, nnz, g.rbits, g.cbits, M, value, dtype, dedup, reduce, out_row, out_col, out_value,
                                out_rowptr, dedup ? host_nnz_out : &ignored, &used, s));
    if (used) return PS_OK;
  }
  Scratch scratch(s);
  const bool need_perm = (value && !fast) || out_perm;
  PS_SCRATCH(k0, scratch, KeyT, nnz);
  PS_SCRATCH(k1, scratch, KeyT, nnz);
  PS_SCRATCH(p0, scratch, uint32_t, nnz);
  PS_SCRATCH(p1, scratch, uint32_t, nnz);
  const int passes = rs_num_passes<KeyT>(g.bits);
  PaySrc psrc = fast ? PaySrc{value, pay_src_mode(dtype)} : PaySrc{nullptr, PSRC_IOTA};

  if (!dedup) {
    // last pass unpacks keys into out_row/out_col and drops the payload where it belongs
    KeyDst kd{out_row, out_col, KDST_UNPACK_COO, g.cbits};
    PayDst pd{nullptr, PDST_NONE};
    uint32_t *perm32 = nullptr;
    if (fast) pd = PayDst{out_value, pay_dst_mode(dtype)};
    else if (out_perm) pd = PayDst{out_perm, PDST_I64};
    else if (need_perm) {
      perm32 = (passes & 1) ? p0 : p1;  // the buffer the last pass would have written
      pd = PayDst{perm32, PDST_U32};
    }
    PS_TRY(radix_sort_pairs<KeyT>(ksrc, psrc, nnz, g.bits, k0, k1, p0, p1, kd, pd, scratch, s));
    if (value && !fast) {
      if (out_perm) PS_TRY(gather_rows_launch<int64_t>(value, width * (int64_t)dtype_size(dtype), out_perm, nnz, out_value, s));
      else PS_TRY(gather_rows_launch<uint32_t>(value, width * (int64_t)dtype_size(dtype), perm32, nnz, out_value, s));
    }
    return PS_OK;
  }

  // dedup: the sorted pairs land in the scratch buffers, then one pass compacts + reduces
  KeyT *ks = (passes & 1) ? k0 : k1;
  uint32_t *psorted = (passes & 1) ? p0 : p1;
  PS_TRY(radix_sort_pairs<KeyT>(ksrc, psrc, nnz, g.bits, k0, k1, p0, p1, KeyDst{ks, nullptr, KDST_KEYS, 0},
                                PayDst{psorted, PDST_U32}, scratch, s));
  PS_SCRATCH(d_nnz, scratch, long long, 1);
  KeySrc sk{ks, nullptr, KSRC_KEYS, 0};
  PaySrc sp{psorted, PSRC_U32};
  long long h_nnz = 0;
  if (!value) {
    PS_TRY(dedup_dispatch<KeyT>(sk, sp, nnz, g.cbits, M, DD_INDEX_ONLY, dtype, reduce, out_row, out_col, nullptr, nullptr,
                                out_rowptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else if (fast) {
    PS_TRY(dedup_dispatch<KeyT>(sk, sp, nnz, g.cbits, M, DD_FAST, dtype, reduce, out_row, out_col, out_value, nullptr,
                                out_rowptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else {
    PS_SCRATCH(run_start, scratch, int64_t, nnz + 1);
    PS_TRY(dedup_dispatch<KeyT>(sk, sp, nnz, g.cbits, M, DD_PTR, dtype, reduce, out_row, out_col, nullptr, run_start,
                                out_rowptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
    PS_TRY(segment_reduce_launch(value, dtype, width, run_start, h_nnz, reduce, out_value, psorted, nnz, s));
  }
  *host_nnz_out = (int64_t)h_nnz;
  return PS_OK;
}

}  // namespace ps

using namespace ps;

extern "C" {

// debug builds (make TIMELINE=1): cycles thread 0 of every radix-pass CTA of the COO sort spent per phase; resets
int ps_debug_timeline(unsigned long long *out16) {
#ifdef PS_RS_TIMELINE
  unsigned long long z[16] = {0};
  if (cudaMemcpyFromSymbol(out16, ps::g_rs_timeline, sizeof(z)) != cudaSuccess) return PS_ERR_CUDA;
  if (cudaMemcpyToSymbol(ps::g_rs_timeline, z, sizeof(z)) != cudaSuccess) return PS_ERR_CUDA;
  return PS_OK;
#else
  (void)out16;
  return PS_ERR_ARG;
#endif
}

int ps_sort_coo(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N, const void *value,
                int dtype, int64_t width, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *out_perm,
                void *stream) {
  PS_CHECK(nnz >= 0 && M >= 0 && N >= 0, "sort_coo: negative size");
  PS_CHECK(nnz < (int64_t(1) << 32), "sort_coo: nnz=%lld exceeds the 2^32-1 entries per call this build supports",
           (long long)nnz);
  PS_TRY(check_value_args(value, out_value, dtype, width));
  if (nnz == 0) return PS_OK;
  KeyGeom g;
  PS_TRY(key_geometry(M, N, &g));
  cudaStream_t s = (cudaStream_t)stream;
  if (g.bits <= 32)
    return sort_impl<uint32_t>(row, col, nnz, g, value, dtype, width, false, PS_SUM, out_row, out_col, out_value,
                               out_perm, nullptr, M, nullptr, s);
  return sort_impl<uint64_t>(row, col, nnz, g, value, dtype, width, false, PS_SUM, out_row, out_col, out_value, out_perm,
                             nullptr, M, nullptr, s);
}

int64_t ps_coo_inspect_bytes(void) { return (int64_t)sizeof(LsInspect); }

int ps_coo_inspect(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N, void *state,
                   int64_t *host_out, void *stream) {
  PS_CHECK(nnz >= 0 && M >= 0 && N >= 0, "coo_inspect: negative size");
  PS_CHECK(host_out != nullptr, "coo_inspect: host_out is NULL");
  host_out[0] = host_out[1] = host_out[2] = -1;
  if (nnz == 0) return PS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  KeyGeom g;
  HybridPlan hp;
  if (state == nullptr || nnz >= (int64_t(1) << 32) || key_geometry(M, N, &g) != PS_OK || g.bits <= 32 ||
      !hybrid_plan(nnz, M, g.rbits, g.cbits, &hp))
    return ps_index_max(row, nnz, col, nnz, host_out, stream);  // the plain validation pass; host_out[2] stays -1
  LsInspect *st = reinterpret_cast<LsInspect *>(state);
  PS_TRY(hybrid_inspect(row, col, nnz, g.cbits, hp, st, true, s));
  struct { long long max_rc[2]; unsigned int max_bucket[2]; } h;
  PS_TRY(read_back(&h, st, sizeof(h), s));
  host_out[0] = (int64_t)h.max_rc[0];
  host_out[1] = (int64_t)h.max_rc[1];
  host_out[2] = (int64_t)h.max_bucket[0];
  return PS_OK;
}

int ps_coalesce(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N, const void *value, int dtype,
                int64_t width, int reduce, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *out_rowptr,
                int64_t *host_nnz_out, const void *inspect, int64_t inspect_max_bucket, void *stream) {
  PS_CHECK(nnz >= 0 && M >= 0 && N >= 0, "coalesce: negative size");
  PS_CHECK(nnz < (int64_t(1) << 32), "coalesce: nnz=%lld exceeds the 2^32-1 entries per call this build supports",
           (long long)nnz);
  PS_CHECK(host_nnz_out != nullptr, "coalesce: host_nnz_out is NULL");
  PS_CHECK(reduce >= PS_SUM && reduce <= PS_MAX, "coalesce: bad reduce %d", reduce);
  PS_TRY(check_value_args(value, out_value, dtype, width));
  *host_nnz_out = 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (nnz == 0) {
    if (out_rowptr) PS_CUDA(cudaMemsetAsync(out_rowptr, 0, sizeof(int64_t) * (size_t)(M + 1), s));
    return PS_OK;
  }
  KeyGeom g;
  PS_TRY(key_geometry(M, N, &g));
  if (g.bits <= 32)
    return sort_impl<uint32_t>(row, col, nnz, g, value, dtype, width, true, reduce, out_row, out_col, out_value, nullptr,
                               host_nnz_out, M, out_rowptr, s);
  return sort_impl<uint64_t>(row, col, nnz, g, value, dtype, width, true, reduce, out_row, out_col, out_value, nullptr,
                             host_nnz_out, M, out_rowptr, s, reinterpret_cast<const LsInspect *>(inspect),
                             inspect ? inspect_max_bucket : -1);
}

int ps_coalesce_u32(const uint32_t *row, const uint32_t *col, int64_t row_offset, int64_t nnz, int64_t M, int64_t N,
                    const void *value, int dtype, int64_t width, int reduce, int64_t *out_row, int64_t *out_col,
                    void *out_value, int64_t *out_rowptr, int64_t *host_nnz_out, void *stream) {
  PS_CHECK(nnz >= 0 && M >= 0 && N >= 0 && row_offset >= 0, "coalesce_u32: negative size");
  PS_CHECK(nnz < (int64_t(1) << 32), "coalesce_u32: nnz=%lld exceeds the 2^32-1 entries per call this build supports",
           (long long)nnz);
  PS_CHECK(row_offset + M <= (int64_t(1) << 32) && N <= (int64_t(1) << 32),
           "coalesce_u32: row_offset + M = %lld or N = %lld exceeds the 2^32 ids a 32-bit word carries",
           (long long)(row_offset + M), (long long)N);
  PS_CHECK(host_nnz_out != nullptr, "coalesce_u32: host_nnz_out is NULL");
  PS_CHECK(reduce >= PS_SUM && reduce <= PS_MAX, "coalesce_u32: bad reduce %d", reduce);
  PS_TRY(check_value_args(value, out_value, dtype, width));
  *host_nnz_out = 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (nnz == 0) {
    if (out_rowptr) PS_CUDA(cudaMemsetAsync(out_rowptr, 0, sizeof(int64_t) * (size_t)(M + 1), s));
    return PS_OK;
  }
  KeyGeom g;
  PS_TRY(key_geometry(M, N, &g));
  KeySrc src{row, col, KSRC_PACK_COO32, g.cbits, row_offset};
  if (g.bits <= 32)
    return sort_impl<uint32_t>(nullptr, nullptr, nnz, g, value, dtype, width, true, reduce, out_row, out_col, out_value,
                               nullptr, host_nnz_out, M, out_rowptr, s, nullptr, -1, &src);
  return sort_impl<uint64_t>(nullptr, nullptr, nnz, g, value, dtype, width, true, reduce, out_row, out_col, out_value,
                             nullptr, host_nnz_out, M, out_rowptr, s, nullptr, -1, &src);
}

int ps_coalesce_sorted(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N, const void *value,
                       int dtype, int64_t width, int reduce, int64_t *out_row, int64_t *out_col, void *out_value,
                       int64_t *host_nnz_out, void *stream) {
  PS_CHECK(nnz >= 0, "coalesce_sorted: negative size");
  // adjacent entries are compared as (row << 32 | col): ids of 2^32 and beyond would merge falsely / be masked
  PS_CHECK(M >= 0 && N >= 0 && M <= (int64_t(1) << 32) && N <= (int64_t(1) << 32),
           "coalesce_sorted: sparse sizes (%lld, %lld) exceed the 2^32 ids per dimension of the packed comparison",
           (long long)M, (long long)N);
  PS_CHECK(host_nnz_out != nullptr, "coalesce_sorted: host_nnz_out is NULL");
  PS_CHECK(reduce >= PS_SUM && reduce <= PS_MAX, "coalesce_sorted: bad reduce %d", reduce);
  PS_TRY(check_value_args(value, out_value, dtype, width));
  *host_nnz_out = 0;
  if (nnz == 0) return PS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  const int cbits = 32;  // row, col < 2^32: checked above
  KeySrc sk{row, col, KSRC_PACK_COO, cbits};
  PS_SCRATCH(d_nnz, scratch, long long, 1);
  long long h_nnz = 0;
  if (!value) {
    PS_TRY(dedup_dispatch<uint64_t>(sk, PaySrc{nullptr, PSRC_IOTA}, nnz, cbits, 0, DD_INDEX_ONLY, dtype, reduce, out_row,
                                    out_col, nullptr, nullptr, nullptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else if (fast_value(dtype, width)) {
    PS_TRY(dedup_dispatch<uint64_t>(sk, PaySrc{value, pay_src_mode(dtype)}, nnz, cbits, 0, DD_FAST, dtype, reduce, out_row,
                                    out_col, out_value, nullptr, nullptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else {
    PS_SCRATCH(run_start, scratch, int64_t, nnz + 1);
    PS_TRY(dedup_dispatch<uint64_t>(sk, PaySrc{nullptr, PSRC_IOTA}, nnz, cbits, 0, DD_PTR, dtype, reduce, out_row, out_col,
                                    nullptr, run_start, nullptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
    PS_TRY(segment_reduce_launch(value, dtype, width, run_start, h_nnz, reduce, out_value, nullptr, nnz, s));
  }
  *host_nnz_out = (int64_t)h_nnz;
  return PS_OK;
}

}  // extern "C"
