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
#include "common.cuh"
#include "radix_sort.cuh"
#include "segment.cuh"

namespace ps {

enum { DD_INDEX_ONLY = 0, DD_FAST = 1, DD_PTR = 2 };

template <typename T> __device__ __forceinline__ T pay_as(uint32_t p);
template <> __device__ __forceinline__ float pay_as<float>(uint32_t p) { return __uint_as_float(p); }
template <> __device__ __forceinline__ int32_t pay_as<int32_t>(uint32_t p) { return (int32_t)p; }
template <> __device__ __forceinline__ __nv_bfloat16 pay_as<__nv_bfloat16>(uint32_t p) {
  return __ushort_as_bfloat16((unsigned short)p);
}
template <> __device__ __forceinline__ __half pay_as<__half>(uint32_t p) { return __ushort_as_half((unsigned short)p); }

// One tile = THREADS*ITEMS consecutive sorted entries; thread t owns ITEMS consecutive ones.
template <typename KeyT, typename T, int RED, int MODE, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
dedup_kernel(KeySrc ksrc, PaySrc psrc, int64_t n, int cbits, int64_t *__restrict__ out_row,
             int64_t *__restrict__ out_col, T *__restrict__ out_value, int64_t *__restrict__ run_start,
             uint64_t *state /*[0]=ticket,[1..]=tiles*/, long long *d_nnz_out) {
  constexpr int TILE = THREADS * ITEMS;
  using A = typename Acc<T>::type;
  __shared__ KeyT s_key[TILE + 1];  // s_key[0] = key preceding the tile
  __shared__ uint32_t s_pay[MODE == DD_FAST ? TILE : 1];
  __shared__ int64_t warp_sums[33];
  __shared__ int64_t s_tile, s_excl;
  if (threadIdx.x == 0) s_tile = (int64_t)atomicAdd((unsigned long long *)state, 1ull);
  __syncthreads();
  const int64_t tile = s_tile;
  const int64_t tile_base = tile * TILE;
  const int tile_n = (int)((n - tile_base) < TILE ? (n - tile_base) : TILE);

  for (int j = threadIdx.x; j < tile_n; j += THREADS) {
    s_key[j + 1] = load_key<KeyT>(ksrc, tile_base + j);
    if (MODE == DD_FAST) s_pay[j] = load_pay(psrc, tile_base + j);
  }
  if (threadIdx.x == 0) s_key[0] = tile_base > 0 ? load_key<KeyT>(ksrc, tile_base - 1) : KeyT(0);
  __syncthreads();

  const int j0 = threadIdx.x * ITEMS;
  unsigned heads = 0;
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = j0 + k;
    const bool h = j < tile_n && ((tile_base + j == 0) || s_key[j + 1] != s_key[j]);
    heads |= (unsigned)h << k;
    cnt += h;
  }
  int64_t total;
  int64_t excl = block_exclusive_scan<int64_t, THREADS>((int64_t)cnt, warp_sums, &total);
  if (threadIdx.x < 32) {
    const uint64_t p = warp_lookback(state + 1, tile, (uint64_t)total);
    if (threadIdx.x == 0) s_excl = (int64_t)p;
  }
  __syncthreads();
  excl += s_excl;
  if (tile_base + tile_n == n && threadIdx.x == 0) {
    *d_nnz_out = (long long)(s_excl + total);
    if (MODE == DD_PTR) run_start[s_excl + total] = n;
  }

  const KeyT cmask = (KeyT)((uint64_t(1) << cbits) - 1);
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    if (!((heads >> k) & 1u)) continue;
    const int j = j0 + k;
    const KeyT key = s_key[j + 1];
    const int64_t o = excl++;
    out_row[o] = (int64_t)((uint64_t)key >> cbits);
    out_col[o] = (int64_t)(key & cmask);
    if (MODE == DD_PTR) run_start[o] = tile_base + j;
    if (MODE == DD_FAST) {
      A acc = to_acc<T>(pay_as<T>(s_pay[j]));
      int jj = j + 1;
      while (jj < tile_n && s_key[jj + 1] == key) {
        acc = Red<RED, A>::op(acc, to_acc<T>(pay_as<T>(s_pay[jj])));
        ++jj;
      }
      int64_t len = jj - j;
      if (jj == tile_n) {  // the run may continue past the tile: finish it from global memory
        int64_t g = tile_base + tile_n;
        while (g < n && load_key<KeyT>(ksrc, g) == key) {
          acc = Red<RED, A>::op(acc, to_acc<T>(pay_as<T>(load_pay(psrc, g))));
          ++g;
          ++len;
        }
      }
      out_value[o] = from_acc<T>(Red<RED, A>::finish(acc, len));
    }
  }
}

template <typename KeyT, typename T, int RED, int MODE>
int dedup_launch(KeySrc ksrc, PaySrc psrc, int64_t n, int cbits, int64_t *out_row, int64_t *out_col, void *out_value,
                 int64_t *run_start, long long *d_nnz, Scratch &scratch, cudaStream_t s) {
  constexpr int THREADS = 256, ITEMS = 8;
  const int64_t tiles = ceil_div(n, THREADS * ITEMS);
  PS_SCRATCH(state, scratch, uint64_t, tiles + 1);
  PS_CUDA(cudaMemsetAsync(state, 0, sizeof(uint64_t) * (size_t)(tiles + 1), s));
  {
    ProfScope prof(PROF_DEDUP, s);
    dedup_kernel<KeyT, T, RED, MODE, THREADS, ITEMS><<<(unsigned)tiles, THREADS, 0, s>>>(
        ksrc, psrc, n, cbits, out_row, out_col, reinterpret_cast<T *>(out_value), run_start, state, d_nnz);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

// values that ride along as the 32-bit payload: one element of <= 4 bytes per entry
static bool fast_value(int dtype, int64_t width) {
  return width == 1 && (dtype == PS_F32 || dtype == PS_I32 || dtype == PS_BF16 || dtype == PS_F16);
}
static int pay_src_mode(int dtype) { return dtype_size(dtype) == 4 ? PSRC_U32 : PSRC_U16; }
static int pay_dst_mode(int dtype) { return dtype_size(dtype) == 4 ? PDST_U32 : PDST_U16; }

template <typename KeyT>
int dedup_dispatch(KeySrc ksrc, PaySrc psrc, int64_t n, int cbits, int mode, int dtype, int reduce, int64_t *out_row,
                   int64_t *out_col, void *out_value, int64_t *run_start, long long *d_nnz, Scratch &scratch,
                   cudaStream_t s) {
  if (mode == DD_INDEX_ONLY)
    return dedup_launch<KeyT, float, PS_SUM, DD_INDEX_ONLY>(ksrc, psrc, n, cbits, out_row, out_col, nullptr, nullptr,
                                                            d_nnz, scratch, s);
  if (mode == DD_PTR)
    return dedup_launch<KeyT, float, PS_SUM, DD_PTR>(ksrc, psrc, n, cbits, out_row, out_col, nullptr, run_start, d_nnz,
                                                     scratch, s);
#define PS_DD(T)                                                                                              \
  PS_DISPATCH_REDUCE(reduce, RED, return (dedup_launch<KeyT, T, RED, DD_FAST>(ksrc, psrc, n, cbits, out_row,   \
                                                                             out_col, out_value, nullptr, d_nnz, \
                                                                             scratch, s)))
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
// sort + (optionally) dedup, shared by ps_sort_coo and ps_coalesce
template <typename KeyT>
int sort_impl(const int64_t *row, const int64_t *col, int64_t nnz, KeyGeom g, const void *value, int dtype,
              int64_t width, bool dedup, int reduce, int64_t *out_row, int64_t *out_col, void *out_value,
              int64_t *out_perm, int64_t *host_nnz_out, cudaStream_t s) {
  Scratch scratch(s);
  const bool fast = value && fast_value(dtype, width) && !out_perm;
  const bool need_perm = (value && !fast) || out_perm;
  PS_SCRATCH(k0, scratch, KeyT, nnz);
  PS_SCRATCH(k1, scratch, KeyT, nnz);
  PS_SCRATCH(p0, scratch, uint32_t, nnz);
  PS_SCRATCH(p1, scratch, uint32_t, nnz);
  const int passes = rs_num_passes<KeyT>(g.bits);
  KeySrc ksrc{row, col, KSRC_PACK_COO, g.cbits};
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
    PS_TRY(dedup_dispatch<KeyT>(sk, sp, nnz, g.cbits, DD_INDEX_ONLY, dtype, reduce, out_row, out_col, nullptr, nullptr,
                                d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else if (fast) {
    PS_TRY(dedup_dispatch<KeyT>(sk, sp, nnz, g.cbits, DD_FAST, dtype, reduce, out_row, out_col, out_value, nullptr,
                                d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else {
    PS_SCRATCH(run_start, scratch, int64_t, nnz + 1);
    PS_TRY(dedup_dispatch<KeyT>(sk, sp, nnz, g.cbits, DD_PTR, dtype, reduce, out_row, out_col, nullptr, run_start,
                                d_nnz, scratch, s));
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
                               out_perm, nullptr, s);
  return sort_impl<uint64_t>(row, col, nnz, g, value, dtype, width, false, PS_SUM, out_row, out_col, out_value, out_perm,
                             nullptr, s);
}

int ps_coalesce(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N, const void *value, int dtype,
                int64_t width, int reduce, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *host_nnz_out,
                void *stream) {
  PS_CHECK(nnz >= 0 && M >= 0 && N >= 0, "coalesce: negative size");
  PS_CHECK(nnz < (int64_t(1) << 32), "coalesce: nnz=%lld exceeds the 2^32-1 entries per call this build supports",
           (long long)nnz);
  PS_CHECK(host_nnz_out != nullptr, "coalesce: host_nnz_out is NULL");
  PS_CHECK(reduce >= PS_SUM && reduce <= PS_MAX, "coalesce: bad reduce %d", reduce);
  PS_TRY(check_value_args(value, out_value, dtype, width));
  *host_nnz_out = 0;
  if (nnz == 0) return PS_OK;
  KeyGeom g;
  PS_TRY(key_geometry(M, N, &g));
  cudaStream_t s = (cudaStream_t)stream;
  if (g.bits <= 32)
    return sort_impl<uint32_t>(row, col, nnz, g, value, dtype, width, true, reduce, out_row, out_col, out_value, nullptr,
                               host_nnz_out, s);
  return sort_impl<uint64_t>(row, col, nnz, g, value, dtype, width, true, reduce, out_row, out_col, out_value, nullptr,
                             host_nnz_out, s);
}

int ps_coalesce_sorted(const int64_t *row, const int64_t *col, int64_t nnz, const void *value, int dtype, int64_t width,
                       int reduce, int64_t *out_row, int64_t *out_col, void *out_value, int64_t *host_nnz_out,
                       void *stream) {
  PS_CHECK(nnz >= 0, "coalesce_sorted: negative size");
  PS_CHECK(host_nnz_out != nullptr, "coalesce_sorted: host_nnz_out is NULL");
  PS_CHECK(reduce >= PS_SUM && reduce <= PS_MAX, "coalesce_sorted: bad reduce %d", reduce);
  PS_TRY(check_value_args(value, out_value, dtype, width));
  *host_nnz_out = 0;
  if (nnz == 0) return PS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  // keys are compared as (row << 32 | col): requires row, col < 2^32 (true for any storage
  // whose sparse sizes fit the packed-key sort above)
  const int cbits = 32;
  KeySrc sk{row, col, KSRC_PACK_COO, cbits};
  PS_SCRATCH(d_nnz, scratch, long long, 1);
  long long h_nnz = 0;
  if (!value) {
    PS_TRY(dedup_dispatch<uint64_t>(sk, PaySrc{nullptr, PSRC_IOTA}, nnz, cbits, DD_INDEX_ONLY, dtype, reduce, out_row,
                                    out_col, nullptr, nullptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else if (fast_value(dtype, width)) {
    PS_TRY(dedup_dispatch<uint64_t>(sk, PaySrc{value, pay_src_mode(dtype)}, nnz, cbits, DD_FAST, dtype, reduce, out_row,
                                    out_col, out_value, nullptr, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
  } else {
    PS_SCRATCH(run_start, scratch, int64_t, nnz + 1);
    PS_TRY(dedup_dispatch<uint64_t>(sk, PaySrc{nullptr, PSRC_IOTA}, nnz, cbits, DD_PTR, dtype, reduce, out_row, out_col,
                                    nullptr, run_start, d_nnz, scratch, s));
    PS_TRY(read_back(&h_nnz, d_nnz, sizeof(h_nnz), s));
    PS_TRY(segment_reduce_launch(value, dtype, width, run_start, h_nnz, reduce, out_value, nullptr, nnz, s));
  }
  *host_nnz_out = (int64_t)h_nnz;
  return PS_OK;
}

}  // extern "C"
