// ind2ptr / ptr2ind and the small pointer/count utilities around them.
//
// Reference behaviour restated (not ported):
//   ind2ptr  csrc/convert.cpp:13-23, csrc/cpu/convert_cpu.cpp:6-30, csrc/cuda/convert_cuda.cu:6-40
//   ptr2ind  csrc/convert.cpp:46-56, csrc/cpu/convert_cpu.cpp:32-48, csrc/cuda/convert_cuda.cu:42-68
// The reference kernels are thread-per-element with a serial head/tail fill on one
// thread (ind2ptr) and thread-per-row serial expansion (ptr2ind).  Here:
//   ind2ptr  = 128-bit coalesced boundary detection; gaps longer than a warp are filled
//              warp-cooperatively; head [0, ind[0]] and tail (ind[last], M] are filled by
//              dedicated blocks of the same launch with coalesced grid-stride stores.
//   ptr2ind  = load-balanced over OUTPUT tiles: 32-ary warp search for the first/last row
//              of the tile, row starts scattered into shared memory, block max-scan,
//              coalesced 128-bit stores.  Degree skew cannot serialise a thread.
#include <mutex>

#include "common.cuh"
#include "convert.cuh"

namespace ps {

// ============================================================================= ind2ptr
template <typename IndT>
__device__ __forceinline__ int64_t ind_load(const IndT *p) { return (int64_t)*p; }

constexpr int I2P_ITEMS = 4;  // consecutive elements per thread: 16 B of uint32 keys / 32 B of int64 indices

template <typename IndT, int THREADS>
__global__ void __launch_bounds__(THREADS)
ind2ptr_kernel(const IndT *__restrict__ ind, int64_t numel, int64_t M, int64_t *__restrict__ out,
               int elem_blocks, int fill_blocks, bool aligned16) {
  constexpr int ITEMS = I2P_ITEMS;
  if ((int)blockIdx.x >= elem_blocks) {
    // ---- head / tail fill blocks
    const int64_t first = (int64_t)ind[0], last = (int64_t)ind[numel - 1];
    const int64_t fb = (int)blockIdx.x - elem_blocks;
    const int64_t stride = (int64_t)fill_blocks * THREADS;
    for (int64_t r = fb * THREADS + threadIdx.x; r <= first && r <= M; r += stride) out[r] = 0;
    for (int64_t r = last + 1 + fb * THREADS + threadIdx.x; r <= M; r += stride) out[r] = numel;
    return;
  }
  // ---- boundary detection: each thread owns ITEMS consecutive elements (128-bit loads)
  const int64_t e0 = ((int64_t)blockIdx.x * THREADS + threadIdx.x) * ITEMS;
  const int64_t left = numel - e0;
  const int have = left >= ITEMS ? ITEMS : (left > 0 ? (int)left : 0);
  int64_t v[ITEMS];
  if (have == ITEMS && aligned16) {
    if (sizeof(IndT) == 8) {
#pragma unroll
      for (int k = 0; k < ITEMS; k += 2) {
        const longlong2 t = ld_stream2(reinterpret_cast<const int64_t *>(ind) + e0 + k);
        v[k] = t.x;
        v[k + 1] = t.y;
      }
    } else {
      uint32_t w[4];
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3])
                   : "l"(reinterpret_cast<const uint32_t *>(ind) + e0));
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) v[k] = (int64_t)w[k];
    }
  } else {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) v[k] = k < have ? (int64_t)ind[e0 + k] : (k > 0 ? v[k - 1] : 0);
  }
  // previous element: neighbour lane's last one, lane 0 reads it from memory
  const int64_t up = __shfl_up_sync(0xffffffffu, v[ITEMS - 1], 1);
  int64_t prev = v[0];
  if (have > 0 && e0 > 0) prev = lane_id() == 0 ? (int64_t)ind[e0 - 1] : up;
  // rows r in (p_k, v_k] get e0 + k, p_0 = prev, p_k = v_{k-1}   (head handled by the fill blocks);
  // absent elements repeat their predecessor, so their interval is empty
  bool any_long = false;
  int64_t p = prev;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int64_t gap = v[k] - p;
    if (gap > 32) any_long = true;
    else
      for (int64_t r = p + 1; r <= v[k]; ++r) out[r] = e0 + k;
    p = v[k];
  }
  if (!__any_sync(0xffffffffu, any_long)) return;
  p = prev;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {  // warp-cooperative fill of every gap longer than a warp
    unsigned m = __ballot_sync(0xffffffffu, v[k] - p > 32);
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const int64_t l = __shfl_sync(0xffffffffu, p + 1, src);
      const int64_t h = __shfl_sync(0xffffffffu, v[k], src);
      const int64_t val = __shfl_sync(0xffffffffu, e0 + k, src);
      for (int64_t r = l + lane_id(); r <= h; r += 32) out[r] = val;
    }
    p = v[k];
  }
}

template <typename IndT>
int ind2ptr_launch(const IndT *ind, int64_t numel, int64_t M, int64_t *out, cudaStream_t s) {
  if (M < 0 || numel < 0) { set_error("ind2ptr: negative size"); return PS_ERR_ARG; }
  if (numel == 0) {  // convert_cpu.cpp:9-11: empty input -> zeros(M+1)
    PS_CUDA(cudaMemsetAsync(out, 0, sizeof(int64_t) * (size_t)(M + 1), s));
    return PS_OK;
  }
  constexpr int THREADS = 256;
  const int64_t elem_blocks = ceil_div(numel, THREADS * I2P_ITEMS);
  int64_t fill_blocks = ceil_div(M + 1, THREADS * 8);
  const int64_t cap = 2 * (int64_t)sm_count();
  if (fill_blocks > cap) fill_blocks = cap;
  if (fill_blocks < 1) fill_blocks = 1;
  if (elem_blocks + fill_blocks > 2147483647LL) { set_error("ind2ptr: input too large"); return PS_ERR_ARG; }
  {
    ProfScope prof(PROF_IND2PTR, s);
    ind2ptr_kernel<IndT, THREADS><<<(unsigned)(elem_blocks + fill_blocks), THREADS, 0, s>>>(
        ind, numel, M, out, (int)elem_blocks, (int)fill_blocks, (reinterpret_cast<uintptr_t>(ind) & 15) == 0);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}
template int ind2ptr_launch<int64_t>(const int64_t *, int64_t, int64_t, int64_t *, cudaStream_t);
template int ind2ptr_launch<uint32_t>(const uint32_t *, int64_t, int64_t, int64_t *, cudaStream_t);

// ============================================================================= ptr2ind
template <int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
ptr2ind_kernel(const int64_t *__restrict__ ptr, int64_t M, int64_t E, int64_t *__restrict__ out,
               const int32_t *__restrict__ bounds) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = (int64_t)blockIdx.x * TILE;
  const int tile_n = (int)((E - e0) < TILE ? (E - e0) : TILE);
  expand_tile<THREADS, ITEMS>(ptr, M, e0, tile_n, seg, aux, bounds);
  for (int j = threadIdx.x; j < tile_n; j += THREADS) out[e0 + j] = (int64_t)seg[j];
}

__global__ void __launch_bounds__(256)
tile_bounds_kernel(const int64_t *__restrict__ ptr, int64_t S, int64_t tile, int64_t tiles,
                   int32_t *__restrict__ bounds) {
  const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (t > tiles) return;  // warp-uniform
  int64_t c = S;
  if (t < tiles) c = warp_upper_bound(ptr, S, t * tile);
  if (lane_id() == 0) bounds[t] = (int32_t)(c > 0 ? c - 1 : 0);
}

int tile_bounds_launch(const int64_t *ptr, int64_t S, int64_t E, int tile, int32_t *bounds, cudaStream_t s) {
  const int64_t tiles = ceil_div(E, tile);
  tile_bounds_kernel<<<(unsigned)ceil_div((tiles + 1) * 32, 256), 256, 0, s>>>(ptr, S, tile, tiles, bounds);
  PS_LAUNCH_OK();
  return PS_OK;
}

// ============================================================================= small utilities
__global__ void ptr2count_kernel(const int64_t *__restrict__ ptr, int64_t M, int64_t *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < M) out[i] = ptr[i + 1] - ptr[i];
}

// Single-pass inclusive scan (decoupled look-back) writing out_ptr[i+1]; out_ptr[0] = 0.
template <int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
count2ptr_kernel(const int64_t *__restrict__ count, const int64_t *__restrict__ idx, int64_t n,
                 int64_t *__restrict__ out, uint64_t *state /* [0]=ticket, [1..]=tiles */) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int64_t warp_sums[33];
  __shared__ int64_t s_tile, s_excl;
  if (threadIdx.x == 0) s_tile = (int64_t)atomicAdd((unsigned long long *)state, 1ull);
  __syncthreads();
  const int64_t tile = s_tile;
  const int64_t base = tile * TILE + (int64_t)threadIdx.x * ITEMS;
  int64_t v[ITEMS];
  int64_t sum = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int64_t i = base + k;
    v[k] = i < n ? (idx ? count[idx[i]] : count[i]) : 0;
    sum += v[k];
  }
  int64_t total;
  int64_t excl = block_exclusive_scan<int64_t, THREADS>(sum, warp_sums, &total);
  if (threadIdx.x < 32) {
    const uint64_t p = warp_lookback(state + 1, tile, (uint64_t)total);
    if (threadIdx.x == 0) s_excl = (int64_t)p;
  }
  __syncthreads();
  excl += s_excl;
  if (tile == 0 && threadIdx.x == 0) out[0] = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    excl += v[k];
    if (base + k < n) out[base + k + 1] = excl;
  }
}

__global__ void histogram_kernel(const int64_t *__restrict__ ind, int64_t n, int64_t *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    atomicAdd((unsigned long long *)&out[ld_stream(ind + i)], 1ull);
}

__global__ void index_max_kernel(const int64_t *__restrict__ a, int64_t na, const int64_t *__restrict__ b,
                                 int64_t nb, long long *out2) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  long long ma = -1, mb = -1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < na; i += stride) {
    const long long v = ld_stream(a + i);
    ma = v > ma ? v : ma;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += stride) {
    const long long v = ld_stream(b + i);
    mb = v > mb ? v : mb;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const long long oa = __shfl_xor_sync(0xffffffffu, ma, d), ob = __shfl_xor_sync(0xffffffffu, mb, d);
    ma = oa > ma ? oa : ma;
    mb = ob > mb ? ob : mb;
  }
  if (lane_id() == 0) {
    if (ma >= 0) atomicMax(&out2[0], ma);
    if (mb >= 0) atomicMax(&out2[1], mb);
  }
}

// flags[0] &= non-decreasing, flags[1] &= strictly increasing (lexicographic on (row,col))
__global__ void coo_order_kernel(const int64_t *__restrict__ row, const int64_t *__restrict__ col, int64_t nnz,
                                 int *flags) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  bool inv = false, eq = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < nnz; i += stride) {
    const int64_t r0 = row[i - 1], r1 = row[i], c0 = col[i - 1], c1 = col[i];
    inv |= (r1 < r0) || (r1 == r0 && c1 < c0);
    eq |= (r1 == r0 && c1 == c0);
  }
  if (__any_sync(0xffffffffu, inv) && lane_id() == 0) { flags[0] = 0; flags[1] = 0; }
  if (__any_sync(0xffffffffu, eq) && lane_id() == 0) flags[1] = 0;
}

__global__ void add_scalar_kernel(const int64_t *__restrict__ in, int64_t n, int64_t delta, int64_t *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i] + delta;
}

int count2ptr_launch(const int64_t *count, const int64_t *idx, int64_t n, int64_t *out_ptr, cudaStream_t s) {
  if (n == 0) {
    PS_CUDA(cudaMemsetAsync(out_ptr, 0, sizeof(int64_t), s));
    return PS_OK;
  }
  constexpr int THREADS = 256, ITEMS = 8;
  const int64_t tiles = ceil_div(n, THREADS * ITEMS);
  Scratch scratch(s);
  PS_SCRATCH(state, scratch, uint64_t, tiles + 1);
  PS_CUDA(cudaMemsetAsync(state, 0, sizeof(uint64_t) * (size_t)(tiles + 1), s));
  count2ptr_kernel<THREADS, ITEMS><<<(unsigned)tiles, THREADS, 0, s>>>(count, idx, n, out_ptr, state);
  PS_LAUNCH_OK();
  return PS_OK;
}

static inline unsigned stream_grid(int64_t n, int threads) {
  int64_t b = ceil_div(n, threads);
  const int64_t cap = 16 * (int64_t)sm_count();
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_ind2ptr(const int64_t *ind, int64_t numel, int64_t M, int64_t *out_ptr, void *stream) {
  return ind2ptr_launch<int64_t>(ind, numel, M, out_ptr, (cudaStream_t)stream);
}

int ps_ptr2ind(const int64_t *ptr, int64_t M, int64_t E, int64_t *out_ind, void *stream) {
  PS_CHECK(M >= 0 && E >= 0, "ptr2ind: negative size");
  PS_CHECK(M < 2147483647LL, "ptr2ind: M=%lld exceeds the 2^31-1 rows this build supports", (long long)M);
  if (E == 0) return PS_OK;
  PS_CHECK(M > 0, "ptr2ind: E > 0 with no rows");
  constexpr int THREADS = 256, ITEMS = 8;
  const int64_t tiles = ceil_div(E, THREADS * ITEMS);
  Scratch scratch((cudaStream_t)stream);
  PS_SCRATCH(bounds, scratch, int32_t, tiles + 1);
  {
    ProfScope prof(PROF_PTR2IND, (cudaStream_t)stream);
    if (int rc = tile_bounds_launch(ptr, M, E, THREADS * ITEMS, bounds, (cudaStream_t)stream)) return rc;
    ptr2ind_kernel<THREADS, ITEMS><<<(unsigned)tiles, THREADS, 0, (cudaStream_t)stream>>>(ptr, M, E, out_ind, bounds);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_ptr2count(const int64_t *ptr, int64_t M, int64_t *out_count, void *stream) {
  if (M <= 0) return PS_OK;
  ptr2count_kernel<<<(unsigned)ceil_div(M, 256), 256, 0, (cudaStream_t)stream>>>(ptr, M, out_count);
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_count2ptr(const int64_t *count, const int64_t *idx, int64_t n, int64_t *out_ptr, int64_t *host_total,
                 void *stream) {
  PS_CHECK(n >= 0, "count2ptr: negative size");
  PS_TRY(count2ptr_launch(count, idx, n, out_ptr, (cudaStream_t)stream));
  if (host_total) PS_TRY(read_back(host_total, out_ptr + n, sizeof(int64_t), (cudaStream_t)stream));
  return PS_OK;
}

int ps_histogram(const int64_t *ind, int64_t numel, int64_t bins, int64_t *out_count, void *stream) {
  PS_CHECK(numel >= 0 && bins >= 0, "histogram: negative size");
  cudaStream_t s = (cudaStream_t)stream;
  if (bins > 0) PS_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int64_t) * (size_t)bins, s));
  if (numel == 0) return PS_OK;
  histogram_kernel<<<stream_grid(numel, 256), 256, 0, s>>>(ind, numel, out_count);
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_index_max(const int64_t *a, int64_t na, const int64_t *b, int64_t nb, int64_t *host_max, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  PS_SCRATCH(d, scratch, long long, 2);
  PS_CUDA(cudaMemsetAsync(d, 0xff, 2 * sizeof(long long), s));  // -1, -1
  if (!b) nb = 0;
  const int64_t n = na > nb ? na : nb;
  if (n > 0) {
    index_max_kernel<<<stream_grid(n, 256), 256, 0, s>>>(a, na, b, nb, d);
    PS_LAUNCH_OK();
  }
  return read_back(host_max, d, 2 * sizeof(long long), s);
}

int ps_coo_order(const int64_t *row, const int64_t *col, int64_t nnz, int32_t *host_flags, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  PS_SCRATCH(d, scratch, int, 2);
  const int ones[2] = {1, 1};
  PS_CUDA(cudaMemcpyAsync(d, ones, sizeof(ones), cudaMemcpyHostToDevice, s));
  if (nnz > 1) {
    coo_order_kernel<<<stream_grid(nnz, 256), 256, 0, s>>>(row, col, nnz, d);
    PS_LAUNCH_OK();
  }
  return read_back(host_flags, d, 2 * sizeof(int), s);
}

int ps_add_scalar(const int64_t *in, int64_t n, int64_t delta, int64_t *out, void *stream) {
  if (n <= 0) return PS_OK;
  add_scalar_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(in, n, delta, out);
  PS_LAUNCH_OK();
  return PS_OK;
}

// ----------------------------------------------------------------------------- ptr2ind without any allocation
// ps_ptr2ind needs one int32 per output tile (the first segment touching the tile) and takes it from the
// stream-ordered pool.  Callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc) get the same two kernels with that table in a fixed array that lives in the library's device image:
// every GPU that loads the library holds PS_PTR2IND_NOALLOC_SCRATCH_BYTES of it until the process exits.
// In host code `g_ptr2ind_bounds` is only the host-side stand-in of the symbol, NOT a device address: the address a
// kernel may be given comes from cudaGetSymbolAddress, asked once per device with that device current.
}  // extern "C" (the array and its address table have internal linkage)

namespace {
constexpr int64_t NOALLOC_MAX_TILES = 1 << 20;                   // tiles of 2,048 entries: E <= 2^31
constexpr int64_t NOALLOC_WORDS = NOALLOC_MAX_TILES + 64;        // tiles + 1 in use, padded to whole 256-byte lines
static __device__ __align__(256) int32_t g_ptr2ind_bounds[NOALLOC_WORDS];  // 4 MiB + 256 B per device
static_assert(sizeof(int32_t) * NOALLOC_WORDS == PS_PTR2IND_NOALLOC_SCRATCH_BYTES, "ps_abi.h states the size");

std::mutex g_noalloc_mu;
int32_t *g_noalloc_addr[64] = {nullptr};  // per device ordinal: the array's address on THAT device

// the device address of g_ptr2ind_bounds on the current device (resolved on the first call per device)
int noalloc_bounds(int32_t **out) {
  int dev = -1;
  PS_CUDA(cudaGetDevice(&dev));
  PS_CHECK(dev >= 0 && dev < 64, "ptr2ind_noalloc: device ordinal %d outside this build's table of 64", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_addr[dev]) {
    void *p = nullptr;
    PS_CUDA(cudaGetSymbolAddress(&p, g_ptr2ind_bounds));
    PS_CHECK(p != nullptr, "ptr2ind_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
    g_noalloc_addr[dev] = static_cast<int32_t *>(p);
  }
  *out = g_noalloc_addr[dev];
  return PS_OK;
}
}  // namespace

extern "C" {

int ps_ptr2ind_noalloc(const int64_t *ptr, int64_t M, int64_t E, int64_t *out_ind, void *stream) {
  PS_CHECK(M >= 0 && E >= 0, "ptr2ind_noalloc: negative size");
  PS_CHECK(M < 2147483647LL, "ptr2ind_noalloc: M=%lld exceeds the 2^31-1 rows this build supports", (long long)M);
  PS_CHECK(E <= (int64_t)PS_PTR2IND_NOALLOC_MAX_E,
           "ptr2ind_noalloc: E=%lld exceeds the %lld entries the fixed scratch array covers; use ps_ptr2ind",
           (long long)E, (long long)PS_PTR2IND_NOALLOC_MAX_E);
  if (E == 0) return PS_OK;
  PS_CHECK(M > 0, "ptr2ind_noalloc: E > 0 with no rows");
  constexpr int THREADS = 256, ITEMS = 8;
  static_assert(THREADS * ITEMS * NOALLOC_MAX_TILES == PS_PTR2IND_NOALLOC_MAX_E, "ps_abi.h states the limit");
  const int64_t tiles = ceil_div(E, THREADS * ITEMS);
  int32_t *bounds = nullptr;
  PS_TRY(noalloc_bounds(&bounds));
  // nothing has to start at zero: tile_bounds_kernel writes bounds[0 .. tiles] and ptr2ind_kernel reads only those
  // words, later on the same stream.  Calls that overlap on the GPU (two streams) would race on the array.
  PS_TRY(tile_bounds_launch(ptr, M, E, THREADS * ITEMS, bounds, (cudaStream_t)stream));
  ptr2ind_kernel<THREADS, ITEMS><<<(unsigned)tiles, THREADS, 0, (cudaStream_t)stream>>>(ptr, M, E, out_ind, bounds);
  PS_LAUNCH_OK();
  return PS_OK;
}

}  // extern "C"
