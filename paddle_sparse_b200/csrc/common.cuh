// paddle_sparse_b200 — shared device/host helpers for the sm_100a kernels.
//
// Everything in csrc/ is compiled with
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo
// into ONE shared library (libpaddle_sparse_b200.so) whose only exported symbols are
// the extern "C" entry points declared in include/ps_abi.h.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/ps_abi.h"

namespace ps {

// ----------------------------------------------------------------------------- errors
void set_error(const char *fmt, ...);
int sm_count();  // SMs of the current device (148 on B200), cached
// cudaFuncSetAttribute is per DEVICE: true the first time (kernel, current device) is asked about
bool func_needs_config(const void *kernel);
#define PS_FUNC_SMEM(kern, bytes)                                                                        \
  do {                                                                                                   \
    if (ps::func_needs_config((const void *)(kern)))                                                      \
      PS_CUDA(cudaFuncSetAttribute((kern), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));    \
  } while (0)

#define PS_CUDA(expr)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (expr);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      ps::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); \
      return PS_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

#define PS_CHECK(cond, ...)                \
  do {                                     \
    if (!(cond)) {                         \
      ps::set_error(__VA_ARGS__);          \
      return PS_ERR_ARG;                   \
    }                                      \
  } while (0)

#define PS_TRY(expr)          \
  do {                        \
    int rc_ = (expr);         \
    if (rc_ != PS_OK) return rc_; \
  } while (0)

// every kernel launch is followed by PS_LAUNCH_OK(): it checks the launch and counts it
// (ps_launch_count(), reported by bench.py as gpu_launches)
void count_launch();
#define PS_LAUNCH_OK()            \
  do {                            \
    ps::count_launch();           \
    PS_CUDA(cudaGetLastError());  \
  } while (0)

// ----------------------------------------------------------------------------- live kernel timing
// Optional (ps_prof_enable): CUDA events recorded on the launching stream around the launches of
// one kernel family, so bench.py can report the dominant kernel's average duration measured
// inside the timed region.  Off by default: no events, no overhead.
enum ProfId { PROF_RS_HIST = 0, PROF_RS_PASS, PROF_DEDUP, PROF_IND2PTR, PROF_PTR2IND, PROF_INVERT_PERM,
              PROF_TRANSPOSE_GATHER, PROF_SEGMENT_REDUCE, PROF_GATHER, PROF_LOCAL_SORT, PROF_BUCKET_HIST, PROF_CSC_LOCAL,
              PROF_COUNT };
struct ProfScope {
  ProfScope(int id, cudaStream_t s);
  ~ProfScope();
  int id_;
  cudaStream_t s_;
  void *slot_;
};

// ----------------------------------------------------------------------------- scratch
// Stream-ordered scratch: cudaMallocAsync from the device's default pool (release
// threshold raised once so blocks are retained), freed on the same stream when the
// object leaves scope.  No hidden synchronisation.
class Scratch {
 public:
  explicit Scratch(cudaStream_t s) : s_(s) {}
  ~Scratch() {
    for (void *p : ptrs_) cudaFreeAsync(p, s_);
  }
  Scratch(const Scratch &) = delete;
  Scratch &operator=(const Scratch &) = delete;

  template <typename T>
  T *get(size_t n, cudaError_t *err) {
    void *p = nullptr;
    size_t bytes = (n ? n : 1) * sizeof(T);
    bytes = (bytes + 255) & ~size_t(255);
    *err = alloc_(&p, bytes);
    if (*err == cudaSuccess) ptrs_.push_back(p);
    return static_cast<T *>(p);
  }

 private:
  cudaError_t alloc_(void **p, size_t bytes);
  cudaStream_t s_;
  std::vector<void *> ptrs_;
};

#define PS_SCRATCH(var, scratch, T, n)               \
  T *var;                                            \
  do {                                               \
    cudaError_t e_;                                  \
    var = (scratch).get<T>((n), &e_);                \
    if (e_ != cudaSuccess) {                         \
      ps::set_error("%s:%d: scratch alloc of %zu x %zu B failed: %s", __FILE__, __LINE__, \
                    (size_t)(n), sizeof(T), cudaGetErrorString(e_));               \
      return PS_ERR_CUDA;                            \
    }                                                \
  } while (0)

// Pinned one-slot mailbox for the single device->host count read an entry point may do.
int read_back(void *host_dst, const void *dev_src, size_t bytes, cudaStream_t s);

// ----------------------------------------------------------------------------- dtypes
inline size_t dtype_size(int dt) {
  switch (dt) {
    case PS_F32: return 4;
    case PS_BF16: return 2;
    case PS_F16: return 2;
    case PS_F64: return 8;
    case PS_I32: return 4;
    case PS_I64: return 8;
    case PS_U8: return 1;
    default: return 0;
  }
}

// accumulator type: bf16/f16 accumulate in fp32 and are rounded once on store
template <typename T> struct Acc { using type = T; };
template <> struct Acc<__nv_bfloat16> { using type = float; };
template <> struct Acc<__half> { using type = float; };

template <typename T> __device__ __forceinline__ typename Acc<T>::type to_acc(T v) { return v; }
template <> __device__ __forceinline__ float to_acc<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_acc<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_acc(typename Acc<T>::type v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_acc<__half>(float v) { return __float2half_rn(v); }

template <typename A> struct Limits;
template <> struct Limits<float> {
  __device__ static float lowest() { return -__int_as_float(0x7f800000); }
  __device__ static float highest() { return __int_as_float(0x7f800000); }
};
template <> struct Limits<double> {
  __device__ static double lowest() { return -__longlong_as_double(0x7ff0000000000000LL); }
  __device__ static double highest() { return __longlong_as_double(0x7ff0000000000000LL); }
};
template <> struct Limits<int32_t> {
  __device__ static int32_t lowest() { return INT32_MIN; }
  __device__ static int32_t highest() { return INT32_MAX; }
};
template <> struct Limits<int64_t> {
  __device__ static int64_t lowest() { return INT64_MIN; }
  __device__ static int64_t highest() { return INT64_MAX; }
};

// a <= 4-byte value as it rides through the sorts: the bits of one 32-bit payload word
template <typename T> __device__ __forceinline__ T pay_as(uint32_t p);
template <> __device__ __forceinline__ float pay_as<float>(uint32_t p) { return __uint_as_float(p); }
template <> __device__ __forceinline__ int32_t pay_as<int32_t>(uint32_t p) { return (int32_t)p; }
template <> __device__ __forceinline__ __nv_bfloat16 pay_as<__nv_bfloat16>(uint32_t p) {
  return __ushort_as_bfloat16((unsigned short)p);
}
template <> __device__ __forceinline__ __half pay_as<__half>(uint32_t p) { return __ushort_as_half((unsigned short)p); }

// reduce operator on accumulator values; RED is a PS_SUM/PS_MEAN/PS_MIN/PS_MAX constant
template <int RED, typename A>
struct Red {
  __device__ static A init() {
    if (RED == PS_MIN) return Limits<A>::highest();
    if (RED == PS_MAX) return Limits<A>::lowest();
    return A(0);
  }
  __device__ static A op(A a, A b) {
    if (RED == PS_MIN) return b < a ? b : a;
    if (RED == PS_MAX) return b > a ? b : a;
    return a + b;
  }
  // value for a segment of `count` elements (count == 0 -> 0, the torch_scatter convention)
  __device__ static A finish(A a, int64_t count) {
    if (count == 0) return A(0);
    if (RED == PS_MEAN) return mean_(a, count);
    return a;
  }
  __device__ static float mean_(float a, int64_t c) { return a / (float)c; }
  __device__ static double mean_(double a, int64_t c) { return a / (double)c; }
  // integer mean: segment_csr divides in C++ (truncation, torch_scatter segment_csr_cpu.cpp: val / count)
  __device__ static int32_t mean_(int32_t a, int64_t c) { return (int32_t)((int64_t)a / c); }
  __device__ static int64_t mean_(int64_t a, int64_t c) { return a / c; }
  // scatter(reduce="mean") divides with rounding_mode="floor" (torch_scatter scatter.py: scatter_mean)
  __device__ static A finish_floor(A a, int64_t count) {
    if (count == 0) return A(0);
    if (RED == PS_MEAN) return floor_mean_(a, count);
    return a;
  }
  __device__ static float floor_mean_(float a, int64_t c) { return a / (float)c; }
  __device__ static double floor_mean_(double a, int64_t c) { return a / (double)c; }
  __device__ static int32_t floor_mean_(int32_t a, int64_t c) { return (int32_t)floor_mean_((int64_t)a, c); }
  __device__ static int64_t floor_mean_(int64_t a, int64_t c) {
    const int64_t q = a / c, r = a % c;
    return (r != 0 && ((r < 0) != (c < 0))) ? q - 1 : q;
  }
};

// Host-side dispatch over the value dtypes of paddle_sparse/testing.py:12-17.
#define PS_DISPATCH_DTYPE(dt, T, ...)                                   \
  switch (dt) {                                                         \
    case PS_F32: { using T = float; __VA_ARGS__; break; }               \
    case PS_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }      \
    case PS_F16: { using T = __half; __VA_ARGS__; break; }              \
    case PS_F64: { using T = double; __VA_ARGS__; break; }              \
    case PS_I32: { using T = int32_t; __VA_ARGS__; break; }             \
    case PS_I64: { using T = int64_t; __VA_ARGS__; break; }             \
    default: ps::set_error("unsupported value dtype %d", (int)(dt)); return PS_ERR_ARG; \
  }

#define PS_DISPATCH_REDUCE(red, RED, ...)                               \
  switch (red) {                                                        \
    case PS_SUM: { constexpr int RED = PS_SUM; __VA_ARGS__; break; }    \
    case PS_MEAN: { constexpr int RED = PS_MEAN; __VA_ARGS__; break; }  \
    case PS_MIN: { constexpr int RED = PS_MIN; __VA_ARGS__; break; }    \
    case PS_MAX: { constexpr int RED = PS_MAX; __VA_ARGS__; break; }    \
    default: ps::set_error("unsupported reduce %d", (int)(red)); return PS_ERR_ARG; \
  }

// ----------------------------------------------------------------------------- device helpers
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// streaming (read-once) loads: bypass L1 allocation so the sort tiles do not thrash it
__device__ __forceinline__ int64_t ld_stream(const int64_t *p) {
  long long v;
  asm volatile("ld.global.nc.L1::no_allocate.s64 %0, [%1];" : "=l"(v) : "l"(p));
  return (int64_t)v;
}
__device__ __forceinline__ uint64_t ld_stream(const uint64_t *p) {
  unsigned long long v;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return (uint64_t)v;
}
__device__ __forceinline__ uint32_t ld_stream(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ longlong2 ld_stream2(const int64_t *p) {  // 128-bit, p 16-B aligned
  longlong2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.s64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
  return v;
}

// 128-bit evict-first load through the coherent path: safe on a buffer the same thread overwrites afterwards
// (the in-place mul_ / add_), unlike the .nc forms above
__device__ __forceinline__ uint4 ld_once(const uint4 *p) {
  uint4 v;
  asm volatile("ld.global.cs.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline int bits_for(int64_t n) {  // bits needed to represent values in [0, n)
  int b = 0;
  while (b < 63 && (int64_t(1) << b) < n) ++b;
  return b;
}

// ----------------------------------------------------------------------------- block primitives
// Exclusive block scan of one value per thread (THREADS multiple of 32, <= 1024).
// `total` receives the block sum.  `warp_sums` is 33 elements of shared memory.
template <typename T, int THREADS>
__device__ __forceinline__ T block_exclusive_scan(T v, T *warp_sums, T *total) {
  constexpr int WARPS = THREADS / 32;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  T inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= (unsigned)d) inc += o;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    T w = lane < WARPS ? warp_sums[lane] : T(0);
    T winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      T o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= (unsigned)d) winc += o;
    }
    if (lane < WARPS) warp_sums[lane] = winc - w;
    if (lane == WARPS - 1) warp_sums[32] = winc;
  }
  __syncthreads();
  T res = warp_sums[warp] + inc - v;
  *total = warp_sums[32];
  __syncthreads();  // warp_sums may be reused by the caller
  return res;
}

// ----------------------------------------------------------------------------- decoupled look-back
// Single-pass chained scan state (Merrill & Garland): one 64-bit word per tile holding a
// 2-bit status and a 62-bit value, so flag and value are published by one store.
struct TileState {
  static constexpr uint64_t EMPTY = 0, AGG = 1, PREFIX = 2;
  __device__ static uint64_t pack(uint64_t status, uint64_t v) { return (status << 62) | v; }
  __device__ static uint64_t status(uint64_t w) { return w >> 62; }
  __device__ static uint64_t value(uint64_t w) { return w & ((uint64_t(1) << 62) - 1); }
  __device__ static void store(uint64_t *p, uint64_t w) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
  }
  __device__ static uint64_t load(const uint64_t *p) {
    uint64_t w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
  }
};

// Warp-parallel look-back (called by ALL 32 lanes of one warp): lane l inspects tile
// (base - l); returns the exclusive prefix of `tile` in every lane and publishes the
// tile's inclusive prefix.  Tiles must be numbered by an atomic ticket so that every
// predecessor is already resident (forward progress).
__device__ __forceinline__ uint64_t warp_lookback(uint64_t *state, int64_t tile, uint64_t aggregate) {
  const unsigned lane = lane_id();
  if (tile == 0) {
    if (lane == 0) TileState::store(state, TileState::pack(TileState::PREFIX, aggregate));
    return 0;
  }
  if (lane == 0) TileState::store(state + tile, TileState::pack(TileState::AGG, aggregate));
  uint64_t excl = 0;
  int64_t base = tile - 1;
  while (true) {
    const int64_t t = base - (int64_t)lane;
    const uint64_t w = t >= 0 ? TileState::load(state + t) : TileState::pack(TileState::PREFIX, 0);
    const uint64_t st = TileState::status(w);
    const unsigned pmask = __ballot_sync(0xffffffffu, st == TileState::PREFIX);
    const unsigned emask = __ballot_sync(0xffffffffu, st == TileState::EMPTY);
    unsigned need = 0xffffffffu;
    if (pmask) {
      const int fp = __ffs(pmask) - 1;
      need = fp == 31 ? 0xffffffffu : ((2u << fp) - 1u);
    }
    if (emask & need) continue;  // a needed predecessor is not published yet: poll again
    uint64_t c = ((need >> lane) & 1u) ? TileState::value(w) : 0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    excl += c;
    if (pmask) break;
    base -= 32;
  }
  if (lane == 0)
    TileState::store(state + tile, TileState::pack(TileState::PREFIX, excl + aggregate));
  return excl;
}

// Same protocol, W x 32 predecessors per round trip (all W loads of a lane in flight together).  For chains whose
// tiles publish their aggregates almost simultaneously (equal work per tile, e.g. one sorted bucket per CTA round):
// the walk back to the nearest published prefix is then up to #resident-CTAs long, and its cost is round trips.
template <int W>
__device__ __forceinline__ uint64_t warp_lookback_wide(uint64_t *state, int64_t tile, uint64_t aggregate) {
  const unsigned lane = lane_id();
  if (tile == 0) {
    if (lane == 0) TileState::store(state, TileState::pack(TileState::PREFIX, aggregate));
    return 0;
  }
  if (lane == 0) TileState::store(state + tile, TileState::pack(TileState::AGG, aggregate));
  uint64_t excl = 0;
  int64_t base = tile - 1;
  bool done = false;
  while (!done) {
    uint64_t w[W];
#pragma unroll
    for (int q = 0; q < W; ++q) {
      const int64_t t = base - (int64_t)lane - 32 * q;
      w[q] = t >= 0 ? TileState::load(state + t) : TileState::pack(TileState::PREFIX, 0);
    }
    int used = 0;
#pragma unroll
    for (int q = 0; q < W; ++q) {
      if (done || used != q) continue;  // uniform: stop at the first window with an unpublished needed word
      const uint64_t st = TileState::status(w[q]);
      const unsigned pmask = __ballot_sync(0xffffffffu, st == TileState::PREFIX);
      const unsigned emask = __ballot_sync(0xffffffffu, st == TileState::EMPTY);
      unsigned need = 0xffffffffu;
      if (pmask) {
        const int fp = __ffs(pmask) - 1;
        need = fp == 31 ? 0xffffffffu : ((2u << fp) - 1u);
      }
      if (emask & need) continue;  // poll this window again in the next round
      uint64_t c = ((need >> lane) & 1u) ? TileState::value(w[q]) : 0;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
      excl += c;
      used = q + 1;
      if (pmask) done = true;
    }
    base -= 32 * used;
  }
  if (lane == 0)
    TileState::store(state + tile, TileState::pack(TileState::PREFIX, excl + aggregate));
  return excl;
}

}  // namespace ps
