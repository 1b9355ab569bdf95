// In-repo replacements for the paddle_scatter calls on the hot path:
//   segment_csr  storage.py:471, reduce.py:51, tensor.py:437
//   gather_csr   mul.py:16,79, index_select.py:29,65
//   scatter      reduce.py:42
// plus the fused broadcast multiplies of mul.py:76-93 and the dense reduction of
// reduce.py:17-27.  paddle_scatter's source is not part of the reference tree; semantics are
// restated from its documented behaviour (empty segment / untouched row -> 0).
#include <cstdlib>

#include "common.cuh"
#include "convert.cuh"
#include "segment.cuh"

namespace ps {

// ============================================================================= segment reduce
// G lanes cooperate on one output (segment s, feature f); o = s*width + f.
// W1: width == 1 known at compile time (the [nnz] value vector of reduce.py:51) - no 64-bit divide per output,
// and the segment walk is unrolled so that each lane keeps several loads in flight.
// Long segments (hub rows of power-law graphs) do not ride on G lanes: a group that meets a segment of more than
// `long_thresh` entries appends its output index to `long_list` and moves on; segment_reduce_long_kernel then gives
// every listed output a whole CTA.  The reference's segment_csr (paddle_scatter, a port of torch_scatter) walks a
// segment with a fixed group of lanes whatever its length.
constexpr int SR_SPLIT = 16;
constexpr int64_t SR_PART_MIN = 4096;
struct alignas(16) SrTask {
  int64_t lo, hi;   // the part's range of entries
  int64_t gid;      // output index (segment * width + feature)
  uint32_t slot;    // position in the list: partial[slot * SR_SPLIT + q], long_arrive[slot]
  uint16_t q, parts;
};
static_assert(sizeof(SrTask) == 32, "two 16-byte loads per task");
__device__ __forceinline__ int sr_parts(int64_t len) {
  const int64_t p = (len + SR_PART_MIN - 1) / SR_PART_MIN;
  return (int)(p < 1 ? 1 : (p > SR_SPLIT ? SR_SPLIT : p));
}

template <typename T, int RED, int G, bool W1>
__global__ void __launch_bounds__(256, (W1 && G == 32) ? 8 : 4)  // long [nnz] rows: 32 registers, full occupancy hides the latency
segment_reduce_kernel(const T *__restrict__ src, int64_t width, const int64_t *__restrict__ ptr, int64_t S,
                      T *__restrict__ out, const uint32_t *__restrict__ perm, int64_t long_thresh,
                      unsigned long long *__restrict__ long_list /*[0] = count, [1] = tasks, [2..] = output indices*/,
                      unsigned int *__restrict__ long_arrive /*[slot]: parts of that output finished so far*/,
                      SrTask *__restrict__ long_tasks /*one record per part of a listed output*/,
                      int vec16 /*src is 16-byte aligned*/) {
  using A = typename Acc<T>::type;
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int sub = threadIdx.x % G;
  const int64_t total = W1 ? S : S * width;
  const bool live = gid < total;
  const int64_t sgm = live ? (W1 ? gid : gid / width) : 0;
  const int64_t f = (live && !W1) ? gid - sgm * width : 0;
  int64_t lo = 0, hi = 0;
  if (live) {
    lo = ptr[sgm];
    hi = ptr[sgm + 1];
  }
  const bool deferred = hi - lo > long_thresh;  // uniform inside the group
  if (deferred) {
    if (sub == 0) {
      const unsigned long long slot = atomicAdd(long_list, 1ull);
      long_list[2 + slot] = (unsigned long long)gid;
      long_arrive[slot] = 0u;
      const int parts = sr_parts(hi - lo);
      const unsigned long long t0 = atomicAdd(long_list + 1, (unsigned long long)parts);
      for (int q = 0; q < parts; ++q)  // everything a part needs in one 32-byte record: no pointer chasing later
        long_tasks[t0 + q] = SrTask{lo + (hi - lo) * q / parts, lo + (hi - lo) * (q + 1) / parts, gid, (uint32_t)slot,
                                    (uint16_t)q, (uint16_t)parts};
    }
    hi = lo;  // nothing to walk; the store below is suppressed
  }
  A acc = Red<RED, A>::init();
  if (W1) {
    int64_t e = lo + sub;
    // 16-byte loads for a full warp on a long enough segment of a <= 4-byte type: a lane takes 4 (f32 / i32) or 8
    // (bf16 / f16) consecutive entries per load instead of one (the 2-byte reductions were issue-bound at one
    // element per load: 72 % issue-active, 0.37 of HBM); the unaligned head and tail are one entry per lane
    constexpr int V = sizeof(T) <= 4 ? 16 / (int)sizeof(T) : 1;
    if (V > 1 && G == 32 && !perm && vec16 && hi - lo >= 2 * V * G) {
      const int64_t a0 = (lo + V - 1) / V * V, a1 = hi / V * V;
      if (e < a0) acc = Red<RED, A>::op(acc, to_acc<T>(src[e]));
      const uint4 *p16 = reinterpret_cast<const uint4 *>(src);
      const int64_t vend = a1 / V;
      int64_t v = a0 / V + sub;
      auto fold = [&](const uint4 &x) {
        const T *t = reinterpret_cast<const T *>(&x);
#pragma unroll
        for (int k = 0; k < V; ++k) acc = Red<RED, A>::op(acc, to_acc<T>(t[k]));
      };
      for (; v + G < vend; v += 2 * G) {
        const uint4 x = __ldg(p16 + v), y = __ldg(p16 + v + G);
        fold(x);
        fold(y);
      }
      for (; v < vend; v += G) fold(__ldg(p16 + v));
      e = a1 + sub;
      if (e < hi) acc = Red<RED, A>::op(acc, to_acc<T>(src[e]));
      e = hi;
    }
    if (!perm) {
      for (; e + 3 * G < hi; e += 4 * G) {
        const T v0 = src[e], v1 = src[e + G], v2 = src[e + 2 * G], v3 = src[e + 3 * G];
        acc = Red<RED, A>::op(acc, to_acc<T>(v0));
        acc = Red<RED, A>::op(acc, to_acc<T>(v1));
        acc = Red<RED, A>::op(acc, to_acc<T>(v2));
        acc = Red<RED, A>::op(acc, to_acc<T>(v3));
      }
    }
    for (; e < hi; e += G) acc = Red<RED, A>::op(acc, to_acc<T>(src[perm ? (int64_t)perm[e] : e]));
  } else {
    for (int64_t e = lo + sub; e < hi; e += G) {
      const int64_t r = perm ? (int64_t)perm[e] : e;
      acc = Red<RED, A>::op(acc, to_acc<T>(src[r * width + f]));
    }
  }
#pragma unroll
  for (int d = G / 2; d > 0; d >>= 1) acc = Red<RED, A>::op(acc, __shfl_xor_sync(0xffffffffu, acc, d));
  if (live && sub == 0 && !deferred) out[gid] = from_acc<T>(Red<RED, A>::finish(acc, hi - lo));
}

// Listed (long) outputs: a segment is cut into up to SR_SPLIT equal parts of at least SR_PART_MIN entries; the lister
// appends one task per part, the CTAs of the long kernel walk the task list (balanced whatever the mix of lengths,
// and an empty list costs one small launch), each reducing its part into partial[slot * SR_SPLIT + part]; the part
// that finishes last (arrival counter per listed output, zeroed by the lister) folds the partials in part order.  256 threads walk a part with four loads in flight each -
// 16-byte loads (4 or 8 entries) when the values are a contiguous [nnz] vector of a <= 4-byte type - and their
// results are combined in a fixed order (lane tree, then warp 0 over the 8 warp results): deterministic for a given
// segment.  One CTA per segment (round 2's first version) left the hub rows of a power-law graph on the critical
// path: a 230 K-entry row took ~170 us on its own.
template <typename T, int RED>
__global__ void __launch_bounds__(256)
segment_reduce_long_kernel(const T *__restrict__ src, int64_t width, const int64_t *__restrict__ ptr, T *__restrict__ out,
                           const uint32_t *__restrict__ perm, const unsigned long long *__restrict__ long_list,
                           unsigned int *__restrict__ long_arrive, const SrTask *__restrict__ long_tasks,
                           typename Acc<T>::type *partial, int vec16) {
  using A = typename Acc<T>::type;
  __shared__ A part[8];
  const unsigned long long ntasks = long_list[1];
  if (blockIdx.x >= ntasks) return;
  SrTask nxt = long_tasks[blockIdx.x];
  for (unsigned long long t = blockIdx.x; t < ntasks; t += gridDim.x) {
    const SrTask task = nxt;
    if (t + gridDim.x < ntasks) nxt = long_tasks[t + gridDim.x];  // requested before this part's data
    const unsigned long long li = task.slot;
    const int py = task.q, parts = task.parts;
    const int64_t gid = task.gid;
    const int64_t f = width == 1 ? 0 : gid % width;
    const int64_t lo = task.lo, hi = task.hi;
    A acc = Red<RED, A>::init();
    int64_t e = lo + threadIdx.x;
    constexpr int V = sizeof(T) <= 4 ? 16 / (int)sizeof(T) : 1;
    if (V > 1 && width == 1 && !perm && vec16 && hi - lo >= 2 * V * 256) {
      const int64_t a0 = (lo + V - 1) / V * V, a1 = hi / V * V;
      if (e < a0) acc = Red<RED, A>::op(acc, to_acc<T>(src[e]));
      const uint4 *p16 = reinterpret_cast<const uint4 *>(src);
      const int64_t vend = a1 / V;
      int64_t v = a0 / V + threadIdx.x;
      auto fold = [&](const uint4 &x) {
        const T *t = reinterpret_cast<const T *>(&x);
#pragma unroll
        for (int k = 0; k < V; ++k) acc = Red<RED, A>::op(acc, to_acc<T>(t[k]));
      };
      for (; v + 3 * 256 < vend; v += 4 * 256) {
        uint4 x[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) x[q] = __ldg(p16 + v + q * 256);
#pragma unroll
        for (int q = 0; q < 4; ++q) fold(x[q]);
      }
      for (; v < vend; v += 256) fold(__ldg(p16 + v));
      e = a1 + threadIdx.x;
      if (e < hi) acc = Red<RED, A>::op(acc, to_acc<T>(src[e]));
      e = hi;
    }
    for (; e + 3 * 256 < hi; e += 4 * 256) {
      T v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t r = perm ? (int64_t)perm[e + q * 256] : e + q * 256;
        v[q] = src[r * width + f];
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) acc = Red<RED, A>::op(acc, to_acc<T>(v[q]));
    }
    for (; e < hi; e += 256) acc = Red<RED, A>::op(acc, to_acc<T>(src[(perm ? (int64_t)perm[e] : e) * width + f]));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc = Red<RED, A>::op(acc, __shfl_xor_sync(0xffffffffu, acc, d));
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
      A a = threadIdx.x < 8 ? part[threadIdx.x] : Red<RED, A>::init();
#pragma unroll
      for (int d = 4; d > 0; d >>= 1) a = Red<RED, A>::op(a, __shfl_xor_sync(0xffffffffu, a, d));
      if (threadIdx.x == 0) {
        if (parts == 1) {
          out[gid] = from_acc<T>(Red<RED, A>::finish(a, hi - lo));
        } else {
          // the part that arrives last folds all partials IN PART ORDER: the result does not depend on which one it is
          volatile A *pv = partial + li * SR_SPLIT;
          pv[py] = a;
          __threadfence();
          if (atomicAdd(&long_arrive[li], 1u) == (unsigned)parts - 1u) {
            __threadfence();
            A r = pv[0];
            for (int q = 1; q < parts; ++q) r = Red<RED, A>::op(r, pv[q]);
            const int64_t sgm = gid / width;
            out[gid] = from_acc<T>(Red<RED, A>::finish(r, ptr[sgm + 1] - ptr[sgm]));
          }
        }
      }
    }
    __syncthreads();
  }
}

// ============================================================================= segment reduce, load-balanced over ENTRIES
// The kernels above give every segment a group of lanes: a warp's life time (two dependent pointer loads, a few data
// round trips, the shuffle tree, one store: ~4 us) times the number of warp rounds bounds them, not bytes - which is
// why bf16 rows took the time of f32 rows (0.37 against 0.67 of HBM).  This one is the merge-path form: a CTA takes a
// tile of 4,096 CONSECUTIVE entries whatever the segments look like (16 per thread, 16-byte loads), every thread folds
// the runs of equal segment inside its 16 entries, runs that begin and end inside one thread are written at once, the
// open runs at the threads' edges are combined by a segmented scan over the 256 threads (left to right: a fixed
// order), and a segment that crosses tiles leaves one partial per tile: `carry` where it leaves a tile, `head` where
// it ends; a small second kernel folds them in tile order.  [nnz] vectors of <= 4-byte types, no permutation.
constexpr int ST_THREADS = 256, ST_ITEMS = 16, ST_TILE = ST_THREADS * ST_ITEMS;
struct StTile {       // what a tile leaves for the fix-up kernel
  long long head_seg;  // segment that began in an earlier tile and ENDS in this one (-1: none); its part here: head
  int carry;           // 0: nothing leaves the tile; 1: the last segment goes on; 2: and it spans the whole tile
  int pad;
};

template <typename T, int RED>
__global__ void __launch_bounds__(ST_THREADS)
segment_tile_kernel(const T *__restrict__ src, const int64_t *__restrict__ ptr, int64_t S, int64_t E,
                    const int32_t *__restrict__ bounds, T *__restrict__ out, StTile *__restrict__ tinfo,
                    typename Acc<T>::type *__restrict__ thead, typename Acc<T>::type *__restrict__ tcarry, int vec16) {
  using A = typename Acc<T>::type;
  using R = Red<RED, A>;
  constexpr int THREADS = ST_THREADS, ITEMS = ST_ITEMS, TILE = ST_TILE, WARPS = THREADS / 32;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  __shared__ int32_t sh_first[THREADS], sh_last[THREADS];
  __shared__ A sh_val[THREADS];
  __shared__ A w_val[WARPS];
  __shared__ int w_flag[WARPS];
  __shared__ int sh_before;  // the tile's first segment began in an earlier tile
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const int64_t tile = blockIdx.x;
  const int64_t e0 = tile * TILE;
  const int tile_n = (int)((E - e0) < TILE ? (E - e0) : TILE);
  const int p0 = (int)tid * ITEMS;
  const int nvalid = tile_n - p0 < ITEMS ? (tile_n - p0 > 0 ? tile_n - p0 : 0) : ITEMS;
  // loads that do not depend on the look-up are requested first (a CTA's life is a chain of dependent round trips)
  const int64_t pbeg = ptr[0], pend = ptr[S];
  int64_t first_start = 0;  // thread 0: where the tile's first segment starts
  if (tid == 0) first_start = ptr[bounds[tile]];
  // ---- the thread's values: requested FIRST, so that their latency overlaps the segment look-up below (one after
  // the other the two dependent chains made a CTA live ~4.5 us for its 16 KB)
  constexpr int V = 16 / (int)sizeof(T);  // entries per 16-byte load
  constexpr int NV = ITEMS / V;
  const T *base = src + e0 + p0;
  const bool vec = vec16 && nvalid == ITEMS;
  uint4 raw[NV];  // only LOADED here: the conversions wait until the look-up is done
  if (vec) {
#pragma unroll
    for (int q = 0; q < NV; ++q) raw[q] = __ldg(reinterpret_cast<const uint4 *>(base) + q);
  }
  // ---- segments of my entries.  The usual case - all 16 in ONE segment (rows are long) - is found by two binary
  // searches of the tile's sorted table of segment starts and needs no per-entry work at all: resolving a segment id
  // for every entry (expand_tile_owned) cost ~48 instructions per entry and made the kernel issue-bound at 2 TB/s.
  int32_t own[ITEMS];
  bool uniform = false;  // all my entries lie in segment own[0]
  bool masked = false;   // segment starts given by `starts` (else by own[])
  uint32_t starts = 0;
  int b_lo = 0, b_hi = 0;  // table entries [b_lo, b_hi) are the starts inside my entries
  const bool inside = e0 >= pbeg && e0 + TILE <= pend;  // (else some entries may belong to no segment)
  if (tid < 2) aux[32 + tid] = bounds[tile + tid];
  __syncthreads();
  const int32_t s_lo = aux[32], s_hi = aux[33];
  const int nseg = s_hi - s_lo;
  if (nseg <= TILE && inside) {  // uniform over the CTA
    for (int i = tid; i < nseg; i += THREADS) {
      const int64_t p = ptr[(int64_t)s_lo + 1 + i] - e0;
      seg[i] = p < 0 ? 0 : (p > TILE ? TILE : (int32_t)p);
    }
    __syncthreads();
    auto count_le = [&](int lo, int hi, int p) {  // first index in [lo, hi] whose start exceeds p
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (seg[mid] <= p) lo = mid + 1;
        else hi = mid;
      }
      return lo;
    };
    b_lo = count_le(0, nseg, p0);
    b_hi = count_le(b_lo, nseg, p0 + ITEMS - 1);
    own[0] = s_lo + b_lo;
    // Segment starts inside my entries as a 16-bit mask (a warp runs the slow path below as soon as ONE lane has a
    // boundary - two warps in three with 492-entry rows - so that path must be cheap for the lanes that have none)
    for (int b = b_lo; b < b_hi; ++b) starts |= 1u << (seg[b] - p0);  // seg[b] - p0 in [1, ITEMS - 1]
    masked = true;
    uniform = starts == 0u;
  } else {
    __syncthreads();  // (aux is rewritten by the general look-up)
    expand_tile_owned<THREADS, ITEMS>(ptr, S, e0, tile_n, seg, aux, bounds, own);
    if (!inside) {  // entries outside [ptr[0], ptr[S]) belong to no segment (ptr need not cover src): id -1
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        const int64_t e = e0 + p0 + k;
        if (e < pbeg || e >= pend) own[k] = -1;
      }
    }
  }
  A v[ITEMS];
  if (vec) {
#pragma unroll
    for (int q = 0; q < NV; ++q) {
      const T *t = reinterpret_cast<const T *>(&raw[q]);
#pragma unroll
      for (int k = 0; k < V; ++k) v[q * V + k] = to_acc<T>(t[k]);
    }
  } else {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) v[k] = k < nvalid ? to_acc<T>(base[k]) : R::init();
  }
  // ---- runs of equal segment inside the thread
  int32_t first_seg = -2 - (int32_t)tid, last_seg = -2 - (int32_t)tid;  // an idle thread matches nobody
  A first_acc = R::init(), acc = R::init();
  bool single = true;
  if (uniform && nvalid > 0) {  // one run: a plain sum of my values (entries past the tile's end hold the neutral value)
    first_seg = last_seg = own[0];
    A t[ITEMS / 2];
#pragma unroll
    for (int k = 0; k < ITEMS / 2; ++k) t[k] = R::op(v[2 * k], v[2 * k + 1]);
#pragma unroll
    for (int w = ITEMS / 4; w > 0; w >>= 1)
#pragma unroll
      for (int k = 0; k < w; ++k) t[k] = R::op(t[2 * k], t[2 * k + 1]);
    acc = first_acc = t[0];
  } else if (masked && nvalid > 0) {
    int32_t cur = own[0];
    int b = b_lo, run0 = 0;
    first_seg = cur;
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      if (k < nvalid) {
        if ((starts >> k) & 1u) {  // the run of `cur` ended at k - 1
          if (single) {
            first_acc = acc;
            single = false;
          } else {  // it began and ended inside this thread: complete
            out[cur] = from_acc<T>(R::finish(acc, (int64_t)(k - run0)));
          }
          while (b < b_hi && seg[b] <= p0 + k) ++b;  // (empty segments share a start: the last one holds the entries)
          cur = s_lo + b;
          run0 = k;
          acc = R::init();
        }
        acc = R::op(acc, v[k]);
      }
    }
    if (single) first_acc = acc;
    last_seg = cur;
  } else if (nvalid > 0) {
    int32_t cur = own[0];
    int run0 = 0;
    first_seg = cur;
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      if (k < nvalid) {
        if (own[k] != cur) {  // the run of `cur` ended at k - 1
          if (single) {
            first_acc = acc;
            single = false;
          } else if (cur >= 0) {  // it began and ended inside this thread: complete
            out[cur] = from_acc<T>(R::finish(acc, (int64_t)(k - run0)));
          }
          cur = own[k];
          run0 = k;
          acc = R::init();
        }
        acc = R::op(acc, v[k]);
      }
    }
    if (single) first_acc = acc;
    last_seg = cur;
  }
  sh_first[tid] = first_seg;
  sh_last[tid] = last_seg;
  // (the tile's first entry lies in segment bounds[tile] or, past empty segments, in a later one that starts AT e0)
  if (tid == 0) sh_before = (nvalid > 0 && first_seg >= 0 && first_seg == bounds[tile] && first_start < e0) ? 1 : 0;
  __syncthreads();
  const bool cont = tid > 0 && nvalid > 0 && sh_last[tid - 1] == first_seg;  // my first run began in the thread before
  // ---- segmented inclusive scan over the threads: x = the run that is open at the thread's right edge
  A sc = acc;  // (single: the only run; else the last one)
  int fl = (single && cont) ? 0 : 1;  // 1: the open run starts in this thread
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const A o = __shfl_up_sync(0xffffffffu, sc, d);
    const int of = __shfl_up_sync(0xffffffffu, fl, d);
    if (lane >= (unsigned)d) {
      if (!fl) sc = R::op(o, sc);
      fl |= of;
    }
  }
  if (lane == 31) {
    w_val[warp] = sc;
    w_flag[warp] = fl;
  }
  __syncthreads();
  if (warp == 0) {
    A ws = lane < WARPS ? w_val[lane] : R::init();
    int wf = lane < WARPS ? w_flag[lane] : 1;
#pragma unroll
    for (int d = 1; d < WARPS; d <<= 1) {
      const A o = __shfl_up_sync(0xffffffffu, ws, d);
      const int of = __shfl_up_sync(0xffffffffu, wf, d);
      if (lane >= (unsigned)d) {
        if (!wf) ws = R::op(o, ws);
        wf |= of;
      }
    }
    if (lane < WARPS) w_val[lane] = ws;
  }
  __syncthreads();
  if (warp > 0 && !fl) sc = R::op(w_val[warp - 1], sc);
  sh_val[tid] = sc;
  __syncthreads();
  if (nvalid == 0) return;
  const int32_t tile_first = sh_first[0];
  const bool before = sh_before != 0;
  // ---- (a) my first run ended inside me
  if (!single && first_seg >= 0) {
    const A total = cont ? R::op(sh_val[tid - 1], first_acc) : first_acc;
    if (first_seg == tile_first && before) {
      thead[tile] = total;
      tinfo[tile].head_seg = first_seg;
    } else {
      out[first_seg] = from_acc<T>(R::finish(total, RED == PS_MEAN ? ptr[first_seg + 1] - ptr[first_seg] : 1));
    }
  }
  // ---- (b) the run open at my right edge: does it go on in the next thread?
  const bool last_thread = p0 + ITEMS >= tile_n;
  if (!last_thread && sh_first[tid + 1] == last_seg) return;
  if (last_seg < 0) return;
  // (only the thread at the tile's end has to look where its segment ends; a count is needed for the mean alone)
  const int64_t seg_end = (last_thread || RED == PS_MEAN) ? ptr[last_seg + 1] : 0;
  if (last_thread && seg_end > e0 + tile_n) {  // it leaves the tile
    tcarry[tile] = sc;
    tinfo[tile].carry = (last_seg == tile_first && before) ? 2 : 1;
    return;
  }
  if (last_seg == tile_first && before) {
    thead[tile] = sc;
    tinfo[tile].head_seg = last_seg;
  } else {
    out[last_seg] = from_acc<T>(R::finish(sc, RED == PS_MEAN ? seg_end - ptr[last_seg] : 1));
  }
}

// segments that crossed tiles: the tile where one ENDS folds the partials of the tiles before it, in tile order
template <typename T, int RED>
__global__ void __launch_bounds__(256)
segment_tile_fix_kernel(const int64_t *__restrict__ ptr, int64_t tiles, const StTile *__restrict__ tinfo,
                        const typename Acc<T>::type *__restrict__ thead,
                        const typename Acc<T>::type *__restrict__ tcarry, T *__restrict__ out) {
  using A = typename Acc<T>::type;
  using R = Red<RED, A>;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < tiles; t += (int64_t)gridDim.x * blockDim.x) {
    const long long sgm = tinfo[t].head_seg;
    if (sgm < 0) continue;
    // the parts are t0 (carry, not whole), t0 + 1 .. t - 1 (carry, whole), t (head): find t0, then fold left to right
    int64_t t0 = t - 1;
    while (t0 > 0 && tinfo[t0].carry == 2) --t0;
    A acc = tcarry[t0];
    for (int64_t u = t0 + 1; u < t; ++u) acc = R::op(acc, tcarry[u]);
    acc = R::op(acc, thead[t]);
    out[sgm] = from_acc<T>(R::finish(acc, ptr[sgm + 1] - ptr[sgm]));
  }
}

// PS_SEG_TILE (bench knob, read once): 1 = the tile kernel above on long [nnz] vectors.  NOT the default: measured on
// the configs[2] shape (tools/segbench.py, same box) it takes 199 us (f32) / 168 us (bf16) against 104 / 94 us for
// lanes-per-segment, and 143 / 119 us against 127 / 132 us on the R-MAT stream - four block barriers, two binary
// searches and a 256-thread segmented scan per 16 KB tile cost more than they balance (first version, which resolved
// a segment id per entry: 240 us, issue-bound at 2.4 instructions per SM-cycle).
static int seg_tile() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_SEG_TILE");
    v = e ? atoi(e) : 0;
  }
  return v;
}

template <typename T, int RED>
static int segment_tile_typed(const T *src, const int64_t *ptr, int64_t S, int64_t E, T *out, cudaStream_t s) {
  using A = typename Acc<T>::type;
  const int64_t tiles = ceil_div(E, ST_TILE);
  Scratch scratch(s);
  PS_SCRATCH(bounds, scratch, int32_t, tiles + 1);
  PS_SCRATCH(tinfo, scratch, StTile, tiles);
  PS_SCRATCH(thead, scratch, A, tiles);
  PS_SCRATCH(tcarry, scratch, A, tiles);
  PS_CUDA(cudaMemsetAsync(out, 0, sizeof(T) * (size_t)S, s));  // empty segments: 0 (the torch_scatter convention)
  PS_CUDA(cudaMemsetAsync(tinfo, 0xff, sizeof(StTile) * (size_t)tiles, s));  // head_seg = -1; carry is set below
  ProfScope prof(PROF_SEGMENT_REDUCE, s);
  PS_TRY(tile_bounds_launch(ptr, S, E, ST_TILE, bounds, s));
  segment_tile_kernel<T, RED><<<(unsigned)tiles, ST_THREADS, 0, s>>>(src, ptr, S, E, bounds, out, tinfo, thead, tcarry,
                                                                     (reinterpret_cast<uintptr_t>(src) & 15u) == 0);
  PS_LAUNCH_OK();
  segment_tile_fix_kernel<T, RED><<<(unsigned)ceil_div(tiles, 256), 256, 0, s>>>(ptr, tiles, tinfo, thead, tcarry, out);
  PS_LAUNCH_OK();
  return PS_OK;
}

// PS_SEG_VEC (bench knob, read once): 0 = element-wise loads in the main kernel, 1 = 16-byte loads
static int seg_vec() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_SEG_VEC");
    v = e ? atoi(e) : 1;
  }
  return v;
}

template <typename T, int RED>
static int segment_reduce_typed(const T *src, int64_t width, const int64_t *ptr, int64_t S, T *out,
                                const uint32_t *perm, int64_t E_hint, cudaStream_t s) {
  const int64_t outputs = S * width;
  if (outputs == 0) return PS_OK;
  const double avg = S > 0 ? (double)E_hint / (double)S : 0.0;
  if constexpr (sizeof(T) <= 4) {
    // long [nnz] vectors with rows of 16 entries and more: load-balanced over the entries (E_hint is the length of src)
    if (seg_tile() && width == 1 && !perm && E_hint >= (int64_t(1) << 16) && avg >= 16.0 && S < 2147483647LL)
      return segment_tile_typed<T, RED>(src, ptr, S, E_hint, out, s);
  }
  int G = 32;
  if (width >= 4 || avg <= 2.0) G = 1;
  else if (avg <= 12.0) G = 4;
  else if (avg <= 48.0) G = 8;
  const int64_t threads = outputs * G;
  const unsigned blocks = (unsigned)ceil_div(threads, 256);
  // segments a group of G lanes should not walk alone: more than 128 entries per lane
  const int64_t long_thresh = 128 * (int64_t)G;
  const size_t list_cap = (size_t)(E_hint > 0 ? (E_hint / long_thresh + 1) * (width > 0 ? width : 1) : outputs) + 2;
  Scratch scratch(s);
  PS_SCRATCH(long_list, scratch, unsigned long long, (list_cap < (size_t)outputs + 2 ? list_cap : (size_t)outputs + 2) + 2);
  PS_CUDA(cudaMemsetAsync(long_list, 0, 2 * sizeof(unsigned long long), s));
  using A = typename Acc<T>::type;
  const size_t n_list = list_cap < (size_t)outputs + 2 ? list_cap : (size_t)outputs + 2;
  PS_SCRATCH(partial, scratch, A, n_list * SR_SPLIT);
  PS_SCRATCH(long_arrive, scratch, unsigned int, n_list);
  PS_SCRATCH(long_tasks, scratch, SrTask, n_list * SR_SPLIT);
  ProfScope prof(PROF_SEGMENT_REDUCE, s);
#define PS_SR(G_, W1_) \
  segment_reduce_kernel<T, RED, G_, W1_><<<blocks, 256, 0, s>>>(src, width, ptr, S, out, perm, long_thresh, long_list, \
                                                                 long_arrive, long_tasks, seg_vec() && (reinterpret_cast<uintptr_t>(src) & 15u) == 0)
  if (width == 1) {
    switch (G) {
      case 1: PS_SR(1, true); break;
      case 4: PS_SR(4, true); break;
      case 8: PS_SR(8, true); break;
      default: PS_SR(32, true); break;
    }
  } else {
    switch (G) {
      case 1: PS_SR(1, false); break;
      case 4: PS_SR(4, false); break;
      case 8: PS_SR(8, false); break;
      default: PS_SR(32, false); break;
    }
  }
#undef PS_SR
  PS_LAUNCH_OK();
  segment_reduce_long_kernel<T, RED><<<(unsigned)sm_count() * 8, 256, 0, s>>>(
      src, width, ptr, out, perm, long_list, long_arrive, long_tasks, partial,
      (reinterpret_cast<uintptr_t>(src) & 15u) == 0);
  PS_LAUNCH_OK();
  return PS_OK;
}

int segment_reduce_launch(const void *src, int dtype, int64_t width, const int64_t *ptr, int64_t S, int reduce,
                          void *out, const uint32_t *perm32, int64_t E_hint, cudaStream_t s) {
  PS_DISPATCH_DTYPE(dtype, T,
                    PS_DISPATCH_REDUCE(reduce, RED,
                                       return (segment_reduce_typed<T, RED>(reinterpret_cast<const T *>(src), width, ptr,
                                                                            S, reinterpret_cast<T *>(out), perm32,
                                                                            E_hint, s))));
  return PS_OK;
}

// ============================================================================= row gathers
template <typename U, typename IndexT>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const U *__restrict__ src, int64_t upr, const IndexT *__restrict__ index, int64_t n_units,
                   U *__restrict__ out) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_units) return;
  const int64_t r = q / upr, u = q - r * upr;
  out[q] = src[(int64_t)index[r] * upr + u];
}

template <typename IndexT>
int gather_rows_launch(const void *src, int64_t row_bytes, const IndexT *index, int64_t n_out, void *out,
                       cudaStream_t s) {
  if (n_out <= 0 || row_bytes <= 0) return PS_OK;
  const int unit = copy_unit(row_bytes, src, out);
  const int64_t upr = row_bytes / unit, n_units = n_out * upr;
  const unsigned blocks = (unsigned)ceil_div(n_units, 256);
  ProfScope prof(PROF_GATHER, s);
#define PS_G(U) gather_rows_kernel<U, IndexT><<<blocks, 256, 0, s>>>((const U *)src, upr, index, n_units, (U *)out)
  switch (unit) {
    case 16: PS_G(uint4); break;
    case 8: PS_G(uint64_t); break;
    case 4: PS_G(uint32_t); break;
    case 2: PS_G(uint16_t); break;
    default: PS_G(uint8_t); break;
  }
#undef PS_G
  PS_LAUNCH_OK();
  return PS_OK;
}
template int gather_rows_launch<int64_t>(const void *, int64_t, const int64_t *, int64_t, void *, cudaStream_t);
template int gather_rows_launch<uint32_t>(const void *, int64_t, const uint32_t *, int64_t, void *, cudaStream_t);

// gather_csr: out[e,:] = src[segment(e),:], segment ids resolved per output tile
template <typename U, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
gather_csr_kernel(const U *__restrict__ src, int64_t upr, const int64_t *__restrict__ ptr, int64_t S, int64_t E,
                  U *__restrict__ out, const int32_t *__restrict__ bounds) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = (int64_t)blockIdx.x * TILE;
  const int tile_n = (int)((E - e0) < TILE ? (E - e0) : TILE);
  expand_tile<THREADS, ITEMS>(ptr, S, e0, tile_n, seg, aux, bounds);
  const int64_t units = (int64_t)tile_n * upr;
  for (int64_t q = threadIdx.x; q < units; q += THREADS) {
    const int64_t j = q / upr, u = q - j * upr;
    out[(e0 + j) * upr + u] = src[(int64_t)seg[j] * upr + u];
  }
}

// ============================================================================= broadcast multiply / add
// OP: 0 = multiply (mul.py), 1 = add (add.py) - same traversal, one arithmetic instruction apart
enum { BC_MUL = 0, BC_ADD = 1 };
template <typename T, int OP>
__device__ __forceinline__ T bc_values(T a, T b) {
  return OP == BC_MUL ? from_acc<T>(to_acc<T>(a) * to_acc<T>(b)) : from_acc<T>(to_acc<T>(a) + to_acc<T>(b));
}

template <typename T, int OP, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
mul_row_kernel(const T *value, int64_t width, const int64_t *__restrict__ rowptr, int64_t M, int64_t nnz,
               const T *__restrict__ other, T *out, const int32_t *__restrict__ bounds) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = (int64_t)blockIdx.x * TILE;
  const int tile_n = (int)((nnz - e0) < TILE ? (nnz - e0) : TILE);
  expand_tile<THREADS, ITEMS>(rowptr, M, e0, tile_n, seg, aux, bounds);
  const int64_t units = (int64_t)tile_n * width;
  for (int64_t q = threadIdx.x; q < units; q += THREADS) {
    const int64_t j = q / width, f = q - j * width;
    const int64_t o = (e0 + j) * width + f;
    out[o] = bc_values<T, OP>(value[o], other[(int64_t)seg[j] * width + f]);
  }
}

// width == 1: the [nnz] value vector times an [M,1] operand; no divide, ITEMS independent loads per thread
template <typename T, int OP, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
mul_row_w1_kernel(const T *value, const int64_t *__restrict__ rowptr, int64_t M, int64_t nnz,
                  const T *__restrict__ other, T *out, const int32_t *__restrict__ bounds) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = (int64_t)blockIdx.x * TILE;
  const int tile_n = (int)((nnz - e0) < TILE ? (nnz - e0) : TILE);
  T v[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = k * THREADS + threadIdx.x;
    if (j < tile_n) v[k] = value[e0 + j];
  }
  expand_tile<THREADS, ITEMS>(rowptr, M, e0, tile_n, seg, aux, bounds);
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = k * THREADS + threadIdx.x;
    if (j < tile_n) out[e0 + j] = bc_values<T, OP>(v[k], other[seg[j]]);
  }
}

// the element-wise body for one given tile (the partial last tile after the vector kernel)
template <typename T, int OP, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
mul_row_w1_tail_kernel(const T *value, const int64_t *__restrict__ rowptr, int64_t M, int64_t nnz, int64_t tile,
                       const T *__restrict__ other, T *out, const int32_t *__restrict__ bounds) {
  constexpr int TILE = THREADS * ITEMS;
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = tile * TILE;
  const int tile_n = (int)((nnz - e0) < TILE ? (nnz - e0) : TILE);
  T v[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = k * THREADS + threadIdx.x;
    if (j < tile_n) v[k] = value[e0 + j];
  }
  expand_tile<THREADS, ITEMS>(rowptr, M, e0, tile_n, seg, aux, bounds);
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int j = k * THREADS + threadIdx.x;
    if (j < tile_n) out[e0 + j] = bc_values<T, OP>(v[k], other[seg[j]]);
  }
}

// width == 1 on 16-byte-aligned vectors: a thread owns ITEMS consecutive entries, moved as 128-bit loads / stores
// (ITEMS * sizeof(T) is 16, 32 or 64 bytes), and reads their rows out of expand_tile_owned's registers; the partial
// last tile takes the element-wise kernel above.
template <typename T, int OP, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
mul_row_w1_vec_kernel(const T *value, const int64_t *__restrict__ rowptr, int64_t M, int64_t full_tiles,
                      const T *__restrict__ other, T *out, const int32_t *__restrict__ bounds) {
  constexpr int TILE = THREADS * ITEMS;
  constexpr int VECS = ITEMS * (int)sizeof(T) / 16;
  static_assert(ITEMS * sizeof(T) % 16 == 0, "a thread's entries are whole 16-byte vectors");
  __shared__ int32_t seg[TILE];
  __shared__ int32_t aux[40];
  const int64_t e0 = (int64_t)blockIdx.x * TILE;
  const int64_t mine = e0 + (int64_t)threadIdx.x * ITEMS;
  union alignas(16) Pack {
    uint4 q[VECS];
    T v[ITEMS];
    __device__ Pack() {}
  } in, res;
  const uint4 *src = reinterpret_cast<const uint4 *>(value + mine);
#pragma unroll
  for (int k = 0; k < VECS; ++k) in.q[k] = ld_once(src + k);
  int32_t own[ITEMS];
  expand_tile_owned<THREADS, ITEMS>(rowptr, M, e0, TILE, seg, aux, bounds, own);
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) res.v[k] = bc_values<T, OP>(in.v[k], other[own[k]]);
  uint4 *dst = reinterpret_cast<uint4 *>(out + mine);
#pragma unroll
  for (int k = 0; k < VECS; ++k) dst[k] = res.q[k];
}

template <typename T, int OP, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
mul_col_w1_kernel(const T *value, const int64_t *__restrict__ col, int64_t nnz, const T *__restrict__ other, T *out) {
  const int64_t e0 = (int64_t)blockIdx.x * (THREADS * ITEMS) + threadIdx.x;
  int64_t c[ITEMS];
  T v[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int64_t e = e0 + k * THREADS;
    if (e < nnz) {
      c[k] = ld_stream(col + e);
      v[k] = value[e];
    }
  }
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int64_t e = e0 + k * THREADS;
    if (e < nnz) out[e] = bc_values<T, OP>(v[k], other[c[k]]);
  }
}

template <typename T, int OP>
__global__ void __launch_bounds__(256)
mul_col_kernel(const T *value, int64_t width, const int64_t *__restrict__ col, int64_t n_units,
               const T *__restrict__ other, T *out) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_units) return;
  const int64_t e = q / width, f = q - e * width;
  out[q] = bc_values<T, OP>(value[q], other[col[e] * width + f]);
}

template <typename T, int OP>
__global__ void __launch_bounds__(256)
mul_nnz_kernel(const T *value, int64_t n, const T *__restrict__ other, T *out) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < n) out[q] = bc_values<T, OP>(value[q], other[q]);
}

// ============================================================================= scatter reduce (by col)
template <typename A> __device__ __forceinline__ void atomic_red_add(A *p, A v) { atomicAdd(p, v); }
template <> __device__ __forceinline__ void atomic_red_add<int64_t>(int64_t *p, int64_t v) {
  atomicAdd(reinterpret_cast<unsigned long long *>(p), (unsigned long long)v);
}
__device__ __forceinline__ void atomic_red_min(int32_t *p, int32_t v) { atomicMin(p, v); }
__device__ __forceinline__ void atomic_red_max(int32_t *p, int32_t v) { atomicMax(p, v); }
__device__ __forceinline__ void atomic_red_min(int64_t *p, int64_t v) { atomicMin((long long *)p, (long long)v); }
__device__ __forceinline__ void atomic_red_max(int64_t *p, int64_t v) { atomicMax((long long *)p, (long long)v); }
// IEEE order == signed-int order for non-negative floats and reversed unsigned order for negatives
// floats ordered through their bit patterns: the SIGN BIT picks the branch (-0.0 belongs with the negatives: `v >= 0`
// sent it down the signed-integer path, where its pattern is the smallest of all and stuck)
__device__ __forceinline__ void atomic_red_min(float *p, float v) {
  if (__float_as_int(v) >= 0) atomicMin(reinterpret_cast<int *>(p), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned int *>(p), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_red_max(float *p, float v) {
  if (__float_as_int(v) >= 0) atomicMax(reinterpret_cast<int *>(p), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned int *>(p), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_red_min(double *p, double v) {
  if (__double_as_longlong(v) >= 0) atomicMin(reinterpret_cast<long long *>(p), __double_as_longlong(v));
  else atomicMax(reinterpret_cast<unsigned long long *>(p), (unsigned long long)__double_as_longlong(v));
}
__device__ __forceinline__ void atomic_red_max(double *p, double v) {
  if (__double_as_longlong(v) >= 0) atomicMax(reinterpret_cast<long long *>(p), __double_as_longlong(v));
  else atomicMin(reinterpret_cast<unsigned long long *>(p), (unsigned long long)__double_as_longlong(v));
}

template <typename A, int RED>
__global__ void fill_init_kernel(A *p, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = Red<RED, A>::init();
}

template <typename T, int RED, bool W1>
__global__ void __launch_bounds__(256)
scatter_reduce_kernel(const T *__restrict__ src, int64_t width, const int64_t *__restrict__ index, int64_t n_units,
                      typename Acc<T>::type *__restrict__ acc, int *__restrict__ cnt) {
  using A = typename Acc<T>::type;
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_units) return;
  if (W1) width = 1;
  const int64_t e = W1 ? q : q / width, f = W1 ? 0 : q - e * width;
  const int64_t r = index[e];
  const A v = to_acc<T>(src[q]);
  if (RED == PS_SUM || RED == PS_MEAN) atomic_red_add<A>(&acc[r * width + f], v);
  else if (RED == PS_MIN) atomic_red_min(&acc[r * width + f], v);
  else atomic_red_max(&acc[r * width + f], v);
  if (RED != PS_SUM && cnt != nullptr && f == 0) atomicAdd(&cnt[r], 1);
}

template <typename T, int RED>
__global__ void __launch_bounds__(256)
scatter_finish_kernel(const typename Acc<T>::type *__restrict__ acc, const int *__restrict__ cnt,
                      const int64_t *__restrict__ count64, int64_t width, int64_t n_units, T *__restrict__ out) {
  using A = typename Acc<T>::type;
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_units) return;
  const int64_t c = RED == PS_SUM ? 1 : (count64 ? count64[q / width] : (int64_t)cnt[q / width]);
  out[q] = from_acc<T>(Red<RED, A>::finish_floor(acc[q], c));
}

template <typename T, int RED>
static int scatter_reduce_typed(const T *src, int64_t width, const int64_t *index, int64_t n, int64_t dim_size, T *out,
                                const int64_t *count64, cudaStream_t s) {
  using A = typename Acc<T>::type;
  const int64_t out_units = dim_size * width;
  if (out_units == 0) return PS_OK;
  Scratch scratch(s);
  constexpr bool same = sizeof(A) == sizeof(T);  // accumulate in place unless T is a 16-bit float
  A *acc = reinterpret_cast<A *>(out);
  if (!same) {
    PS_SCRATCH(tmp, scratch, A, out_units);
    acc = tmp;
  }
  int *cnt = nullptr;
  if (RED != PS_SUM && count64 == nullptr) {  // a caller that knows the per-output counts saves one atomic per entry
    PS_SCRATCH(c, scratch, int, dim_size);
    PS_CUDA(cudaMemsetAsync(c, 0, sizeof(int) * (size_t)dim_size, s));
    cnt = c;
  }
  fill_init_kernel<A, RED><<<(unsigned)ceil_div(out_units, 256), 256, 0, s>>>(acc, out_units);
  PS_LAUNCH_OK();
  const int64_t n_units = n * width;
  if (n_units > 0) {
    const unsigned blocks = (unsigned)ceil_div(n_units, 256);
    if (width == 1) scatter_reduce_kernel<T, RED, true><<<blocks, 256, 0, s>>>(src, width, index, n_units, acc, cnt);
    else scatter_reduce_kernel<T, RED, false><<<blocks, 256, 0, s>>>(src, width, index, n_units, acc, cnt);
    PS_LAUNCH_OK();
  }
  if (RED != PS_SUM || !same) {
    scatter_finish_kernel<T, RED><<<(unsigned)ceil_div(out_units, 256), 256, 0, s>>>(acc, cnt, count64, width, out_units,
                                                                                      out);
    PS_LAUNCH_OK();
  }
  return PS_OK;
}

// ============================================================================= dense reduction (dim=None)
template <typename T> struct Wide { using type = double; };
template <> struct Wide<int32_t> { using type = int64_t; };
template <> struct Wide<int64_t> { using type = int64_t; };
template <typename T> __device__ __forceinline__ typename Wide<T>::type to_wide(T v) { return (typename Wide<T>::type)to_acc<T>(v); }
template <typename T> __device__ __forceinline__ T from_wide(typename Wide<T>::type v) {
  return from_acc<T>((typename Acc<T>::type)v);
}

template <typename T, int RED, int THREADS>
__global__ void __launch_bounds__(THREADS)
reduce_all_kernel(const T *__restrict__ src, int64_t n, typename Wide<T>::type *__restrict__ partial) {
  using W = typename Wide<T>::type;
  __shared__ W sh[THREADS / 32];
  W acc = Red<RED, W>::init();
  const int64_t stride = (int64_t)gridDim.x * THREADS;
  for (int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride)
    acc = Red<RED, W>::op(acc, to_wide<T>(src[i]));
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc = Red<RED, W>::op(acc, __shfl_xor_sync(0xffffffffu, acc, d));
  if (lane_id() == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = threadIdx.x < THREADS / 32 ? sh[threadIdx.x] : Red<RED, W>::init();
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc = Red<RED, W>::op(acc, __shfl_xor_sync(0xffffffffu, acc, d));
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
  }
}

template <typename T, int RED>
__global__ void reduce_all_finish_kernel(const typename Wide<T>::type *__restrict__ partial, int nblocks, int64_t n,
                                         T *__restrict__ out) {
  using W = typename Wide<T>::type;
  W acc = Red<RED, W>::init();
  for (int i = threadIdx.x; i < nblocks; i += 32) acc = Red<RED, W>::op(acc, partial[i]);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc = Red<RED, W>::op(acc, __shfl_xor_sync(0xffffffffu, acc, d));
  if (threadIdx.x == 0) out[0] = from_wide<T>(Red<RED, W>::finish(acc, n));
}

template <typename T, int RED>
static int reduce_all_typed(const T *src, int64_t n, T *out, cudaStream_t s) {
  using W = typename Wide<T>::type;
  constexpr int THREADS = 256;
  int64_t blocks = ceil_div(n > 0 ? n : 1, THREADS * 8);
  const int64_t cap = 4 * (int64_t)sm_count();
  if (blocks > cap) blocks = cap;
  Scratch scratch(s);
  PS_SCRATCH(partial, scratch, W, blocks);
  reduce_all_kernel<T, RED, THREADS><<<(unsigned)blocks, THREADS, 0, s>>>(src, n, partial);
  PS_LAUNCH_OK();
  reduce_all_finish_kernel<T, RED><<<1, 32, 0, s>>>(partial, (int)blocks, n, out);
  PS_LAUNCH_OK();
  return PS_OK;
}

}  // namespace ps

namespace ps {

template <int OP>
static int bc_row(const void *value, int dtype, int64_t width, const int64_t *rowptr, int64_t M, int64_t nnz,
                  const void *other, void *out, cudaStream_t s) {
  PS_CHECK(width >= 1 && M >= 0 && nnz >= 0, "mul/add_row: bad sizes");
  PS_CHECK(M < 2147483647LL, "mul/add_row: too many rows");
  if (nnz == 0) return PS_OK;
  constexpr int THREADS = 256, ITEMS = 8;
  const bool aligned = (reinterpret_cast<uintptr_t>(value) | reinterpret_cast<uintptr_t>(out)) % 16 == 0;
  // the vector kernel moves 32 bytes per thread at least: 16 entries of a 2-byte type, 8 otherwise
  const bool vec = width == 1 && aligned;
  const int tile = THREADS * (vec && dtype_size(dtype) == 2 ? 2 * ITEMS : ITEMS);
  const unsigned blocks = (unsigned)ceil_div(nnz, tile);
  Scratch scratch(s);
  PS_SCRATCH(bounds, scratch, int32_t, (int64_t)blocks + 1);
  if (int rc = tile_bounds_launch(rowptr, M, nnz, tile, bounds, s)) return rc;
  const int64_t full_tiles = nnz / tile;
  if (vec) {
    if (full_tiles > 0) {
      PS_DISPATCH_DTYPE(dtype, T,
                        (mul_row_w1_vec_kernel<T, OP, THREADS, (sizeof(T) == 2 ? 2 * ITEMS : ITEMS)>
                         <<<(unsigned)full_tiles, THREADS, 0, s>>>(
                             reinterpret_cast<const T *>(value), rowptr, M, full_tiles,
                             reinterpret_cast<const T *>(other), reinterpret_cast<T *>(out), bounds)));
      PS_LAUNCH_OK();
    }
    if (full_tiles * tile < nnz) {  // the partial last tile: same tile index, element-wise
      PS_DISPATCH_DTYPE(dtype, T,
                        (mul_row_w1_tail_kernel<T, OP, THREADS, (sizeof(T) == 2 ? 2 * ITEMS : ITEMS)>
                         <<<1, THREADS, 0, s>>>(
                             reinterpret_cast<const T *>(value), rowptr, M, nnz, full_tiles,
                             reinterpret_cast<const T *>(other), reinterpret_cast<T *>(out), bounds)));
    }
  } else if (width == 1) {
    PS_DISPATCH_DTYPE(dtype, T,
                      (mul_row_w1_kernel<T, OP, THREADS, ITEMS><<<blocks, THREADS, 0, s>>>(
                          reinterpret_cast<const T *>(value), rowptr, M, nnz, reinterpret_cast<const T *>(other),
                          reinterpret_cast<T *>(out), bounds)));
  } else {
    PS_DISPATCH_DTYPE(dtype, T,
                      (mul_row_kernel<T, OP, THREADS, ITEMS><<<blocks, THREADS, 0, s>>>(
                          reinterpret_cast<const T *>(value), width, rowptr, M, nnz,
                          reinterpret_cast<const T *>(other), reinterpret_cast<T *>(out), bounds)));
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

template <int OP>
static int bc_col(const void *value, int dtype, int64_t width, const int64_t *col, int64_t nnz, const void *other,
                  void *out, cudaStream_t s) {
  PS_CHECK(width >= 1 && nnz >= 0, "mul/add_col: bad sizes");
  const int64_t n_units = nnz * width;
  if (n_units == 0) return PS_OK;
  if (width == 1) {
    constexpr int THREADS = 256, ITEMS = 4;
    const unsigned wblocks = (unsigned)ceil_div(nnz, THREADS * ITEMS);
    PS_DISPATCH_DTYPE(dtype, T,
                      (mul_col_w1_kernel<T, OP, THREADS, ITEMS><<<wblocks, THREADS, 0, s>>>(
                          reinterpret_cast<const T *>(value), col, nnz, reinterpret_cast<const T *>(other),
                          reinterpret_cast<T *>(out))));
  } else {
    const unsigned blocks = (unsigned)ceil_div(n_units, 256);
    PS_DISPATCH_DTYPE(dtype, T,
                      (mul_col_kernel<T, OP><<<blocks, 256, 0, s>>>(reinterpret_cast<const T *>(value), width, col,
                                                                    n_units, reinterpret_cast<const T *>(other),
                                                                    reinterpret_cast<T *>(out))));
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

template <int OP>
static int bc_nnz(const void *value, int dtype, int64_t n, const void *other, void *out, cudaStream_t s) {
  PS_CHECK(n >= 0, "mul/add_nnz: negative size");
  if (n == 0) return PS_OK;
  const unsigned blocks = (unsigned)ceil_div(n, 256);
  PS_DISPATCH_DTYPE(dtype, T,
                    (mul_nnz_kernel<T, OP><<<blocks, 256, 0, s>>>(reinterpret_cast<const T *>(value), n,
                                                                  reinterpret_cast<const T *>(other),
                                                                  reinterpret_cast<T *>(out))));
  PS_LAUNCH_OK();
  return PS_OK;
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_segment_reduce(const void *src, int dtype, int64_t width, int64_t E, const int64_t *ptr, int64_t S, int reduce,
                      void *out, void *stream) {
  PS_CHECK(S >= 0 && width >= 1 && E >= 0, "segment_reduce: bad sizes");
  if (S == 0) return PS_OK;
  // E only steers the lanes-per-segment choice (average segment length)
  return segment_reduce_launch(src, dtype, width, ptr, S, reduce, out, nullptr, E, (cudaStream_t)stream);
}

int ps_gather_rows(const void *src, int64_t row_bytes, const int64_t *index, int64_t n_out, void *out, void *stream) {
  PS_CHECK(row_bytes >= 0 && n_out >= 0, "gather_rows: negative size");
  return gather_rows_launch<int64_t>(src, row_bytes, index, n_out, out, (cudaStream_t)stream);
}

int ps_gather_csr(const void *src, int64_t row_bytes, const int64_t *ptr, int64_t S, int64_t E, void *out,
                  void *stream) {
  PS_CHECK(row_bytes >= 0 && S >= 0 && E >= 0, "gather_csr: negative size");
  PS_CHECK(S < 2147483647LL, "gather_csr: too many segments");
  if (E == 0 || row_bytes == 0) return PS_OK;
  PS_CHECK(S > 0, "gather_csr: E > 0 with no segments");
  constexpr int THREADS = 256, ITEMS = 8;
  const unsigned blocks = (unsigned)ceil_div(E, THREADS * ITEMS);
  const int unit = copy_unit(row_bytes, src, out);
  const int64_t upr = row_bytes / unit;
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  PS_SCRATCH(bounds, scratch, int32_t, (int64_t)blocks + 1);
  if (int rc = tile_bounds_launch(ptr, S, E, THREADS * ITEMS, bounds, s)) return rc;
#define PS_G(U) \
  gather_csr_kernel<U, THREADS, ITEMS><<<blocks, THREADS, 0, s>>>((const U *)src, upr, ptr, S, E, (U *)out, bounds)
  switch (unit) {
    case 16: PS_G(uint4); break;
    case 8: PS_G(uint64_t); break;
    case 4: PS_G(uint32_t); break;
    case 2: PS_G(uint16_t); break;
    default: PS_G(uint8_t); break;
  }
#undef PS_G
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_scatter_reduce(const void *src, int dtype, int64_t width, const int64_t *index, int64_t n, int64_t dim_size,
                      int reduce, void *out, const int64_t *count, void *stream) {
  PS_CHECK(n >= 0 && dim_size >= 0 && width >= 1, "scatter_reduce: bad sizes");
  cudaStream_t s = (cudaStream_t)stream;
  PS_DISPATCH_DTYPE(dtype, T,
                    PS_DISPATCH_REDUCE(reduce, RED,
                                       return (scatter_reduce_typed<T, RED>(reinterpret_cast<const T *>(src), width,
                                                                            index, n, dim_size,
                                                                            reinterpret_cast<T *>(out), count, s))));
  return PS_OK;
}

int ps_reduce_all(const void *src, int dtype, int64_t n, int reduce, void *out, void *stream) {
  PS_CHECK(n >= 0, "reduce_all: negative size");
  cudaStream_t s = (cudaStream_t)stream;
  PS_DISPATCH_DTYPE(dtype, T,
                    PS_DISPATCH_REDUCE(reduce, RED,
                                       return (reduce_all_typed<T, RED>(reinterpret_cast<const T *>(src), n,
                                                                        reinterpret_cast<T *>(out), s))));
  return PS_OK;
}

int ps_mul_row(const void *value, int dtype, int64_t width, const int64_t *rowptr, int64_t M, int64_t nnz,
               const void *other, void *out, void *stream) {
  return bc_row<BC_MUL>(value, dtype, width, rowptr, M, nnz, other, out, (cudaStream_t)stream);
}
int ps_add_row(const void *value, int dtype, int64_t width, const int64_t *rowptr, int64_t M, int64_t nnz,
               const void *other, void *out, void *stream) {
  return bc_row<BC_ADD>(value, dtype, width, rowptr, M, nnz, other, out, (cudaStream_t)stream);
}
int ps_mul_col(const void *value, int dtype, int64_t width, const int64_t *col, int64_t nnz, const void *other,
               void *out, void *stream) {
  return bc_col<BC_MUL>(value, dtype, width, col, nnz, other, out, (cudaStream_t)stream);
}
int ps_add_col(const void *value, int dtype, int64_t width, const int64_t *col, int64_t nnz, const void *other,
               void *out, void *stream) {
  return bc_col<BC_ADD>(value, dtype, width, col, nnz, other, out, (cudaStream_t)stream);
}
int ps_mul_nnz(const void *value, int dtype, int64_t n, const void *other, void *out, void *stream) {
  return bc_nnz<BC_MUL>(value, dtype, n, other, out, (cudaStream_t)stream);
}
int ps_add_nnz(const void *value, int dtype, int64_t n, const void *other, void *out, void *stream) {
  return bc_nnz<BC_ADD>(value, dtype, n, other, out, (cudaStream_t)stream);
}

}  // extern "C"
