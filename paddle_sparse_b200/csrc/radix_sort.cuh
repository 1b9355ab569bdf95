// Stable LSD radix sort of (key, 32-bit payload) pairs for sm_100a — the workhorse behind
// sort-on-construct / coalesce (storage.py:158-171, 454-486: the reference calls
// Tensor.argsort on 64-bit keys, utils.py:21-23) and csr2csc (storage.py:425-434).
//
// Design (one-sweep: every pass reads and writes each pair exactly once):
//   1. hist kernel   — one read of the keys builds the digit histograms of ALL passes in
//                      shared memory; a 256-thread scan turns them into bin bases.
//   2. pass kernel   — per tile: keys loaded warp-striped (coalesced), ranked with
//                      __match_any_sync + per-warp digit counters (no atomics), tile digit
//                      totals chained to earlier tiles by decoupled look-back (one 64-bit
//                      status word per tile x digit, epoch-tagged so a single memset serves
//                      every pass), pairs staged in shared memory in tile-sorted order and
//                      written out as contiguous per-digit runs (coalesced).
// The first pass can read its keys straight from the int64 (row, col) arrays (packing
// row << cbits | col on the fly) and the last pass can widen the payload to int64, so no
// separate pack / unpack passes touch HBM.
#pragma once
#include "common.cuh"

namespace ps {

constexpr int RS_RADIX = 256;
constexpr int RS_MAX_PASSES = 8;

struct SortPlan {
  int passes;
  int begin[RS_MAX_PASSES];
  int bits[RS_MAX_PASSES];
};

inline SortPlan make_sort_plan(int total_bits) {
  SortPlan p;
  if (total_bits < 1) total_bits = 1;
  p.passes = (total_bits + 7) / 8;
  const int base = total_bits / p.passes, rem = total_bits % p.passes;
  int b = 0;
  for (int i = 0; i < p.passes; ++i) {
    p.begin[i] = b;
    p.bits[i] = base + (i < rem ? 1 : 0);
    b += p.bits[i];
  }
  return p;
}

// ---- where the first pass finds its keys / payload, and where the last one puts them
enum { KSRC_KEYS = 0, KSRC_PACK_COO = 1, KSRC_I64 = 2 };
enum { PSRC_U32 = 0, PSRC_IOTA = 1, PSRC_U16 = 2, PSRC_U8 = 3 };
enum { PDST_U32 = 0, PDST_I64 = 1, PDST_U16 = 2, PDST_U8 = 3, PDST_NONE = 4 };
enum { KDST_NONE = 0, KDST_KEYS = 1, KDST_UNPACK_COO = 2 };

struct KeySrc {
  const void *a;  // keys (KSRC_KEYS), row (KSRC_PACK_COO), the int64 array (KSRC_I64)
  const void *b;  // col (KSRC_PACK_COO)
  int mode;
  int cbits;
};
struct PaySrc {
  const void *p;
  int mode;
};
struct PayDst {
  void *p;
  int mode;
};
struct KeyDst {
  void *a;  // keys (KDST_KEYS) or int64 row (KDST_UNPACK_COO)
  void *b;  // int64 col (KDST_UNPACK_COO)
  int mode;
  int cbits;
};

template <typename KeyT>
__device__ __forceinline__ KeyT load_key(const KeySrc &s, int64_t i) {
  if (s.mode == KSRC_KEYS) return ld_stream(reinterpret_cast<const KeyT *>(s.a) + i);
  if (s.mode == KSRC_PACK_COO) {
    const uint64_t r = (uint64_t)ld_stream(reinterpret_cast<const int64_t *>(s.a) + i);
    const uint64_t c = (uint64_t)ld_stream(reinterpret_cast<const int64_t *>(s.b) + i);
    return (KeyT)((r << s.cbits) | c);
  }
  return (KeyT)ld_stream(reinterpret_cast<const int64_t *>(s.a) + i);
}
__device__ __forceinline__ uint32_t load_pay(const PaySrc &s, int64_t i) {
  if (s.mode == PSRC_U32) return ld_stream(reinterpret_cast<const uint32_t *>(s.p) + i);
  if (s.mode == PSRC_IOTA) return (uint32_t)i;
  if (s.mode == PSRC_U16) return (uint32_t) reinterpret_cast<const uint16_t *>(s.p)[i];
  return (uint32_t) reinterpret_cast<const uint8_t *>(s.p)[i];
}
__device__ __forceinline__ void store_pay(const PayDst &d, int64_t i, uint32_t v) {
  if (d.mode == PDST_U32) reinterpret_cast<uint32_t *>(d.p)[i] = v;
  else if (d.mode == PDST_I64) reinterpret_cast<int64_t *>(d.p)[i] = (int64_t)v;
  else if (d.mode == PDST_U16) reinterpret_cast<uint16_t *>(d.p)[i] = (uint16_t)v;
  else if (d.mode == PDST_U8) reinterpret_cast<uint8_t *>(d.p)[i] = (uint8_t)v;
}
template <typename KeyT>
__device__ __forceinline__ void store_key(const KeyDst &d, int64_t i, KeyT k) {
  if (d.mode == KDST_KEYS) reinterpret_cast<KeyT *>(d.a)[i] = k;
  else if (d.mode == KDST_UNPACK_COO) {
    reinterpret_cast<int64_t *>(d.a)[i] = (int64_t)((uint64_t)k >> d.cbits);
    reinterpret_cast<int64_t *>(d.b)[i] = (int64_t)((uint64_t)k & ((uint64_t(1) << d.cbits) - 1));
  }
}

// ----------------------------------------------------------------------------- histogram of all passes
template <typename KeyT>
__global__ void __launch_bounds__(512)
rs_hist_kernel(KeySrc ksrc, int64_t n, SortPlan plan, unsigned long long *__restrict__ ghist /*[passes][256]*/) {
  __shared__ unsigned int sh[RS_MAX_PASSES * RS_RADIX];
  for (int j = threadIdx.x; j < plan.passes * RS_RADIX; j += blockDim.x) sh[j] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const KeyT k = load_key<KeyT>(ksrc, i);
#pragma unroll
    for (int p = 0; p < RS_MAX_PASSES; ++p) {
      if (p < plan.passes) {
        const unsigned d = (unsigned)(k >> plan.begin[p]) & ((1u << plan.bits[p]) - 1u);
        atomicAdd(&sh[p * RS_RADIX + d], 1u);
      }
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < plan.passes * RS_RADIX; j += blockDim.x)
    if (sh[j]) atomicAdd(&ghist[j], (unsigned long long)sh[j]);
}

// exclusive scan of each pass's histogram, in place: grid = passes, block = 256
static __global__ void rs_scan_kernel(unsigned long long *ghist) {
  __shared__ unsigned long long warp_sums[33];
  unsigned long long *h = ghist + (size_t)blockIdx.x * RS_RADIX;
  const unsigned long long v = h[threadIdx.x];
  unsigned long long total;
  const unsigned long long e = block_exclusive_scan<unsigned long long, RS_RADIX>(v, warp_sums, &total);
  h[threadIdx.x] = e;
}

// ----------------------------------------------------------------------------- one-sweep pass
// status word: [63:58] epoch (pass+1), [57:56] status, [55:0] count
struct RsState {
  static constexpr uint64_t AGG = 1, PREFIX = 2;
  __device__ static uint64_t pack(uint64_t epoch, uint64_t st, uint64_t v) { return (epoch << 58) | (st << 56) | v; }
  __device__ static uint64_t epoch(uint64_t w) { return w >> 58; }
  __device__ static uint64_t status(uint64_t w) { return (w >> 56) & 3u; }
  __device__ static uint64_t value(uint64_t w) { return w & ((uint64_t(1) << 56) - 1); }
};

template <typename KeyT, int THREADS, int ITEMS>
struct RsSmem {
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr int WARPS = THREADS / 32;
  KeyT keys[TILE];
  uint32_t pay[TILE];
  uint32_t warp_cnt[WARPS][RS_RADIX];
  uint64_t gbase[RS_RADIX];
  uint32_t dig_excl[RS_RADIX];
  uint32_t scan_tmp[34];
  long long tile;
};

// Tile loads / stores with the (uniform) mode switch hoisted out of the unrolled loops.
template <typename KeyT, int ITEMS>
__device__ __forceinline__ void load_tile_keys(const KeySrc &s, int64_t tile_base, int warp_base, unsigned lane,
                                               int tile_n, KeyT (&key)[ITEMS]) {
  if (s.mode == KSRC_KEYS) {
    const KeyT *p = reinterpret_cast<const KeyT *>(s.a) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      key[i] = j < tile_n ? ld_stream(p + j) : ~KeyT(0);
    }
  } else if (s.mode == KSRC_PACK_COO) {
    const int64_t *r = reinterpret_cast<const int64_t *>(s.a) + tile_base;
    const int64_t *c = reinterpret_cast<const int64_t *>(s.b) + tile_base;
    int64_t rv[ITEMS], cv[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      rv[i] = j < tile_n ? ld_stream(r + j) : -1;
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      cv[i] = j < tile_n ? ld_stream(c + j) : -1;
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      key[i] = j < tile_n ? (KeyT)(((uint64_t)rv[i] << s.cbits) | (uint64_t)cv[i]) : ~KeyT(0);
    }
  } else {
    const int64_t *p = reinterpret_cast<const int64_t *>(s.a) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      key[i] = j < tile_n ? (KeyT)ld_stream(p + j) : ~KeyT(0);
    }
  }
}

template <int ITEMS>
__device__ __forceinline__ void load_tile_pay(const PaySrc &s, int64_t tile_base, int warp_base, unsigned lane,
                                              int tile_n, uint32_t (&pay)[ITEMS]) {
  if (s.mode == PSRC_U32) {
    const uint32_t *p = reinterpret_cast<const uint32_t *>(s.p) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      pay[i] = j < tile_n ? ld_stream(p + j) : 0u;
    }
  } else if (s.mode == PSRC_IOTA) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) pay[i] = (uint32_t)(tile_base + warp_base + i * 32 + lane);
  } else if (s.mode == PSRC_U16) {
    const uint16_t *p = reinterpret_cast<const uint16_t *>(s.p) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      pay[i] = j < tile_n ? (uint32_t)p[j] : 0u;
    }
  } else {
    const uint8_t *p = reinterpret_cast<const uint8_t *>(s.p) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      pay[i] = j < tile_n ? (uint32_t)p[j] : 0u;
    }
  }
}

// One tile: THREADS * ITEMS pairs.  Phases:
//   1 load keys (all loads in flight)
//   2 per item round: match lanes with equal digits, rank inside the warp, warp-private digit counts
//   3 digit totals -> the tile's aggregate is PUBLISHED EARLY (successors rarely wait on it), block scan
//   4 position of every item inside the sorted tile
//   5 stage pairs in tile-sorted order; digit threads finish the decoupled look-back meanwhile
//   6 coalesced scatter of contiguous per-digit runs
template <typename KeyT, int THREADS, int ITEMS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
rs_pass_kernel(KeySrc ksrc, PaySrc psrc, KeyDst kdst, PayDst pdst, int64_t n, int begin_bit, int num_bits, int epoch,
               const unsigned long long *__restrict__ bin_base /*[256]*/,
               uint64_t *__restrict__ tile_state /*[tiles][256]*/, unsigned long long *ticket) {
  constexpr int TILE = THREADS * ITEMS;
  constexpr int WARPS = THREADS / 32;
  static_assert(THREADS >= RS_RADIX, "one thread per digit is required");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RsSmem<KeyT, THREADS, ITEMS> &sm = *reinterpret_cast<RsSmem<KeyT, THREADS, ITEMS> *>(smem_raw);

  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  if (tid == 0) sm.tile = (long long)atomicAdd(ticket, 1ull);
  for (int j = tid; j < WARPS * RS_RADIX; j += THREADS) (&sm.warp_cnt[0][0])[j] = 0;
  __syncthreads();
  const int64_t tile = sm.tile;
  const int64_t tile_base = tile * TILE;
  const int tile_n = (int)((n - tile_base) < TILE ? (n - tile_base) : TILE);
  const unsigned mask = (1u << num_bits) - 1u;
  const int warp_base = warp * (32 * ITEMS);

  // ---- 1: keys, digits, match masks (padding keys are all-ones: the largest digit, ranked last)
  KeyT key[ITEMS];
  load_tile_keys<KeyT, ITEMS>(ksrc, tile_base, warp_base, lane, tile_n, key);
  // ---- 2: rank inside the warp, item round by item round (stable), with warp-private digit counters:
  // plain LDS / STS (shared-memory atomics cost 2 cycles per lane on this part) ordered by __syncwarp
  unsigned peers[ITEMS];  // becomes the rank of the item inside its (warp, digit) group
  uint32_t *my_cnt = sm.warp_cnt[warp];
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const unsigned d = (unsigned)(key[i] >> begin_bit) & mask;
    const unsigned m = __match_any_sync(0xffffffffu, d);
    const uint32_t before = my_cnt[d];
    __syncwarp();
    if ((m & lt) == 0) my_cnt[d] = before + __popc(m);  // lowest peer lane updates
    __syncwarp();
    peers[i] = before + __popc(m & lt);
  }
  __syncthreads();

  // ---- 3: per digit: exclusive prefix over warps, tile total; publish the aggregate right away
  uint32_t total = 0;
  uint64_t *st = tile_state + tid;  // word of (tile t, digit tid) at st[t * 256]
  if (tid < RS_RADIX) {
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = sm.warp_cnt[w][tid];
      sm.warp_cnt[w][tid] = total;
      total += c;
    }
    TileState::store(st + tile * RS_RADIX, RsState::pack(epoch, tile == 0 ? RsState::PREFIX : RsState::AGG, total));
  }
  uint32_t tile_total;
  const uint32_t dexcl = block_exclusive_scan<uint32_t, THREADS>(total, sm.scan_tmp, &tile_total);
  if (tid < RS_RADIX) sm.dig_excl[tid] = dexcl;
  __syncthreads();

  // ---- 4: position inside the sorted tile
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const unsigned d = (unsigned)(key[i] >> begin_bit) & mask;
    peers[i] += sm.dig_excl[d] + my_cnt[d];
  }

  // ---- 5: stage in tile-sorted order (payloads are loaded late: registers, see -Xptxas -v)
  {
    uint32_t pay[ITEMS];
    load_tile_pay<ITEMS>(psrc, tile_base, warp_base, lane, tile_n, pay);
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) sm.keys[peers[i]] = key[i];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) sm.pay[peers[i]] = pay[i];
  }
  // look-back: by now the predecessors have usually published their inclusive prefix
  if (tid < RS_RADIX) {
    uint64_t excl = 0;
    if (tile > 0) {
      // walk back LB tiles per step with all LB loads in flight (the walk is latency-bound: a serial
      // walk costs one L2 round trip per predecessor that has only published its aggregate)
      constexpr int LB = 8;
      int64_t t = tile - 1;
      bool done = false;
      while (!done) {
        uint64_t w[LB];
#pragma unroll
        for (int q = 0; q < LB; ++q) w[q] = (t - q >= 0) ? TileState::load(st + (t - q) * RS_RADIX) : 0;
#pragma unroll
        for (int q = 0; q < LB; ++q) {
          if (done) break;
          if (RsState::epoch(w[q]) != (uint64_t)epoch) {  // not published in this pass yet: resume here
            t -= q;
            goto next_round;
          }
          excl += RsState::value(w[q]);
          if (RsState::status(w[q]) == RsState::PREFIX) done = true;
        }
        t -= LB;
      next_round:;
      }
      TileState::store(st + tile * RS_RADIX, RsState::pack(epoch, RsState::PREFIX, excl + total));
    }
    sm.gbase[tid] = (uint64_t)bin_base[tid] + excl - (uint64_t)dexcl;
  }
  __syncthreads();

  // ---- 6: coalesced scatter: consecutive staged slots of one digit go to consecutive addresses
  const bool full = tile_n == TILE;
  if (kdst.mode == KDST_KEYS && pdst.mode == PDST_U32) {
    KeyT *ko = reinterpret_cast<KeyT *>(kdst.a);
    uint32_t *po = reinterpret_cast<uint32_t *>(pdst.p);
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const int j = k * THREADS + tid;
      if (full || j < tile_n) {
        const KeyT kk = sm.keys[j];
        const unsigned d = (unsigned)(kk >> begin_bit) & mask;
        const int64_t dst = (int64_t)(sm.gbase[d] + (uint64_t)j);
        ko[dst] = kk;
        po[dst] = sm.pay[j];
      }
    }
  } else {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const int j = k * THREADS + tid;
      if (full || j < tile_n) {
        const KeyT kk = sm.keys[j];
        const unsigned d = (unsigned)(kk >> begin_bit) & mask;
        const int64_t dst = (int64_t)(sm.gbase[d] + (uint64_t)j);
        store_key<KeyT>(kdst, dst, kk);
        store_pay(pdst, dst, sm.pay[j]);
      }
    }
  }
}

// ----------------------------------------------------------------------------- host driver
// Tile geometry per key type (measured on B200, profiles/): 8-byte keys 256 x 12 with 4 CTAs per SM,
// 4-byte keys 256 x 16.  PS_RS_VARIANT (bench knob, read once) overrides: 0 = default.
int rs_variant();

template <typename KeyT, int THREADS, int ITEMS, int MIN_BLOCKS>
int rs_launch_pass(KeySrc ks, PaySrc ps_, KeyDst ko, PayDst pd, int64_t n, int begin, int bits, int epoch,
                   const unsigned long long *bin_base, uint64_t *state, unsigned long long *ticket, cudaStream_t s) {
  constexpr int TILE = THREADS * ITEMS;
  const size_t smem = sizeof(RsSmem<KeyT, THREADS, ITEMS>);
  PS_CUDA(cudaFuncSetAttribute(rs_pass_kernel<KeyT, THREADS, ITEMS, MIN_BLOCKS>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t tiles = ceil_div(n, TILE);
  {
    ProfScope prof(PROF_RS_PASS, s);
    rs_pass_kernel<KeyT, THREADS, ITEMS, MIN_BLOCKS><<<(unsigned)tiles, THREADS, smem, s>>>(
        ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

template <typename KeyT> struct RsGeom;
template <> struct RsGeom<uint64_t> { static constexpr int T0 = 256, I0 = 12, B0 = 4; };
template <> struct RsGeom<uint32_t> { static constexpr int T0 = 256, I0 = 16, B0 = 4; };

template <typename KeyT>
int rs_tile_size() {
  switch (rs_variant()) {
    case 1: return 512 * 12;
    case 2: return 256 * 16;
    case 3: return 384 * 16;
    case 4: return 256 * 8;
    default: return RsGeom<KeyT>::T0 * RsGeom<KeyT>::I0;
  }
}

template <typename KeyT>
int rs_dispatch_pass(KeySrc ks, PaySrc ps_, KeyDst ko, PayDst pd, int64_t n, int begin, int bits, int epoch,
                     const unsigned long long *bin_base, uint64_t *state, unsigned long long *ticket, cudaStream_t s) {
  switch (rs_variant()) {
    case 1: return rs_launch_pass<KeyT, 512, 12, 2>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
    case 2: return rs_launch_pass<KeyT, 256, 16, 3>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
    case 3: return rs_launch_pass<KeyT, 384, 16, 2>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
    case 4: return rs_launch_pass<KeyT, 256, 8, 6>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
    default:
      return rs_launch_pass<KeyT, RsGeom<KeyT>::T0, RsGeom<KeyT>::I0, RsGeom<KeyT>::B0>(ks, ps_, ko, pd, n, begin, bits,
                                                                                        epoch, bin_base, state, ticket, s);
  }
}

// Sorts n pairs on key bits [0, total_bits).  First-pass sources are ksrc/psrc.  The last
// pass writes keys to keys_final (KDST_NONE: keys dropped) and payloads to pay_final.
// keys_tmp/pay_tmp are two scratch buffers of n elements each (unused when 1 pass).
template <typename KeyT>
int radix_sort_pairs(KeySrc ksrc, PaySrc psrc, int64_t n, int total_bits, KeyT *keys_tmp0, KeyT *keys_tmp1,
                     uint32_t *pay_tmp0, uint32_t *pay_tmp1, KeyDst keys_final, PayDst pay_final,
                     Scratch &scratch, cudaStream_t s) {
  if (n <= 0) return PS_OK;
  const SortPlan plan = make_sort_plan(total_bits);
  if (plan.passes > RS_MAX_PASSES) { set_error("radix sort: %d key bits unsupported", total_bits); return PS_ERR_ARG; }
  const int64_t tiles = ceil_div(n, rs_tile_size<KeyT>());

  // one allocation: [passes*256 hist | passes tickets (padded to 32) | tiles*256 state words]
  const size_t hist_words = (size_t)plan.passes * RS_RADIX;
  const size_t total_words = hist_words + 32 + (size_t)tiles * RS_RADIX;
  PS_SCRATCH(words, scratch, uint64_t, total_words);
  PS_CUDA(cudaMemsetAsync(words, 0, total_words * sizeof(uint64_t), s));
  unsigned long long *ghist = reinterpret_cast<unsigned long long *>(words);
  unsigned long long *tickets = ghist + hist_words;
  uint64_t *state = words + hist_words + 32;

  int64_t hb = ceil_div(n, 512 * 16);
  const int64_t cap = 8 * (int64_t)sm_count();
  if (hb > cap) hb = cap;
  {
    ProfScope prof(PROF_RS_HIST, s);
    rs_hist_kernel<KeyT><<<(unsigned)hb, 512, 0, s>>>(ksrc, n, plan, ghist);
  }
  PS_LAUNCH_OK();
  rs_scan_kernel<<<plan.passes, RS_RADIX, 0, s>>>(ghist);
  PS_LAUNCH_OK();

  KeyT *kt[2] = {keys_tmp0, keys_tmp1};
  uint32_t *pt[2] = {pay_tmp0, pay_tmp1};
  for (int p = 0; p < plan.passes; ++p) {
    KeySrc ks = ksrc;
    PaySrc ps_ = psrc;
    if (p > 0) {
      ks = KeySrc{kt[(p - 1) & 1], nullptr, KSRC_KEYS, 0};
      ps_ = PaySrc{pt[(p - 1) & 1], PSRC_U32};
    }
    KeyDst ko{kt[p & 1], nullptr, KDST_KEYS, 0};
    PayDst pd{pt[p & 1], PDST_U32};
    if (p == plan.passes - 1) {
      ko = keys_final;
      pd = pay_final;
    }
    PS_TRY(rs_dispatch_pass<KeyT>(ks, ps_, ko, pd, n, plan.begin[p], plan.bits[p], p + 1,
                                  ghist + (size_t)p * RS_RADIX, state, tickets + p, s));
  }
  return PS_OK;
}

}  // namespace ps
