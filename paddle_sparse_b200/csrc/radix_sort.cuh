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

constexpr int RS_MAX_RBITS = 9;                  // widest digit any pass may use
constexpr int RS_MAX_RADIX = 1 << RS_MAX_RBITS;  // histogram rows are padded to this many bins
constexpr int RS_MAX_PASSES = 8;

struct SortPlan {
  int passes;
  int begin[RS_MAX_PASSES];
  int bits[RS_MAX_PASSES];
};

// fewest passes of at most `rbits` bits each, the bits spread evenly over them
inline SortPlan make_sort_plan(int total_bits, int rbits) {
  SortPlan p;
  if (total_bits < 1) total_bits = 1;
  p.passes = (total_bits + rbits - 1) / rbits;
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
enum { PSRC_U32 = 0, PSRC_IOTA = 1, PSRC_U16 = 2, PSRC_U8 = 3, PSRC_I64 = 4 };
enum { PDST_U32 = 0, PDST_I64 = 1, PDST_U16 = 2, PDST_U8 = 3, PDST_NONE = 4 };
enum { KDST_NONE = 0, KDST_KEYS = 1, KDST_UNPACK_COO = 2, KDST_I64 = 3 };

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
  if (s.mode == PSRC_I64) return (uint32_t)ld_stream(reinterpret_cast<const int64_t *>(s.p) + i);
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
  else if (d.mode == KDST_I64) reinterpret_cast<int64_t *>(d.a)[i] = (int64_t)k;
  else if (d.mode == KDST_UNPACK_COO) {
    reinterpret_cast<int64_t *>(d.a)[i] = (int64_t)((uint64_t)k >> d.cbits);
    reinterpret_cast<int64_t *>(d.b)[i] = (int64_t)((uint64_t)k & ((uint64_t(1) << d.cbits) - 1));
  }
}

// ----------------------------------------------------------------------------- histogram of all passes
template <typename KeyT>
__global__ void __launch_bounds__(512)
rs_hist_kernel(KeySrc ksrc, int64_t n, SortPlan plan, unsigned long long *__restrict__ ghist /*[passes][512]*/) {
  __shared__ unsigned int sh[RS_MAX_PASSES * RS_MAX_RADIX];
  for (int j = threadIdx.x; j < plan.passes * RS_MAX_RADIX; j += blockDim.x) sh[j] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const KeyT k = load_key<KeyT>(ksrc, i);
#pragma unroll
    for (int p = 0; p < RS_MAX_PASSES; ++p) {
      if (p < plan.passes) {
        const unsigned d = (unsigned)(k >> plan.begin[p]) & ((1u << plan.bits[p]) - 1u);
        atomicAdd(&sh[p * RS_MAX_RADIX + d], 1u);
      }
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < plan.passes * RS_MAX_RADIX; j += blockDim.x)
    if (sh[j]) atomicAdd(&ghist[j], (unsigned long long)sh[j]);
}

// exclusive scan of each pass's histogram, in place: grid = passes, block = 512
static __global__ void rs_scan_kernel(unsigned long long *ghist) {
  __shared__ unsigned long long warp_sums[33];
  unsigned long long *h = ghist + (size_t)blockIdx.x * RS_MAX_RADIX;
  const unsigned long long v = h[threadIdx.x];
  unsigned long long total;
  const unsigned long long e = block_exclusive_scan<unsigned long long, RS_MAX_RADIX>(v, warp_sums, &total);
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

// up to three 32-bit payload words travel with every key (structure of arrays in global memory)
template <int NPAY> struct PaySrcN { PaySrc s[NPAY > 0 ? NPAY : 1]; };
template <int NPAY> struct PayDstN { PayDst d[NPAY > 0 ? NPAY : 1]; };

template <typename KeyT, int NPAY, int RBITS, int THREADS, int ITEMS>
struct RsSmem {
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr int WARPS = THREADS / 32;
  static constexpr int RADIX = 1 << RBITS;
  KeyT keys[TILE];                         // tile in sorted order
  uint32_t pay[NPAY > 0 ? NPAY : 1][TILE];  // arrival order (cp.async prefetch), then sorted order
  uint32_t cnt[WARPS][RADIX];              // items of the (warp, digit) group
  uint16_t base[WARPS][RADIX];             // start of the (warp, digit) group inside the sorted tile
  uint32_t gbase[RADIX];                   // global position of the digit's run minus its tile offset
  uint32_t scan_tmp[32];
  long long tile;
};

__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gsrc) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Tile loads with the (uniform) mode switch hoisted out of the unrolled loops.
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
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      key[i] = j < tile_n ? (KeyT)ld_stream(r + j) : ~KeyT(0);
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      if (j < tile_n) key[i] = (KeyT)(((uint64_t)key[i] << s.cbits) | (uint64_t)ld_stream(c + j));
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

// payload words whose source is not a plain u32 array (those are prefetched with cp.async)
template <int ITEMS>
__device__ __forceinline__ void load_tile_pay_sync(const PaySrc &s, int64_t tile_base, int warp_base, unsigned lane,
                                                   int tile_n, uint32_t (&pay)[ITEMS]) {
  if (s.mode == PSRC_IOTA) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) pay[i] = (uint32_t)(tile_base + warp_base + i * 32 + lane);
  } else if (s.mode == PSRC_I64) {
    const int64_t *p = reinterpret_cast<const int64_t *>(s.p) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      pay[i] = j < tile_n ? (uint32_t)ld_stream(p + j) : 0u;
    }
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

// ---- decoupled look-back for one digit: w[] holds the words of tiles t0, t0-1, ... (loaded early so
// the round trip overlaps the staging phase); tiles before 0 read as "prefix 0"
template <int LB>
__device__ __forceinline__ void rs_lb_load(const uint64_t *stw, int radix, int64_t t0, int epoch, uint64_t (&w)[LB]) {
#pragma unroll
  for (int q = 0; q < LB; ++q)
    w[q] = (t0 - q >= 0) ? TileState::load(stw + (t0 - q) * radix) : RsState::pack(epoch, RsState::PREFIX, 0);
}
// The walk continues LB2 tiles per round trip: the prefix frontier trails the newest aggregate by roughly
// (L2 round trip / tile period) tiles, so later rounds must cover many tiles at once to keep up.
template <int LB, int LB2>
__device__ __forceinline__ uint32_t rs_lb_finish(const uint64_t *stw, int radix, int64_t tile, int epoch,
                                                 uint64_t (&w)[LB]) {
  uint64_t excl = 0;
  int64_t t = tile - 1;
  {
    int q = 0;
#pragma unroll
    for (; q < LB; ++q) {
      if (RsState::epoch(w[q]) != (uint64_t)epoch) break;  // not published in this pass yet: resume here
      excl += RsState::value(w[q]);
      if (RsState::status(w[q]) == RsState::PREFIX) return (uint32_t)excl;
    }
    t -= q;
  }
  uint64_t v[LB2];
  while (true) {
    rs_lb_load<LB2>(stw, radix, t, epoch, v);
    int q = 0;
#pragma unroll
    for (; q < LB2; ++q) {
      if (RsState::epoch(v[q]) != (uint64_t)epoch) break;
      excl += RsState::value(v[q]);
      if (RsState::status(v[q]) == RsState::PREFIX) return (uint32_t)excl;
    }
    t -= q;
  }
}

// One tile = THREADS * ITEMS pairs; CTAs are persistent and draw tiles from an atomic ticket (the next
// ticket is requested while the current tile is processed).  Phases per tile:
//   1 keys -> registers (all loads in flight); u32 payload words start moving to shared memory (cp.async)
//   2 rank inside the warp: one shared-memory atomicAdd per item on warp-private digit counters
//   3 digit totals -> the tile's aggregate is PUBLISHED EARLY and the first look-back words are requested;
//     block scan of the totals gives every (warp, digit) group its start inside the sorted tile
//   4 position of every item inside the sorted tile
//   5 stage pairs in tile-sorted order; finish the decoupled look-back
//   6 coalesced scatter of contiguous per-digit runs
// PS_RS_TIMELINE (debug builds only): thread 0 of every CTA accumulates the SM cycles it spends between
// phase markers; ps_debug_timeline() reads the totals.
#ifdef PS_RS_TIMELINE
static __device__ unsigned long long g_rs_timeline[16];  // one copy per translation unit
#define PS_TL_DECL long long tl_prev = clock64(); unsigned long long tl_acc[12] = {0}
#define PS_TL(k) do { if (tid == 0) { const long long now_ = clock64(); tl_acc[k] += (unsigned long long)(now_ - tl_prev); tl_prev = now_; } } while (0)
#define PS_TL_FLUSH do { if (tid == 0) for (int k_ = 0; k_ < 12; ++k_) atomicAdd(&g_rs_timeline[k_], tl_acc[k_]); } while (0)
#else
#define PS_TL_DECL
#define PS_TL(k)
#define PS_TL_FLUSH
#endif

template <typename KeyT, int NPAY, int RBITS, int THREADS, int ITEMS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
rs_pass_kernel(KeySrc ksrc, PaySrcN<NPAY> psrc, KeyDst kdst, PayDstN<NPAY> pdst, int64_t n, int64_t tiles,
               int begin_bit, int num_bits, int epoch, const unsigned long long *__restrict__ bin_base,
               uint64_t *__restrict__ tile_state /*[tiles][RADIX]*/, unsigned long long *ticket) {
  using Smem = RsSmem<KeyT, NPAY, RBITS, THREADS, ITEMS>;
  constexpr int TILE = Smem::TILE, WARPS = Smem::WARPS, RADIX = Smem::RADIX;
  constexpr int DPT = RADIX > THREADS ? RADIX / THREADS : 1;  // consecutive digits owned by a digit thread
  constexpr int LB = DPT == 1 ? 4 : 2, LB2 = 16 / DPT;
  static_assert(RADIX % (THREADS < RADIX ? THREADS : RADIX) == 0, "THREADS must divide the radix (or exceed it)");
  static_assert(TILE < 65536, "group starts are kept in 16 bits");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem &sm = *reinterpret_cast<Smem *>(smem_raw);

  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned mask = (1u << num_bits) - 1u;
  const unsigned lt = lanemask_lt();
  const int warp_base = warp * (32 * ITEMS);
  uint32_t *my_cnt = sm.cnt[warp];
  const uint16_t *my_base = sm.base[warp];
  const bool digit_thread = (int)tid * DPT < RADIX;
  const int d0 = digit_thread ? tid * DPT : 0;  // first digit this thread owns
  uint64_t *stw = tile_state + d0;               // word of (tile t, digit d0 + q) at stw[t * RADIX + q]
  uint32_t bin0[DPT];
#pragma unroll
  for (int q = 0; q < DPT; ++q) bin0[q] = (uint32_t)bin_base[d0 + q];

  long long next_tile = 0;
  if (tid == 0) next_tile = (long long)atomicAdd(ticket, 1ull);
  PS_TL_DECL;
  while (true) {
    if (tid == 0) sm.tile = next_tile;
    PS_TL(0);
    __syncthreads();  // also: every read of the previous tile's staged pairs is done
    PS_TL(1);
    const int64_t tile = sm.tile;
    if (tile >= tiles) break;
    if (tid == 0) next_tile = (long long)atomicAdd(ticket, 1ull);  // consumed at the top of the next round
    const int64_t tile_base = tile * TILE;
    const int tile_n = (int)((n - tile_base) < TILE ? (n - tile_base) : TILE);

    // ---- 1: keys (padding keys are all-ones: the largest digit, ranked last); payload prefetch
    KeyT key[ITEMS];
    load_tile_keys<KeyT, ITEMS>(ksrc, tile_base, warp_base, lane, tile_n, key);
    if (NPAY > 0) {
#pragma unroll
      for (int p = 0; p < NPAY; ++p) {
        if (psrc.s[p].mode == PSRC_U32) {
          const uint32_t *src = reinterpret_cast<const uint32_t *>(psrc.s[p].p) + tile_base;
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) {
            const int j = warp_base + i * 32 + lane;
            if (j < tile_n) cp_async4(&sm.pay[p][j], src + j);
          }
        }
      }
      cp_async_commit();
    }
#pragma unroll
    for (int q = lane; q < RADIX; q += 32) my_cnt[q] = 0;
    __syncwarp();

    // ---- 2: rank inside the warp (stable): set s covers rounds [s*SI, (s+1)*SI) and owns byte s
    // One shared-memory atomic per item returns the item's rank inside its (warp, digit) group.  Stability
    // needs the lanes of ONE instruction that hit the same counter to be served in ascending lane order:
    // that is how ATOMS behaves on sm_100a (tools/microbench/atom_order.cu: 6e10 lane-instances under every
    // collision / bank-conflict pattern, none out of order; rank_bench.cu: 2 SM-cycles per warp instruction
    // against 40 for MATCH.ANY and 17 for eight ballots) and ps_selftest() re-checks it on the device in use.
    unsigned rank[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) rank[i] = atomicAdd(&my_cnt[(unsigned)(key[i] >> begin_bit) & mask], 1u);
    PS_TL(2);
    __syncthreads();
    PS_TL(3);

    // ---- 3: per digit: exclusive prefix over warps, tile total; publish the aggregate right away
    uint32_t tot[DPT];
    uint64_t lbw[DPT][LB];
#pragma unroll
    for (int q = 0; q < DPT; ++q) tot[q] = 0;
    if (digit_thread) {
#pragma unroll
      for (int q = 0; q < DPT; ++q) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) t += sm.cnt[w][d0 + q];
        tot[q] = t;
        TileState::store(stw + tile * RADIX + q, RsState::pack(epoch, tile == 0 ? RsState::PREFIX : RsState::AGG, t));
      }
#pragma unroll
      for (int q = 0; q < DPT; ++q) rs_lb_load<LB>(stw + q, RADIX, tile - 1, epoch, lbw[q]);
    }
    uint32_t dexcl;  // tile-exclusive start of digit d0
    {
      uint32_t tsum = 0;
#pragma unroll
      for (int q = 0; q < DPT; ++q) tsum += tot[q];
      uint32_t inc = tsum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d) inc += o;
      }
      if (lane == 31) sm.scan_tmp[warp] = inc;
      __syncthreads();
      uint32_t wb = 0;
#pragma unroll
      for (int w = 0; w < WARPS; ++w) wb += (w < (int)warp) ? sm.scan_tmp[w] : 0u;
      dexcl = wb + inc - tsum;
    }
    if (digit_thread) {
      uint32_t dx = dexcl;  // second sweep over the warp counters (cheaper than holding the prefixes in registers)
#pragma unroll
      for (int q = 0; q < DPT; ++q) {
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
          sm.base[w][d0 + q] = (uint16_t)dx;
          dx += sm.cnt[w][d0 + q];
        }
      }
    }
    PS_TL(4);
    __syncthreads();
    PS_TL(5);

    // ---- 4: position inside the sorted tile = group start + rank inside the group
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) rank[i] += (unsigned)my_base[(unsigned)(key[i] >> begin_bit) & mask];

    // ---- 5: stage in tile-sorted order
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) sm.keys[rank[i]] = key[i];
    if (NPAY > 0) {
      cp_async_wait_all();
#pragma unroll
      for (int p = 0; p < NPAY; ++p) {
        uint32_t pv[ITEMS];
        const bool prefetched = psrc.s[p].mode == PSRC_U32;
        if (prefetched) {
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) pv[i] = sm.pay[p][warp_base + i * 32 + lane];
          __syncthreads();  // every arrival-order slot is read before sorted order overwrites it
        } else {
          load_tile_pay_sync<ITEMS>(psrc.s[p], tile_base, warp_base, lane, tile_n, pv);
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) sm.pay[p][rank[i]] = pv[i];
      }
    }
    PS_TL(6);
    // look-back: the first words were requested in phase 3 and have usually arrived by now
    if (digit_thread) {
#pragma unroll
      for (int q = 0; q < DPT; ++q) {
        uint32_t excl = 0;
        if (tile > 0) {
          excl = rs_lb_finish<LB, LB2>(stw + q, RADIX, tile, epoch, lbw[q]);
          TileState::store(stw + tile * RADIX + q, RsState::pack(epoch, RsState::PREFIX, (uint64_t)excl + tot[q]));
        }
        sm.gbase[d0 + q] = bin0[q] + excl - dexcl;
        dexcl += tot[q];
      }
    }
    PS_TL(7);
    __syncthreads();
    PS_TL(8);

    // ---- 6: coalesced scatter: consecutive staged slots of one digit go to consecutive addresses
    const bool full = tile_n == TILE;
    bool plain = kdst.mode == KDST_KEYS;
#pragma unroll
    for (int p = 0; p < NPAY; ++p) plain = plain && pdst.d[p].mode == PDST_U32;
    if (plain) {
      KeyT *ko = reinterpret_cast<KeyT *>(kdst.a);
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        if (full || j < tile_n) {
          const KeyT kk = sm.keys[j];
          const unsigned d = (unsigned)(kk >> begin_bit) & mask;
          const size_t dst = (size_t)(uint32_t)(sm.gbase[d] + (uint32_t)j);
          ko[dst] = kk;
#pragma unroll
          for (int p = 0; p < NPAY; ++p) reinterpret_cast<uint32_t *>(pdst.d[p].p)[dst] = sm.pay[p][j];
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        if (full || j < tile_n) {
          const KeyT kk = sm.keys[j];
          const unsigned d = (unsigned)(kk >> begin_bit) & mask;
          const int64_t dst = (int64_t)(uint32_t)(sm.gbase[d] + (uint32_t)j);
          store_key<KeyT>(kdst, dst, kk);
#pragma unroll
          for (int p = 0; p < NPAY; ++p) store_pay(pdst.d[p], dst, sm.pay[p][j]);
        }
      }
    }
    PS_TL(9);
  }
  PS_TL_FLUSH;
}

// ----------------------------------------------------------------------------- host driver
// Geometry per key type (measured on B200, profiles/).  PS_RS_VARIANT (bench knob, read once) overrides.
int rs_variant();

struct RsGeomRt {
  int rbits, tile;
};

template <typename KeyT, int NPAY, int RBITS, int THREADS, int ITEMS, int MIN_BLOCKS>
int rs_launch_pass(KeySrc ks, PaySrcN<NPAY> ps_, KeyDst ko, PayDstN<NPAY> pd, int64_t n, int begin, int bits,
                   int epoch, const unsigned long long *bin_base, uint64_t *state, unsigned long long *ticket,
                   cudaStream_t s) {
  using Smem = RsSmem<KeyT, NPAY, RBITS, THREADS, ITEMS>;
  auto kern = rs_pass_kernel<KeyT, NPAY, RBITS, THREADS, ITEMS, MIN_BLOCKS>;
  const size_t smem = sizeof(Smem);
  static int resident = 0;  // CTAs per SM of this instantiation (same on every B200 of the box)
  if (resident == 0) {
    PS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int r = 0;
    PS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r, kern, THREADS, smem));
    resident = r > 0 ? r : 1;
  }
  const int64_t tiles = ceil_div(n, Smem::TILE);
  int64_t grid = (int64_t)sm_count() * resident;
  if (grid > tiles) grid = tiles;
  {
    ProfScope prof(PROF_RS_PASS, s);
    kern<<<(unsigned)grid, THREADS, smem, s>>>(ks, ps_, ko, pd, n, tiles, begin, bits, epoch, bin_base, state, ticket);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

// variant table: {rbits, threads, items, min blocks}; 0 is the default of the key type
#define PS_RS_VARIANTS(KeyT, X)                                  \
  switch (rs_variant()) {                                        \
    case 1:                                                      \
      if constexpr (sizeof(KeyT) == 8) { X(8, 512, 12, 2); }     \
      else { X(8, 512, 16, 2); }                                 \
      break;                                                     \
    case 2: X(9, 512, 12, 1); break;                             \
    case 3: X(8, 1024, 12, 1); break;                            \
    default:                                                     \
      if constexpr (sizeof(KeyT) == 8) { X(8, 256, 12, 4); }     \
      else { X(8, 256, 16, 4); }                                 \
      break;                                                     \
  }

template <typename KeyT>
RsGeomRt rs_geometry() {
  RsGeomRt g{8, 256 * 12};
#define PS_X(RB, T, I, B) g = RsGeomRt{RB, (T) * (I)}
  PS_RS_VARIANTS(KeyT, PS_X)
#undef PS_X
  return g;
}

template <typename KeyT, int NPAY>
int rs_dispatch_pass(KeySrc ks, PaySrcN<NPAY> ps_, KeyDst ko, PayDstN<NPAY> pd, int64_t n, int begin, int bits,
                     int epoch, const unsigned long long *bin_base, uint64_t *state, unsigned long long *ticket,
                     cudaStream_t s) {
#define PS_X(RB, T, I, B) \
  return rs_launch_pass<KeyT, NPAY, RB, T, I, B>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s)
  PS_RS_VARIANTS(KeyT, PS_X)
#undef PS_X
  return PS_ERR_ARG;
}

// Sorts n (key, NPAY payload words) tuples on key bits [0, total_bits).  First-pass sources are
// ksrc/psrc.  The last pass writes keys to keys_final (KDST_NONE: keys dropped) and payloads to
// pay_final.  keys_tmp*/pay_tmp* are scratch buffers of n elements each (pay_tmp*[p] for word p;
// unused when one pass suffices).
template <typename KeyT, int NPAY>
int radix_sort_tuples(KeySrc ksrc, PaySrcN<NPAY> psrc, int64_t n, int total_bits, KeyT *keys_tmp0, KeyT *keys_tmp1,
                      uint32_t *const *pay_tmp0, uint32_t *const *pay_tmp1, KeyDst keys_final,
                      PayDstN<NPAY> pay_final, Scratch &scratch, cudaStream_t s) {
  if (n <= 0) return PS_OK;
  const RsGeomRt geom = rs_geometry<KeyT>();
  const SortPlan plan = make_sort_plan(total_bits, geom.rbits);
  if (plan.passes > RS_MAX_PASSES) { set_error("radix sort: %d key bits unsupported", total_bits); return PS_ERR_ARG; }
  const int radix = 1 << geom.rbits;
  const int64_t tiles = ceil_div(n, geom.tile);

  // one allocation: [passes*512 hist | passes tickets (padded to 32) | tiles*radix state words]
  const size_t hist_words = (size_t)plan.passes * RS_MAX_RADIX;
  const size_t total_words = hist_words + 32 + (size_t)tiles * radix;
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
  rs_scan_kernel<<<plan.passes, RS_MAX_RADIX, 0, s>>>(ghist);
  PS_LAUNCH_OK();

  KeyT *kt[2] = {keys_tmp0, keys_tmp1};
  for (int p = 0; p < plan.passes; ++p) {
    KeySrc ks = ksrc;
    PaySrcN<NPAY> ps_ = psrc;
    if (p > 0) {
      ks = KeySrc{kt[(p - 1) & 1], nullptr, KSRC_KEYS, 0};
      for (int w = 0; w < NPAY; ++w) ps_.s[w] = PaySrc{((p - 1) & 1) ? pay_tmp1[w] : pay_tmp0[w], PSRC_U32};
    }
    KeyDst ko{kt[p & 1], nullptr, KDST_KEYS, 0};
    PayDstN<NPAY> pd;
    for (int w = 0; w < NPAY; ++w) pd.d[w] = PayDst{(p & 1) ? pay_tmp1[w] : pay_tmp0[w], PDST_U32};
    if (p == plan.passes - 1) {
      ko = keys_final;
      pd = pay_final;
    }
    PS_TRY((rs_dispatch_pass<KeyT, NPAY>(ks, ps_, ko, pd, n, plan.begin[p], plan.bits[p], p + 1,
                                         ghist + (size_t)p * RS_MAX_RADIX, state, tickets + p, s)));
  }
  return PS_OK;
}

// number of passes radix_sort_* will run for this key type (callers pick the ping-pong buffer with it)
template <typename KeyT>
int rs_num_passes(int total_bits) { return make_sort_plan(total_bits, rs_geometry<KeyT>().rbits).passes; }

// the common single-payload form
template <typename KeyT>
int radix_sort_pairs(KeySrc ksrc, PaySrc psrc, int64_t n, int total_bits, KeyT *keys_tmp0, KeyT *keys_tmp1,
                     uint32_t *pay_tmp0, uint32_t *pay_tmp1, KeyDst keys_final, PayDst pay_final, Scratch &scratch,
                     cudaStream_t s) {
  PaySrcN<1> ps_;
  ps_.s[0] = psrc;
  PayDstN<1> pd;
  pd.d[0] = pay_final;
  uint32_t *t0[1] = {pay_tmp0}, *t1[1] = {pay_tmp1};
  return radix_sort_tuples<KeyT, 1>(ksrc, ps_, n, total_bits, keys_tmp0, keys_tmp1, t0, t1, keys_final, pd, scratch, s);
}

}  // namespace ps
