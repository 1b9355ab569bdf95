// Stable LSD radix sort of (key, 32-bit payload) pairs for sm_100a — the workhorse behind
// sort-on-construct / coalesce (storage.py:158-171, 454-486: the reference calls
// Tensor.argsort on 64-bit keys, utils.py:21-23) and csr2csc (storage.py:425-434).
//
// Design (one-sweep: every pass reads and writes each tuple exactly once):
//   1. hist kernel   - one read of the keys builds the digit histograms of ALL passes in shared memory; a scan
//                      turns them into bin bases (callers that know them pass them in: hybrid coalesce, inverse).
//   2. pass kernel   - one persistent 1024-thread CTA per SM, 12-16 K-tuple tiles from an atomic ticket: keys in
//                      registers, ranked by ONE shared-memory atomicAdd per item on warp-private 16-bit counters
//                      (ATOMS serves lanes in order on sm_100a: stable), tile digit totals published early and
//                      chained to earlier tiles by a warp-cooperative decoupled look-back (one 64-bit status word
//                      per tile x digit, epoch-tagged so a single memset serves every pass), tuples staged in
//                      shared memory in tile-sorted order and written out as contiguous per-digit runs.
// The first pass can read its keys straight from the int64 (row, col) arrays (packing row << cbits | col on the
// fly), the last pass can widen keys / payload to int64 or keep only the low key bits, so no separate pack /
// unpack passes touch HBM.  Up to three u32 payload words travel with a key.
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
enum { KSRC_KEYS = 0, KSRC_PACK_COO = 1, KSRC_I64 = 2, KSRC_PACK_COO32 = 3, KSRC_U8 = 4 /* one byte per key */ };
enum { PSRC_U32 = 0, PSRC_IOTA = 1, PSRC_U16 = 2, PSRC_U8 = 3, PSRC_I64 = 4, PSRC_NONE = 5 /* word unused */ };
enum { PDST_U32 = 0, PDST_I64 = 1, PDST_U16 = 2, PDST_U8 = 3, PDST_NONE = 4,
       // p = device array of one base pointer PER DIGIT (single-pass partitions whose bins live in different
       // buffers, e.g. the receive buffers of peer GPUs): the word goes to p[digit][bin_base[digit] + rank in bin]
       PDST_I64_BY_DIGIT = 5, PDST_U32_BY_DIGIT = 6 };
enum { KDST_NONE = 0, KDST_KEYS = 1, KDST_UNPACK_COO = 2, KDST_I64 = 3, KDST_LOW32 = 4 /* low cbits bits as u32 */ };

struct KeySrc {
  const void *a;  // keys (KSRC_KEYS), row (KSRC_PACK_COO: int64, KSRC_PACK_COO32: uint32), the int64 array (KSRC_I64)
  const void *b;  // col (KSRC_PACK_COO / KSRC_PACK_COO32)
  int mode;
  int cbits;
  int64_t off = 0;  // KSRC_PACK_COO32: subtracted from row (tuples received from peers carry GLOBAL row ids)
  int lo32 = 0;     // KSRC_PACK_COO: every row and col fits 32 bits, so a pass may read only their low words
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
  if (s.mode == KSRC_U8) return (KeyT) reinterpret_cast<const uint8_t *>(s.a)[i];
  if (s.mode == KSRC_PACK_COO32) {
    const uint64_t r = (uint64_t)ld_stream(reinterpret_cast<const uint32_t *>(s.a) + i) - (uint64_t)s.off;
    const uint64_t c = (uint64_t)ld_stream(reinterpret_cast<const uint32_t *>(s.b) + i);
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
__device__ __forceinline__ void store_pay(const PayDst &d, int64_t i, uint32_t v, unsigned digit) {
  if (d.mode == PDST_I64_BY_DIGIT) reinterpret_cast<int64_t *const *>(d.p)[digit][i] = (int64_t)v;
  else if (d.mode == PDST_U32_BY_DIGIT) reinterpret_cast<uint32_t *const *>(d.p)[digit][i] = v;
  else store_pay(d, i, v);
}
template <typename KeyT>
__device__ __forceinline__ void store_key(const KeyDst &d, int64_t i, KeyT k) {
  if (d.mode == KDST_KEYS) reinterpret_cast<KeyT *>(d.a)[i] = k;
  else if (d.mode == KDST_I64) reinterpret_cast<int64_t *>(d.a)[i] = (int64_t)k;
  else if (d.mode == KDST_LOW32)
    reinterpret_cast<uint32_t *>(d.a)[i] = (uint32_t)((uint64_t)k & ((uint64_t(1) << d.cbits) - 1));
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
  auto count = [&](KeyT k) {
#pragma unroll
    for (int p = 0; p < RS_MAX_PASSES; ++p) {
      if (p < plan.passes) {
        const unsigned d = (unsigned)(k >> plan.begin[p]) & ((1u << plan.bits[p]) - 1u);
        atomicAdd(&sh[p * RS_MAX_RADIX + d], 1u);
      }
    }
  };
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {  // four keys in flight per thread: the loop is latency-bound otherwise
    KeyT k[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) k[q] = load_key<KeyT>(ksrc, i + q * stride);
#pragma unroll
    for (int q = 0; q < 4; ++q) count(k[q]);
  }
  for (; i < n; i += stride) count(load_key<KeyT>(ksrc, i));
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

// Shared memory of one CTA.  The register file holds the tile in ARRIVAL order (ITEMS keys per thread), shared
// memory holds it in SORTED order; u32 payload words of the NEXT tile land in pay_in (cp.async) meanwhile.
template <typename KeyT, int NPAY, int RBITS, int THREADS, int ITEMS, bool PREFETCH>
struct RsSmem {
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr int WARPS = THREADS / 32;
  static constexpr int RADIX = 1 << RBITS;
  static constexpr int ROW = RADIX / 2 + 1;  // counter row in 32-bit words: odd stride, so a column read is conflict-free
  KeyT keys[TILE];
  uint32_t pay[NPAY > 0 ? NPAY : 1][TILE];
  // landing zone of the bulk asynchronous copies (TMA, cp.async.bulk): 16-byte aligned, completion on `mbar`
  alignas(16) uint32_t pay_in[(PREFETCH && NPAY > 0) ? NPAY : 1][(PREFETCH && NPAY > 0) ? TILE : 4];
  alignas(8) uint64_t mbar;
  uint32_t cnt[WARPS][ROW];  // two 16-bit counters per word: items of (warp, digit); later the group's start in the tile
  uint32_t gbase[RADIX];     // global position of the digit's run minus its offset inside the tile
  uint16_t part_tot[THREADS / RADIX > 0 ? THREADS / RADIX : 1][RADIX];  // digit totals of each group of warps
  uint32_t dtot[RADIX];      // items of the digit in this tile
  uint32_t dstart[RADIX];    // offset of the digit inside the sorted tile
  uint32_t scan_tmp[32];
  long long tile_next;
};

// ---- bulk asynchronous copy global -> shared (the 1-D form of TMA; SASS: UBLKCP) completing on an mbarrier.  One
// elected thread arms the barrier with the byte count and issues the copies; the data path involves no thread, no
// register and no LSU slot.  Source, destination and size must be multiples of 16 bytes.
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void *smem_dst, const void *gsrc, unsigned bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(__cvta_generic_to_global(gsrc)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}

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
  } else if (s.mode == KSRC_U8) {
    const uint8_t *p = reinterpret_cast<const uint8_t *>(s.a) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      key[i] = j < tile_n ? (KeyT)__ldg(p + j) : ~KeyT(0);
    }
  } else if (s.mode == KSRC_PACK_COO32) {
    const uint32_t *r = reinterpret_cast<const uint32_t *>(s.a) + tile_base;
    const uint32_t *c = reinterpret_cast<const uint32_t *>(s.b) + tile_base;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = warp_base + i * 32 + lane;
      key[i] = j < tile_n ? (KeyT)((uint64_t)ld_stream(r + j) - (uint64_t)s.off) : ~KeyT(0);
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

// 64-bit keys packed from two 32-bit words (KSRC_PACK_COO32, or the low words of KSRC_PACK_COO's int64 arrays): the
// REQUEST only loads - row word into the key's high half, col word into its low half, no arithmetic on the loaded
// values, so nothing waits for them at the request site - and the key is put together when the tile is ranked, one
// whole scatter phase later (rs_finish_raw_keys).  With the shift / or at the request site 26 % of a first pass's
// warp samples sat there waiting for the column loads, 44 % when the source was 32-bit.
template <int ITEMS>
__device__ __forceinline__ void load_tile_raw(const KeySrc &s, int64_t tile_base, int warp_base, unsigned lane,
                                              int tile_n, uint64_t (&key)[ITEMS]) {
  const int w = s.mode == KSRC_PACK_COO32 ? 1 : 2;  // 32-bit words per element (little endian: low word first)
  const uint32_t *r = reinterpret_cast<const uint32_t *>(s.a) + tile_base * w;
  const uint32_t *c = reinterpret_cast<const uint32_t *>(s.b) + tile_base * w;
  uint32_t hi[ITEMS], lo[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int j = warp_base + i * 32 + lane;
    hi[i] = j < tile_n ? ld_stream(r + (size_t)j * w) : 0xffffffffu;
  }
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int j = warp_base + i * 32 + lane;
    lo[i] = j < tile_n ? ld_stream(c + (size_t)j * w) : 0xffffffffu;
  }
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(key[i]) : "r"(lo[i]), "r"(hi[i]));  // register pairing
}
template <int ITEMS>
__device__ __forceinline__ void rs_finish_raw_keys(const KeySrc &s, int warp_base, unsigned lane, int tile_n,
                                                   uint64_t (&key)[ITEMS]) {
  const uint32_t off = (uint32_t)s.off;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int j = warp_base + i * 32 + lane;
    const uint64_t r = (uint64_t)((uint32_t)(key[i] >> 32) - off), c = key[i] & 0xffffffffull;
    key[i] = j < tile_n ? ((r << s.cbits) | c) : ~uint64_t(0);
  }
}
template <int ITEMS>
__device__ __forceinline__ void load_tile_raw(const KeySrc &, int64_t, int, unsigned, int, uint32_t (&)[ITEMS]) {}
template <int ITEMS>
__device__ __forceinline__ void rs_finish_raw_keys(const KeySrc &, int, unsigned, int, uint32_t (&)[ITEMS]) {}

// payload words whose source is not a plain u32 array (those are prefetched with cp.async)
template <int ITEMS>
__device__ __forceinline__ void load_tile_pay_sync(const PaySrc &s, int64_t tile_base, int warp_base, unsigned lane,
                                                   int tile_n, uint32_t (&pay)[ITEMS]) {
  if (s.mode == PSRC_IOTA) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) pay[i] = (uint32_t)(tile_base + warp_base + i * 32 + lane);
  } else if (s.mode == PSRC_NONE) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) pay[i] = 0u;
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

// ---- decoupled look-back, warp-cooperative: the 32 / DPW lanes {k, k + DPW, ...} of a warp serve digit k of the
// warp together.  Lane (k, r) loads the word of tile t - r (one register pair per lane instead of a window per
// digit thread); the words are exchanged by shuffles and every lane of the group tracks the same walk state.
// The first words are loaded EARLY (phase 3) so their round trip overlaps the staging phase; later rounds read
// two words per lane.  Tiles before 0 read as "prefix 0".
__device__ __forceinline__ uint64_t rs_lb_word(const uint64_t *stw, int radix, int64_t t, int epoch) {
  return t >= 0 ? TileState::load(stw + t * radix) : RsState::pack(epoch, RsState::PREFIX, 0);
}
constexpr int RS_LB_WORDS = 8;  // status words one lane requests per round trip (GS * RS_LB_WORDS tiles per round)
template <int DPW>
__device__ __forceinline__ uint32_t rs_lb_finish(const uint64_t *stw, int radix, int64_t tile, int epoch, uint64_t w0,
                                                 unsigned lane) {
  constexpr int GS = 32 / DPW;
  const unsigned lk = lane % DPW, lr = lane / DPW;
  uint64_t excl = 0;
  int64_t t = tile - 1;  // next tile to consume
  bool done = false, stuck = false;
  auto consume = [&](uint64_t w) {
    if (done || stuck) return;
    if (RsState::epoch(w) != (uint64_t)epoch) { stuck = true; return; }  // not published in this pass yet
    excl += RsState::value(w);
    --t;
    if (RsState::status(w) == RsState::PREFIX) done = true;
  };
#pragma unroll
  for (int r = 0; r < GS; ++r) consume(__shfl_sync(0xffffffffu, w0, lk + r * DPW));
  // With one big tile per SM the CTAs run almost in lock step: when a tile looks back, its ~SM-count predecessors
  // have published aggregates but no prefix yet, so the walk is long and its cost is round trips to L2.  Every lane
  // therefore keeps RS_LB_WORDS loads in flight per round (round 2 timeline: the look-back was 23 % of a tile's
  // time with two).
  while (!__all_sync(0xffffffffu, done)) {
    uint64_t u[RS_LB_WORDS];
#pragma unroll
    for (int w = 0; w < RS_LB_WORDS; ++w) u[w] = done ? 0 : rs_lb_word(stw, radix, t - (int64_t)lr - w * GS, epoch);
    stuck = false;
#pragma unroll
    for (int w = 0; w < RS_LB_WORDS; ++w) {
#pragma unroll
      for (int r = 0; r < GS; ++r) consume(__shfl_sync(0xffffffffu, u[w], lk + r * DPW));
    }
  }
  return (uint32_t)excl;
}

// PS_RS_TIMELINE (debug builds, make TIMELINE=1): thread 0 of every CTA accumulates the SM cycles it spends
// between phase markers; ps_debug_timeline() reads the totals.
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

// One tile = THREADS * ITEMS tuples, ONE persistent CTA per SM.  Why so large: the scatter of a pass writes
// RADIX interleaved streams in runs of TILE / RADIX elements, and the HBM/L2 path of a B200 only approaches
// streaming speed when those runs are >= ~48 elements (tools/microbench/scatter_bw.cu: runs of 12 -> 2.0 TB/s,
// 24 -> 3.0, 48 -> 4.1, 96 -> 4.5, streaming copy 5.5).  The register file holds the tile in arrival order, shared
// memory holds it in sorted order.  Tiles come from an atomic ticket; the loop is software-pipelined: the
// NEXT tile's key loads (and cp.async payload) are issued as soon as the current tile is staged, so their
// latency overlaps the look-back and the scatter.  Phases per tile:
//   2 rank: ONE shared-memory atomicAdd per item on warp-private 16-bit digit counters returns the item's rank
//     inside its (warp, digit) group
//   3 every warp scans RADIX / WARPS digit columns across the warps (lane = warp index); lanes < RADIX / WARPS
//     own one digit each: they PUBLISH the tile's aggregate early and request the first look-back words; a block
//     scan of the digit totals turns the counters into group starts inside the sorted tile
//   4 position of every item; stage tuples in tile-sorted order; request the next tile
//   5 finish the decoupled look-back
//   6 coalesced scatter of contiguous per-digit runs
template <typename KeyT, int NPAY, int RBITS, int THREADS, int ITEMS, bool PREFETCH, int MINB = 1024 / THREADS>
__global__ void __launch_bounds__(THREADS, MINB)
rs_pass_kernel(KeySrc ksrc, PaySrcN<NPAY> psrc, KeyDst kdst, PayDstN<NPAY> pdst, int64_t n, int64_t tiles,
               int begin_bit, int num_bits, int epoch, const unsigned long long *__restrict__ bin_base,
               uint64_t *__restrict__ tile_state /*[tiles][RADIX]*/, unsigned long long *ticket) {
  using Smem = RsSmem<KeyT, NPAY, RBITS, THREADS, ITEMS, PREFETCH>;
  constexpr int TILE = Smem::TILE, WARPS = Smem::WARPS, RADIX = Smem::RADIX;
  constexpr int DPW = RADIX / WARPS;  // digit columns scanned by one warp; lane k < DPW owns digit warp*DPW + k
  static_assert(RADIX % WARPS == 0 && DPW <= 32, "one warp scans RADIX / WARPS digit columns");
  static_assert(32 * ITEMS < 65536 && TILE < 65536, "counters and group starts are kept in 16 bits");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem &sm = *reinterpret_cast<Smem *>(smem_raw);

  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned mask = (1u << num_bits) - 1u;
  const int warp_base = warp * (32 * ITEMS);
  uint32_t *my_cnt = sm.cnt[warp];
  const bool digit_thread = lane < (unsigned)DPW;  // lane k < DPW owns digit warp*DPW + k; lanes k + r*DPW assist
  const int my_digit = warp * DPW + lane % DPW;
  uint64_t *stw = tile_state + my_digit;  // word of (tile t, my digit) at stw[t * RADIX]
  const uint32_t bin0 = (uint32_t)bin_base[my_digit];

  // request one tile: keys into registers, u32 payload words towards pay_in
  KeyT key[ITEMS];
  // u32 payload words of a FULL tile whose source is 16-byte aligned arrive by bulk asynchronous copy (one thread
  // issues, the mbarrier counts the bytes); other tiles / sources are read by the threads in the staging phase
  bool pay_bulk[NPAY > 0 ? NPAY : 1];
#pragma unroll
  for (int p = 0; p < (NPAY > 0 ? NPAY : 1); ++p)
    pay_bulk[p] = PREFETCH && NPAY > 0 && psrc.s[p].mode == PSRC_U32 &&
                  (reinterpret_cast<uintptr_t>(psrc.s[p].p) & 15u) == 0;
  const bool raw_keys = sizeof(KeyT) == 8 && (ksrc.mode == KSRC_PACK_COO32 || (ksrc.mode == KSRC_PACK_COO && ksrc.lo32));
  auto bulk_tile = [&](int64_t t) { return PREFETCH && NPAY > 0 && t < tiles && (n - t * TILE) >= TILE; };
  auto request_keys = [&](int64_t t) {
    const int64_t tb = t * TILE;
    const int tn = (int)((n - tb) < TILE ? (n - tb) : TILE);
    if (raw_keys) load_tile_raw<ITEMS>(ksrc, tb, warp_base, lane, tn, key);
    else load_tile_keys<KeyT, ITEMS>(ksrc, tb, warp_base, lane, tn, key);
  };
  auto request_pay = [&](int64_t t) {  // thread 0 only, after every read of pay_in of the previous tile
    if (!bulk_tile(t)) return;
    unsigned bytes = 0;
#pragma unroll
    for (int p = 0; p < NPAY; ++p) bytes += pay_bulk[p] ? (unsigned)(TILE * sizeof(uint32_t)) : 0u;
    if (bytes == 0) return;
    mbar_expect_tx(&sm.mbar, bytes);
#pragma unroll
    for (int p = 0; p < NPAY; ++p)
      if (pay_bulk[p])
        bulk_copy_g2s(&sm.pay_in[p][0], reinterpret_cast<const uint32_t *>(psrc.s[p].p) + t * TILE,
                      (unsigned)(TILE * sizeof(uint32_t)), &sm.mbar);
  };
  unsigned mbar_phase = 0;
  bool any_bulk = false;
#pragma unroll
  for (int p = 0; p < (NPAY > 0 ? NPAY : 1); ++p) any_bulk = any_bulk || pay_bulk[p];

  long long ticket_ahead = 0;  // thread 0: the ticket drawn one tile ahead
  if (tid == 0) {
    sm.tile_next = (long long)atomicAdd(ticket, 1ull);
    if (PREFETCH && NPAY > 0) mbar_init(&sm.mbar, 1);
  }
  __syncthreads();
  int64_t tile = sm.tile_next;
  if (tile < tiles) request_keys(tile);
  if (tid == 0) {
    if (any_bulk) request_pay(tile);
    ticket_ahead = (long long)atomicAdd(ticket, 1ull);
  }
  PS_TL_DECL;

  while (tile < tiles) {
    const int64_t tile_base = tile * TILE;
    const int tile_n = (int)((n - tile_base) < TILE ? (n - tile_base) : TILE);
    PS_TL(0);
#pragma unroll
    for (int q = lane; q < Smem::ROW; q += 32) my_cnt[q] = 0;
    if (raw_keys) rs_finish_raw_keys<ITEMS>(ksrc, warp_base, lane, tile_n, key);  // the words requested a tile ago
    __syncwarp();

    // ---- 2: rank.  Stability needs the lanes of ONE atomic instruction that hit the same counter to be served
    // in ascending lane order: that is how ATOMS behaves on sm_100a (tools/microbench/atom_order.cu: 6e10
    // lane-instances under every collision / bank-conflict pattern, none out of order; rank_bench.cu: ~2 SM
    // cycles per warp instruction against 40 for MATCH.ANY and 17 for eight ballots).  Padding keys of a partial
    // tile are all-ones: the largest digit, and ranked last inside it.
    // 1024-thread CTAs are capped at 64 registers: their ranks / slots (< TILE <= 65535) are kept two to a register
    constexpr bool PACK = THREADS >= 1024;
    constexpr int NRANK = PACK ? (ITEMS + 1) / 2 : ITEMS;
    unsigned rank[NRANK];
    auto slot_of = [&](int i) -> unsigned {
      if constexpr (PACK) return (rank[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
      else return rank[i];
    };
    if constexpr (PACK) {
#pragma unroll
      for (int i = 0; i < NRANK; ++i) rank[i] = 0;
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const unsigned d = (unsigned)(key[i] >> begin_bit) & mask;
      const unsigned sh = (d & 1u) * 16u;
      const unsigned r = (atomicAdd(&my_cnt[d >> 1], 1u << sh) >> sh) & 0xffffu;
      if constexpr (PACK) rank[i >> 1] |= r << ((i & 1) * 16);
      else rank[i] = r;
    }
    PS_TL(1);
    __syncthreads();
    PS_TL(2);

    // ---- 3: prefix over the warps, per digit.  Thread (digit, part) walks the counters of its part of the warps
    // row-wise (conflict-free, no shuffles: round 1 scanned the digit columns with 5-step shuffle scans, 16 % of a
    // tile's time), the parts are combined, a block scan over the digit totals gives every digit its offset inside the
    // sorted tile, and that offset is folded back into the counters, which thereby become the start of every
    // (warp, digit) group in the tile.
    constexpr int PARTS = THREADS / RADIX > 0 ? THREADS / RADIX : 1;
    constexpr int WPP = WARPS / PARTS;            // warps per part
    constexpr int DW = RADIX / 32;                // warps whose lanes cover the digits once
    static_assert(THREADS % RADIX == 0 && WARPS % PARTS == 0 && DW <= 32, "digit / part ownership");
    const unsigned dg = tid % RADIX, part = tid / RADIX;
    uint16_t *c16 = reinterpret_cast<uint16_t *>(&sm.cnt[part * WPP][0]) + dg;
    {
      uint32_t run = 0;
#pragma unroll
      for (int w = 0; w < WPP; ++w) {
        const uint32_t c = c16[w * (Smem::ROW * 2)];
        c16[w * (Smem::ROW * 2)] = (uint16_t)run;
        run += c;
      }
      sm.part_tot[part][dg] = (uint16_t)run;
    }
    __syncthreads();
    {
      uint32_t dt = 0, before = 0;
#pragma unroll
      for (int h = 0; h < PARTS; ++h) {
        const uint32_t t = sm.part_tot[h][dg];
        before += h < (int)part ? t : 0u;
        dt += t;
      }
      uint32_t inc = dt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d) inc += o;
      }
      if (lane == 31) sm.scan_tmp[warp] = inc;
      __syncthreads();
      uint32_t wb = 0;
#pragma unroll
      for (int w = 0; w < DW; ++w) wb += (w < (int)(warp % DW)) ? sm.scan_tmp[part * DW + w] : 0u;
      const uint32_t ds = wb + inc - dt;  // offset of digit dg inside the sorted tile
      if (part == 0) {
        sm.dtot[dg] = dt;
        sm.dstart[dg] = ds;
      }
      const uint32_t add = ds + before;
#pragma unroll
      for (int w = 0; w < WPP; ++w) c16[w * (Smem::ROW * 2)] = (uint16_t)(c16[w * (Smem::ROW * 2)] + add);
    }
    if (tid == 0) sm.tile_next = ticket_ahead;
    PS_TL(3);
    __syncthreads();
    PS_TL(4);
    // digit threads publish the tile's aggregate and request the first look-back words
    const uint32_t tot = sm.dtot[my_digit];
    const uint32_t dexcl = sm.dstart[my_digit];
    if (digit_thread)
      TileState::store(stw + tile * RADIX, RsState::pack(epoch, tile == 0 ? RsState::PREFIX : RsState::AGG, tot));
    const uint64_t lbw = rs_lb_word(stw, RADIX, tile - 1 - (int64_t)(lane / DPW), epoch);

    // ---- 4: position inside the sorted tile = group start + rank inside the group; stage
    {
      const uint16_t *my_start = reinterpret_cast<const uint16_t *>(my_cnt);
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const unsigned st = (unsigned)my_start[(unsigned)(key[i] >> begin_bit) & mask];
        if constexpr (PACK) rank[i >> 1] += st << ((i & 1) * 16);  // start + rank < TILE: no carry into the neighbour
        else rank[i] += st;
      }
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) sm.keys[slot_of(i)] = key[i];
    PS_TL(10);
    if (NPAY > 0) {
      const bool landed = any_bulk && bulk_tile(tile);  // uniform: this tile's u32 words came by bulk copy
      if (landed) {
        mbar_wait(&sm.mbar, mbar_phase);
        mbar_phase ^= 1u;
      }
#pragma unroll
      for (int p = 0; p < NPAY; ++p) {
        uint32_t pv[ITEMS];
        if (landed && pay_bulk[p]) {
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) pv[i] = sm.pay_in[p][warp_base + i * 32 + lane];
        } else if (psrc.s[p].mode == PSRC_U32) {
          const uint32_t *src = reinterpret_cast<const uint32_t *>(psrc.s[p].p) + tile_base;
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) {
            const int j = warp_base + i * 32 + lane;
            pv[i] = j < tile_n ? ld_stream(src + j) : 0u;
          }
        } else {
          load_tile_pay_sync<ITEMS>(psrc.s[p], tile_base, warp_base, lane, tile_n, pv);
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) sm.pay[p][slot_of(i)] = pv[i];
      }
    }
    PS_TL(5);
    // ---- 5: look-back (the first words were requested in phase 3 and have usually arrived by now)
    {
      const uint32_t excl = rs_lb_finish<DPW>(stw, RADIX, tile, epoch, lbw, lane);  // tile 0: "prefix 0" at once
      if (digit_thread) {
        if (tile > 0) TileState::store(stw + tile * RADIX, RsState::pack(epoch, RsState::PREFIX, (uint64_t)excl + tot));
        sm.gbase[my_digit] = bin0 + excl - dexcl;
      }
    }
    PS_TL(9);
    // the registers are free again: request the next tile now, its latency hides behind look-back and scatter
    const int64_t next = sm.tile_next;
    if (next < tiles) request_keys(next);
    if (tid == 0) ticket_ahead = (long long)atomicAdd(ticket, 1ull);

    PS_TL(6);
    __syncthreads();
    PS_TL(7);
    if (tid == 0 && any_bulk) request_pay(next);  // every thread is past its reads of pay_in: the buffer is free

    // ---- 6: coalesced scatter: consecutive staged slots of one digit go to consecutive addresses
    const bool full = tile_n == TILE;
    bool plain = kdst.mode == KDST_KEYS;
#pragma unroll
    for (int p = 0; p < NPAY; ++p) plain = plain && (pdst.d[p].mode == PDST_U32 || pdst.d[p].mode == PDST_NONE);
    if (plain) {
      KeyT *ko = reinterpret_cast<KeyT *>(kdst.a);
#pragma unroll 4
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        if (full || j < tile_n) {
          const KeyT kk = sm.keys[j];
          const unsigned d = (unsigned)(kk >> begin_bit) & mask;
          const size_t dst = (size_t)(uint32_t)(sm.gbase[d] + (uint32_t)j);
          ko[dst] = kk;
#pragma unroll
          for (int p = 0; p < NPAY; ++p)
            if (pdst.d[p].mode == PDST_U32) reinterpret_cast<uint32_t *>(pdst.d[p].p)[dst] = sm.pay[p][j];
        }
      }
    } else {
#pragma unroll 2
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        if (full || j < tile_n) {
          const KeyT kk = sm.keys[j];
          const unsigned d = (unsigned)(kk >> begin_bit) & mask;
          const int64_t dst = (int64_t)(uint32_t)(sm.gbase[d] + (uint32_t)j);
          store_key<KeyT>(kdst, dst, kk);
#pragma unroll
          for (int p = 0; p < NPAY; ++p) store_pay(pdst.d[p], dst, sm.pay[p][j], d);
        }
      }
    }
    PS_TL(8);
    tile = next;
  }
  PS_TL_FLUSH;
}

// ----------------------------------------------------------------------------- host driver
// Geometry: 1024 threads, one CTA per SM; ITEMS per (key type, payload words) chosen so the sorted tile, the
// payload prefetch buffer and the counters fit the 227 KB of shared memory.  PS_RS_VARIANT (bench knob, read
// once): 1 = 9-bit digits.
int rs_variant();

template <typename KeyT, int NPAY> struct RsGeom;
template <> struct RsGeom<uint64_t, 1> { static constexpr int ITEMS = 12; static constexpr bool PREFETCH = true; };
template <> struct RsGeom<uint32_t, 1> { static constexpr int ITEMS = 16; static constexpr bool PREFETCH = true; };
template <> struct RsGeom<uint32_t, 3> { static constexpr int ITEMS = 12; static constexpr bool PREFETCH = false; };
template <> struct RsGeom<uint64_t, 3> { static constexpr int ITEMS = 8; static constexpr bool PREFETCH = false; };

struct RsGeomRt {
  int rbits, tile;
};
// Launch shape (PS_RS_VARIANT): 0 (default) = 512 threads holding 2 x ITEMS tuples each, 1 = 1024 threads holding
// ITEMS each; one CTA per SM and the same tile either way.  With 1024 threads the register file caps a thread at 64
// registers and the key / rank / payload arrays spill (56-72 B per thread); 512 threads get 128 registers and no
// spills: 7 % faster passes (0.606 -> 0.561 ms on the 61.9 M-tuple u64 passes).  9-bit digits and two
// 512-thread CTAs per SM on half tiles were measured too and brought nothing.
template <typename KeyT, int NPAY>
RsGeomRt rs_geometry() {
  return RsGeomRt{8, 1024 * RsGeom<KeyT, NPAY>::ITEMS};
}

// Digit width of a sort that computes its own histograms: 9 bits where that saves a whole pass over HBM (64-bit keys
// of 41-45 or 49-54 bits: 5 or 6 passes instead of 6 or 7; a 9-bit pass costs what an 8-bit one does), else 8.
// Callers that bring their own bin bases (hybrid coalesce, inverse permutation) planned with 8 and stay there.
template <typename KeyT, int NPAY>
inline int rs_auto_rbits(int total_bits) {
  if (sizeof(KeyT) == 8 && NPAY == 1 && (total_bits + 8) / 9 < (total_bits + 7) / 8) return 9;
  return 8;
}

template <typename KeyT, int NPAY, int RBITS, int THREADS, int IMUL = 1>
int rs_launch_pass(KeySrc ks, PaySrcN<NPAY> ps_, KeyDst ko, PayDstN<NPAY> pd, int64_t n, int begin, int bits,
                   int epoch, const unsigned long long *bin_base, uint64_t *state, unsigned long long *ticket,
                   cudaStream_t s) {
  using G = RsGeom<KeyT, NPAY>;
  using Smem = RsSmem<KeyT, NPAY, RBITS, THREADS, G::ITEMS * IMUL, G::PREFETCH>;
  auto kern = rs_pass_kernel<KeyT, NPAY, RBITS, THREADS, G::ITEMS * IMUL, G::PREFETCH, 1024 / (THREADS * IMUL)>;
  const size_t smem = sizeof(Smem);
  PS_FUNC_SMEM(kern, smem);
  const int64_t tiles = ceil_div(n, Smem::TILE);
  int64_t grid = (int64_t)sm_count() * (1024 / (THREADS * IMUL));
  if (grid > tiles) grid = tiles;
  {
    ProfScope prof(PROF_RS_PASS, s);
    kern<<<(unsigned)grid, THREADS, smem, s>>>(ks, ps_, ko, pd, n, tiles, begin, bits, epoch, bin_base, state, ticket);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

template <typename KeyT, int NPAY>
int rs_dispatch_pass(int rbits, KeySrc ks, PaySrcN<NPAY> ps_, KeyDst ko, PayDstN<NPAY> pd, int64_t n, int begin,
                     int bits, int epoch, const unsigned long long *bin_base, uint64_t *state,
                     unsigned long long *ticket, cudaStream_t s) {
  if constexpr (sizeof(KeyT) == 8 && NPAY == 1) {
    if (rbits == 9)
      return rs_launch_pass<KeyT, NPAY, 9, 512, 2>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
  }
  if (rs_variant() == 1)
    return rs_launch_pass<KeyT, NPAY, 8, 1024, 1>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
  return rs_launch_pass<KeyT, NPAY, 8, 512, 2>(ks, ps_, ko, pd, n, begin, bits, epoch, bin_base, state, ticket, s);
}

// Sorts n (key, NPAY payload words) tuples on key bits [first_bit, first_bit + total_bits).  First-pass sources are
// ksrc/psrc.  The last pass writes keys to keys_final (KDST_NONE: keys dropped) and payloads to
// pay_final.  keys_tmp*/pay_tmp* are scratch buffers of n elements each (pay_tmp*[p] for word p;
// unused when one pass suffices).
template <typename KeyT, int NPAY>
int radix_sort_tuples(KeySrc ksrc, PaySrcN<NPAY> psrc, int64_t n, int total_bits, KeyT *keys_tmp0, KeyT *keys_tmp1,
                      uint32_t *const *pay_tmp0, uint32_t *const *pay_tmp1, KeyDst keys_final,
                      PayDstN<NPAY> pay_final, Scratch &scratch, cudaStream_t s, int first_bit = 0,
                      const unsigned long long *scanned_hist = nullptr /* [passes][512] bin bases, if known */) {
  if (n <= 0) return PS_OK;
  const RsGeomRt geom = rs_geometry<KeyT, NPAY>();
  const int rbits = scanned_hist ? geom.rbits : rs_auto_rbits<KeyT, NPAY>(total_bits);
  SortPlan plan = make_sort_plan(total_bits, rbits);
  for (int i = 0; i < plan.passes; ++i) plan.begin[i] += first_bit;  // sort on key bits [first_bit, first_bit + total_bits)
  if (plan.passes > RS_MAX_PASSES) { set_error("radix sort: %d key bits unsupported", total_bits); return PS_ERR_ARG; }
  const int radix = 1 << rbits;
  const int64_t tiles = ceil_div(n, geom.tile);

  // one allocation: [passes*512 hist | passes tickets (padded to 32) | tiles*radix state words]
  const size_t hist_words = (size_t)plan.passes * RS_MAX_RADIX;
  const size_t total_words = hist_words + 32 + (size_t)tiles * radix;
  PS_SCRATCH(words, scratch, uint64_t, total_words);
  PS_CUDA(cudaMemsetAsync(words, 0, total_words * sizeof(uint64_t), s));
  unsigned long long *ghist = reinterpret_cast<unsigned long long *>(words);
  unsigned long long *tickets = ghist + hist_words;
  uint64_t *state = words + hist_words + 32;

  if (scanned_hist) {
    ghist = const_cast<unsigned long long *>(scanned_hist);
  } else {
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
  }

  KeyT *kt[2] = {keys_tmp0, keys_tmp1};
  for (int p = 0; p < plan.passes; ++p) {
    KeySrc ks = ksrc;
    PaySrcN<NPAY> ps_ = psrc;
    if (p > 0) {
      ks = KeySrc{kt[(p - 1) & 1], nullptr, KSRC_KEYS, 0};
      for (int w = 0; w < NPAY; ++w)
        ps_.s[w] = psrc.s[w].mode == PSRC_NONE ? psrc.s[w] : PaySrc{((p - 1) & 1) ? pay_tmp1[w] : pay_tmp0[w], PSRC_U32};
    }
    KeyDst ko{kt[p & 1], nullptr, KDST_KEYS, 0};
    PayDstN<NPAY> pd;
    for (int w = 0; w < NPAY; ++w)
      pd.d[w] = psrc.s[w].mode == PSRC_NONE ? PayDst{nullptr, PDST_NONE} : PayDst{(p & 1) ? pay_tmp1[w] : pay_tmp0[w], PDST_U32};
    if (p == plan.passes - 1) {
      ko = keys_final;
      pd = pay_final;
    }
    PS_TRY((rs_dispatch_pass<KeyT, NPAY>(rbits, ks, ps_, ko, pd, n, plan.begin[p], plan.bits[p], p + 1,
                                         ghist + (size_t)p * RS_MAX_RADIX, state, tickets + p, s)));
  }
  return PS_OK;
}

// number of passes radix_sort_* will run for this key type (callers pick the ping-pong buffer with it)
template <typename KeyT>
int rs_num_passes(int total_bits) { return make_sort_plan(total_bits, rs_auto_rbits<KeyT, 1>(total_bits)).passes; }

// the common single-payload form
template <typename KeyT>
int radix_sort_pairs(KeySrc ksrc, PaySrc psrc, int64_t n, int total_bits, KeyT *keys_tmp0, KeyT *keys_tmp1,
                     uint32_t *pay_tmp0, uint32_t *pay_tmp1, KeyDst keys_final, PayDst pay_final, Scratch &scratch,
                     cudaStream_t s, int first_bit = 0, const unsigned long long *scanned_hist = nullptr) {
  PaySrcN<1> ps_;
  ps_.s[0] = psrc;
  PayDstN<1> pd;
  pd.d[0] = pay_final;
  uint32_t *t0[1] = {pay_tmp0}, *t1[1] = {pay_tmp1};
  return radix_sort_tuples<KeyT, 1>(ksrc, ps_, n, total_bits, keys_tmp0, keys_tmp1, t0, t1, keys_final, pd, scratch, s,
                                    first_bit, scanned_hist);
}

}  // namespace ps
