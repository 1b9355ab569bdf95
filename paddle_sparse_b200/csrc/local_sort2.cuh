// Hybrid coalesce, second-generation bucket kernel (round 2).
//
// local_sort.cuh sorts a bucket with four stable LSD passes inside shared memory: ~120 warp instructions per 32
// entries and pass (ncu, round 1: 923 M instructions, 1.44 ms, issue-bound).  A bucket's low keys are ~3,200 random
// 29-bit values, so ONE counting pass on their top 12 bits already separates them into 4,096 bins of ~0.8 entries;
// what is left is to order the few entries that share a bin, which each entry does by COMPARING itself with its
// bin mates: final slot = bin start + #{mates with a smaller (key, arrival index)}.  The arrival index breaks ties,
// so equal keys keep their input order (the stable order the oracle's sort gives, and the order duplicates are
// summed in).  One counting pass + one short compare loop replace four ranked passes; 256 threads and ~50 KB per
// CTA keep four CTAs in different phases on an SM.  Then, as before: run heads, duplicates folded in sorted order,
// the bucket's output offset chained over the buckets (aggregate published EARLY, prefix consumed LATE), int64
// row / col / value and the rowptr entries of the bucket's own rows written coalesced.
//   reference behaviour: storage.py:158-171 (sort-on-construct) + :454-486 (coalesce) + :211-222 (rowptr)
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

namespace ps {

constexpr int L2_THREADS = 512;
constexpr int L2_ITEMS = 8;
constexpr int L2_CAP = L2_THREADS * L2_ITEMS;  // 4,096 entries per bucket
constexpr int L2_WARPS = L2_THREADS / 32;
constexpr int L2_BINBITS = 12;
constexpr int L2_BINS = 1 << L2_BINBITS;
constexpr int L2_HEAVY = 8;  // a bin beyond this many entries: bin mates are ranked by position, not by owner
static_assert(L2_BINS == L2_THREADS * 8, "the bin scan gives every thread eight bins (four counter words)");

struct L2Meta {
  long long bucket;
  unsigned int start, end;
  int heads;
  // what the round needs to know about its bucket (filled by l2_fetch_meta)
  unsigned int orig;      // the top-T-bit bucket the local keys are relative to
  unsigned int buf;       // which source buffer holds the bucket (refined buckets alternate)
  int binshift;           // the 12 bin bits of the counting pass start here (applied to the 32-bit key part)
  long long rlo, rhi;     // rows whose rowptr entries this bucket writes: [rlo, rhi)
  unsigned int nmerge;    // original buckets merged into this one (descriptor buckets of light row ranges) ...
  unsigned int bnd[7];    // ... and where the 2nd .. 8th of them start: an entry's local key is extended by the
                          // number of these at or before its position, shifted by R
};
// Buckets given by descriptors (refine.cuh): sorted by start; entry nb of `start` is the total number of entries.
struct L2Desc {
  const uint32_t *start, *klo, *khi, *orig;  // orig: original bucket | buffer << 31
  const void *keys1;                         // buffer 1 (buffer 0 is the kernel's lowkeys / lowvals)
  const uint32_t *vals1;
};
template <bool DESC>
__device__ __forceinline__ L2Meta l2_fetch_meta(long long b, int nb, const unsigned int *__restrict__ bucket_off,
                                                const L2Desc &desc, int R, int cbits, int64_t M, int binshift0) {
  L2Meta m;
  m.bucket = b;
  m.start = m.end = 0u;
  m.heads = 0;
  m.orig = 0u;
  m.buf = 0u;
  m.binshift = binshift0;
  m.rlo = m.rhi = 0;
  m.nmerge = 1u;
  if (b >= nb) return m;
  const uint64_t rowmask = (uint64_t(1) << cbits) - 1;
  if (!DESC) {
    m.start = bucket_off[b];
    m.end = bucket_off[b + 1];
    m.orig = (unsigned)b;
    m.rlo = (long long)(((uint64_t)b << R) >> cbits);
    m.rhi = (long long)((((uint64_t)b + 1) << R) >> cbits);
    if (b == nb - 1 || m.rhi > M + 1) m.rhi = M + 1;
  } else {
    m.start = desc.start[b];
    m.end = desc.start[b + 1];
    const uint32_t ob = desc.orig[b], klo = desc.klo[b], khi = desc.khi[b];
    m.orig = ob & 0xffffu;
    m.nmerge = ((ob >> 16) & 7u) + 1u;
    m.buf = ob >> 31;
    for (unsigned q = 1; q < m.nmerge; ++q) m.bnd[q - 1] = bucket_off[m.orig + q];
    const uint32_t diff = klo ^ khi;  // every key of the bucket shares the bits above the highest set bit of this
    const int h = diff ? 32 - __clz(diff) : 0;  // number of low bits in which keys of the bucket may differ
    m.binshift = h > L2_BINBITS ? h - L2_BINBITS : 0;
    // rows whose first possible key lies in [this bucket's lowest key, the next listed bucket's lowest key)
    const uint64_t lo = ((uint64_t)m.orig << R) | klo;
    m.rlo = b == 0 ? 0 : (long long)((lo + rowmask) >> cbits);
    if (b + 1 < nb) {
      const uint64_t nx = ((uint64_t)(desc.orig[b + 1] & 0xffffu) << R) | desc.klo[b + 1];
      m.rhi = (long long)((nx + rowmask) >> cbits);
      if (m.rhi > M + 1) m.rhi = M + 1;
    } else {
      m.rhi = M + 1;
    }
  }
  return m;
}
template <bool HAS_VALUE>
struct L2Smem {
  uint32_t skeys[2][L2_CAP];                   // sorted low keys (their top 32 bits): [slot] = bucket being sorted /
                                               // awaiting its offset
  uint8_t slo[2][L2_CAP];                      // ... and their low LO <= 4 bits (bucket-local keys of up to 36 bits)
  uint32_t svals[2][HAS_VALUE ? L2_CAP : 4];   // their values
  uint64_t wk[L2_CAP];                         // by bin (arrival order inside a bin): key's 32-bit part << 16 | its low
                                               // bits << 12 | arrival index - ONE word, so that "(key, arrival index)
                                               // lexicographic" is one 64-bit compare in the bin-mate loop
  uint32_t bins[L2_BINS / 2 + 2];              // two 16-bit counters per word; after the scan: bin starts (+ sentinel)
  uint32_t wcnt[L2_ITEMS * L2_WARPS];
  uint32_t scan_tmp[33];
  unsigned int maxbin, nheads;
  long long excl;
  L2Meta meta[2];
  L2Meta meta_next;  // fetched mid-round by thread 0, moved into meta[] when the round ends
};

// DEDUP = false: sort only (sort-on-construct): every entry is kept, the bucket's output offset is its input offset.
// The loop is software-pipelined over the buckets a CTA draws: bucket k is sorted and its number of distinct keys
// PUBLISHED, then bucket k-1 - whose sorted form waits in the other half of the double buffer - reads its output
// offset from the chain and is written out.  The chain (a decoupled look-back over the buckets) is thereby consumed
// one whole bucket-sort after it was fed: in round 2's first version, which consumed it at once, 38 % of all warp
// samples sat at the barrier behind the look-back warp.
// KeyT = uint32_t: `lowkeys` holds the bucket-local keys (R <= 32 bits); uint64_t: the full packed keys, of which the
// low R <= 36 bits are the bucket-local key (shards of very wide matrices: 51- to 53-bit keys whose local part does
// not fit 32 bits).  The local key is handled as (top 32 bits, low LO = R - 32 bits); the low bits share a 16-bit
// word with the 12-bit arrival index, so "(key, arrival index) lexicographic" stays two integer compares.
template <typename T, int RED, bool HAS_VALUE, bool DEDUP, typename KeyT, bool DESC = false>
__global__ void __launch_bounds__(L2_THREADS, 2)
ls2_sort_dedup_kernel(const KeyT *__restrict__ lowkeys, const uint32_t *__restrict__ lowvals,
                      const unsigned int *__restrict__ bucket_off, int nb, int R, int cbits, int64_t M,
                      int64_t *__restrict__ out_row, int64_t *__restrict__ out_col, T *__restrict__ out_value,
                      int64_t *__restrict__ out_rowptr, uint64_t *state /*[0]=ticket,[1..]=buckets*/,
                      long long *d_nnz_out, L2Desc desc) {
  static_assert(!DESC || sizeof(KeyT) == 4, "descriptor buckets are built for 32-bit local keys");
  using A = typename Acc<T>::type;
  constexpr int THREADS = L2_THREADS, ITEMS = L2_ITEMS, WARPS = L2_WARPS;
  extern __shared__ __align__(16) unsigned char l2_raw[];
  L2Smem<HAS_VALUE> &sm = *reinterpret_cast<L2Smem<HAS_VALUE> *>(l2_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned lt = lanemask_lt();
  const int LO = R > 32 ? R - 32 : 0;                       // low key bits kept beside the 32-bit part
  const int binshift0 = (R > L2_BINBITS ? R - L2_BINBITS : 0) - LO;  // applied to the 32-bit part
  const uint32_t lomask = (1u << LO) - 1u;
  const uint64_t rmask = R >= 64 ? ~uint64_t(0) : ((uint64_t(1) << R) - 1);
  const uint64_t cmask = (uint64_t(1) << cbits) - 1;
  const uint16_t *bstart16 = reinterpret_cast<const uint16_t *>(sm.bins);

  if (tid == 0) {  // the first bucket and its range
    const long long b = (long long)atomicAdd((unsigned long long *)state, 1ull);
    sm.meta[0] = l2_fetch_meta<DESC>(b, nb, bucket_off, desc, R, cbits, M, binshift0);
  }
  int slot = 0;
  bool pending = false;
  while (true) {
    __syncthreads();  // the previous round is completely written out; meta[slot] is filled
    const int64_t bucket = sm.meta[slot].bucket;
    const bool valid = bucket < nb;
    long long nxt_bucket = 0;  // thread 0: the bucket after this one
    if (valid) {
      const uint32_t start = sm.meta[slot].start;
      const int cnt = (int)(sm.meta[slot].end - start);
      const int binshift = sm.meta[slot].binshift;
      const unsigned nmerge = DESC ? sm.meta[slot].nmerge : 1u;
      const KeyT *kin = lowkeys;
      const uint32_t *vin = lowvals;
      if (DESC && sm.meta[slot].buf) {
        kin = reinterpret_cast<const KeyT *>(desc.keys1);
        vin = desc.vals1;
      }
      const int ie = (cnt + THREADS - 1) / THREADS;  // item rows in use (uniform)
      uint32_t *skeys = sm.skeys[slot], *svals = sm.svals[slot];
      uint8_t *slo = sm.slo[slot];
      // ---- A: keys and values into registers (entry j = i * THREADS + tid: its arrival index)
      uint32_t key[ITEMS], val[HAS_VALUE ? ITEMS : 1];
      uint32_t klo = 0;  // the low LO bits of the ITEMS keys, 4 bits each
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const int j = i * THREADS + (int)tid;
        key[i] = 0xffffffffu;
        if (i < ie && j < cnt) {
          if (sizeof(KeyT) == 8) {
            const uint64_t low = (uint64_t)ld_stream(kin + start + j) & rmask;
            key[i] = (uint32_t)(low >> LO);
            klo |= ((uint32_t)low & lomask) << (4 * i);
          } else {
            key[i] = (uint32_t)ld_stream(kin + start + j);
            if (DESC && nmerge > 1) {
              uint32_t delta = 0;
#pragma unroll
              for (int q = 0; q < 7; ++q) delta += (q + 1 < (int)nmerge && start + (uint32_t)j >= sm.meta[slot].bnd[q]) ? 1u : 0u;
              key[i] |= delta << R;
            }
          }
        }
      }
      if (HAS_VALUE) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          const int j = i * THREADS + (int)tid;
          val[i] = (i < ie && j < cnt) ? ld_stream(vin + start + j) : 0u;
        }
      }
      for (int q = tid; q < L2_BINS / 2 + 2; q += THREADS) sm.bins[q] = 0;
      if (tid == 0) {
        sm.maxbin = 0;
        sm.nheads = 0;
        nxt_bucket = (long long)atomicAdd((unsigned long long *)state, 1ull);
      }
      __syncthreads();
      // ---- B: counting pass on the top bits of the low key; the atomic's return value is the arrival rank in the bin
      uint32_t br[ITEMS];  // bin | rank << 16, later the entry's slot
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        br[i] = 0;
        if (i < ie && i * THREADS + (int)tid < cnt) {
          const unsigned b = DESC ? ((key[i] >> binshift) & (unsigned)(L2_BINS - 1)) : (key[i] >> binshift);
          const unsigned sh = (b & 1u) * 16u;
          const unsigned r = (atomicAdd(&sm.bins[b >> 1], 1u << sh) >> sh) & 0xffffu;
          br[i] = b | (r << 16);
        }
      }
      __syncthreads();
      // ---- C: exclusive scan of the 4,096 bin counts (8 per thread), largest bin
      {
        uint32_t w[4], c[8], sum = 0, mx = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) w[q] = sm.bins[tid * 4 + q];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint32_t t = (w[q >> 1] >> ((q & 1) * 16)) & 0xffffu;
          mx = t > mx ? t : mx;
          c[q] = sum;
          sum += t;
        }
        uint32_t total;
        const uint32_t base = block_exclusive_scan<uint32_t, THREADS>(sum, sm.scan_tmp, &total);
#pragma unroll
        for (int q = 0; q < 4; ++q) sm.bins[tid * 4 + q] = ((c[2 * q] + base) & 0xffffu) | ((c[2 * q + 1] + base) << 16);
        if (tid == THREADS - 1) sm.bins[L2_BINS / 2] = total & 0xffffu;  // sentinel: start of "bin 4096" = cnt
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          const uint32_t o = __shfl_xor_sync(0xffffffffu, mx, d);
          mx = o > mx ? o : mx;
        }
        if (lane == 0 && mx > 1) atomicMax(&sm.maxbin, mx);
      }
      if (tid == 0) sm.meta_next = l2_fetch_meta<DESC>(nxt_bucket, nb, bucket_off, desc, R, cbits, M, binshift0);
      __syncthreads();
      // ---- D / E: no bin holds two entries -> the bin start IS the sorted slot.  Otherwise entries go into their bins
      // in arrival order first, and every entry finds its slot = bin start + number of bin mates that sort before it
      // ((key, arrival index) lexicographic: equal keys keep their input order).
      if (sm.maxbin > 1) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          if (i < ie && i * THREADS + (int)tid < cnt) {
            const unsigned pos = (unsigned)bstart16[br[i] & 0xffffu] + (br[i] >> 16);
            sm.wk[pos] = ((uint64_t)key[i] << 16) | ((((klo >> (4 * i)) & 15u) << 12) | (unsigned)(i * THREADS + tid));
          }
        }
        __syncthreads();
        if (sm.maxbin > (unsigned)L2_HEAVY) {
          // Skewed keys (power-law columns inside a hub row): a bin may hold a hundred entries, and with the entries of a
          // thread spread at random every warp would walk the longest bin of its lanes for every item.  Here the work
          // is dealt by POSITION in the bin-ordered array instead: the lanes of a warp sit in the same bin (or in
          // neighbouring small ones), the walk over the bin mates is a broadcast read of the same words, and a bin of m
          // entries costs m * m / 32 warp iterations instead of m per warp and item.  The slots go back to the entries'
          // owners through a 16-bit map laid over the bin counters, which are no longer needed.
          uint32_t res[ITEMS];
#pragma unroll
          for (int k = 0; k < ITEMS; ++k) {
            const int p = k * THREADS + (int)tid;
            res[k] = 0xffffffffu;
            if (k < ie && p < cnt) {
              const uint64_t me = sm.wk[p];
              const uint32_t k32 = (uint32_t)(me >> 16);
              const unsigned b = DESC ? ((k32 >> binshift) & (unsigned)(L2_BINS - 1)) : (k32 >> binshift);
              const unsigned s_ = bstart16[b], e_ = bstart16[b + 1];
              unsigned less = 0;
              for (unsigned t = s_; t < e_; ++t) less += sm.wk[t] < me ? 1u : 0u;
              res[k] = ((s_ + less) << 12) | ((uint32_t)me & 0xfffu);
            }
          }
          __syncthreads();  // every read of the bin starts is done
          uint16_t *slotmap = reinterpret_cast<uint16_t *>(sm.bins);
#pragma unroll
          for (int k = 0; k < ITEMS; ++k)
            if (res[k] != 0xffffffffu) slotmap[res[k] & 0xfffu] = (uint16_t)(res[k] >> 12);
          __syncthreads();
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) {
            if (i < ie && i * THREADS + (int)tid < cnt) {
              const unsigned pos = slotmap[i * THREADS + (int)tid];
              skeys[pos] = key[i];
              if (sizeof(KeyT) == 8) slo[pos] = (uint8_t)((klo >> (4 * i)) & 15u);
              if (HAS_VALUE) svals[pos] = val[i];
            }
          }
        } else
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          if (i < ie && i * THREADS + (int)tid < cnt) {
            const unsigned b = br[i] & 0xffffu;
            const unsigned s_ = bstart16[b], e_ = bstart16[b + 1];
            const uint64_t me =
                ((uint64_t)key[i] << 16) | ((((klo >> (4 * i)) & 15u) << 12) | ((unsigned)(i * THREADS) + tid));
            unsigned less = 0;
            if (e_ - s_ > 1) {
              for (unsigned t = s_; t < e_; ++t) less += sm.wk[t] < me ? 1u : 0u;
            }
            skeys[s_ + less] = key[i];
            if (sizeof(KeyT) == 8) slo[s_ + less] = (uint8_t)((klo >> (4 * i)) & 15u);
            if (HAS_VALUE) svals[s_ + less] = val[i];
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          if (i < ie && i * THREADS + (int)tid < cnt) {
            const unsigned pos = (unsigned)bstart16[br[i] & 0xffffu];
            skeys[pos] = key[i];
            if (sizeof(KeyT) == 8) slo[pos] = (uint8_t)((klo >> (4 * i)) & 15u);
            if (HAS_VALUE) svals[pos] = val[i];
          }
        }
      }
      __syncthreads();
      // ---- F: number of distinct keys, published at once (the chain is fed as early as it can be)
      if (DEDUP) {
        unsigned mine = 0;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          if (i < ie) {
            const int j = i * THREADS + (int)tid;
            const bool h = j < cnt && (j == 0 || skeys[j] != skeys[j - 1] || (sizeof(KeyT) == 8 && slo[j] != slo[j - 1]));
            mine += __popc(__ballot_sync(0xffffffffu, h));
          }
        }
        if (lane == 0 && mine) atomicAdd(&sm.nheads, mine);
        __syncthreads();
        if (tid == 0) {
          const unsigned total = sm.nheads;
          sm.meta[slot].heads = (int)total;
          TileState::store(state + 1 + bucket,
                           TileState::pack(bucket == 0 ? TileState::PREFIX : TileState::AGG, (uint64_t)total));
        }
      }
    }
    // ---- the bucket sorted one round ago: offset from the chain, heads ranked, output
    if (pending) {
      const int ps_ = slot ^ 1;
      const int64_t pbucket = sm.meta[ps_].bucket;
      const uint32_t start = sm.meta[ps_].start;
      const int cnt = (int)(sm.meta[ps_].end - start);
      const int ie = (cnt + THREADS - 1) / THREADS;
      const uint32_t *skeys = sm.skeys[ps_], *svals = sm.svals[ps_];
      const uint8_t *slo = sm.slo[ps_];
      auto lowkey = [&](int j) -> uint64_t {  // the bucket-local key of sorted slot j
        return sizeof(KeyT) == 8 ? (((uint64_t)skeys[j] << LO) | slo[j]) : (uint64_t)skeys[j];
      };
      unsigned heads = 0, below[ITEMS];
      if (DEDUP) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          below[i] = 0;
          if (i < ie) {  // uniform
            const int j = i * THREADS + (int)tid;
            const bool h = j < cnt && (j == 0 || skeys[j] != skeys[j - 1] || (sizeof(KeyT) == 8 && slo[j] != slo[j - 1]));
            const unsigned bal = __ballot_sync(0xffffffffu, h);
            heads |= (unsigned)h << i;
            below[i] = __popc(bal & lt);
            if (lane == 0) sm.wcnt[i * WARPS + warp] = __popc(bal);
          } else if (lane == 0) {
            sm.wcnt[i * WARPS + warp] = 0;
          }
        }
        __syncthreads();
        if (warp == 0) {
          constexpr int PER = ITEMS * WARPS / 32;
          uint32_t v[PER], sum = 0;
#pragma unroll
          for (int q = 0; q < PER; ++q) {
            v[q] = sm.wcnt[lane * PER + q];
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
            sm.wcnt[lane * PER + q] = run;
            run += v[q];
          }
          // the chain: walk back over the published counts to the nearest prefix (wide: 128 buckets per round trip)
          uint64_t pfx = 0;
          if (pbucket > 0) {
            uint64_t *st = state + 1;
            int64_t base = pbucket - 1;
            bool done = false;
            while (!done) {
              constexpr int W = 4;
              uint64_t w[W];
#pragma unroll
              for (int q = 0; q < W; ++q) {
                const int64_t t = base - (int64_t)lane - 32 * q;
                w[q] = t >= 0 ? TileState::load(st + t) : TileState::pack(TileState::PREFIX, 0);
              }
              int used = 0;
#pragma unroll
              for (int q = 0; q < W; ++q) {
                if (done || used != q) continue;
                const uint64_t s_ = TileState::status(w[q]);
                const unsigned pmask = __ballot_sync(0xffffffffu, s_ == TileState::PREFIX);
                const unsigned emask = __ballot_sync(0xffffffffu, s_ == TileState::EMPTY);
                unsigned need = 0xffffffffu;
                if (pmask) {
                  const int fp = __ffs(pmask) - 1;
                  need = fp == 31 ? 0xffffffffu : ((2u << fp) - 1u);
                }
                if (emask & need) continue;
                uint64_t c = ((need >> lane) & 1u) ? TileState::value(w[q]) : 0;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
                pfx += c;
                used = q + 1;
                if (pmask) done = true;
              }
              base -= 32 * used;
            }
            if (lane == 0)
              TileState::store(st + pbucket, TileState::pack(TileState::PREFIX, pfx + (uint64_t)sm.meta[ps_].heads));
          }
          if (lane == 0) sm.excl = (long long)pfx;
        }
        __syncthreads();
      } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          const int j = i * THREADS + (int)tid;
          below[i] = (unsigned)j;  // every entry is kept: its output slot is its sorted slot
          if (i < ie && j < cnt) heads |= 1u << i;
        }
      }
      const int64_t excl = DEDUP ? (int64_t)sm.excl : (int64_t)start;
      const int nheads = DEDUP ? sm.meta[ps_].heads : cnt;
      // output: consecutive lanes hold consecutive heads, coalesced wherever nothing was merged
      const uint64_t hi_bits = (uint64_t)sm.meta[ps_].orig << R;
      const int64_t rbase = (int64_t)(hi_bits >> cbits);  // first row of the original bucket (local keys count from it)
      const int64_t rlo = sm.meta[ps_].rlo, rhi = sm.meta[ps_].rhi;  // rows owned by this bucket: [rlo, rhi)
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        if (!((heads >> i) & 1u)) continue;
        const int j = i * THREADS + (int)tid;
        const uint64_t lk = lowkey(j);
        const int64_t o = excl + below[i] + (DEDUP ? sm.wcnt[i * WARPS + warp] : 0u);
        // R >= cbits (a bucket is a range of whole rows): the row inside the bucket and the column are fields of the
        // local key itself - 32-bit arithmetic for 32-bit local keys
        int64_t r;
        if (sizeof(KeyT) == 8) {
          const uint64_t full = hi_bits | lk;
          r = (int64_t)(full >> cbits);
          out_col[o] = (int64_t)(full & cmask);
        } else {
          r = rbase + (int64_t)((uint32_t)lk >> cbits);
          out_col[o] = (int64_t)((uint32_t)lk & (uint32_t)cmask);
        }
        out_row[o] = r;
        if (HAS_VALUE && !DEDUP) out_value[o] = pay_as<T>(svals[j]);
        if (HAS_VALUE && DEDUP) {
          A acc = to_acc<T>(pay_as<T>(svals[j]));
          int jj = j + 1;
          while (jj < cnt && lowkey(jj) == lk) {
            acc = Red<RED, A>::op(acc, to_acc<T>(pay_as<T>(svals[jj])));
            ++jj;
          }
          out_value[o] = from_acc<T>(Red<RED, A>::finish(acc, (int64_t)(jj - j)));
        }
        if (out_rowptr) {
          const int64_t pr = j == 0 ? rlo - 1
                                    : (sizeof(KeyT) == 8 ? (int64_t)((hi_bits | lowkey(j - 1)) >> cbits)
                                                         : rbase + (int64_t)(skeys[j - 1] >> cbits));
          if (r != pr && r <= M) {  // this head opens row r ...
            out_rowptr[r] = o;
            if (r - pr > 1)  // ... and the empty rows before it (rare)
              for (int64_t q = pr + 1; q < r; ++q) out_rowptr[q] = o;
          }
        }
      }
      if (out_rowptr) {  // rows of this bucket after its last entry (all of them when the bucket is empty)
        // (+, not |: the extended keys of a merged bucket carry bits at and above R)
        int64_t first = cnt > 0 ? (int64_t)((hi_bits + lowkey(cnt - 1)) >> cbits) + 1 : rlo;
        if (first < rlo) first = rlo;  // (a refined bucket may start inside a row that an earlier bucket opened)
        for (int64_t q = first + tid; q < rhi; q += THREADS) out_rowptr[q] = excl + nheads;
      }
      if (DEDUP && pbucket == nb - 1 && tid == 0) *d_nnz_out = (long long)(excl + nheads);
    }
    if (!valid) break;
    pending = true;
    __syncthreads();  // every reader of meta[slot ^ 1] (the bucket just written out) is done
    if (tid == 0) sm.meta[slot ^ 1] = sm.meta_next;
    slot ^= 1;
  }
}

}  // namespace ps
