// Hybrid coalesce on SKEWED rows: an adaptive second partition level.
//
// After the global passes the entries are grouped by the top T bits of the key (local_sort.cuh / coalesce.cu), and a
// bucket CTA of local_sort2.cuh finishes a bucket of at most 4,096 entries.  Power-law graphs break that bound: on the
// R-MAT stream of the bench 76 % of the entries sit in ~1,500 of the 16,384 buckets (heaviest 1.3 M entries, heaviest
// single row 148 K).  Round 1 sent the WHOLE input to the plain LSD sort when one bucket overflowed.  Here only the
// oversized buckets are refined, level by level, by a SEGMENTED stable counting pass on the next 8 bits of the
// bucket-local key (the three-kernel scheme csc_local.cuh uses for hub columns: per-chunk digit histograms -> per-
// segment scan -> per-chunk scatter, every chunk a balanced 4,096-entry unit of work whatever the segment's size):
//   * a segment's 256 digit groups are packed greedily into FINAL buckets of <= 4,096 entries (adjacent small groups
//     share a bucket, so a thin tail does not become thousands of tiny buckets);
//   * a group that is still too large becomes a segment of the next level (fixed depth does not work: the heaviest
//     row needs 13 more key bits, most buckets none);
//   * final buckets are described by (start, lowest local key, highest local key, original bucket, buffer): the
//     bucket kernel picks its 12 bin bits right below the bucket's common key prefix and writes the rowptr entries of
//     the rows whose FIRST possible key falls between this bucket's lowest key and the next bucket's.
// Levels alternate between two buffers; entries of buckets that needed no refinement never move.
// All levels are enqueued unconditionally (a level with no segments is three empty launches): no host round trip
// between them.  A single key with more than 4,096 duplicates cannot be split: `overflow` is raised and the caller
// falls back to the LSD sort.
//   reference behaviour: storage.py:158-171 (sort-on-construct) + :454-486 (coalesce) - argsort has no cliff either
#pragma once
#include "common.cuh"

namespace ps {

constexpr int RF_CH = 4096;  // entries per chunk (the unit of work of the histogram and scatter kernels)
constexpr int RF_THREADS = 256;
constexpr int RF_ITEMS = RF_CH / RF_THREADS;
constexpr int RF_WARPS = RF_THREADS / 32;
constexpr int RF_MAX_LEVELS = 4;  // 8 bits each: local keys of up to 32 bits

struct RfSeg {
  uint32_t start, end;  // entries [start, end) of the level's source buffer
  uint32_t chunk0;      // first chunk of the segment in the level's chunk list
  uint32_t orig;        // original bucket (top T key bits)
  uint32_t klo;         // lowest possible bucket-local key of the segment
  uint32_t pad;
};
struct RfCtl {
  unsigned int nseg[RF_MAX_LEVELS + 2];    // [l]: segments of level l (1-based); [RF_MAX_LEVELS + 1] catches overflow
  unsigned int nchunk[RF_MAX_LEVELS + 2];
  unsigned int nfinal;
  unsigned int overflow;
};
struct RfLists {  // device arrays (capacities checked on the device: `overflow`)
  RfSeg *seg[RF_MAX_LEVELS + 2];
  uint32_t *chunk_seg[RF_MAX_LEVELS + 2];  // chunk -> segment of its level
  unsigned int seg_cap, chunk_cap, final_cap;
  // final buckets, unordered: start, lowest / highest local key, original bucket | buffer << 31
  uint32_t *f_start, *f_klo, *f_khi, *f_orig;
};

__device__ __forceinline__ void rf_push_final(const RfLists &L, RfCtl *ctl, uint32_t start, uint32_t klo, uint32_t khi,
                                              uint32_t orig, int buf) {
  const unsigned f = atomicAdd(&ctl->nfinal, 1u);
  if (f >= L.final_cap) {
    ctl->overflow = 1u;
    return;
  }
  L.f_start[f] = start;
  L.f_klo[f] = klo;
  L.f_khi[f] = khi;
  L.f_orig[f] = orig | ((uint32_t)buf << 31);
}
__device__ __forceinline__ void rf_push_seg(const RfLists &L, RfCtl *ctl, int level, int max_level, uint32_t start,
                                            uint32_t end, uint32_t orig, uint32_t klo) {
  if (level > max_level) {  // every key bit is consumed: one key with more duplicates than a bucket holds
    ctl->overflow = 1u;
    return;
  }
  const unsigned s = atomicAdd(&ctl->nseg[level], 1u);
  const unsigned nch = (end - start + RF_CH - 1) / RF_CH;
  const unsigned c0 = atomicAdd(&ctl->nchunk[level], nch);
  if (s >= L.seg_cap || c0 + nch > L.chunk_cap) {
    ctl->overflow = 1u;
    return;
  }
  L.seg[level][s] = RfSeg{start, end, c0, orig, klo, 0u};
  for (unsigned q = 0; q < nch; ++q) L.chunk_seg[level][c0 + q] = s;
}

// original buckets: oversized ones become the segments of level 1; the others are final, NEIGHBOURS MERGED while they
// fit one bucket and the bucket-local key has spare bits (a power-law graph has thousands of light row ranges of a
// few hundred entries each).  A merged bucket's local keys are extended by the original bucket's distance from the
// first one: (delta << R) | key, which must fit 32 bits - groups of 2^gbits <= 8 aligned neighbours, gbits <= 32 - R.
// f_orig: first original bucket | (originals merged - 1) << 16 | buffer << 31.
static __global__ void __launch_bounds__(256)
rf_select_kernel(const unsigned int *__restrict__ bucket_off, int nb, int R, int gbits, unsigned cap, RfLists L,
                 RfCtl *ctl) {
  const uint32_t kmax = R >= 32 ? 0xffffffffu : ((1u << R) - 1u);
  const int gsz = 1 << gbits;
  const int groups = (nb + gsz - 1) >> gbits;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += gridDim.x * blockDim.x) {
    int first = -1, last = -1;
    unsigned size = 0;
    for (int q = 0; q <= gsz; ++q) {
      const int b = (g << gbits) + q;
      const bool in = q < gsz && b < nb;
      const unsigned s = in ? bucket_off[b] : 0u, e = in ? bucket_off[b + 1] : 0u;
      const unsigned sz = e - s;
      const bool big = sz > cap;
      if (first >= 0 && (!in || big || size + sz > cap)) {  // close the open run of originals [first, last]
        rf_push_final(L, ctl, bucket_off[first], 0u, (uint32_t)(((uint64_t)(last - first) << R) | kmax),
                      (uint32_t)first | ((uint32_t)(last - first) << 16), 0);
        first = -1;
        size = 0;
      }
      if (!in || sz == 0) continue;  // an empty bucket's rows belong to the bucket listed before it
      if (big) {
        rf_push_seg(L, ctl, 1, RF_MAX_LEVELS, s, e, (uint32_t)b, 0u);
      } else {
        if (first < 0) first = b;
        last = b;
        size += sz;
      }
    }
  }
}

// per-chunk histogram of the level's digit
static __global__ void __launch_bounds__(RF_THREADS)
rf_hist_kernel(const uint32_t *__restrict__ keys, const RfSeg *__restrict__ segs, const uint32_t *__restrict__ chunk_seg,
               const RfCtl *__restrict__ ctl, int level, int shift, unsigned dmask, unsigned chunk_cap,
               unsigned int *__restrict__ chunk_hist /*[chunks][256]*/) {
  if (ctl->overflow) return;
  __shared__ unsigned int h[256];
  unsigned chunks = ctl->nchunk[level];
  if (chunks > chunk_cap) chunks = chunk_cap;
  for (unsigned c = blockIdx.x; c < chunks; c += gridDim.x) {
    const RfSeg sg = segs[chunk_seg[c]];
    h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned lo = sg.start + (c - sg.chunk0) * RF_CH;
    const unsigned hi = lo + RF_CH < sg.end ? lo + RF_CH : sg.end;
    for (unsigned i = lo + threadIdx.x; i < hi; i += RF_THREADS) atomicAdd(&h[(keys[i] >> shift) & dmask], 1u);
    __syncthreads();
    chunk_hist[(size_t)c * 256 + threadIdx.x] = h[threadIdx.x];
    __syncthreads();
  }
}

// per segment: every chunk's offset of every digit inside the segment; the digit groups packed into final buckets,
// oversized groups handed to the next level
static __global__ void __launch_bounds__(256)
rf_scan_kernel(const RfSeg *__restrict__ segs, const RfCtl *__restrict__ ctl_in, RfCtl *ctl, int level, int shift,
               unsigned dmask, unsigned cap, unsigned seg_cap, unsigned int *__restrict__ chunk_hist,
               unsigned int *__restrict__ seg_dstart /*[segments][256]*/, RfLists L, int max_level) {
  if (ctl_in->overflow) return;
  __shared__ unsigned int warp_sums[33];
  __shared__ unsigned int tot[256], dst[256], b_chunk[256], plan[6];
  __shared__ uint16_t g_first[256], g_last[256], b_digit[256];
  unsigned nseg = ctl_in->nseg[level];
  if (nseg > seg_cap) nseg = seg_cap;
  const unsigned d = threadIdx.x;
  for (unsigned s = blockIdx.x; s < nseg; s += gridDim.x) {
    const RfSeg sg = segs[s];
    const unsigned nch = (sg.end - sg.start + RF_CH - 1) / RF_CH;
    // column d of the segment's chunk histograms, eight chunks in flight (a hub bucket has hundreds of chunks: one
    // dependent load per chunk made this kernel the slowest of a level)
    unsigned int *col = chunk_hist + (size_t)sg.chunk0 * 256 + d;
    unsigned int run = 0;
    for (unsigned c = 0; c < nch; c += 8) {
      unsigned int t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = c + u < nch ? col[(size_t)(c + u) * 256] : 0u;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (c + u < nch) col[(size_t)(c + u) * 256] = run;
        run += t[u];
      }
    }
    unsigned int total;
    const unsigned int dstart = block_exclusive_scan<unsigned int, 256>(run, warp_sums, &total);
    tot[d] = run;
    dst[d] = dstart;  // (the scatter kernel adds the digit's start to the chunk's offset)
    seg_dstart[(size_t)s * 256 + d] = dstart;
    __syncthreads();
    // greedy packing of the 256 groups, in key order: thread 0 walks the totals in shared memory and only RECORDS the
    // buckets and sub-segments; their list slots come from three atomics for the whole segment and the records are
    // written by all threads (one global atomic per bucket made this kernel the slowest of a level)
    if (d == 0) {
      int first = -1, ng = 0, nbg = 0;
      unsigned size = 0, nck = 0;
      for (int q = 0; q <= (int)dmask + 1; ++q) {
        const unsigned sz = q <= (int)dmask ? tot[q] : 0u;
        const bool big = sz > cap;
        if (first >= 0 && (q > (int)dmask || big || size + sz > cap)) {  // close the open bucket: digits [first, last]
          int last = q - 1;
          while (tot[last] == 0) --last;
          g_first[ng] = (uint16_t)first;
          g_last[ng] = (uint16_t)last;
          ++ng;
          first = -1;
          size = 0;
        }
        if (q > (int)dmask || sz == 0) continue;
        if (big) {
          b_digit[nbg] = (uint16_t)q;
          b_chunk[nbg] = nck;
          nck += (sz + RF_CH - 1) / RF_CH;
          ++nbg;
        } else {
          if (first < 0) first = q;
          size += sz;
        }
      }
      plan[0] = (unsigned)ng;
      plan[1] = (unsigned)nbg;
      plan[2] = nck;
      plan[3] = ng ? atomicAdd(&ctl->nfinal, (unsigned)ng) : 0u;
      if (nbg) {
        if (level + 1 > max_level) {  // every key bit is consumed: one key with more duplicates than a bucket holds
          ctl->overflow = 1u;
          plan[1] = 0u;
        } else {
          plan[4] = atomicAdd(&ctl->nseg[level + 1], (unsigned)nbg);
          plan[5] = atomicAdd(&ctl->nchunk[level + 1], nck);
        }
      }
      if (plan[3] + (unsigned)ng > L.final_cap) {
        ctl->overflow = 1u;
        plan[0] = 0u;
      }
      if (plan[1] && (plan[4] + plan[1] > L.seg_cap || plan[5] + nck > L.chunk_cap)) {
        ctl->overflow = 1u;
        plan[1] = 0u;
      }
    }
    __syncthreads();
    {
      const int buf = level & 1;
      const unsigned span = 1u << shift;  // local keys per digit (shift <= 24)
      if (d < plan[0]) {
        const unsigned f = plan[3] + d, fi = g_first[d], la = g_last[d];
        L.f_start[f] = sg.start + dst[fi];
        L.f_klo[f] = sg.klo + fi * span;
        L.f_khi[f] = sg.klo + la * span + (span - 1u);
        L.f_orig[f] = sg.orig | ((uint32_t)buf << 31);
      }
      const unsigned nbg = plan[1];
      if (d < nbg) {
        const unsigned q = b_digit[d];
        L.seg[level + 1][plan[4] + d] =
            RfSeg{sg.start + dst[q], sg.start + dst[q] + tot[q], plan[5] + b_chunk[d], sg.orig, sg.klo + q * span, 0u};
      }
      for (unsigned m = 0; m < nbg; ++m) {  // chunk -> segment map of the sub-segments, all threads
        const unsigned nc = (tot[b_digit[m]] + RF_CH - 1) / RF_CH;
        for (unsigned q = d; q < nc; q += 256) L.chunk_seg[level + 1][plan[5] + b_chunk[m] + q] = plan[4] + m;
      }
    }
    __syncthreads();
  }
}

// per chunk: stable counting pass on the digit, staged in shared memory, written coalesced to the other buffer
struct RfSmem {
  uint32_t k[RF_CH];
  uint32_t v[RF_CH];
  uint32_t cnt[RF_WARPS][128];  // two 16-bit counters per word
  uint16_t dbase[256];          // start of digit d inside the sorted chunk
  uint32_t gb[256];             // global position of sorted slot j of digit d = gb[d] + j
  uint32_t scan_tmp[33];
};
template <bool HAS_VALUE>
static __global__ void __launch_bounds__(RF_THREADS, 3)  // 80 registers: no spills (28 B at four CTAs per SM)
rf_scatter_kernel(const uint32_t *__restrict__ src_k, const uint32_t *__restrict__ src_v, uint32_t *__restrict__ dst_k,
                  uint32_t *__restrict__ dst_v, const RfSeg *__restrict__ segs, const uint32_t *__restrict__ chunk_seg,
                  const RfCtl *__restrict__ ctl, int level, int shift, unsigned dmask, unsigned chunk_cap,
                  const unsigned int *__restrict__ chunk_hist, const unsigned int *__restrict__ seg_dstart) {
  if (ctl->overflow) return;
  extern __shared__ __align__(16) unsigned char rf_raw[];
  RfSmem &sm = *reinterpret_cast<RfSmem *>(rf_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  uint32_t *my_cnt = sm.cnt[warp];
  unsigned chunks = ctl->nchunk[level];
  if (chunks > chunk_cap) chunks = chunk_cap;
  for (unsigned c = blockIdx.x; c < chunks; c += gridDim.x) {
    const unsigned sid = chunk_seg[c];
    const RfSeg sg = segs[sid];
    const unsigned lo = sg.start + (c - sg.chunk0) * RF_CH;
    const int cnt = (int)((lo + RF_CH < sg.end ? lo + RF_CH : sg.end) - lo);
    const unsigned int goff = chunk_hist[(size_t)c * 256 + tid] + seg_dstart[(size_t)sid * 256 + tid];  // digit tid
    // warp-blocked layout: ranking order (warp, item, lane) == entry order, which keeps the pass stable
    const int ie = (cnt + RF_THREADS - 1) / RF_THREADS;
    const int wb0 = warp * 32 * ie;
#pragma unroll
    for (int q = lane; q < 128; q += 32) my_cnt[q] = 0;
    __syncwarp();
    uint32_t key[RF_ITEMS], val[HAS_VALUE ? RF_ITEMS : 1], dr[RF_ITEMS];
#pragma unroll
    for (int i = 0; i < RF_ITEMS; ++i) {
      const int j = wb0 + i * 32 + (int)lane;
      const bool ok = i < ie && j < cnt;
      key[i] = ok ? ld_stream(src_k + lo + j) : 0u;
      if (HAS_VALUE) val[i] = ok ? ld_stream(src_v + lo + j) : 0u;
      dr[i] = ok ? ((key[i] >> shift) & dmask) : 0xffffffffu;
    }
#pragma unroll
    for (int i = 0; i < RF_ITEMS; ++i) {
      if (dr[i] != 0xffffffffu) {
        const unsigned d = dr[i];
        const unsigned sh = (d & 1u) * 16u;
        const unsigned r = (atomicAdd(&my_cnt[d >> 1], 1u << sh) >> sh) & 0xffffu;  // lanes served in order
        dr[i] = d | (r << 8);
      }
    }
    __syncthreads();
    {  // prefix over the warps for digit tid, block scan over the digit totals
      uint16_t *c16 = reinterpret_cast<uint16_t *>(&sm.cnt[0][0]) + tid;
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < RF_WARPS; ++w) {
        const uint32_t t = c16[w * 256];
        c16[w * 256] = (uint16_t)tot;
        tot += t;
      }
      uint32_t total;
      const uint32_t dstart = block_exclusive_scan<uint32_t, RF_THREADS>(tot, sm.scan_tmp, &total);
      sm.dbase[tid] = (uint16_t)dstart;
      sm.gb[tid] = sg.start + goff - dstart;
    }
    __syncthreads();
    {
      const uint16_t *my_start = reinterpret_cast<const uint16_t *>(my_cnt);
#pragma unroll
      for (int i = 0; i < RF_ITEMS; ++i) {
        if (dr[i] != 0xffffffffu) {
          const unsigned d = dr[i] & 255u;
          const unsigned pos = (dr[i] >> 8) + (unsigned)my_start[d] + (unsigned)sm.dbase[d];
          sm.k[pos] = key[i];
          if (HAS_VALUE) sm.v[pos] = val[i];
        }
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int q = 0; q < RF_ITEMS; ++q) {
      const int j = q * RF_THREADS + (int)tid;
      if (j < cnt) {
        const uint32_t kk = sm.k[j];
        const size_t dst = (size_t)(sm.gb[(kk >> shift) & dmask] + (uint32_t)j);
        dst_k[dst] = kk;
        if (HAS_VALUE) dst_v[dst] = sm.v[j];
      }
    }
    __syncthreads();
  }
}

}  // namespace ps
