// CSC side as a hybrid sort: the entries are in CSR order, so csr2csc (storage.py:425-434: argsort(M*col + row)) is a
// STABLE sort on col alone.  Round 1 ran it as three global LSD passes over 16-byte tuples.  Here the TOP col bits
// are sorted by one or two one-sweep passes (radix_sort.cuh), which leaves buckets of a few thousand entries that
// cover 2^L <= 256 consecutive columns each; one CTA then finishes a bucket with ONE stable counting pass on the
// low L bits inside shared memory and writes everything the CSC side owes, streaming and coalesced, straight to
// its final place:
//   csr2csc (storage.py:425-434), colptr / colcount (:386-420; the bucket's digit histogram IS its columns'
//   counts, so the separate ind2ptr pass over the sorted columns is gone) and, when asked, the transposed arrays
//   col[csr2csc], row[csr2csc], value[csr2csc] (transpose.py:9-33).
// Skew: a bucket larger than one CTA's capacity (hub columns) is cut into chunks; a small pre-pass gives every
// chunk its per-digit offsets, so chunks of one bucket are independent and the work stays balanced.
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"
#include "local_sort.cuh"

namespace ps {

constexpr int CL_THREADS = 256;
constexpr int CL_ITEMS = 16;
constexpr int CL_CAP = CL_THREADS * CL_ITEMS;  // entries one CTA finishes in shared memory
constexpr int CL_WARPS = CL_THREADS / 32;
constexpr int CL_MAX_T = 15;   // 2^15 bucket counters = 128 KB of shared memory in the histogram
constexpr int CL_MAX_L = 8;    // low col bits left to the local pass
#ifndef PS_CL_MINB
#define PS_CL_MINB 3
#endif
constexpr int CL_MINB = PS_CL_MINB;  // CTAs per SM the local pass is compiled for: 3 = 80 registers and no spills (4 = 64
                                     // registers with 44-52 B of spills: fill_cache_ 2.729 against 2.665 ms)
constexpr int CL_ROW = 256 / 2;  // counter row in words (two 16-bit counters per word)

// ---- bucket sizes: histogram of col >> shift (one read of the int64 column array)
static __global__ void __launch_bounds__(1024)
cl_bucket_hist_kernel(const int64_t *__restrict__ col, int64_t n, int shift, unsigned nb,
                      unsigned int *__restrict__ gcount /*[nb]*/) {
  extern __shared__ unsigned int cl_sh[];
  for (unsigned j = threadIdx.x; j < nb; j += blockDim.x) cl_sh[j] = 0;
  __syncthreads();
  const int64_t pairs = n >> 1;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const bool aligned = (reinterpret_cast<uintptr_t>(col) & 15u) == 0;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (aligned) {
    for (; i + 3 * stride < pairs; i += 4 * stride) {
      longlong2 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = ld_stream2(col + 2 * (i + q * stride));
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint64_t a = (uint64_t)v[q].x >> shift, b = (uint64_t)v[q].y >> shift;
        if (a < nb) atomicAdd(&cl_sh[(unsigned)a], 1u);
        if (b < nb) atomicAdd(&cl_sh[(unsigned)b], 1u);
      }
    }
    for (; i < pairs; i += stride) {
      const longlong2 v = ld_stream2(col + 2 * i);
      const uint64_t a = (uint64_t)v.x >> shift, b = (uint64_t)v.y >> shift;
      if (a < nb) atomicAdd(&cl_sh[(unsigned)a], 1u);
      if (b < nb) atomicAdd(&cl_sh[(unsigned)b], 1u);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const uint64_t a = (uint64_t)col[n - 1] >> shift;
      if (a < nb) atomicAdd(&cl_sh[(unsigned)a], 1u);
    }
  } else {
    for (; i < n; i += stride) {
      const uint64_t a = (uint64_t)ld_stream(col + i) >> shift;
      if (a < nb) atomicAdd(&cl_sh[(unsigned)a], 1u);
    }
  }
  __syncthreads();
  for (unsigned j = threadIdx.x; j < nb; j += blockDim.x)
    if (cl_sh[j]) atomicAdd(&gcount[j], cl_sh[j]);
}

// bucket offsets, the bin bases of the global passes and the chunk table come from ls_derive_kernel (local_sort.cuh)
struct ClInfo {          // device-side facts the later kernels read (LsChunkOut::info)
  unsigned int chunks;   // total number of chunks
  unsigned int n_multi;  // buckets cut into more than one chunk
};

// ---- chunks of oversized buckets only: per-chunk digit histogram ...
static __global__ void __launch_bounds__(256)
cl_chunk_hist_kernel(const uint32_t *__restrict__ scol, const unsigned int *__restrict__ bucket_off,
                     const unsigned int *__restrict__ chunk_start, const unsigned int *__restrict__ chunk_bucket,
                     const ClInfo *__restrict__ info, unsigned lmask, unsigned int *__restrict__ chunk_hist /*[chunks][256]*/) {
  if (info->n_multi == 0) return;
  __shared__ unsigned int h[256];
  const unsigned chunks = info->chunks;
  for (unsigned c = blockIdx.x; c < chunks; c += gridDim.x) {
    const unsigned b = chunk_bucket[c];
    const unsigned first = chunk_start[b];
    if (chunk_start[b + 1] - first < 2) continue;  // uniform
    h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned lo = bucket_off[b] + (c - first) * CL_CAP;
    const unsigned end = bucket_off[b + 1];
    const unsigned hi = lo + CL_CAP < end ? lo + CL_CAP : end;
    for (unsigned i = lo + threadIdx.x; i < hi; i += 256) atomicAdd(&h[scol[i] & lmask], 1u);
    __syncthreads();
    chunk_hist[(size_t)c * 256 + threadIdx.x] = h[threadIdx.x];
    __syncthreads();
  }
}

// ... turned into every chunk's per-digit offset inside its bucket; also the bucket's colptr / colcount
static __global__ void __launch_bounds__(256)
cl_chunk_scan_kernel(const unsigned int *__restrict__ bucket_off, const unsigned int *__restrict__ chunk_start,
                     const unsigned int *__restrict__ multi_list, const ClInfo *__restrict__ info, int L, int64_t N,
                     unsigned int *__restrict__ chunk_hist, int64_t *__restrict__ out_colptr,
                     int64_t *__restrict__ out_colcount) {
  const unsigned n_multi = info->n_multi;
  if (n_multi == 0) return;
  __shared__ unsigned int warp_sums[33];
  const unsigned d = threadIdx.x;
  for (unsigned m = blockIdx.x; m < n_multi; m += gridDim.x) {
    const unsigned b = multi_list[m];
    const unsigned c0 = chunk_start[b], c1 = chunk_start[b + 1];
    // column d of the bucket's chunk histograms, eight chunks in flight (a hub column's bucket has hundreds of chunks:
    // one dependent load per chunk made this kernel take 146 us on the R-MAT stream)
    unsigned int *colp = chunk_hist + (size_t)c0 * 256 + d;
    const unsigned nch = c1 - c0;
    unsigned int run = 0;
    for (unsigned c = 0; c < nch; c += 8) {
      unsigned int t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = c + u < nch ? colp[(size_t)(c + u) * 256] : 0u;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (c + u < nch) colp[(size_t)(c + u) * 256] = run;
        run += t[u];
      }
    }
    unsigned int total;
    const unsigned int dstart = block_exclusive_scan<unsigned int, 256>(run, warp_sums, &total);
    for (unsigned c = 0; c < nch; c += 8) {
      unsigned int t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = c + u < nch ? colp[(size_t)(c + u) * 256] : 0u;
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (c + u < nch) colp[(size_t)(c + u) * 256] = t[u] + dstart;
    }
    const int64_t colid = ((int64_t)b << L) + d;
    if (d < (1u << L) && colid <= N) {
      if (out_colptr) out_colptr[colid] = (int64_t)bucket_off[b] + dstart;
      if (out_colcount && colid < N) out_colcount[colid] = (int64_t)run;
    }
    if (d == 0 && out_colptr && colid + ((int64_t)1 << L) == N) out_colptr[N] = (int64_t)bucket_off[b + 1];
  }
}

// ---- the local pass: one chunk per CTA round.  256 threads and ~42 KB of shared memory per CTA, so that four or
// five CTAs in different phases share an SM (the pass is a chain of short latency-bound phases: metadata, column
// loads, ranking, scans, payload loads, write-out).  The payload is staged in TWO rounds through the same two
// buffers - (entry index) -> csr2csc, t_row; then (row, value) -> t_col, t_value - and the next chunk's metadata and
// the second round's loads are requested one phase early.
// csc2csr (storage.py:439-447: argsort(csr2csc), i.e. the inverse permutation) rides along: the kernel knows, for every
// entry, its CSR position p (the carried index) and its CSC position q (where it is written).
//   mode 1 (the output fits L2 comfortably): out[p] = q directly.
//   mode 2: the FIRST partition pass of the bucketed inverse (transpose.cu: invert_perm_impl) is done here - the pairs
//   (p, q) of a chunk are grouped by the digit `(p >> shift) & mask` in shared memory and appended, one run per digit,
//   to that digit's region of (tp, tq), whose start is known in closed form (p is a permutation) and whose fill level
//   is a global cursor.  Order inside a region does not matter: the last stage drops the pairs of a 4,096-entry window
//   into a shared-memory table.  Saves one full read of csr2csc and one global pass.
struct ClInverse {
  int mode;  // 0 = none
  int shift, bits;
  unsigned long long *cursor;  // [2^bits], initialised with the regions' starts
  uint32_t *tp, *tq;
  int64_t *out;  // mode 1
};
struct ClMeta {
  long long chunk;
  unsigned b, boff, bend, start, multi;
};
struct ClSmem {
  uint32_t a[CL_CAP];
  uint32_t v[CL_CAP];
  uint8_t dig[CL_CAP];    // low col bits of the sorted slot
  uint16_t ix[CL_CAP];    // inverse round: sorted slot of the pair that goes to this partition slot
  uint32_t cnt[CL_WARPS][CL_ROW];  // two 16-bit counters per word; later the group's offset inside the digit;
                                   // inverse round: [0,256) bin counts / starts, [256,512) global bases
  uint16_t dbase[256];             // start of digit d inside the sorted chunk
  uint32_t gb[256];                // global position of sorted slot j of digit d = gb[d] + j
  uint32_t scan_tmp[40];
  ClMeta meta[2];
};

__device__ __forceinline__ ClMeta cl_fetch_meta(unsigned long long *ticket, unsigned chunks,
                                                const unsigned int *__restrict__ bucket_off,
                                                const unsigned int *__restrict__ chunk_start,
                                                const unsigned int *__restrict__ chunk_bucket) {
  ClMeta m;
  m.chunk = (long long)atomicAdd(ticket, 1ull);
  m.b = m.boff = m.bend = m.start = m.multi = 0;
  if (m.chunk < (long long)chunks) {
    m.b = chunk_bucket[m.chunk];
    const unsigned cfirst = chunk_start[m.b];
    m.multi = chunk_start[m.b + 1] - cfirst > 1 ? 1u : 0u;
    m.boff = bucket_off[m.b];
    m.bend = bucket_off[m.b + 1];
    m.start = m.boff + (unsigned)(m.chunk - cfirst) * CL_CAP;
  }
  return m;
}

template <bool CARRY, bool HAS_VALUE, int INV>
__global__ void __launch_bounds__(CL_THREADS, CL_MINB)
cl_local_kernel(const uint32_t *__restrict__ scol, const uint32_t *__restrict__ sp, const uint32_t *__restrict__ srow,
                const uint32_t *__restrict__ sval, const unsigned int *__restrict__ bucket_off,
                const unsigned int *__restrict__ chunk_start, const unsigned int *__restrict__ chunk_bucket,
                const unsigned int *__restrict__ chunk_hist, const ClInfo *__restrict__ info, int nb, int L, int64_t N,
                int64_t *__restrict__ out_csr2csc, int64_t *__restrict__ out_colptr, int64_t *__restrict__ out_colcount,
                int64_t *__restrict__ out_t_row, int64_t *__restrict__ out_t_col, uint32_t *__restrict__ out_t_val,
                unsigned long long *ticket, ClInverse inv) {
  constexpr int THREADS = CL_THREADS, ITEMS = CL_ITEMS, WARPS = CL_WARPS;
  static_assert(THREADS == 256 && WARPS == 8, "thread d owns digit d in the scans below");
  extern __shared__ __align__(16) unsigned char cl_raw[];
  ClSmem &sm = *reinterpret_cast<ClSmem *>(cl_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  uint32_t *my_cnt = sm.cnt[warp];
  const unsigned lmask = (1u << L) - 1u;
  const unsigned chunks = info->chunks;

  if (tid == 0) sm.meta[0] = cl_fetch_meta(ticket, chunks, bucket_off, chunk_start, chunk_bucket);
  int cur = 0;
  while (true) {
    __syncthreads();  // the previous chunk is completely written out; meta[cur] is there
    const ClMeta me = sm.meta[cur];
    if (me.chunk >= (long long)chunks) break;
    const int64_t chunk = me.chunk;
    const unsigned b = me.b, boff = me.boff, start = me.start;
    const bool multi = me.multi != 0;
    const int cnt = (int)((me.bend - start) < (unsigned)CL_CAP ? (me.bend - start) : (unsigned)CL_CAP);
    ClMeta nxt;
    if (tid == 0) nxt = cl_fetch_meta(ticket, chunks, bucket_off, chunk_start, chunk_bucket);  // consumed at the end
    // warp-blocked layout: warp w holds entries [w*32*ie, (w+1)*32*ie); ranking order (warp, item, lane) == entry
    // order, which keeps the pass stable (ties keep CSR = row order)
    const int ie = (cnt + THREADS - 1) / THREADS;
    const int wb0 = warp * 32 * ie;
#pragma unroll
    for (int q = lane; q < CL_ROW; q += 32) my_cnt[q] = 0;
    __syncwarp();
    uint32_t dr[ITEMS];  // digit | rank << 8, later the sorted slot
    uint32_t pa[ITEMS];  // first-round payload (entry index), requested before the scans
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = wb0 + i * 32 + lane;
      const bool ok = i < ie && j < cnt;
      dr[i] = ok ? (ld_stream(scol + start + j) & lmask) : 0xffffffffu;
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = wb0 + i * 32 + lane;
      pa[i] = (i < ie && j < cnt) ? ld_stream(sp + start + j) : 0u;
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      if (dr[i] != 0xffffffffu) {
        const unsigned d = dr[i];
        const unsigned sh = (d & 1u) * 16u;
        const unsigned r = (atomicAdd(&my_cnt[d >> 1], 1u << sh) >> sh) & 0xffffu;  // lanes served in order
        dr[i] = d | (r << 8);
      }
    }
    __syncthreads();
    // prefix over the warps for digit tid, then a block scan over the 256 digit totals
    {
      uint16_t *c16 = reinterpret_cast<uint16_t *>(&sm.cnt[0][0]) + tid;
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < WARPS; ++w) {
        const uint32_t c = c16[w * (CL_ROW * 2)];
        c16[w * (CL_ROW * 2)] = (uint16_t)tot;
        tot += c;
      }
      uint32_t inc = tot;
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
      const uint32_t dstart = wb + inc - tot;  // start of digit tid inside the sorted chunk
      sm.dbase[tid] = (uint16_t)dstart;
      // where the chunk's run of this digit goes: a bucket that is one chunk is written out as one streaming range
      const uint32_t rel = multi ? chunk_hist[(size_t)chunk * 256 + tid] : dstart;
      sm.gb[tid] = boff + rel - dstart;
      if (!multi) {  // the digit histogram IS the columns' counts: colptr / colcount of this bucket's columns
        const int64_t colid = ((int64_t)b << L) + tid;
        if (tid <= lmask && colid <= N) {
          if (out_colptr) out_colptr[colid] = (int64_t)boff + dstart;
          if (out_colcount && colid < N) out_colcount[colid] = (int64_t)tot;
        }
      }
    }
    __syncthreads();
    // sorted slots; stage round 1 (digit, entry index)
    {
      const uint16_t *my_start = reinterpret_cast<const uint16_t *>(my_cnt);
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        if (dr[i] != 0xffffffffu) {
          const unsigned d = dr[i] & 255u;
          const unsigned pos = (dr[i] >> 8) + (unsigned)my_start[d] + (unsigned)sm.dbase[d];
          dr[i] = pos;
          sm.dig[pos] = (uint8_t)d;
          sm.a[pos] = pa[i];
        }
      }
    }
    // round 2 is requested now when nothing comes in between: its latency hides behind the write-out of round 1
    uint32_t pb[CARRY && HAS_VALUE ? ITEMS : 1];
    if (CARRY && INV != 2) {
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const int j = wb0 + i * 32 + lane;
        const bool ok = i < ie && j < cnt;
        pa[i] = ok ? ld_stream(srow + start + j) : 0u;
        if (HAS_VALUE) pb[i] = ok ? ld_stream(sval + start + j) : 0u;
      }
    }
    uint32_t *ibin = &sm.cnt[0][0];   // inverse round: [0,256) counts, then starts
    uint32_t *igb = ibin + 256;       //                [256,512) global base of the digit's run minus its start
    __syncthreads();
    if (INV == 2) ibin[tid] = 0;  // every thread is past its reads of the rank counters
    // coalesced write-out: sorted slot j of digit d goes to gb[d] + j (one streaming range when the bucket is one
    // chunk)
    const int64_t colbase = (int64_t)b << L;
#pragma unroll 4
    for (int k = 0; k < ITEMS; ++k) {
      const int j = k * THREADS + tid;
      if (j < cnt) {
        const unsigned d = sm.dig[j];
        // (a bucket that is one chunk goes out as one streaming range: no per-digit base to look up)
        const size_t dst = (size_t)((multi ? sm.gb[d] : boff) + (uint32_t)j);
        const uint32_t p = sm.a[j];
        out_csr2csc[dst] = (int64_t)p;
        if (CARRY) out_t_row[dst] = colbase + d;
        if (INV == 1) inv.out[p] = (int64_t)dst;
      }
    }
    if (INV == 2) {
      // ---- inverse round: pairs (p, q) grouped by the digit of p, one run per digit appended to its region
      const unsigned imask = (1u << inv.bits) - 1u;
      uint32_t ir[ITEMS];  // digit | rank << 8
      __syncthreads();     // the counters are zero
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        ir[k] = 0;
        if (j < cnt) {
          const unsigned d = (sm.a[j] >> inv.shift) & imask;
          ir[k] = d | (atomicAdd(&ibin[d], 1u) << 8);  // any order inside a run will do
        }
      }
      __syncthreads();
      {
        const uint32_t c = ibin[tid];
        uint32_t total;
        const uint32_t st = block_exclusive_scan<uint32_t, THREADS>(c, sm.scan_tmp, &total);
        const unsigned long long g = c ? atomicAdd(inv.cursor + tid, (unsigned long long)c) : 0ull;
        ibin[tid] = st;
        igb[tid] = (uint32_t)g - st;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        if (j < cnt) sm.ix[ibin[ir[k] & 255u] + (ir[k] >> 8)] = (uint16_t)j;
      }
      __syncthreads();
#pragma unroll 4
      for (int k = 0; k < ITEMS; ++k) {
        const int j2 = k * THREADS + tid;
        if (j2 < cnt) {
          const unsigned j = sm.ix[j2];
          const uint32_t p = sm.a[j];
          const uint32_t q = (multi ? sm.gb[sm.dig[j]] : boff) + j;
          const size_t dst = (size_t)(igb[(p >> inv.shift) & imask] + (uint32_t)j2);
          inv.tp[dst] = p;
          inv.tq[dst] = q;
        }
      }
      if (CARRY) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          const int j = wb0 + i * 32 + lane;
          const bool ok = i < ie && j < cnt;
          pa[i] = ok ? ld_stream(srow + start + j) : 0u;
          if (HAS_VALUE) pb[i] = ok ? ld_stream(sval + start + j) : 0u;
        }
      }
    }
    if (CARRY) {
      __syncthreads();  // round 1 (and the inverse round) is read out of the buffers
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        if (dr[i] != 0xffffffffu) {
          sm.a[dr[i]] = pa[i];
          if (HAS_VALUE) sm.v[dr[i]] = pb[i];
        }
      }
      __syncthreads();
#pragma unroll 4
      for (int k = 0; k < ITEMS; ++k) {
        const int j = k * THREADS + tid;
        if (j < cnt) {
          const size_t dst = (size_t)((multi ? sm.gb[sm.dig[j]] : boff) + (uint32_t)j);
          out_t_col[dst] = (int64_t)sm.a[j];
          if (HAS_VALUE) out_t_val[dst] = sm.v[j];
        }
      }
    }
    if (tid == 0) {
      if (b == (unsigned)nb - 1 && !multi && out_colptr && (((int64_t)nb) << L) == N)
        out_colptr[N] = (int64_t)me.bend;  // N is the first column past the last bucket: nobody else owns colptr[N]
      sm.meta[cur ^ 1] = nxt;
    }
    cur ^= 1;
  }
}

}  // namespace ps
