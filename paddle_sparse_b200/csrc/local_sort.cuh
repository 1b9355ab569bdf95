// Hybrid coalesce: bucket by the TOP bits with one-sweep passes, then sort every bucket INSIDE shared memory.
//
// An LSD sort of the 44-bit (row, col) key of an ogbn-products-shaped graph needs six passes over HBM.  Here the
// top T <= 14 key bits are sorted globally (one or two passes of the one-sweep kernel, radix_sort.cuh), which
// leaves 2^T buckets of a few thousand entries that are already in bucket order; one CTA then loads a bucket
// (32-bit low key + 32-bit value), sorts it with LSD passes that never leave the SM (ranking by shared-memory
// atomics, the same column scans as the global pass), merges duplicates, and writes the final int64 row / col,
// the reduced values and the rowptr of its own row range - coalesced, straight to their final place.
//   reference behaviour: storage.py:158-171 (sort-on-construct) + :454-486 (coalesce) + :211-222 (rowptr)
// Needs: key bits - T <= 32, T <= row bits (a bucket is a whole range of rows), every bucket <= LS_CAP entries
// (checked on the histogram, otherwise the caller falls back to the plain LSD sort).
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

namespace ps {

constexpr int LS_THREADS = 1024;
constexpr int LS_ITEMS = 8;
constexpr int LS_CAP = LS_THREADS * LS_ITEMS;  // entries one CTA sorts in shared memory
constexpr int LS_MAX_T = 15;                    // 2^15 bucket counters = 128 KB of shared memory in the histogram

// ---- histogram of the top T key bits (bucket sizes); keys are packed on the fly from the int64 arrays.
// With `max_rc` it is also the constructor's validation pass (storage.py:65-89: max(row), max(col)) - the same
// 16 bytes per entry are read once for both; entries whose bucket id falls outside [0, 2^T) (out-of-range
// indices the caller is about to reject) are not counted.
static __global__ void __launch_bounds__(1024)
ls_bucket_hist_kernel(const int64_t *__restrict__ row, const int64_t *__restrict__ col, int64_t n, int cbits, int R,
                      int T, unsigned int *__restrict__ gcount /*[2^T]*/, long long *__restrict__ max_rc /*[2] | null*/) {
  extern __shared__ unsigned int ls_sh[];
  const unsigned nb = 1u << T;
  for (unsigned j = threadIdx.x; j < nb; j += blockDim.x) ls_sh[j] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  long long mr = -1, mc = -1;
  for (; i + 3 * stride < n; i += 4 * stride) {
    uint64_t r[4], c[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) r[q] = (uint64_t)ld_stream(row + i + q * stride);
#pragma unroll
    for (int q = 0; q < 4; ++q) c[q] = (uint64_t)ld_stream(col + i + q * stride);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t b = ((r[q] << cbits) | c[q]) >> R;
      if (b < nb && (c[q] >> cbits) == 0) atomicAdd(&ls_sh[(unsigned)b], 1u);
      mr = (long long)r[q] > mr ? (long long)r[q] : mr;
      mc = (long long)c[q] > mc ? (long long)c[q] : mc;
    }
  }
  for (; i < n; i += stride) {
    const uint64_t r = (uint64_t)ld_stream(row + i), c = (uint64_t)ld_stream(col + i);
    const uint64_t b = ((r << cbits) | c) >> R;
    if (b < nb && (c >> cbits) == 0) atomicAdd(&ls_sh[(unsigned)b], 1u);
    mr = (long long)r > mr ? (long long)r : mr;
    mc = (long long)c > mc ? (long long)c : mc;
  }
  if (max_rc) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      const long long a = __shfl_xor_sync(0xffffffffu, mr, d), b = __shfl_xor_sync(0xffffffffu, mc, d);
      mr = a > mr ? a : mr;
      mc = b > mc ? b : mc;
    }
    if ((threadIdx.x & 31u) == 0) {
      if (mr >= 0) atomicMax(&max_rc[0], mr);
      if (mc >= 0) atomicMax(&max_rc[1], mc);
    }
  }
  __syncthreads();
  for (unsigned j = threadIdx.x; j < nb; j += blockDim.x)
    if (ls_sh[j]) atomicAdd(&gcount[j], ls_sh[j]);
}

// What one read of (row, col) leaves behind for the hybrid coalesce (ps_coo_inspect fills it at construction time,
// ps_coalesce consumes it): the validation maxima, the largest bucket, the bin bases of the global passes, the
// bucket offsets.  The first 24 bytes are what the host reads back.
struct LsInspect {
  long long max_rc[2];
  unsigned int max_bucket[2];
  unsigned long long ghist[2 * RS_MAX_RADIX];
  unsigned int bucket_off[(1 << LS_MAX_T) + 1];
  unsigned int gcount[1 << LS_MAX_T];
};

// ---- one CTA: bucket offsets (exclusive scan), the largest bucket, and the scanned digit histograms of the
// global passes (pass p sorts bucket-id bits [pbegin[p], pbegin[p] + pbits[p]))
struct LsPlan {
  int passes, pbegin[2], pbits[2];
};
// One CTA of 32 warps; warp w owns the contiguous bucket segment [w * nb / 32, (w + 1) * nb / 32) and reads it
// COALESCED (lane l takes buckets seg + k * 32 + l): a per-thread run of consecutive buckets reads with a 128-byte
// stride between lanes, which made this single-CTA kernel take 70 - 120 us on 2^15 buckets.
// `chunk_cap` > 0 (the CSC side, csc_local.cuh) also builds the chunk table: a bucket of more than chunk_cap entries
// is cut into ceil(count / chunk_cap) chunks, every other bucket is one chunk.
struct LsChunkOut {
  unsigned int *chunk_start;   // [nb + 1]
  unsigned int *chunk_bucket;  // [chunks]
  unsigned int *multi_list;    // [nb]
  unsigned int *info;          // [0] = chunks, [1] = buckets cut into several chunks
};
static __global__ void __launch_bounds__(1024)
ls_derive_kernel(const unsigned int *__restrict__ gcount, int T, LsPlan plan, unsigned int *__restrict__ bucket_off,
                 unsigned long long *__restrict__ ghist /*[passes][512]*/, unsigned int *__restrict__ max_bucket,
                 int chunk_cap, LsChunkOut co) {
  __shared__ unsigned int dig[2][RS_MAX_RADIX];  // 32-bit shared atomics are native, 64-bit ones are CAS loops
  __shared__ unsigned int wtot[32], wctot[32];
  __shared__ unsigned int smax, s_multi;
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  for (int j = tid; j < 2 * RS_MAX_RADIX; j += 1024) (&dig[0][0])[j] = 0;
  if (tid == 0) smax = s_multi = 0;
  __syncthreads();
  const int nb = 1 << T;
  const int seg = (nb + 31) / 32;               // buckets per warp
  const int rounds = (seg + 31) / 32;           // <= 32 for T <= 15
  const int seg0 = (int)warp * seg;
  // Real loops, not unrolled: this kernel runs once on one CTA, so straight-line code costs an instruction-cache miss
  // per line (a fully unrolled version took 90 us, 4x this one); the second sweep re-reads the counts (cache hits).
  unsigned int mx = 0, wsum = 0, wcsum = 0;
  // sweep 1: totals of the warp's segment, largest bucket, digit histograms of the global passes
#pragma unroll 1
  for (int k = 0; k < rounds; ++k) {
    const int b = seg0 + k * 32 + (int)lane;
    const bool in = k * 32 + (int)lane < seg && b < nb;
    const unsigned int c = in ? gcount[b] : 0u;
    wsum += c;
    wcsum += !in ? 0u : (chunk_cap > 0 && c > (unsigned)chunk_cap ? (c + chunk_cap - 1) / chunk_cap : 1u);
    mx = c > mx ? c : mx;
    if (c) {
      atomicAdd(&dig[0][(b >> plan.pbegin[0]) & ((1 << plan.pbits[0]) - 1)], c);
      if (plan.passes > 1) atomicAdd(&dig[1][(b >> plan.pbegin[1]) & ((1 << plan.pbits[1]) - 1)], c);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const unsigned int o = __shfl_xor_sync(0xffffffffu, mx, d);
    mx = o > mx ? o : mx;
    wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
    wcsum += __shfl_xor_sync(0xffffffffu, wcsum, d);
  }
  if (lane == 0) {
    wtot[warp] = wsum;
    wctot[warp] = wcsum;
    atomicMax(&smax, mx);
  }
  __syncthreads();
  unsigned int base = 0, cbase = 0, total = 0, ctotal = 0;
  for (int w = 0; w < 32; ++w) {
    if (w < (int)warp) { base += wtot[w]; cbase += wctot[w]; }
    total += wtot[w];
    ctotal += wctot[w];
  }
  // sweep 2: per round a warp scan of the counts (and of the chunk counts) on top of the running carries
  unsigned int carry = base, ccarry = cbase;
#pragma unroll 1
  for (int k = 0; k < rounds; ++k) {
    const int b = seg0 + k * 32 + (int)lane;
    const bool in = k * 32 + (int)lane < seg && b < nb;
    const unsigned int c = in ? gcount[b] : 0u;
    const unsigned int ch = !in ? 0u : (chunk_cap > 0 && c > (unsigned)chunk_cap ? (c + chunk_cap - 1) / chunk_cap : 1u);
    unsigned int inc = c, cinc = ch;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d), oc = __shfl_up_sync(0xffffffffu, cinc, d);
      if (lane >= (unsigned)d) { inc += o; cinc += oc; }
    }
    if (in) {
      bucket_off[b] = carry + inc - c;
      if (chunk_cap > 0) {
        const unsigned int cs = ccarry + cinc - ch;
        co.chunk_start[b] = cs;
        for (unsigned int j = 0; j < ch; ++j) co.chunk_bucket[cs + j] = (unsigned)b;
        if (ch > 1) co.multi_list[atomicAdd(&s_multi, 1u)] = (unsigned)b;
      }
    }
    carry += __shfl_sync(0xffffffffu, inc, 31);
    ccarry += __shfl_sync(0xffffffffu, cinc, 31);
  }
  if (tid == 1023) {
    bucket_off[nb] = total;
    if (chunk_cap > 0) co.chunk_start[nb] = ctotal;
  }
  __syncthreads();
  if (tid == 0) {
    if (max_bucket) *max_bucket = smax;
    if (chunk_cap > 0) {
      co.info[0] = ctotal;
      co.info[1] = s_multi;
    }
  }
  // exclusive scans of the digit histograms: warp p scans pass p (16 digits per lane)
  if (warp < (unsigned)plan.passes) {
    unsigned long long v[16], sum = 0;
    for (int q = 0; q < 16; ++q) {
      v[q] = dig[warp][lane * 16 + q];
      sum += v[q];
    }
    unsigned long long inc = sum;
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= (unsigned)d) inc += o;
    }
    unsigned long long r = inc - sum;
    for (int q = 0; q < 16; ++q) {
      ghist[warp * RS_MAX_RADIX + lane * 16 + q] = r;
      r += v[q];
    }
  }
}

// ---- the in-shared-memory sort + dedup of one bucket per CTA round
constexpr int LS_SORT_THREADS = 512;   // two CTAs per SM
constexpr int LS_SORT_ITEMS = 16;      // LS_CAP = 512 * 16
constexpr int LS_WARPS = LS_SORT_THREADS / 32;
constexpr int LS_ROW = 256 / 2;        // counter row in words (two 16-bit counters per word)
static_assert(LS_SORT_THREADS * LS_SORT_ITEMS == LS_CAP, "bucket capacity");

template <bool HAS_VALUE>
struct LsSmem {
  uint32_t keys[LS_CAP];
  uint32_t vals[HAS_VALUE ? LS_CAP : 1];
  uint32_t cnt[LS_WARPS][LS_ROW];  // two 16-bit counters per word, later the group's offset inside its half
  uint16_t half_tot[2][256];       // items of digit d in warps [0, 8) / [8, 16)
  uint16_t dbase[2][256];          // start of digit d's groups of that half inside the sorted bucket
  uint32_t scan_tmp[32];
  uint32_t wcnt[LS_SORT_ITEMS * LS_WARPS];
  long long bucket, excl;
  int heads;
};

// DEDUP = false: sort only (sort-on-construct, storage.py:158-171): every entry is kept, so the bucket's output
// offset is its input offset and no chaining over the buckets is needed.
template <typename T, int RED, bool HAS_VALUE, bool DEDUP>
__global__ void __launch_bounds__(LS_SORT_THREADS, 2)
ls_sort_dedup_kernel(const uint32_t *__restrict__ lowkeys, const uint32_t *__restrict__ lowvals,
                     const unsigned int *__restrict__ bucket_off, int nb, int R, int cbits, int64_t M,
                     int64_t *__restrict__ out_row, int64_t *__restrict__ out_col, T *__restrict__ out_value,
                     int64_t *__restrict__ out_rowptr, uint64_t *state /*[0]=ticket,[1..]=buckets*/,
                     long long *d_nnz_out) {
  using A = typename Acc<T>::type;
  constexpr int THREADS = LS_SORT_THREADS, ITEMS = LS_SORT_ITEMS, WARPS = LS_WARPS;
  static_assert(THREADS == 512 && WARPS == 16, "the digit prefix below is written for two halves of 8 warps");
  extern __shared__ __align__(16) unsigned char ls_raw[];
  LsSmem<HAS_VALUE> &sm = *reinterpret_cast<LsSmem<HAS_VALUE> *>(ls_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned lt = lanemask_lt();
  uint32_t *my_cnt = sm.cnt[warp];
  const int passes = (R + 7) / 8;
  const uint64_t cmask = (uint64_t(1) << cbits) - 1;

  while (true) {
    __syncthreads();  // the previous bucket is completely written out
    if (tid == 0) sm.bucket = (long long)atomicAdd((unsigned long long *)state, 1ull);
    __syncthreads();
    const int64_t bucket = sm.bucket;
    if (bucket >= nb) break;
    const uint32_t start = bucket_off[bucket];
    const int cnt = (int)(bucket_off[bucket + 1] - start);
    // balanced warp-blocked layout: warp w holds entries [w*32*ie, (w+1)*32*ie), item i of lane l is entry
    // w*32*ie + i*32 + l: ranking order (warp, item, lane) == entry order, which keeps every pass stable
    const int ie = (cnt + THREADS - 1) / THREADS;
    const int wb0 = warp * 32 * ie;
    uint32_t key[ITEMS], val[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int j = wb0 + i * 32 + lane;
      key[i] = 0;
      val[i] = 0;
      if (i < ie && j < cnt) {
        key[i] = ld_stream(lowkeys + start + j);
        if (HAS_VALUE) val[i] = ld_stream(lowvals + start + j);
      }
    }
    // ---- LSD passes inside shared memory
    for (int p = 0; p < passes; ++p) {
      const int shift = p * 8;
      const unsigned mask = (R - shift) >= 8 ? 255u : ((1u << (R - shift)) - 1u);
#pragma unroll
      for (int q = lane; q < LS_ROW; q += 32) my_cnt[q] = 0;
      __syncwarp();
      unsigned rank[ITEMS];
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        rank[i] = 0;
        if (i < ie && wb0 + i * 32 + (int)lane < cnt) {
          const unsigned d = (key[i] >> shift) & mask;
          const unsigned sh = (d & 1u) * 16u;
          rank[i] = (atomicAdd(&my_cnt[d >> 1], 1u << sh) >> sh) & 0xffffu;  // lanes served in order: rs_pass_kernel
        }
      }
      __syncthreads();
      // prefix over the warps, per digit: thread (d, h) walks the 8 warps of half h (row-wise, conflict-free);
      // then a block scan over the 256 digit totals, done redundantly by both halves
      const unsigned dg = tid & 255u, hf = tid >> 8;
      {
        uint16_t *c16 = reinterpret_cast<uint16_t *>(&sm.cnt[hf * 8][0]) + dg;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
          const uint32_t c = c16[w * (LS_ROW * 2)];
          c16[w * (LS_ROW * 2)] = (uint16_t)run;
          run += c;
        }
        sm.half_tot[hf][dg] = (uint16_t)run;
      }
      __syncthreads();
      {
        const uint32_t h0 = sm.half_tot[0][dg], h1 = sm.half_tot[1][dg];
        const uint32_t tot = h0 + h1;
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
        for (int w = 0; w < 8; ++w) wb += (w < (int)(warp & 7u)) ? sm.scan_tmp[hf * 8 + w] : 0u;
        sm.dbase[hf][dg] = (uint16_t)(wb + inc - tot + (hf ? h0 : 0u));
      }
      __syncthreads();
      const uint16_t *my_start = reinterpret_cast<const uint16_t *>(my_cnt);
      const uint16_t *my_dbase = sm.dbase[warp >> 3];
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        if (i < ie && wb0 + i * 32 + (int)lane < cnt) {
          const unsigned d = (key[i] >> shift) & mask;
          const unsigned pos = rank[i] + (unsigned)my_start[d] + (unsigned)my_dbase[d];
          sm.keys[pos] = key[i];
          if (HAS_VALUE) sm.vals[pos] = val[i];
        }
      }
      __syncthreads();
      if (p + 1 < passes) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
          const int j = wb0 + i * 32 + lane;
          if (i < ie && j < cnt) {
            key[i] = sm.keys[j];
            if (HAS_VALUE) val[i] = sm.vals[j];
          }
        }
      }
    }
    // ---- the bucket is sorted in shared memory: run heads (entry order = (item, warp, lane) with the striped
    // layout j = i*THREADS + tid), their ranks, the bucket's offset in the output by look-back over the buckets
    unsigned heads = 0, below[ITEMS];
    int64_t excl = (int64_t)start;
    int nheads = cnt;
    if (DEDUP) {
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        below[i] = 0;
        if (i < ie) {  // uniform
          const int j = i * THREADS + tid;
          const bool h = j < cnt && (j == 0 || sm.keys[j] != sm.keys[j - 1]);
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
        const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
        const uint64_t pfx = warp_lookback_wide<4>(state + 1, bucket, (uint64_t)total);
        if (lane == 0) {
          sm.excl = (long long)pfx;
          sm.heads = (int)total;
        }
      }
      __syncthreads();
      excl = sm.excl;
      nheads = sm.heads;
    } else {
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const int j = i * THREADS + tid;
        below[i] = (unsigned)j;  // every entry is kept: its output slot is its sorted slot
        if (i < ie && j < cnt) heads |= 1u << i;
      }
    }
    const uint64_t hi_bits = (uint64_t)bucket << R;
    // rows owned by this bucket: [rlo, rhi) (the last bucket also owns everything up to M)
    const int64_t rlo = (int64_t)(hi_bits >> cbits);
    int64_t rhi = (int64_t)((((uint64_t)bucket + 1) << R) >> cbits);
    if (bucket == nb - 1 || rhi > M + 1) rhi = M + 1;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      if (!((heads >> i) & 1u)) continue;
      const int j = i * THREADS + tid;
      const uint32_t lk = sm.keys[j];
      const int64_t o = excl + below[i] + (DEDUP ? sm.wcnt[i * WARPS + warp] : 0u);
      const uint64_t full = hi_bits | lk;
      const int64_t r = (int64_t)(full >> cbits);
      out_row[o] = r;  // consecutive lanes hold consecutive heads: coalesced wherever nothing was merged
      out_col[o] = (int64_t)(full & cmask);
      if (HAS_VALUE && !DEDUP) out_value[o] = pay_as<T>(sm.vals[j]);
      if (HAS_VALUE && DEDUP) {
        A acc = to_acc<T>(pay_as<T>(sm.vals[j]));
        int jj = j + 1;
        while (jj < cnt && sm.keys[jj] == lk) {
          acc = Red<RED, A>::op(acc, to_acc<T>(pay_as<T>(sm.vals[jj])));
          ++jj;
        }
        out_value[o] = from_acc<T>(Red<RED, A>::finish(acc, (int64_t)(jj - j)));
      }
      if (out_rowptr) {
        const int64_t pr = j == 0 ? rlo - 1 : (int64_t)((hi_bits | sm.keys[j - 1]) >> cbits);
        if (r != pr && r <= M) {  // (with DEDUP = false duplicates of a key share the row: only a row change writes)  // this head opens row r (and the empty rows before it)
          out_rowptr[r] = o;
          for (int64_t q = pr + 1; q < r; ++q) out_rowptr[q] = o;
        }
      }
    }
    if (out_rowptr) {  // rows of this bucket after its last entry (all of them when the bucket is empty)
      const int64_t first = cnt > 0 ? (int64_t)((hi_bits | sm.keys[cnt - 1]) >> cbits) + 1 : rlo;
      for (int64_t q = first + tid; q < rhi; q += THREADS) out_rowptr[q] = excl + nheads;
    }
    if (DEDUP && bucket == nb - 1 && tid == 0) *d_nnz_out = (long long)(excl + nheads);
  }
}

}  // namespace ps
