// sample_adj on the GPU.
//
// Reference behaviour restated (not ported): csrc/sample.cpp:8-24 -> csrc/cpu/sample_cpu.cpp:10-148.
// The reference has NO device path (sample.cpp:13-15 throws on GPU tensors); its CPU loop is
// one thread, one std::unordered_map probe per sampled edge and one tensor allocation per
// random draw (csrc/cpu/utils.h:22-34).  Here the batch is processed by a short chain of
// kernels, a warp per seed row:
//   count    per-seed output degree -> single-pass scan -> out_rowptr
//   pick     all-neighbours / k draws with replacement / Robert Floyd without replacement
//            (sample_cpu.cpp:44-64 / 66-87 / 89-106), draws from Philox keyed by
//            (seed, seed-row position, draw index)
//   relabel  first-occurrence numbering (sample_cpu.cpp:35-40 and the n_id_map probes) done
//            order-preservingly in parallel: an open-addressing table keeps, per global id,
//            the minimum "priority" (duplicate seeds: last position wins, as
//            n_id_map[i] = n overwrites; picks: position in emission order), flags of first
//            occurrences are scanned into ranks
//   emit     per-row sort by (local id, e_id) (sample_cpu.cpp:133-145) — a register bitonic
//            network when every row has <= 32 picks, two stable radix sorts otherwise.
// The emission order inside a sampled row is Floyd's insertion order (the reference iterates a
// std::unordered_set there, i.e. an implementation-defined order; see oracle/oracle_port.c).
#include "common.cuh"
#include "convert.cuh"
#include "philox.cuh"
#include "radix_sort.cuh"

namespace ps {

// number of entries seed row i emits
__device__ __forceinline__ int64_t sa_row_count(int64_t deg, int64_t k, bool replace) {
  if (k < 0 || (!replace && deg <= k)) return deg;
  if (replace) return deg > 0 ? k : 0;
  return k;
}

__global__ void __launch_bounds__(256)
sa_count_kernel(const int64_t *__restrict__ rowptr, const int64_t *__restrict__ idx, int64_t S, int64_t k,
                bool replace, int64_t *__restrict__ cnt) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= S) return;
  const int64_t n = idx[i];
  cnt[i] = sa_row_count(rowptr[n + 1] - rowptr[n], k, replace);
}

// sum of the degrees of the seed rows (ps_sample_adj_bound, num_neighbors < 0)
__global__ void __launch_bounds__(256)
sa_degree_sum_kernel(const int64_t *__restrict__ rowptr, const int64_t *__restrict__ idx, int64_t S,
                     unsigned long long *out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned long long acc = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < S; i += stride) {
    const int64_t n = idx[i];
    acc += (unsigned long long)(rowptr[n + 1] - rowptr[n]);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
  if (lane_id() == 0 && acc) atomicAdd(out, acc);
}

// One warp per seed row: writes the picked edge ids (emission order), their global column ids
// and the seed position of every pick.
__global__ void __launch_bounds__(256)
sa_pick_kernel(const int64_t *__restrict__ rowptr, const int64_t *__restrict__ col, const int64_t *__restrict__ idx,
               int64_t S, int64_t k, bool replace, uint64_t seed, const int64_t *__restrict__ out_rowptr,
               int64_t *e_buf, int64_t *__restrict__ gcol_buf, int64_t *__restrict__ rowid_buf) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= S) return;
  const unsigned lane = lane_id();
  const int64_t n = idx[i];
  const int64_t row_start = rowptr[n], deg = rowptr[n + 1] - row_start;
  const int64_t base = out_rowptr[i], cnt = out_rowptr[i + 1] - base;
  if (cnt == 0) return;
  if (k < 0 || (!replace && deg <= k)) {
    for (int64_t j = lane; j < cnt; j += 32) e_buf[base + j] = row_start + j;
  } else if (replace) {
    for (int64_t j = lane; j < cnt; j += 32)
      e_buf[base + j] = row_start + philox_randint(seed, (uint64_t)i, (uint32_t)j, 0, deg);
  } else {
    // Robert Floyd: for j = deg-k .. deg-1: t = randint[0, j); take t, or j when t was taken.
    volatile int64_t *picks = e_buf + base;  // holds row_start + pick
    for (int64_t u0 = 0; u0 < k; u0 += 32) {
      const int64_t u_mine = u0 + lane;
      int64_t t_mine = 0;
      if (u_mine < k) t_mine = philox_randint(seed, (uint64_t)i, (uint32_t)u_mine, 0, deg - k + u_mine);
      const int lim = (int)((k - u0) < 32 ? (k - u0) : 32);
      for (int q = 0; q < lim; ++q) {
        const int64_t u = u0 + q;
        const int64_t t = __shfl_sync(0xffffffffu, t_mine, q);
        bool dup = false;
        for (int64_t p = lane; p < u; p += 32) dup |= (picks[p] == row_start + t);
        dup = __any_sync(0xffffffffu, dup);
        if (lane == 0) picks[u] = row_start + (dup ? (deg - k + u) : t);
        __syncwarp();
      }
    }
  }
  __syncwarp();
  for (int64_t j = lane; j < cnt; j += 32) {
    const int64_t e = ((volatile int64_t *)e_buf)[base + j];
    gcol_buf[base + j] = col[e];
    rowid_buf[base + j] = i;
  }
}

__device__ __forceinline__ uint64_t sa_mix(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}

// Insert every seed (priority S-1-n: the LAST duplicate wins) and every pick (priority S + p)
// with atomicMin; remember the slot of every pick.
__global__ void __launch_bounds__(256)
sa_insert_kernel(const int64_t *__restrict__ idx, int64_t S, const int64_t *__restrict__ gcol_buf,
                 const int64_t *__restrict__ out_rowptr, long long *keys, long long *vals, uint64_t mask,
                 uint32_t *__restrict__ slot_buf) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t E = out_rowptr[S];
  if (q >= S + E) return;
  const long long c = q < S ? idx[q] : gcol_buf[q - S];
  const long long pri = q < S ? (S - 1 - q) : q;  // picks: S + (q - S)
  uint64_t h = sa_mix((uint64_t)c) & mask;
  while (true) {
    const long long prev = atomicCAS((unsigned long long *)&keys[h], (unsigned long long)-1LL, (unsigned long long)c);
    if (prev == -1LL || prev == c) break;
    h = (h + 1) & mask;
  }
  atomicMin(&vals[h], pri);
  if (q >= S) slot_buf[q - S] = (uint32_t)h;
}

// flag[p] = 1 iff pick p is the first occurrence of a node that is not a seed
__global__ void __launch_bounds__(256)
sa_flag_kernel(int64_t S, int64_t bound, const int64_t *__restrict__ out_rowptr, const long long *__restrict__ vals,
               const uint32_t *__restrict__ slot_buf, int64_t *__restrict__ flag) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= bound) return;
  const int64_t E = out_rowptr[S];
  flag[p] = (p < E && vals[slot_buf[p]] == S + p) ? 1 : 0;
}

__global__ void __launch_bounds__(256)
sa_relabel_kernel(const int64_t *__restrict__ idx, int64_t S, int64_t bound, const int64_t *__restrict__ out_rowptr,
                  const int64_t *__restrict__ gcol_buf, const long long *__restrict__ vals,
                  const uint32_t *__restrict__ slot_buf, const int64_t *__restrict__ rank /*[bound+1]*/,
                  int64_t *__restrict__ local_buf, int64_t *__restrict__ out_n_id, long long *d_sizes) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t E = out_rowptr[S];
  if (q == 0) {
    d_sizes[0] = E;
    d_sizes[1] = S + rank[bound];
  }
  if (q < S) {
    out_n_id[q] = idx[q];
    return;
  }
  const int64_t p = q - S;
  if (p >= E) return;
  const long long v = vals[slot_buf[p]];
  local_buf[p] = v < S ? (S - 1 - v) : (S + rank[v - S]);
  if (v == S + p) out_n_id[S + rank[p]] = gcol_buf[p];
}

// rows of <= 32 picks: one warp per row, bitonic network on (local, e)
__global__ void __launch_bounds__(256)
sa_emit_small_kernel(int64_t S, const int64_t *__restrict__ out_rowptr, const int64_t *__restrict__ local_buf,
                     const int64_t *__restrict__ e_buf, int64_t *__restrict__ out_col,
                     int64_t *__restrict__ out_e_id) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= S) return;
  const unsigned lane = lane_id();
  const int64_t base = out_rowptr[i];
  const int cnt = (int)(out_rowptr[i + 1] - base);
  int64_t a = INT64_MAX, b = INT64_MAX;
  if ((int)lane < cnt) {
    a = local_buf[base + lane];
    b = e_buf[base + lane];
  }
#pragma unroll
  for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int64_t oa = __shfl_xor_sync(0xffffffffu, a, stride);
      const int64_t ob = __shfl_xor_sync(0xffffffffu, b, stride);
      const bool up = ((lane & size) == 0);          // ascending block
      const bool lower = ((lane & stride) == 0);     // this lane keeps the smaller one when ascending
      const bool other_less = (oa < a) || (oa == a && ob < b);
      const bool other_greater = (oa > a) || (oa == a && ob > b);
      const bool take = (up == lower) ? other_less : other_greater;
      if (take) {
        a = oa;
        b = ob;
      }
    }
  }
  if ((int)lane < cnt) {
    out_col[base + lane] = a;
    out_e_id[base + lane] = b;
  }
}

// general path helpers -------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sa_fill_kernel(int64_t *p, int64_t n, int64_t v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// key2[q] = (rowid[p] << lbits) | local[p] with p = perm1[q]; padded picks carry rowid == S
__global__ void __launch_bounds__(256)
sa_key2_kernel(const uint32_t *__restrict__ perm1, int64_t bound, int64_t S, const int64_t *__restrict__ out_rowptr,
               const int64_t *__restrict__ rowid_buf, const int64_t *__restrict__ local_buf, int lbits,
               uint64_t *__restrict__ key2) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= bound) return;
  const int64_t p = perm1[q];
  const int64_t E = out_rowptr[S];
  key2[q] = p < E ? (((uint64_t)rowid_buf[p] << lbits) | (uint64_t)local_buf[p]) : ((uint64_t)S << lbits);
}

__global__ void __launch_bounds__(256)
sa_emit_perm_kernel(const uint32_t *__restrict__ perm2, int64_t S, const int64_t *__restrict__ out_rowptr,
                    const int64_t *__restrict__ local_buf, const int64_t *__restrict__ e_buf,
                    int64_t *__restrict__ out_col, int64_t *__restrict__ out_e_id) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= out_rowptr[S]) return;
  const int64_t p = perm2[q];
  out_col[q] = local_buf[p];
  out_e_id[q] = e_buf[p];
}

static inline unsigned blocks_for(int64_t n, int threads = 256) { return (unsigned)ceil_div(n > 0 ? n : 1, threads); }

}  // namespace ps

using namespace ps;

namespace ps {
// sample.py:10-28 `sample`: out[i, k] = col[rowptr[s_i] + floor(u * deg(s_i))], u uniform in [0, 1) - uniform draws WITH
// replacement, one launch where the reference chains rand / mul_ / cast / add_ / index.  u of (row position i, draw k)
// is the top 24 bits of Philox4x32-10(seed; (i, k)) scaled by 2^-24 (a float32 in [0, 1), like the framework's rand);
// floor(u * deg) is computed in float32 as the reference does and clamped to deg - 1.  A row without entries yields
// the entry at its (empty) start clamped into the array, as indexing col[rowptr[s]] does in the reference.
__global__ void __launch_bounds__(256)
sample_uniform_kernel(const int64_t *__restrict__ rowptr, const int64_t *__restrict__ col, int64_t nnz,
                      const int64_t *__restrict__ subset, int64_t S, int64_t k, uint64_t seed, int64_t *__restrict__ out) {
  const int64_t total = S * k;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = q / k, d = q - i * k;
    const int64_t r = subset ? subset[i] : i;
    const int64_t lo = rowptr[r], deg = rowptr[r + 1] - lo;
    const float u = (float)(uint32_t)(philox_u64(seed, (uint64_t)i, (uint32_t)d) >> 40) * (1.0f / 16777216.0f);
    int64_t off = (int64_t)(u * (float)deg);
    if (off >= deg) off = deg > 0 ? deg - 1 : 0;
    int64_t e = lo + off;
    if (e >= nnz) e = nnz - 1;
    out[q] = col[e];
  }
}
}  // namespace ps

extern "C" {

int ps_sample_adj_bound(const int64_t *rowptr, const int64_t *idx, int64_t S, int64_t num_neighbors, int replace,
                        int64_t *host_bound, void *stream) {
  PS_CHECK(S >= 0 && host_bound, "sample_adj_bound: bad arguments");
  (void)replace;
  if (num_neighbors >= 0 || S == 0) {
    *host_bound = S == 0 ? 0 : S * num_neighbors;
    return PS_OK;
  }
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  PS_SCRATCH(d, scratch, unsigned long long, 1);
  PS_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long), s));
  int64_t b = ceil_div(S, 256);
  if (b > 4 * (int64_t)sm_count()) b = 4 * (int64_t)sm_count();
  sa_degree_sum_kernel<<<(unsigned)b, 256, 0, s>>>(rowptr, idx, S, d);
  PS_LAUNCH_OK();
  unsigned long long h = 0;
  PS_TRY(read_back(&h, d, sizeof(h), s));
  *host_bound = (int64_t)h;
  return PS_OK;
}

int ps_sample_adj(const int64_t *rowptr, const int64_t *col, int64_t M, int64_t nnz, const int64_t *idx, int64_t S,
                  int64_t num_neighbors, int replace, uint64_t seed, int64_t bound, int64_t *out_rowptr,
                  int64_t *out_col, int64_t *out_n_id, int64_t *out_e_id, int64_t *host_sizes, void *stream) {
  PS_CHECK(S >= 0 && M >= 0 && nnz >= 0 && bound >= 0 && host_sizes, "sample_adj: bad arguments");
  PS_CHECK(S + bound < (int64_t(1) << 31), "sample_adj: batch too large (S + bound = %lld)", (long long)(S + bound));
  cudaStream_t s = (cudaStream_t)stream;
  host_sizes[0] = host_sizes[1] = 0;
  if (S == 0) {
    PS_CUDA(cudaMemsetAsync(out_rowptr, 0, sizeof(int64_t), s));
    return PS_OK;
  }
  const bool rep = replace != 0;
  const int64_t k = num_neighbors;
  Scratch scratch(s);
  const int64_t B = bound > 0 ? bound : 1;

  // ---- count -> out_rowptr
  PS_SCRATCH(cnt, scratch, int64_t, S);
  sa_count_kernel<<<blocks_for(S), 256, 0, s>>>(rowptr, idx, S, k, rep, cnt);
  PS_LAUNCH_OK();
  PS_TRY(count2ptr_launch(cnt, nullptr, S, out_rowptr, s));

  // ---- pick
  PS_SCRATCH(e_buf, scratch, int64_t, B);
  PS_SCRATCH(gcol_buf, scratch, int64_t, B);
  PS_SCRATCH(rowid_buf, scratch, int64_t, B);
  PS_SCRATCH(local_buf, scratch, int64_t, B);
  PS_SCRATCH(slot_buf, scratch, uint32_t, B);
  PS_SCRATCH(flag, scratch, int64_t, B);
  PS_SCRATCH(rank, scratch, int64_t, B + 1);
  PS_SCRATCH(d_sizes, scratch, long long, 2);
  const bool small_rows = k >= 0 && k <= 32;
  if (!small_rows) PS_CUDA(cudaMemsetAsync(e_buf, 0, sizeof(int64_t) * (size_t)B, s));
  sa_pick_kernel<<<blocks_for(S * 32), 256, 0, s>>>(rowptr, col, idx, S, k, rep, seed, out_rowptr, e_buf, gcol_buf,
                                                   rowid_buf);
  PS_LAUNCH_OK();

  // ---- relabel
  uint64_t cap = 64;
  while (cap < 2 * (uint64_t)(S + bound)) cap <<= 1;
  PS_SCRATCH(keys, scratch, long long, cap);
  PS_SCRATCH(vals, scratch, long long, cap);
  PS_CUDA(cudaMemsetAsync(keys, 0xff, sizeof(long long) * cap, s));
  PS_CUDA(cudaMemsetAsync(vals, 0x7f, sizeof(long long) * cap, s));
  sa_insert_kernel<<<blocks_for(S + bound), 256, 0, s>>>(idx, S, gcol_buf, out_rowptr, keys, vals, cap - 1, slot_buf);
  PS_LAUNCH_OK();
  sa_flag_kernel<<<blocks_for(bound), 256, 0, s>>>(S, bound, out_rowptr, vals, slot_buf, flag);
  PS_LAUNCH_OK();
  PS_TRY(count2ptr_launch(flag, nullptr, bound, rank, s));
  sa_relabel_kernel<<<blocks_for(S + bound), 256, 0, s>>>(idx, S, bound, out_rowptr, gcol_buf, vals, slot_buf, rank,
                                                         local_buf, out_n_id, d_sizes);
  PS_LAUNCH_OK();

  // ---- emit, every row sorted by (local id, e_id)
  if (bound > 0) {
    if (small_rows) {
      sa_emit_small_kernel<<<blocks_for(S * 32), 256, 0, s>>>(S, out_rowptr, local_buf, e_buf, out_col, out_e_id);
      PS_LAUNCH_OK();
    } else {
      // stable sort by e_id, then stable sort by (seed position, local id)
      PS_SCRATCH(k0, scratch, uint64_t, bound);
      PS_SCRATCH(k1, scratch, uint64_t, bound);
      PS_SCRATCH(p0, scratch, uint32_t, bound);
      PS_SCRATCH(p1, scratch, uint32_t, bound);
      PS_SCRATCH(perm1, scratch, uint32_t, bound);
      PS_SCRATCH(perm2, scratch, uint32_t, bound);
      PS_SCRATCH(key2, scratch, uint64_t, bound);
      PS_TRY(radix_sort_pairs<uint64_t>(KeySrc{e_buf, nullptr, KSRC_I64, 0}, PaySrc{nullptr, PSRC_IOTA}, bound,
                                        bits_for(nnz + 1), k0, k1, p0, p1, KeyDst{nullptr, nullptr, KDST_NONE, 0},
                                        PayDst{perm1, PDST_U32}, scratch, s));
      const int lbits = bits_for(S + bound + 1);
      sa_key2_kernel<<<blocks_for(bound), 256, 0, s>>>(perm1, bound, S, out_rowptr, rowid_buf, local_buf, lbits, key2);
      PS_LAUNCH_OK();
      PS_TRY(radix_sort_pairs<uint64_t>(KeySrc{key2, nullptr, KSRC_KEYS, 0}, PaySrc{perm1, PSRC_U32}, bound,
                                        lbits + bits_for(S + 1), k0, k1, p0, p1,
                                        KeyDst{nullptr, nullptr, KDST_NONE, 0}, PayDst{perm2, PDST_U32}, scratch, s));
      sa_emit_perm_kernel<<<blocks_for(bound), 256, 0, s>>>(perm2, S, out_rowptr, local_buf, e_buf, out_col, out_e_id);
      PS_LAUNCH_OK();
    }
  }
  long long h[2] = {0, 0};
  PS_TRY(read_back(h, d_sizes, sizeof(h), s));
  host_sizes[0] = (int64_t)h[0];
  host_sizes[1] = (int64_t)h[1];
  return PS_OK;
}

int ps_sample_uniform(const int64_t *rowptr, const int64_t *col, int64_t nnz, const int64_t *subset, int64_t S,
                      int64_t num_neighbors, uint64_t seed, int64_t *out, void *stream) {
  PS_CHECK(S >= 0 && num_neighbors >= 0 && nnz >= 0, "sample_uniform: negative size");
  if (S == 0 || num_neighbors == 0) return PS_OK;
  PS_CHECK(nnz > 0, "sample_uniform: the matrix has no entries to draw from");
  const int64_t total = S * num_neighbors;
  int64_t b = ceil_div(total, 256);
  if (b > 16 * (int64_t)sm_count()) b = 16 * (int64_t)sm_count();
  sample_uniform_kernel<<<(unsigned)b, 256, 0, (cudaStream_t)stream>>>(rowptr, col, nnz, subset, S, num_neighbors, seed, out);
  PS_LAUNCH_OK();
  return PS_OK;
}

}  // extern "C"
