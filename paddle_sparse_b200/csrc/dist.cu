// Row-range partitioning for matrices that span several GPUs.  The reference
// has no distributed code at all; this is the device half of the sample sort named by the
// north star: coarse row histogram -> (host/NCCL: all-reduce, splitters) -> stable partition
// of the local tuples by destination rank -> (NCCL all-to-all) -> local ps_coalesce.
#include "common.cuh"
#include "radix_sort.cuh"
#include "segment.cuh"

namespace ps {

constexpr int DH_SMEM_BUCKETS = 32768;  // 128 KB of 32-bit counters per CTA

// out[b] += #{e : row[e] >> shift == b}; counters privatised in shared memory
__global__ void __launch_bounds__(512)
row_hist_smem_kernel(const int64_t *__restrict__ row, int64_t nnz, int shift, int buckets,
                     unsigned long long *__restrict__ out) {
  extern __shared__ unsigned int sh_hist[];
  for (int j = threadIdx.x; j < buckets; j += blockDim.x) sh_hist[j] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 2;
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
  // four 16-byte loads in flight per thread: with 128 KB of counters only one CTA fits an SM and one load at a time
  // left the kernel at 2 TB/s (0.78 ms per 202 M rows)
  for (; i + 3 * stride + 1 < nnz; i += 4 * stride) {
    longlong2 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ld_stream2(row + i + u * stride);  // row is 16-B aligned, i even
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      atomicAdd(&sh_hist[(unsigned)(v[u].x >> shift)], 1u);
      atomicAdd(&sh_hist[(unsigned)(v[u].y >> shift)], 1u);
    }
  }
  for (; i < nnz; i += stride) {
    if (i + 1 < nnz) {
      const longlong2 v = ld_stream2(row + i);
      atomicAdd(&sh_hist[(unsigned)(v.x >> shift)], 1u);
      atomicAdd(&sh_hist[(unsigned)(v.y >> shift)], 1u);
    } else {
      atomicAdd(&sh_hist[(unsigned)(ld_stream(row + i) >> shift)], 1u);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < buckets; j += blockDim.x)
    if (sh_hist[j]) atomicAdd(&out[j], (unsigned long long)sh_hist[j]);
}

__global__ void __launch_bounds__(256)
row_hist_global_kernel(const int64_t *__restrict__ row, int64_t nnz, int shift, unsigned long long *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += stride)
    atomicAdd(&out[ld_stream(row + i) >> shift], 1ull);
}

// dest[e] = g with splitters[g] <= row[e] < splitters[g+1]; counts[g] += 1
template <typename DestT>
__global__ void __launch_bounds__(256)
dest_kernel(const int64_t *__restrict__ row, int64_t nnz, const int64_t *__restrict__ splitters, int G,
            DestT *__restrict__ dest, unsigned long long *__restrict__ counts) {
  __shared__ int64_t sp[65];
  __shared__ unsigned int sc[64];
  if (threadIdx.x <= G) sp[threadIdx.x] = splitters[threadIdx.x];
  if (threadIdx.x < 64) sc[threadIdx.x] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // four rows in flight per thread; the lanes of a warp that go to the same rank count themselves with ONE atomic
  // (with G = 2 every warp-wide atomicAdd used to serialise 16 deep on its two counters)
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < nnz; i0 += 4 * stride) {  // warp-uniform trip count
    int64_t r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride + threadIdx.x;
      r[u] = i < nnz ? ld_stream(row + i) : -1;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride + threadIdx.x;
      int lo = 0, hi = G - 1;  // largest g with sp[g] <= r
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (sp[mid] <= r[u]) lo = mid;
        else hi = mid - 1;
      }
      const int g = i < nnz ? lo : 64;  // 64: nothing to count
      if (i < nnz) dest[i] = (DestT)lo;
      const unsigned peers = __match_any_sync(0xffffffffu, g);
      if (g < 64 && (threadIdx.x & 31u) == (unsigned)(__ffs(peers) - 1)) atomicAdd(&sc[g], (unsigned)__popc(peers));
    }
  }
  __syncthreads();
  if (threadIdx.x < G && sc[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)sc[threadIdx.x]);
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_dist_row_histogram(const int64_t *row, int64_t nnz, int shift, int64_t buckets, int64_t *out_hist,
                          void *stream) {
  PS_CHECK(nnz >= 0 && buckets > 0 && shift >= 0 && shift < 63, "dist_row_histogram: bad arguments");
  PS_CHECK((reinterpret_cast<uintptr_t>(row) & 15) == 0 || nnz == 0, "dist_row_histogram: row must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  PS_CUDA(cudaMemsetAsync(out_hist, 0, sizeof(int64_t) * (size_t)buckets, s));
  if (nnz == 0) return PS_OK;
  unsigned long long *out = reinterpret_cast<unsigned long long *>(out_hist);
  if (buckets <= DH_SMEM_BUCKETS) {
    const size_t smem = sizeof(unsigned int) * (size_t)buckets;
    PS_CUDA(cudaFuncSetAttribute(row_hist_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(sizeof(unsigned int) * DH_SMEM_BUCKETS)));
    int64_t blocks = ceil_div(nnz, 512 * 2 * 8);
    const int64_t cap = (smem > 100 * 1024 ? 1 : 2) * (int64_t)sm_count();
    if (blocks > cap) blocks = cap;
    row_hist_smem_kernel<<<(unsigned)blocks, 512, smem, s>>>(row, nnz, shift, (int)buckets, out);
  } else {
    int64_t blocks = ceil_div(nnz, 256 * 8);
    const int64_t cap = 8 * (int64_t)sm_count();
    if (blocks > cap) blocks = cap;
    row_hist_global_kernel<<<(unsigned)blocks, 256, 0, s>>>(row, nnz, shift, out);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_dist_partition(const int64_t *row, const int64_t *col, const void *value, int64_t row_bytes, int64_t nnz,
                      const int64_t *splitters, int G, int64_t *out_row, int64_t *out_col, void *out_value,
                      int64_t *out_counts, void *stream) {
  PS_CHECK(nnz >= 0 && G >= 1 && G <= 64 && row_bytes >= 0, "dist_partition: bad arguments (G must be in [1, 64])");
  PS_CHECK(nnz < (int64_t(1) << 32), "dist_partition: nnz=%lld exceeds 2^32-1 entries per call", (long long)nnz);
  cudaStream_t s = (cudaStream_t)stream;
  PS_CUDA(cudaMemsetAsync(out_counts, 0, sizeof(int64_t) * (size_t)G, s));
  if (nnz == 0) return PS_OK;
  Scratch scratch(s);
  PS_SCRATCH(dest, scratch, uint32_t, nnz);
  int64_t blocks = ceil_div(nnz, 256 * 8);
  const int64_t cap = 8 * (int64_t)sm_count();
  if (blocks > cap) blocks = cap;
  dest_kernel<uint32_t><<<(unsigned)blocks, 256, 0, s>>>(row, nnz, splitters, G, dest,
                                                        reinterpret_cast<unsigned long long *>(out_counts));
  PS_LAUNCH_OK();
  const int dbits = G > 1 ? bits_for(G) : 1;
  if (row_bytes == 0 || row_bytes == 4) {
    // ONE stable counting pass on the destination id that carries the tuple itself (row, col and the 4-byte value as
    // three 32-bit payload words; indices are < 2^32 on this path) and writes the int64 outputs directly: 24 B read +
    // 20 B written per tuple, where a pass on (destination, entry index) followed by three gathers moved 68 B
    const bool has_v = value && row_bytes > 0;
    uint32_t *none[3] = {nullptr, nullptr, nullptr};
    PaySrcN<3> ps_;
    ps_.s[0] = PaySrc{row, PSRC_I64};
    ps_.s[1] = PaySrc{col, PSRC_I64};
    ps_.s[2] = has_v ? PaySrc{value, PSRC_U32} : PaySrc{nullptr, PSRC_NONE};
    PayDstN<3> pd;
    pd.d[0] = PayDst{out_row, PDST_I64};
    pd.d[1] = PayDst{out_col, PDST_I64};
    pd.d[2] = has_v ? PayDst{out_value, PDST_U32} : PayDst{nullptr, PDST_NONE};
    PS_TRY((radix_sort_tuples<uint32_t, 3>(KeySrc{dest, nullptr, KSRC_KEYS, 0}, ps_, nnz, dbits, nullptr, nullptr, none,
                                           none, KeyDst{nullptr, nullptr, KDST_NONE, 0}, pd, scratch, s)));
    return PS_OK;
  }
  // wide values: one stable counting pass on the destination id with the entry index as payload, then gathers
  PS_SCRATCH(perm, scratch, uint32_t, nnz);
  PS_TRY(radix_sort_pairs<uint32_t>(KeySrc{dest, nullptr, KSRC_KEYS, 0}, PaySrc{nullptr, PSRC_IOTA}, nnz, dbits, nullptr,
                                    nullptr, nullptr, nullptr, KeyDst{nullptr, nullptr, KDST_NONE, 0},
                                    PayDst{perm, PDST_U32}, scratch, s));
  PS_TRY(gather_rows_launch<uint32_t>(row, 8, perm, nnz, out_row, s));
  PS_TRY(gather_rows_launch<uint32_t>(col, 8, perm, nnz, out_col, s));
  if (value && row_bytes > 0) PS_TRY(gather_rows_launch<uint32_t>(value, row_bytes, perm, nnz, out_value, s));
  return PS_OK;
}

int ps_dist_dest(const int64_t *row, int64_t nnz, const int64_t *splitters, int G, uint8_t *out_dest,
                 int64_t *out_counts, void *stream) {
  PS_CHECK(nnz >= 0 && G >= 1 && G <= 64, "dist_dest: bad arguments (G must be in [1, 64])");
  cudaStream_t s = (cudaStream_t)stream;
  PS_CUDA(cudaMemsetAsync(out_counts, 0, sizeof(int64_t) * (size_t)G, s));
  if (nnz == 0) return PS_OK;
  int64_t blocks = ceil_div(nnz, 256 * 8);
  const int64_t cap = 8 * (int64_t)sm_count();
  if (blocks > cap) blocks = cap;
  dest_kernel<uint8_t><<<(unsigned)blocks, 256, 0, s>>>(row, nnz, splitters, G, out_dest,
                                                       reinterpret_cast<unsigned long long *>(out_counts));
  PS_LAUNCH_OK();
  return PS_OK;
}

int ps_dist_scatter_peer(const uint8_t *dest, const int64_t *row, const int64_t *col, const void *value,
                         int64_t row_bytes, int64_t nnz, int G, const int64_t *bases, void *const *peer_row,
                         void *const *peer_col, void *const *peer_value, int narrow, void *stream) {
  PS_CHECK(nnz >= 0 && G >= 1 && G <= 64, "dist_scatter_peer: bad arguments (G must be in [1, 64])");
  PS_CHECK(row_bytes == 0 || row_bytes == 4, "dist_scatter_peer: values must be absent or 4 bytes per entry");
  PS_CHECK(nnz < (int64_t(1) << 32), "dist_scatter_peer: nnz=%lld exceeds 2^32-1 entries per call", (long long)nnz);
  if (nnz == 0) return PS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  // the "histogram" of this single pass is given: bin d starts at bases[d] inside destination d's buffer
  PS_SCRATCH(gh, scratch, unsigned long long, RS_MAX_RADIX);
  PS_CUDA(cudaMemsetAsync(gh, 0, sizeof(unsigned long long) * RS_MAX_RADIX, s));
  PS_CUDA(cudaMemcpyAsync(gh, bases, sizeof(int64_t) * (size_t)G, cudaMemcpyDeviceToDevice, s));
  const bool has_v = value && row_bytes > 0;
  uint32_t *none[3] = {nullptr, nullptr, nullptr};
  PaySrcN<3> ps_;
  ps_.s[0] = PaySrc{row, PSRC_I64};
  ps_.s[1] = PaySrc{col, PSRC_I64};
  ps_.s[2] = has_v ? PaySrc{value, PSRC_U32} : PaySrc{nullptr, PSRC_NONE};
  PayDstN<3> pd;
  // narrow: row and col cross NVLink as the 32-bit words the pass carries anyway (12 instead of 20 bytes per tuple);
  // the receiver's coalesce reads them as such (ps_coalesce_u32)
  pd.d[0] = PayDst{const_cast<void **>(peer_row), narrow ? PDST_U32_BY_DIGIT : PDST_I64_BY_DIGIT};
  pd.d[1] = PayDst{const_cast<void **>(peer_col), narrow ? PDST_U32_BY_DIGIT : PDST_I64_BY_DIGIT};
  pd.d[2] = has_v ? PayDst{const_cast<void **>(peer_value), PDST_U32_BY_DIGIT} : PayDst{nullptr, PDST_NONE};
  PS_TRY((radix_sort_tuples<uint32_t, 3>(KeySrc{dest, nullptr, KSRC_U8, 0}, ps_, nnz, G > 1 ? bits_for(G) : 1, nullptr,
                                         nullptr, none, none, KeyDst{nullptr, nullptr, KDST_NONE, 0}, pd, scratch, s, 0,
                                         gh)));
  return PS_OK;
}

}  // extern "C"
