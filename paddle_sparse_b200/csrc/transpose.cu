// CSC structures and transpose.
//
// Reference behaviour restated (not ported):
//   storage.py:425-434  csr2csc = argsort(M*col + row)      (a second full 64-bit sort)
//   storage.py:439-447  csc2csr = argsort(csr2csc)          (a third full sort)
//   storage.py:386-420  colptr / colcount
//   transpose.py:9-33   t(): three separate gathers through csr2csc
// Here: the entries are already in CSR order, so csr2csc is a STABLE radix sort on the
// col bits alone with the entry index as payload (ties keep row order); colptr falls out
// of the sorted col keys by boundary detection (the ind2ptr kernel on 32-bit keys);
// csc2csr is one inverse-permutation scatter; t() is one fused gather.
#include "common.cuh"
#include "convert.cuh"
#include "radix_sort.cuh"
#include "segment.cuh"
#include "csc_local.cuh"

namespace ps {

int csc_variant();  // PS_CSC_HYBRID=0 disables the hybrid CSC sort (bench knob)
constexpr int INV_WINDOW_BITS = 12;
__global__ void invert_window_kernel(const uint32_t *__restrict__ target, const uint32_t *__restrict__ source, int64_t n,
                                     int64_t *__restrict__ out);
__global__ void dense_key_hist_kernel(long long n, SortPlan plan, unsigned long long *__restrict__ ghist);

// The hybrid CSC sort (csc_local.cuh).  *used = false (nothing written) when the shape rules it out; the caller then
// runs the LSD pipeline.  `t_row / t_col / t_val` null: csr2csc + colptr / colcount only.
static int csc_hybrid(const int64_t *row, const int64_t *col, const void *value, int64_t n, int64_t N,
                      int64_t *out_csr2csc, int64_t *out_colptr, int64_t *out_colcount, int64_t *t_row, int64_t *t_col,
                      void *t_val, int64_t *out_csc2csr, bool *used, cudaStream_t s) {
  *used = false;
  const int cbits = bits_for(N);
  if (!csc_variant() || n < (int64_t(1) << 16) || cbits < 2) return PS_OK;
  // The low L <= 8 col bits are sorted locally, the top T = cbits - L select a bucket of 2^L columns.  L is the largest
  // width whose AVERAGE bucket (n * 2^L / N entries: only N of the 2^cbits key values exist) still fills at most 90 % of
  // a CTA's capacity; larger buckets (skew) are cut into chunks by the kernels.
  int L = CL_MAX_L < cbits - 1 ? CL_MAX_L : cbits - 1;
  while (L > 0 && (double)n * (double)(int64_t(1) << L) / (double)(N > 0 ? N : 1) > 0.9 * CL_CAP) --L;
  const int T = cbits - L;
  if (T > CL_MAX_T || T < 1) return PS_OK;
  const SortPlan gp = make_sort_plan(T, rs_geometry<uint32_t, 3>().rbits);
  if (gp.passes > 2) return PS_OK;
  const bool carry = t_row != nullptr;
  const int nb = 1 << T;
  LsPlan plan;
  plan.passes = gp.passes;
  for (int p = 0; p < gp.passes; ++p) { plan.pbegin[p] = gp.begin[p]; plan.pbits[p] = gp.bits[p]; }

  Scratch scratch(s);
  const size_t max_chunks = (size_t)nb + (size_t)(n / CL_CAP) + 1;
  PS_SCRATCH(gcount, scratch, unsigned int, nb);
  PS_SCRATCH(bucket_off, scratch, unsigned int, (size_t)nb + 1);
  PS_SCRATCH(chunk_start, scratch, unsigned int, (size_t)nb + 1);
  PS_SCRATCH(chunk_bucket, scratch, unsigned int, max_chunks);
  PS_SCRATCH(multi_list, scratch, unsigned int, nb);
  PS_SCRATCH(ghist, scratch, unsigned long long, 2 * RS_MAX_RADIX);
  PS_SCRATCH(info, scratch, ClInfo, 1);
  PS_SCRATCH(ticket, scratch, unsigned long long, 1);
  PS_CUDA(cudaMemsetAsync(gcount, 0, sizeof(unsigned int) * (size_t)nb, s));
  PS_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned long long), s));
  {
    PS_FUNC_SMEM(cl_bucket_hist_kernel, sizeof(unsigned int) << CL_MAX_T);
    ProfScope prof(PROF_BUCKET_HIST, s);
    cl_bucket_hist_kernel<<<sm_count() * (T <= 14 ? 2 : 1), 1024, sizeof(unsigned int) * (size_t)nb, s>>>(col, n, L, (unsigned)nb,
                                                                                                         gcount);
  }
  PS_LAUNCH_OK();
  ls_derive_kernel<<<1, 1024, 0, s>>>(gcount, T, plan, bucket_off, ghist, nullptr, CL_CAP,
                                      LsChunkOut{chunk_start, chunk_bucket, multi_list, reinterpret_cast<unsigned int *>(info)});
  PS_LAUNCH_OK();

  // global passes on col bits [L, cbits): (col, entry index[, row, value]) -> u32 arrays in bucket order
  PS_SCRATCH(scol, scratch, uint32_t, n);
  PS_SCRATCH(sp, scratch, uint32_t, n);
  PS_SCRATCH(k0, scratch, uint32_t, n);
  uint32_t *srow = nullptr, *sval = nullptr;
  if (carry) {
    PS_SCRATCH(a0, scratch, uint32_t, n);
    PS_SCRATCH(a1, scratch, uint32_t, n);
    PS_SCRATCH(srow_, scratch, uint32_t, n);
    srow = srow_;
    uint32_t *t0[3] = {a0, a1, nullptr}, *t1[3] = {nullptr, nullptr, nullptr};
    if (value) {
      PS_SCRATCH(a2, scratch, uint32_t, n);
      PS_SCRATCH(sval_, scratch, uint32_t, n);
      t0[2] = a2;
      sval = sval_;
    }
    PaySrcN<3> ps_;
    ps_.s[0] = PaySrc{nullptr, PSRC_IOTA};
    ps_.s[1] = PaySrc{row, PSRC_I64};
    ps_.s[2] = value ? PaySrc{value, PSRC_U32} : PaySrc{nullptr, PSRC_NONE};
    PayDstN<3> pd;
    pd.d[0] = PayDst{sp, PDST_U32};
    pd.d[1] = PayDst{srow, PDST_U32};
    pd.d[2] = value ? PayDst{sval, PDST_U32} : PayDst{nullptr, PDST_NONE};
    PS_TRY((radix_sort_tuples<uint32_t, 3>(KeySrc{col, nullptr, KSRC_I64, 0}, ps_, n, T, k0, nullptr, t0, t1,
                                           KeyDst{scol, nullptr, KDST_KEYS, 0}, pd, scratch, s, L, ghist)));
  } else {
    PS_SCRATCH(a0, scratch, uint32_t, n);
    PS_TRY(radix_sort_pairs<uint32_t>(KeySrc{col, nullptr, KSRC_I64, 0}, PaySrc{nullptr, PSRC_IOTA}, n, T, k0, nullptr, a0,
                                      nullptr, KeyDst{scol, nullptr, KDST_KEYS, 0}, PayDst{sp, PDST_U32}, scratch, s, L,
                                      ghist));
  }

  // oversized buckets (hub columns) only: per-chunk digit offsets
  PS_SCRATCH(chunk_hist, scratch, unsigned int, max_chunks * 256);
  cl_chunk_hist_kernel<<<sm_count() * 8, 256, 0, s>>>(scol, bucket_off, chunk_start, chunk_bucket, info, (1u << L) - 1u,
                                                     chunk_hist);
  PS_LAUNCH_OK();
  cl_chunk_scan_kernel<<<sm_count() * 2, 256, 0, s>>>(bucket_off, chunk_start, multi_list, info, L, N, chunk_hist,
                                                     out_colptr, out_colcount);
  PS_LAUNCH_OK();

  // csc2csr rides along (ClInverse): direct scatter when the output stays in L2, else the first partition pass of the
  // bucketed inverse is fused into the local kernel
  ClInverse inv{0, 0, 0, nullptr, nullptr, nullptr, out_csc2csr};
  SortPlan iplan{};
  uint32_t *tp = nullptr, *tq = nullptr;
  unsigned long long *ighist = nullptr;
  if (out_csc2csr) {
    const int nbits = bits_for(n);
    if (nbits <= 20) {
      inv.mode = 1;
    } else {
      const int top = nbits - INV_WINDOW_BITS;
      iplan = make_sort_plan(top, rs_geometry<uint32_t, 1>().rbits);
      for (int i = 0; i < iplan.passes; ++i) iplan.begin[i] += INV_WINDOW_BITS;
      PS_SCRATCH(tp_, scratch, uint32_t, n);
      PS_SCRATCH(tq_, scratch, uint32_t, n);
      PS_SCRATCH(gh, scratch, unsigned long long, (size_t)iplan.passes * RS_MAX_RADIX);
      tp = tp_;
      tq = tq_;
      ighist = gh;
      dense_key_hist_kernel<<<iplan.passes, RS_MAX_RADIX, 0, s>>>((long long)n, iplan, ighist);
      PS_LAUNCH_OK();
      inv.mode = 2;
      inv.shift = iplan.begin[0];
      inv.bits = iplan.bits[0];
      inv.cursor = ighist;  // row 0: the regions' starts, consumed as fill cursors
      inv.tp = tp;
      inv.tq = tq;
    }
  }
  const size_t smem = sizeof(ClSmem);
  int grid = sm_count() * CL_MINB;
#define PS_CL(C_, V_, I_)                                                                                          \
  do {                                                                                                             \
    auto kern = cl_local_kernel<C_, V_, I_>;                                                                       \
    PS_FUNC_SMEM(kern, smem);                                                                                      \
    ProfScope prof(PROF_CSC_LOCAL, s);                                                                             \
    kern<<<grid, CL_THREADS, smem, s>>>(scol, sp, srow, sval, bucket_off, chunk_start, chunk_bucket, chunk_hist,  \
                                       info, nb, L, N, out_csr2csc, out_colptr, out_colcount, t_row, t_col,      \
                                       reinterpret_cast<uint32_t *>(t_val), ticket, inv);                         \
  } while (0)
#define PS_CLI(C_, V_)                         \
  do {                                         \
    if (inv.mode == 2) PS_CL(C_, V_, 2);       \
    else if (inv.mode == 1) PS_CL(C_, V_, 1);  \
    else PS_CL(C_, V_, 0);                     \
  } while (0)
  if (carry && value) PS_CLI(true, true);
  else if (carry) PS_CLI(true, false);
  else PS_CLI(false, false);
#undef PS_CLI
#undef PS_CL
  PS_LAUNCH_OK();
  if (inv.mode == 2) {  // the rest of the bucketed inverse: the second partition pass (if any), then the windows
    const uint32_t *wk = tp, *wv = tq;
    if (iplan.passes > 1) {
      PS_SCRATCH(k1, scratch, uint32_t, n);
      PS_SCRATCH(p1, scratch, uint32_t, n);
      const RsGeomRt geom = rs_geometry<uint32_t, 1>();
      const int64_t tiles = ceil_div(n, geom.tile);
      const size_t words = 32 + (size_t)tiles * (size_t)(1 << geom.rbits);
      PS_SCRATCH(st, scratch, uint64_t, words);
      PS_CUDA(cudaMemsetAsync(st, 0, words * sizeof(uint64_t), s));
      PaySrcN<1> ps_;
      ps_.s[0] = PaySrc{tq, PSRC_U32};
      PayDstN<1> pd;
      pd.d[0] = PayDst{p1, PDST_U32};
      PS_TRY((rs_dispatch_pass<uint32_t, 1>(geom.rbits, KeySrc{tp, nullptr, KSRC_KEYS, 0}, ps_, KeyDst{k1, nullptr, KDST_KEYS, 0},
                                            pd, n, iplan.begin[1], iplan.bits[1], 1, ighist + RS_MAX_RADIX, st + 32,
                                            reinterpret_cast<unsigned long long *>(st), s)));
      wk = k1;
      wv = p1;
    }
    {
      ProfScope prof(PROF_INVERT_PERM, s);
      invert_window_kernel<<<(unsigned)(sm_count() * 4), 512, 0, s>>>(wk, wv, n, out_csc2csr);
    }
    PS_LAUNCH_OK();
  }
  *used = true;
  return PS_OK;
}

__global__ void __launch_bounds__(256)
invert_perm_kernel(const int64_t *__restrict__ perm, int64_t n, int64_t *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[ld_stream(perm + i)] = i;
}

// Second half of the bucketed inverse.  `perm` is a permutation, so after the (target, source) pairs have been
// partitioned on target >> INV_WINDOW_BITS the pairs of window w sit exactly at [w << BITS, (w + 1) << BITS): one CTA
// drops the sources into a shared-memory table indexed by the low target bits and writes the window out coalesced.
__global__ void __launch_bounds__(512)
invert_window_kernel(const uint32_t *__restrict__ target, const uint32_t *__restrict__ source, int64_t n,
                     int64_t *__restrict__ out) {
  constexpr int WIN = 1 << INV_WINDOW_BITS;
  __shared__ uint32_t table[WIN];
  for (int64_t w = blockIdx.x; (w << INV_WINDOW_BITS) < n; w += gridDim.x) {
    const int64_t base = w << INV_WINDOW_BITS;
    const int len = (int)((n - base) < WIN ? (n - base) : WIN);
    uint32_t t[WIN / 512], v[WIN / 512];
#pragma unroll
    for (int k = 0; k < WIN / 512; ++k) {
      const int i = k * 512 + threadIdx.x;
      t[k] = i < len ? ld_stream(target + base + i) : 0u;
      v[k] = i < len ? ld_stream(source + base + i) : 0u;
    }
#pragma unroll
    for (int k = 0; k < WIN / 512; ++k)
      if (k * 512 + (int)threadIdx.x < len) table[t[k] & (WIN - 1)] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < WIN / 512; ++k) {
      const int i = k * 512 + threadIdx.x;
      if (i < len) out[base + i] = (int64_t)table[i];
    }
    __syncthreads();
  }
}

// Bin bases of the partition passes without reading anything: the keys are the integers 0 .. n-1, each exactly once,
// so digit d of bits [lo, lo + bits) occurs count(d) = #{x < n : (x >> lo) & (2^bits - 1) == d} times.
__global__ void __launch_bounds__(RS_MAX_RADIX)
dense_key_hist_kernel(long long n, SortPlan plan, unsigned long long *__restrict__ ghist /*[passes][512] scanned*/) {
  __shared__ unsigned long long warp_sums[33];
  const int p = blockIdx.x, d = threadIdx.x;
  const int lo = plan.begin[p], bits = plan.bits[p];
  unsigned long long c = 0;
  if (d < (1 << bits)) {
    const unsigned long long period = 1ull << (lo + bits), block = 1ull << lo;
    const unsigned long long full = (unsigned long long)n / period, rem = (unsigned long long)n % period;
    const unsigned long long start = (unsigned long long)d * block;
    c = full * block + (rem > start ? (rem - start < block ? rem - start : block) : 0ull);
  }
  unsigned long long total;
  ghist[(size_t)p * RS_MAX_RADIX + d] = block_exclusive_scan<unsigned long long, RS_MAX_RADIX>(c, warp_sums, &total);
}

// out[perm[i]] = i.  Large inputs: a stable radix partition of (perm[i], i) on the high bits of perm[i] (two
// one-sweep passes), then the window kernel above; small inputs (the output fits L2 comfortably): direct scatter.
static int invert_perm_impl(const int64_t *perm, int64_t n, int64_t *out, cudaStream_t s) {
  const int nbits = bits_for(n);
  if (nbits <= 20 || n >= (int64_t(1) << 32)) {
    ProfScope prof(PROF_INVERT_PERM, s);
    invert_perm_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(perm, n, out);
    PS_LAUNCH_OK();
    return PS_OK;
  }
  Scratch scratch(s);
  PS_SCRATCH(k0, scratch, uint32_t, n);
  PS_SCRATCH(k1, scratch, uint32_t, n);
  PS_SCRATCH(p0, scratch, uint32_t, n);
  PS_SCRATCH(p1, scratch, uint32_t, n);
  const int top = nbits - INV_WINDOW_BITS;
  const int passes = rs_num_passes<uint32_t>(top);
  uint32_t *tk = (passes & 1) ? k0 : k1, *tp = (passes & 1) ? p0 : p1;
  SortPlan plan = make_sort_plan(top, rs_geometry<uint32_t, 1>().rbits);
  for (int i = 0; i < plan.passes; ++i) plan.begin[i] += INV_WINDOW_BITS;
  PS_SCRATCH(ghist, scratch, unsigned long long, (size_t)plan.passes * RS_MAX_RADIX);
  dense_key_hist_kernel<<<plan.passes, RS_MAX_RADIX, 0, s>>>((long long)n, plan, ghist);
  PS_LAUNCH_OK();
  PS_TRY(radix_sort_pairs<uint32_t>(KeySrc{perm, nullptr, KSRC_I64, 0}, PaySrc{nullptr, PSRC_IOTA}, n, top, k0, k1, p0, p1,
                                    KeyDst{tk, nullptr, KDST_KEYS, 0}, PayDst{tp, PDST_U32}, scratch, s, INV_WINDOW_BITS,
                                    ghist));
  {
    ProfScope prof(PROF_INVERT_PERM, s);
    invert_window_kernel<<<(unsigned)(sm_count() * 4), 512, 0, s>>>(tk, tp, n, out);
  }
  PS_LAUNCH_OK();
  return PS_OK;
}

// out_col[j] = row[p], out_row[j] = col[p] (unless streamed from colptr), out_value[j] = value[p]
template <typename U, bool WITH_ROW, bool WITH_VALUE>
__global__ void __launch_bounds__(256)
transpose_gather_kernel(const int64_t *__restrict__ csr2csc, const int64_t *__restrict__ row,
                        const int64_t *__restrict__ col, const U *__restrict__ value, int64_t nnz,
                        int64_t *__restrict__ out_row, int64_t *__restrict__ out_col, U *__restrict__ out_value) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nnz) return;
  const int64_t p = ld_stream(csr2csc + j);
  out_col[j] = row[p];
  if (WITH_ROW) out_row[j] = col[p];
  if (WITH_VALUE) out_value[j] = value[p];
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_csr2csc(const int64_t *col, int64_t nnz, int64_t N, int64_t *out_csr2csc, int64_t *out_colptr,
               int64_t *out_colcount, void *stream) {
  PS_CHECK(nnz >= 0 && N >= 0, "csr2csc: negative size");
  PS_CHECK(nnz < (int64_t(1) << 32), "csr2csc: nnz=%lld exceeds the 2^32-1 entries per call this build supports",
           (long long)nnz);
  PS_CHECK(N <= (int64_t(1) << 32), "csr2csc: N=%lld exceeds the 2^32 columns this build supports", (long long)N);
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  const bool want_ptr = out_colptr || out_colcount;
  int64_t *colptr = out_colptr;
  if (want_ptr && !colptr) {
    PS_SCRATCH(tmp, scratch, int64_t, N + 1);
    colptr = tmp;
  }
  if (nnz == 0) {
    if (want_ptr) PS_CUDA(cudaMemsetAsync(colptr, 0, sizeof(int64_t) * (size_t)(N + 1), s));
    if (out_colcount && N > 0) PS_CUDA(cudaMemsetAsync(out_colcount, 0, sizeof(int64_t) * (size_t)N, s));
    return PS_OK;
  }
  {
    bool used = false;
    PS_TRY(csc_hybrid(nullptr, col, nullptr, nnz, N, out_csr2csc, colptr, out_colcount, nullptr, nullptr, nullptr, nullptr, &used,
                      s));
    if (used) return PS_OK;
  }
  const int cbits = bits_for(N);
  PS_SCRATCH(k0, scratch, uint32_t, nnz);
  PS_SCRATCH(k1, scratch, uint32_t, nnz);
  PS_SCRATCH(p0, scratch, uint32_t, nnz);
  PS_SCRATCH(p1, scratch, uint32_t, nnz);
  const int passes = rs_num_passes<uint32_t>(cbits);
  uint32_t *sorted_col = (passes & 1) ? k0 : k1;
  KeyDst kd = want_ptr ? KeyDst{sorted_col, nullptr, KDST_KEYS, 0} : KeyDst{nullptr, nullptr, KDST_NONE, 0};
  PS_TRY(radix_sort_pairs<uint32_t>(KeySrc{col, nullptr, KSRC_I64, 0}, PaySrc{nullptr, PSRC_IOTA}, nnz, cbits, k0, k1,
                                    p0, p1, kd, PayDst{out_csr2csc, PDST_I64}, scratch, s));
  if (want_ptr) PS_TRY(ind2ptr_launch<uint32_t>(sorted_col, nnz, N, colptr, s));
  if (out_colcount) PS_TRY(ps_ptr2count(colptr, N, out_colcount, stream));
  return PS_OK;
}

int ps_invert_perm(const int64_t *perm, int64_t n, int64_t *out, void *stream) {
  PS_CHECK(n >= 0, "invert_perm: negative size");
  if (n == 0) return PS_OK;
  return invert_perm_impl(perm, n, out, (cudaStream_t)stream);
}

int ps_csc_build(const int64_t *row, const int64_t *col, const void *value, int64_t value_bytes, int64_t nnz, int64_t N,
                 int64_t *out_csr2csc, int64_t *out_colptr, int64_t *out_colcount, int64_t *out_csc2csr,
                 int64_t *out_t_row, int64_t *out_t_col, void *out_t_value, void *stream) {
  PS_CHECK(nnz >= 0 && N >= 0, "csc_build: negative size");
  PS_CHECK(nnz < (int64_t(1) << 32), "csc_build: nnz=%lld exceeds the 2^32-1 entries per call this build supports",
           (long long)nnz);
  PS_CHECK(N <= (int64_t(1) << 32), "csc_build: N=%lld exceeds the 2^32 columns this build supports", (long long)N);
  cudaStream_t s = (cudaStream_t)stream;
  if (!out_t_row) {  // nothing rides along (row, value may be NULL): csr2csc, colptr, colcount, csc2csr only
    PS_CHECK(col && out_csr2csc && !out_t_col && !out_t_value, "csc_build: NULL argument");
    if (nnz > 0) {
      bool used = false;
      PS_TRY(csc_hybrid(nullptr, col, nullptr, nnz, N, out_csr2csc, out_colptr, out_colcount, nullptr, nullptr, nullptr,
                        out_csc2csr, &used, s));
      if (used) return PS_OK;
    }
    PS_TRY(ps_csr2csc(col, nnz, N, out_csr2csc, out_colptr, out_colcount, stream));
    if (out_csc2csr && nnz > 0) PS_TRY(invert_perm_impl(out_csr2csc, nnz, out_csc2csr, s));
    return PS_OK;
  }
  PS_CHECK(row && col && out_csr2csc && out_t_row && out_t_col, "csc_build: NULL argument");
  PS_CHECK(!value || (value_bytes == 4 && out_t_value), "csc_build: carried values are one 4-byte element per entry");
  Scratch scratch(s);
  const bool want_ptr = out_colptr || out_colcount;
  int64_t *colptr = out_colptr;
  if (want_ptr && !colptr) {
    PS_SCRATCH(tmp, scratch, int64_t, N + 1);
    colptr = tmp;
  }
  if (nnz == 0) {
    if (want_ptr) PS_CUDA(cudaMemsetAsync(colptr, 0, sizeof(int64_t) * (size_t)(N + 1), s));
    if (out_colcount && N > 0) PS_CUDA(cudaMemsetAsync(out_colcount, 0, sizeof(int64_t) * (size_t)N, s));
    return PS_OK;
  }
  {  // hybrid: two global passes on the top col bits, one local pass that writes every output
    bool used = false;
    PS_TRY(csc_hybrid(row, col, value, nnz, N, out_csr2csc, colptr, out_colcount, out_t_row, out_t_col, out_t_value,
                      out_csc2csr, &used, s));
    if (used) return PS_OK;
  }
  // stable sort by col of (col | entry index, row, value bits): the last pass writes the transposed arrays directly
  const int cbits = bits_for(N);
  PS_SCRATCH(k0, scratch, uint32_t, nnz);
  PS_SCRATCH(k1, scratch, uint32_t, nnz);
  uint32_t *t0[3], *t1[3];
  for (int w = 0; w < 3; ++w) {
    if (w == 2 && !value) { t0[w] = t1[w] = nullptr; continue; }
    PS_SCRATCH(a, scratch, uint32_t, nnz);
    PS_SCRATCH(b, scratch, uint32_t, nnz);
    t0[w] = a;
    t1[w] = b;
  }
  PaySrcN<3> ps_;
  ps_.s[0] = PaySrc{nullptr, PSRC_IOTA};
  ps_.s[1] = PaySrc{row, PSRC_I64};
  ps_.s[2] = value ? PaySrc{value, PSRC_U32} : PaySrc{nullptr, PSRC_NONE};
  PayDstN<3> pd;
  pd.d[0] = PayDst{out_csr2csc, PDST_I64};
  pd.d[1] = PayDst{out_t_col, PDST_I64};
  pd.d[2] = value ? PayDst{out_t_value, PDST_U32} : PayDst{nullptr, PDST_NONE};
  PS_TRY((radix_sort_tuples<uint32_t, 3>(KeySrc{col, nullptr, KSRC_I64, 0}, ps_, nnz, cbits, k0, k1, t0, t1,
                                         KeyDst{out_t_row, nullptr, KDST_I64, 0}, pd, scratch, s)));
  if (want_ptr) PS_TRY(ind2ptr_launch<int64_t>(out_t_row, nnz, N, colptr, s));
  if (out_colcount) PS_TRY(ps_ptr2count(colptr, N, out_colcount, stream));
  if (out_csc2csr) PS_TRY(invert_perm_impl(out_csr2csc, nnz, out_csc2csr, s));
  return PS_OK;
}

int ps_transpose_gather(const int64_t *csr2csc, const int64_t *row, const int64_t *col, const void *value,
                        int64_t row_bytes, int64_t nnz, const int64_t *colptr, int64_t N, int64_t *out_row,
                        int64_t *out_col, void *out_value, void *stream) {
  PS_CHECK(nnz >= 0 && N >= 0 && row_bytes >= 0, "transpose_gather: negative size");
  if (nnz == 0) return PS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const bool with_row = colptr == nullptr;
  if (!with_row) PS_TRY(ps_ptr2ind(colptr, N, nnz, out_row, stream));  // col[csr2csc] is just the sorted col
  const unsigned blocks = (unsigned)ceil_div(nnz, 256);
  int unit = value ? copy_unit(row_bytes, value, out_value) : 0;
  const bool fused_value = value && unit == row_bytes;  // one 1/2/4/8/16-byte unit per entry
#define PS_T(U, WR, WV)                                                                                     \
  transpose_gather_kernel<U, WR, WV><<<blocks, 256, 0, s>>>(csr2csc, row, col, (const U *)value, nnz, out_row, \
                                                           out_col, (U *)out_value)
#define PS_TV(U)                     \
  do {                               \
    if (with_row) PS_T(U, true, true); \
    else PS_T(U, false, true);       \
  } while (0)
  ProfScope *prof = new ProfScope(PROF_TRANSPOSE_GATHER, s);
  if (fused_value) {
    switch (unit) {
      case 16: PS_TV(uint4); break;
      case 8: PS_TV(uint64_t); break;
      case 4: PS_TV(uint32_t); break;
      case 2: PS_TV(uint16_t); break;
      default: PS_TV(uint8_t); break;
    }
  } else {
    if (with_row) PS_T(uint32_t, true, false);
    else PS_T(uint32_t, false, false);
  }
#undef PS_TV
#undef PS_T
  delete prof;
  PS_LAUNCH_OK();
  if (value && !fused_value) PS_TRY(gather_rows_launch<int64_t>(value, row_bytes, csr2csc, nnz, out_value, s));
  return PS_OK;
}

}  // extern "C"
