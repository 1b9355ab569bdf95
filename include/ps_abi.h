/* paddle_sparse_b200 — C ABI of libpaddle_sparse_b200.so
 *
 * The drop-in boundary for the SparseStorage hot path of PFCCLab/paddle_sparse
 * (SURVEY.md section 8b).  Every entry point takes plain device pointers, sizes and
 * a CUDA stream (passed as void* so the header needs no CUDA include) and returns an
 * int status (PS_OK = 0).  On failure ps_last_error() holds a thread-local message.
 * The library is GPU-only: there is no CPU fallback, and no torch / paddle type
 * appears in any signature.
 *
 * Conventions
 *   - index arrays are contiguous int64 on the device (the reference asserts int64
 *     everywhere: paddle_sparse/storage.py:61,94,101);
 *   - `value` arrays are contiguous [n, width] of `dtype` (PS_* below); value == NULL
 *     means "no value tensor" (SparseStorage._value is None);
 *   - fixed-size outputs are caller-allocated; data-dependent outputs are
 *     caller-allocated at their upper bound and the true count is returned through a
 *     HOST pointer (the call then synchronises the stream exactly once);
 *   - inputs are never modified unless the function name ends in `_inplace`;
 *   - kernels are enqueued on `stream`; scratch memory is stream-ordered
 *     (cudaMallocAsync) so no call synchronises unless stated.
 *
 * Citations are file:line under the reference repository (PFCCLab/paddle_sparse).
 */
#ifndef PS_ABI_H
#define PS_ABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PS_OK 0
#define PS_ERR_ARG 1  /* bad argument (replaces PD_CHECK, csrc/cpu/utils.h:6-9) */
#define PS_ERR_CUDA 2 /* CUDA runtime error */

/* value dtypes: paddle_sparse/testing.py:12-17 */
enum { PS_F32 = 0, PS_BF16 = 1, PS_F16 = 2, PS_F64 = 3, PS_I32 = 4, PS_I64 = 5, PS_U8 = 6 };
/* reduce ops: paddle_sparse/testing.py:10 ("add" is a synonym of "sum") */
enum { PS_SUM = 0, PS_MEAN = 1, PS_MIN = 2, PS_MAX = 3 };

/* ------------------------------------------------------------------ library */
/* csrc/version.cpp:8-33 `sparse_cuda_version` -> CUDA_VERSION the library was built with */
int64_t ps_cuda_version(void);
const char *ps_last_error(void);
/* number of CUDA kernels this library has launched so far in the process (bench.py: gpu_launches) */
int64_t ps_launch_count(void);
/* Live per-kernel timing for bench.py's roofline: when enabled, CUDA events are recorded on the
 * launching stream around every launch of the main kernel families; ps_prof_read returns the
 * summed duration and launch count of family `id` since the last read. */
void ps_prof_enable(int on);
int ps_prof_kinds(void);
const char *ps_prof_name(int id);
int ps_prof_read(int id, double *total_ms, int64_t *launches);

/* The shim-owned device allocator of the DLPack boundary (paddle_sparse_b200/paddle_sparse_ops.py): the reference's
 * ops allocate their outputs inside the op with paddle::empty / paddle::zeros (csrc/cpu/convert_cpu.cpp:10,13,
 * csrc/cuda/convert_cuda.cu:27,30, csrc/cpu/sample_cpu.cpp:25-26,123-130); a caller that brings no allocator of its
 * own gets stream-ordered blocks (cudaMallocAsync / cudaFreeAsync) owned by the DLPack capsule that wraps them.
 * ps_memcpy: kind 1 = host->device, 2 = device->host (both synchronise), 3 = device->device (asynchronous). */
int ps_alloc(int64_t bytes, void **out_ptr, void *stream);
int ps_free(void *ptr, void *stream);
int ps_memcpy(void *dst, const void *src, int64_t bytes, int kind, void *stream);
/* cudaStreamSynchronize(stream): a DLPack consumer on another stream is ordered behind the producing stream */
int ps_stream_sync(void *stream);
/* current CUDA device of the calling thread (-1 on error) / make `dev` current: the boundary launches on the device
 * its input tensors live on (CHECK_CUDA + the tensor's place, csrc/cuda/convert_cuda.cu:24,34) */
int ps_current_device(void);
int ps_set_device(int dev);

/* Self-test of the hardware behaviour the stable ranking rests on (shared-memory atomics of one warp instruction that
 * hit the same word are served in ascending lane order; csrc/selftest.cu): host_out[0] = lane-instances checked,
 * [1] = of which in collisions, [2] = served OUT of lane order (must be 0: a device that fails must not run the
 * sorts, they would be unstable).  Synchronises.  ps_rank_mode(): which ranking the sorts use; always 0 = one atomic
 * per item (a match-based mode that would not depend on the lane order is NOT built). */
int ps_selftest_atom_order(int64_t *host_out, void *stream);
int ps_rank_mode(void);

/* ------------------------------------------------------------------ convert */
/* csrc/convert.cpp:13-23 `ind2ptr(ind, M)`; kernels csrc/cuda/convert_cuda.cu:6-40,
 * csrc/cpu/convert_cpu.cpp:6-30.  ind sorted ascending, out_ptr[M+1]. */
int ps_ind2ptr(const int64_t *ind, int64_t numel, int64_t M, int64_t *out_ptr, void *stream);
/* csrc/convert.cpp:46-56 `ptr2ind(ptr, E)`; csrc/cuda/convert_cuda.cu:42-68. out_ind[E]. */
int ps_ptr2ind(const int64_t *ptr, int64_t M, int64_t E, int64_t *out_ind, void *stream);
/* ps_ptr2ind for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO counterpart in the reference.  Same arguments
 * and the same output as ps_ptr2ind; the per-tile table that call takes from the stream-ordered pool lives here in
 * a fixed array inside the library's device image (PS_PTR2IND_NOALLOC_SCRATCH_BYTES on every GPU that loads the
 * library, held until the process exits), so the call performs no cudaMalloc / cudaFree / pool request of any kind.
 * E > PS_PTR2IND_NOALLOC_MAX_E is rejected with PS_ERR_ARG (use ps_ptr2ind).  All calls on one device share the
 * array: calls that overlap on the GPU (different streams without an ordering between them) race; calls on one
 * stream, or ordered by events, are safe.  Nothing in the array has to start at zero - every word a call reads it
 * has written itself. */
#define PS_PTR2IND_NOALLOC_MAX_E 2147483648LL            /* 2^20 tiles of 2,048 entries */
#define PS_PTR2IND_NOALLOC_SCRATCH_BYTES (4194304 + 256) /* one int32 per tile + 1, padded */
int ps_ptr2ind_noalloc(const int64_t *ptr, int64_t M, int64_t E, int64_t *out_ind, void *stream);
/* storage.py:373-381 / 405-412: count[i] = ptr[i+1] - ptr[i]. */
int ps_ptr2count(const int64_t *ptr, int64_t M, int64_t *out_count, void *stream);
/* storage.py:395-398 (colptr = [0, cumsum(colcount)]) and index_select.py:21-23
 * (rowptr' from rowcount[idx]): out_ptr[0]=0, out_ptr[i+1]=sum_{j<=i} count[idx ? idx[j] : j].
 * host_total (nullable) receives out_ptr[n] (synchronises when non-NULL). */
int ps_count2ptr(const int64_t *count, const int64_t *idx, int64_t n, int64_t *out_ptr,
                 int64_t *host_total, void *stream);
/* storage.py:413-418 colcount = scatter_add(ones_like(col), col, dim_size=N). */
int ps_histogram(const int64_t *ind, int64_t numel, int64_t bins, int64_t *out_count, void *stream);

/* ------------------------------------------------------------------ construct */
/* storage.py:65-89 `int(row.max())`, `int(col.max())`: host_max[0] = max(a) (or -1 if
 * na == 0), host_max[1] = max(b) (b nullable).  Synchronises. */
int ps_index_max(const int64_t *a, int64_t na, const int64_t *b, int64_t nb, int64_t *host_max,
                 void *stream);
/* storage.py:158-163 sortedness test and storage.py:449-452 `is_coalesced`:
 * host_flags[0] = 1 iff (row,col) is non-decreasing, host_flags[1] = 1 iff strictly
 * increasing.  Synchronises. */
int ps_coo_order(const int64_t *row, const int64_t *col, int64_t nnz, int32_t *host_flags,
                 void *stream);
/* storage.py:158-171 sort-on-construct (index_sort: utils.py:14-23): STABLE sort of the
 * entries by key row*N+col.  out_row/out_col[nnz]; out_value (nullable iff value NULL)
 * [nnz,width]; out_perm (nullable) receives the permutation (sorted[i] = in[perm[i]]). */
int ps_sort_coo(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N,
                const void *value, int dtype, int64_t width, int64_t *out_row, int64_t *out_col,
                void *out_value, int64_t *out_perm, void *stream);
/* ps_coalesce on tuples held as 32-bit words (what ps_dist_scatter_peer delivers with narrow = 1): entry e is
 * (row[e] - row_offset, col[e]); same outputs, order and reduction as ps_coalesce on the widened arrays.
 * row_offset + M <= 2^32, N <= 2^32.  SURVEY.md 8e (no counterpart in the reference: it has no sharded build). */
int ps_coalesce_u32(const uint32_t *row, const uint32_t *col, int64_t row_offset, int64_t nnz, int64_t M,
                    int64_t N, const void *value, int dtype, int64_t width, int reduce, int64_t *out_row,
                    int64_t *out_col, void *out_value, int64_t *out_rowptr, int64_t *host_nnz_out,
                    void *stream);
/* storage.py:454-486 `SparseStorage.coalesce(reduce)` on SORTED input: keep the first
 * (row,col) of every run, value' = segment_csr(value, run_starts, reduce).
 * Outputs sized nnz; *host_nnz_out = nnz'.  Synchronises once.  M, N = the sparse sizes (every
 * row < M, col < N): at most 2^32 each, PS_ERR_ARG beyond (entries are compared as row << 32 | col). */
int ps_coalesce_sorted(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N,
                       const void *value, int dtype, int64_t width, int reduce, int64_t *out_row,
                       int64_t *out_col, void *out_value, int64_t *host_nnz_out, void *stream);
/* storage.py:65-89 (validation maxima) fused with the first read the coalesce needs: ONE pass over
 * (row, col) yields max(row), max(col) and - when the hybrid plan applies to (nnz, M, N) - the bucket
 * histogram, offsets and pass bases, left in `state` (device, ps_coo_inspect_bytes() bytes, caller-owned).
 * host_out[3] = {max row, max col, largest bucket or -1 when nothing was left in `state`}.  Synchronises
 * once.  Hand `state` and host_out[2] to ps_coalesce on the SAME, unmodified arrays: it then skips its
 * own histogram read and one synchronisation.  state may be NULL (validation only). */
int64_t ps_coo_inspect_bytes(void);
int ps_coo_inspect(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N,
                   void *state, int64_t *host_out, void *stream);
/* coalesce.py:8-29 free function = sort-on-construct + coalesce, fused: one radix sort
 * of the packed keys followed by run detection + segmented reduce.  out_rowptr (nullable, [M+1]): the
 * rowptr of the coalesced result (storage.py:211-222 = ind2ptr(row, M)), written by the same pass.
 * Synchronises once. */
int ps_coalesce(const int64_t *row, const int64_t *col, int64_t nnz, int64_t M, int64_t N,
                const void *value, int dtype, int64_t width, int reduce, int64_t *out_row,
                int64_t *out_col, void *out_value, int64_t *out_rowptr, int64_t *host_nnz_out,
                const void *inspect /* nullable, see ps_coo_inspect */, int64_t inspect_max_bucket,
                void *stream);

/* ------------------------------------------------------------------ CSC structures */
/* storage.py:425-434 `csr2csc` = argsort(M*col + row); on CSR-ordered input this is a
 * STABLE sort by col alone.  out_csr2csc[nnz]; out_colptr (nullable) [N+1] =
 * storage.py:386-400; out_colcount (nullable) [N] = storage.py:405-420. */
int ps_csr2csc(const int64_t *col, int64_t nnz, int64_t N, int64_t *out_csr2csc,
               int64_t *out_colptr, int64_t *out_colcount, void *stream);
/* The whole CSC side in one call (storage.py:386-447 + transpose.py:9-33): the stable sort by col carries
 * (entry index, row, value bits) as payload, so its last pass writes csr2csc AND the transposed arrays
 * out_t_row = col[csr2csc] (the sorted col), out_t_col = row[csr2csc], out_t_value = value[csr2csc] directly:
 * no random gathers.  value (nullable) must be one 4-byte element per entry (value_bytes == 4); out_colptr,
 * out_colcount, out_csc2csr (inverse permutation; rides along with the sort's bucket kernel) are nullable.
 * out_t_row == NULL (then out_t_col, out_t_value, row and value may be NULL too): nothing is carried - only
 * csr2csc, colptr, colcount and csc2csr are built (row shards that will not be transposed). */
int ps_csc_build(const int64_t *row, const int64_t *col, const void *value, int64_t value_bytes,
                 int64_t nnz, int64_t N, int64_t *out_csr2csc, int64_t *out_colptr,
                 int64_t *out_colcount, int64_t *out_csc2csr, int64_t *out_t_row, int64_t *out_t_col,
                 void *out_t_value, void *stream);
/* storage.py:439-447 `csc2csr` = argsort(csr2csc) = inverse permutation. */
int ps_invert_perm(const int64_t *perm, int64_t n, int64_t *out, void *stream);
/* x[perm] for [n_src, row_bytes]-byte rows: storage.py:166-169,238-241, transpose.py:15,
 * index_select.py:104-108.  out[n_out, row_bytes]. */
int ps_gather_rows(const void *src, int64_t row_bytes, const int64_t *index, int64_t n_out,
                   void *out, void *stream);
/* transpose.py:9-33 `t()`: out_row = col[csr2csc] (streamed from colptr when given),
 * out_col = row[csr2csc], out_value = value[csr2csc] in one pass. */
int ps_transpose_gather(const int64_t *csr2csc, const int64_t *row, const int64_t *col,
                        const void *value, int64_t row_bytes, int64_t nnz, const int64_t *colptr,
                        int64_t N, int64_t *out_row, int64_t *out_col, void *out_value,
                        void *stream);

/* ------------------------------------------------------------------ paddle_scatter replacements */
/* paddle_scatter.segment_csr(src, indptr, None, reduce): storage.py:471, reduce.py:51,
 * tensor.py:437.  src [E, width], ptr[S+1], out [S, width]; empty segment -> 0. */
int ps_segment_reduce(const void *src, int dtype, int64_t width, int64_t E, const int64_t *ptr,
                      int64_t S, int reduce, void *out, void *stream);
/* paddle_scatter.gather_csr(src, indptr): mul.py:16,79, index_select.py:29,65.
 * src [S, row_bytes], out [E, row_bytes], out[e] = src[s] for e in [ptr[s], ptr[s+1]). */
int ps_gather_csr(const void *src, int64_t row_bytes, const int64_t *ptr, int64_t S, int64_t E,
                  void *out, void *stream);
/* paddle_scatter.scatter(src, index, 0, None, dim_size, reduce): reduce.py:42.
 * src [n, width], out [dim_size, width]; untouched rows -> 0.
 * count (nullable): count[i] = number of entries with index == i when the caller already holds it
 * (the cached colcount of storage.py:405-420); mean / min / max then skip their counting atomics. */
int ps_scatter_reduce(const void *src, int dtype, int64_t width, const int64_t *index, int64_t n,
                      int64_t dim_size, int reduce, void *out, const int64_t *count, void *stream);
/* reduce.py:17-27 dim=None: out[1] = reduce over all n elements of src. */
int ps_reduce_all(const void *src, int dtype, int64_t n, int reduce, void *out, void *stream);

/* ------------------------------------------------------------------ broadcast mul */
/* mul.py:76-93 row-wise branch: out[e,:] = value[e,:] * other[r,:] for e in row r
 * (gather_csr fused with the multiply).  value/out [nnz,width] (may alias), other [M,width]. */
int ps_mul_row(const void *value, int dtype, int64_t width, const int64_t *rowptr, int64_t M,
               int64_t nnz, const void *other, void *out, void *stream);
/* mul.py:81-82 col-wise branch: out[e,:] = value[e,:] * other[col[e],:]. */
int ps_mul_col(const void *value, int dtype, int64_t width, const int64_t *col, int64_t nnz,
               const void *other, void *out, void *stream);
/* mul.py:96-119 `mul_nnz`: out = value * other, both [n]. */
int ps_mul_nnz(const void *value, int dtype, int64_t n, const void *other, void *out, void *stream);
/* add.py:12-28, 53-69 (dense [M,1] / [1,N] operand) and add.py:72-81 `add_nnz`: the three
 * multiplies above with + in place of *; same arguments. */
int ps_add_row(const void *value, int dtype, int64_t width, const int64_t *rowptr, int64_t M,
               int64_t nnz, const void *other, void *out, void *stream);
int ps_add_col(const void *value, int dtype, int64_t width, const int64_t *col, int64_t nnz,
               const void *other, void *out, void *stream);
int ps_add_nnz(const void *value, int dtype, int64_t n, const void *other, void *out, void *stream);

/* ------------------------------------------------------------------ set operations */
/* mul.py:35-73 sparse (.) sparse, the matching step: for every entry e of A (row[e], col[e];
 * n of them) the position of the equal entry in the coalesced CSR (rowptr_b[M_b+1], col_b) of B,
 * or -1; out_mask[e] = 1 when matched.  Replaces the reference's concat + argsort + adjacent-equal
 * scan; both operands are coalesced there (mul.py:38-41). */
int ps_coo_match(const int64_t *row, const int64_t *col, int64_t n, const int64_t *rowptr_b,
                 const int64_t *col_b, int64_t M_b, int64_t *out_pos, uint8_t *out_mask,
                 void *stream);

/* ------------------------------------------------------------------ resize / reshape */
/* storage.py:281-330 `sparse_resize`: out[0:total] = src[0:min(n,total)] followed by `fill` (a pointer array grows by
 * entries equal to nnz, a count array by zeros; shrinking cuts). */
int ps_pad_i64(const int64_t *src, int64_t n, int64_t fill, int64_t total, int64_t *out, void *stream);
/* storage.py:332-368 `sparse_reshape`: idx = old_n * row + col; out_row = idx / new_n, out_col = idx % new_n. */
int ps_reshape_index(const int64_t *row, const int64_t *col, int64_t n, int64_t old_n, int64_t new_n,
                     int64_t *out_row, int64_t *out_col, void *stream);

/* ------------------------------------------------------------------ cat */
/* cat.py:49-276: out = concat_k(src_k[0:lens[k]] + deltas[k]) in one launch per 32 pieces
 * (rows / cols shifted by the running sizes, pointers and permutations by the running nnz).
 * srcs / lens / deltas are HOST arrays of K entries; srcs[k] and out are device pointers. */
int ps_concat_add_i64(const int64_t *const *srcs, const int64_t *lens, const int64_t *deltas,
                      int64_t K, int64_t *out, void *stream);
/* the value concat of cat.py (any dtype / trailing shape): K device-to-device copies on `stream`. */
int ps_concat_bytes(const void *const *srcs, const int64_t *nbytes, int64_t K, void *out, void *stream);

/* ------------------------------------------------------------------ select */
/* masked_select.py:18-24: indices of the set entries of a byte mask, ascending.
 * out_idx sized n; *host_count = number selected.  Synchronises. */
int ps_mask_to_index(const uint8_t *mask, int64_t n, int64_t *out_idx, int64_t *host_count,
                     void *stream);
/* masked_select.py:93-114 `masked_select_nnz`: entry mask via an indirection
 * (mask[index[e]]) compacted to entry ids.  index nullable (then mask[e]).  Synchronises. */
int ps_mask_entries(const uint8_t *mask, const int64_t *index, int64_t nnz, int64_t *out_idx,
                    int64_t *host_count, void *stream);
/* narrow.py:53-62 `narrow(dim=1)`: keep the entries with lo <= col[e] < hi, in order, as
 * (row[e], col[e] - lo, value[e]) -- the reference builds a boolean mask, selects three
 * times and subtracts; here one compaction writes row / col and the value rows follow by one
 * gather.  Outputs sized nnz; value / out_value nullable (row_bytes = bytes per value row).
 * *host_count = entries kept.  Synchronises. */
int ps_narrow_cols(const int64_t *row, const int64_t *col, const void *value, int64_t row_bytes,
                   int64_t nnz, int64_t lo, int64_t hi, int64_t *out_row, int64_t *out_col,
                   void *out_value, int64_t *host_count, void *stream);
/* index_select.py:17-33 (dim 0) and :53-67 (dim 1, on the CSC arrays): given the new
 * pointer out_ptr[K+1] (ps_count2ptr of the old counts at idx) write, for every output
 * entry e of new row r: out_major[e] = r, out_perm[e] = old_ptr[idx[r]] + e - out_ptr[r],
 * out_minor[e] = minor[out_perm[e]], out_value[e] = value[out_perm[e]].
 * out_perm / out_value nullable. */
int ps_index_select_fill(const int64_t *old_ptr, const int64_t *minor, const void *value,
                         int64_t row_bytes, const int64_t *idx, int64_t K, const int64_t *out_ptr,
                         int64_t E, int64_t *out_major, int64_t *out_minor, void *out_value,
                         int64_t *out_perm, void *stream);

/* ------------------------------------------------------------------ sample_adj */
/* csrc/sample.cpp:8-24 -> csrc/cpu/sample_cpu.cpp:10-148 (the reference has NO GPU path:
 * sample.cpp:13-15 throws).  Draw u of seed position p is Philox4x32-10(seed;(p,u))
 * reduced as low + r % (high-low) (the "fixed seed mapping").  Upper bounds for the
 * caller's allocations: ps_sample_adj_bound.  out_rowptr[S+1]; out_col/out_e_id[bound];
 * out_n_id[S+bound]; host_sizes[0] = E, host_sizes[1] = |n_id|.  Synchronises. */
int ps_sample_adj_bound(const int64_t *rowptr, const int64_t *idx, int64_t S, int64_t num_neighbors,
                        int replace, int64_t *host_bound, void *stream);
int ps_sample_adj(const int64_t *rowptr, const int64_t *col, int64_t M, int64_t nnz, const int64_t *idx,
                  int64_t S, int64_t num_neighbors, int replace, uint64_t seed, int64_t bound,
                  int64_t *out_rowptr, int64_t *out_col, int64_t *out_n_id, int64_t *out_e_id,
                  int64_t *host_sizes, void *stream);

/* sample.py:10-28 `sample(src, num_neighbors, subset)`: out[i, k] = col[rowptr[s_i] + floor(u * deg(s_i))] with u
 * uniform in [0, 1) (draws with replacement).  The reference chains rand / mul_ / cast / add_ / index on the
 * framework's generator; here u of (position i, draw k) is the top 24 bits of Philox4x32-10(seed; (i, k)) * 2^-24.
 * subset NULL: rows 0 .. S-1.  out [S, num_neighbors]. */
int ps_sample_uniform(const int64_t *rowptr, const int64_t *col, int64_t nnz, const int64_t *subset, int64_t S,
                      int64_t num_neighbors, uint64_t seed, int64_t *out, void *stream);

/* ------------------------------------------------------------------ row-range partitioning (multi-GPU) */
/* SURVEY.md 8e (the reference has no distributed code).  Coarse row histogram used to
 * choose nnz-balanced, row-aligned splitters: out_hist[b] = #{e : row[e] >> shift == b}
 * (row must be 16-byte aligned). */
int ps_dist_row_histogram(const int64_t *row, int64_t nnz, int shift, int64_t buckets,
                          int64_t *out_hist, void *stream);
/* Stable partition of the local tuples by destination rank (row in
 * [splitters[g], splitters[g+1])), G <= 64.  Outputs are grouped by rank in send order;
 * out_counts[G] (device).  splitters is a DEVICE array of G+1 row ids.
 * CONTRACT: ps_dist_partition and ps_dist_scatter_peer move row and col as 32-bit words: every id must be
 * < 2^32 (sparse sizes <= 2^32 per dimension).  The entry points do not see the sizes; the host side
 * (paddle_sparse_b200/dist.py: route_tuples) refuses larger matrices before calling them. */
int ps_dist_partition(const int64_t *row, const int64_t *col, const void *value, int64_t row_bytes,
                      int64_t nnz, const int64_t *splitters, int G, int64_t *out_row,
                      int64_t *out_col, void *out_value, int64_t *out_counts, void *stream);
/* The same partition fused with the exchange over peer memory (one process per GPU, receive buffers of
 * all ranks mapped into this process, e.g. CUDA IPC / symmetric memory).  ps_dist_dest: destination rank
 * of every tuple (out_dest[nnz], one BYTE each: G <= 64) and the per-destination counts (device, G entries).  After the
 * ranks have exchanged their counts, ps_dist_scatter_peer runs the ONE stable counting pass that carries
 * (row, col, 4-byte value) and stores destination d's run at element bases[d] of peer_row[d] /
 * peer_col[d] / peer_value[d] (device arrays of G base pointers; bases: device, G entries): the tuples
 * cross NVLink as the pass produces them, tile by tile.  row_bytes: 0 (no value) or 4.  narrow != 0: row and
 * col are stored as uint32 (peer_row / peer_col then address uint32 elements): 12 instead of 20 bytes per tuple
 * on the wire; the receiver builds from them with ps_coalesce_u32. */
int ps_dist_dest(const int64_t *row, int64_t nnz, const int64_t *splitters, int G, uint8_t *out_dest,
                 int64_t *out_counts, void *stream);
int ps_dist_scatter_peer(const uint8_t *dest, const int64_t *row, const int64_t *col,
                         const void *value, int64_t row_bytes, int64_t nnz, int G, const int64_t *bases,
                         void *const *peer_row, void *const *peer_col, void *const *peer_value,
                         int narrow, void *stream);
/* out[i] = in[i] + delta (local <-> global row ids of a shard). */
int ps_add_scalar(const int64_t *in, int64_t n, int64_t delta, int64_t *out, void *stream);

/* ------------------------------------------------------------------ synthetic inputs (bench / tests) */
/* SURVEY.md 8d generator: h = splitmix64(seed + i); row = (h >> 32) % M; col = (h & 0xffffffff) % N;
 * value = fp32((splitmix64((seed ^ 0x9e3779b97f4a7c15) + i) >> 40) * 2^-24).  i in [first, first+n). */
int ps_synth_coo(uint64_t seed, int64_t first, int64_t n, int64_t M, int64_t N, int64_t *out_row,
                 int64_t *out_col, float *out_value, void *stream);
/* The skewed variant of the same generator: R-MAT (0.57, 0.19, 0.19, 0.05) descents on the
 * counter stream, reduced mod M / mod N; values as ps_synth_coo.  Same arguments. */
int ps_synth_rmat(uint64_t seed, int64_t first, int64_t n, int64_t M, int64_t N, int64_t *out_row,
                 int64_t *out_col, float *out_value, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PS_ABI_H */
