/* TEST INFRASTRUCTURE ONLY (oracle).
 *
 * Plain-C restatement of the reference's native CPU algorithms for the hot path.
 * It exists so that the parity tests still have a checker on machines where
 * oracle/_ref (the reference's own sources, compiled unmodified) is absent, and
 * it fixes ONE ordering the reference leaves to libstdc++ (see port_sample_adj).
 * It is pinned against oracle/_ref and against the reference's golden vectors in
 * tests/test_oracle.py.  Only tests/, __graft_entry__.smoke() and bench.py's CPU
 * legs may load it; the product path never does.
 *
 *   port_ind2ptr     follows csrc/cpu/convert_cpu.cpp:6-30
 *   port_ptr2ind     follows csrc/cpu/convert_cpu.cpp:32-48
 *   port_sample_adj  follows csrc/cpu/sample_cpu.cpp:10-148, RNG per oracle/philox.h
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "philox.h"

/* convert_cpu.cpp:6-30 — walk the sorted index once; head rows get 0, every row
 * boundary crossed between element i and i+1 gets i+1, tail rows get numel. */
void port_ind2ptr(const int64_t *ind, int64_t numel, int64_t M, int64_t *out) {
  if (numel == 0) { /* convert_cpu.cpp:9-11 */
    for (int64_t i = 0; i <= M; ++i) out[i] = 0;
    return;
  }
  for (int64_t i = 0; i <= ind[0]; ++i) out[i] = 0;
  int64_t idx = ind[0];
  for (int64_t i = 0; i < numel - 1; ++i) {
    const int64_t next = ind[i + 1];
    for (; idx < next; ++idx) out[idx + 1] = i + 1;
  }
  for (int64_t i = ind[numel - 1] + 1; i < M + 1; ++i) out[i] = numel;
}

/* convert_cpu.cpp:32-48 — expand every [ptr[i], ptr[i+1]) to i. */
void port_ptr2ind(const int64_t *ptr, int64_t M, int64_t E, int64_t *out) {
  (void)E;
  for (int64_t i = 0; i < M; ++i)
    for (int64_t e = ptr[i]; e < ptr[i + 1]; ++e) out[e] = i;
}

/* ---- open-addressing int64 -> int64 map (stands in for std::unordered_map) ---- */
typedef struct {
  int64_t *key, *val;
  uint64_t mask;
} imap;
static uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
static void imap_init(imap *m, uint64_t cap_hint) {
  uint64_t cap = 16;
  while (cap < 2 * cap_hint + 2) cap <<= 1;
  m->key = (int64_t *)malloc(sizeof(int64_t) * cap);
  m->val = (int64_t *)malloc(sizeof(int64_t) * cap);
  for (uint64_t i = 0; i < cap; ++i) m->key[i] = -1;
  m->mask = cap - 1;
}
static int64_t *imap_slot(imap *m, int64_t k, int *found) {
  uint64_t h = mix64((uint64_t)k) & m->mask;
  while (m->key[h] != -1 && m->key[h] != k) h = (h + 1) & m->mask;
  *found = (m->key[h] == k);
  if (!*found) m->key[h] = k;
  return &m->val[h];
}
static void imap_free(imap *m) { free(m->key); free(m->val); }

typedef struct { int64_t local, e; } pair_t;
static int pair_cmp(const void *a, const void *b) {
  const pair_t *x = (const pair_t *)a, *y = (const pair_t *)b;
  if (x->local != y->local) return x->local < y->local ? -1 : 1;
  return (x->e > y->e) - (x->e < y->e);
}

/* Upper bound on E so the caller can allocate: sum over seeds of the row count
 * (num_neighbors < 0), else S * num_neighbors. */
int64_t port_sample_adj_bound(const int64_t *rowptr, const int64_t *idx, int64_t S,
                              int64_t num_neighbors) {
  if (num_neighbors >= 0) return S * num_neighbors;
  int64_t tot = 0;
  for (int64_t i = 0; i < S; ++i) tot += rowptr[idx[i] + 1] - rowptr[idx[i]];
  return tot;
}

/* sample_cpu.cpp:10-148.  Outputs: out_rowptr[S+1], out_col[E], out_e_id[E],
 * out_n_id[<= S+E]; sizes[0] = E, sizes[1] = |n_id|.
 *
 * Ordering fixed here (the reference leaves it to std::unordered_set iteration,
 * sample_cpu.cpp:97,108): a sampled row relabels its picks in INSERTION order
 * (Floyd step order j = deg-k .. deg-1; all-neighbour rows in CSR order), and
 * rows are emitted sorted by (local id, e_id) where the reference's std::sort
 * (sample_cpu.cpp:133-140) leaves equal local ids unordered.  Both are instances
 * of what the reference may produce; tests compare against the real reference on
 * canonicalised sets. */
void port_sample_adj(const int64_t *rowptr, const int64_t *col, const int64_t *idx,
                     int64_t S, int64_t num_neighbors, int replace, uint64_t seed,
                     int64_t *out_rowptr, int64_t *out_col, int64_t *out_n_id,
                     int64_t *out_e_id, int64_t *sizes) {
  const int64_t bound = port_sample_adj_bound(rowptr, idx, S, num_neighbors);
  imap map;
  imap_init(&map, (uint64_t)(S + bound));
  int64_t n_count = 0;
  int found;
  for (int64_t n = 0; n < S; ++n) { /* sample_cpu.cpp:35-40: last position wins */
    *imap_slot(&map, idx[n], &found) = n;
    out_n_id[n_count++] = idx[n];
  }
  pair_t *buf = (pair_t *)malloc(sizeof(pair_t) * (size_t)(bound > 0 ? bound : 1));
  int64_t *picks = (int64_t *)malloc(sizeof(int64_t) *
                                     (size_t)(num_neighbors > 0 ? num_neighbors : 1));
  int64_t E = 0;
  out_rowptr[0] = 0;
  for (int64_t i = 0; i < S; ++i) {
    const int64_t n = idx[i];
    const int64_t row_start = rowptr[n], row_count = rowptr[n + 1] - rowptr[n];
    const int64_t first = E;
    if (num_neighbors < 0 || (!replace && row_count <= num_neighbors)) {
      /* sample_cpu.cpp:44-64 and :98-99 */
      for (int64_t j = 0; j < row_count; ++j) { buf[E].e = row_start + j; ++E; }
    } else if (replace) { /* sample_cpu.cpp:66-87 */
      if (row_count > 0)
        for (int64_t j = 0; j < num_neighbors; ++j) {
          buf[E].e = row_start + ps_philox_randint(seed, (uint64_t)i, (uint32_t)j, 0, row_count);
          ++E;
        }
    } else { /* Robert Floyd, sample_cpu.cpp:100-106 */
      int64_t np = 0;
      uint32_t u = 0;
      for (int64_t j = row_count - num_neighbors; j < row_count; ++j, ++u) {
        int64_t t = ps_philox_randint(seed, (uint64_t)i, u, 0, j);
        int dup = 0;
        for (int64_t q = 0; q < np; ++q) dup |= (picks[q] == t);
        picks[np++] = dup ? j : t;
      }
      for (int64_t q = 0; q < np; ++q) { buf[E].e = row_start + picks[q]; ++E; }
    }
    for (int64_t p = first; p < E; ++p) { /* relabel on first sight */
      const int64_t c = col[buf[p].e];
      int64_t *slot = imap_slot(&map, c, &found);
      if (!found) { *slot = n_count; out_n_id[n_count++] = c; }
      buf[p].local = *slot;
    }
    qsort(buf + first, (size_t)(E - first), sizeof(pair_t), pair_cmp);
    out_rowptr[i + 1] = E;
  }
  for (int64_t p = 0; p < E; ++p) { out_col[p] = buf[p].local; out_e_id[p] = buf[p].e; }
  sizes[0] = E;
  sizes[1] = n_count;
  free(buf); free(picks); imap_free(&map);
}

/* Exposed so tests can pin the CUDA Philox against the same arithmetic. */
uint64_t port_philox_u64(uint64_t seed, uint64_t pos, uint32_t draw) {
  return ps_philox_u64(seed, pos, draw);
}
