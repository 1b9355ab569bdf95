/* TEST INFRASTRUCTURE ONLY (oracle).
 * Philox4x32-10 (Salmon et al., SC'11), restated from the published algorithm.
 * This is the "fixed seed mapping" used to compare sample_adj with the
 * reference: the u-th draw of seed-row position p under `seed`
 * is   philox(key = seed, counter = (p_lo, p_hi, u, 0))  ->  x0 | x1 << 32,
 * reduced to [low, high) by  low + r % (high - low).
 * The CUDA kernel carries its own copy of the same arithmetic
 * (paddle_sparse_b200/csrc/philox.cuh); nothing in the product includes this file. */
#ifndef PS_ORACLE_PHILOX_H
#define PS_ORACLE_PHILOX_H
#include <stdint.h>

static inline void ps_philox_round(uint32_t c[4], const uint32_t k[2]) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
  const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
  const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
  const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
  const uint32_t n0 = hi1 ^ c[1] ^ k[0];
  const uint32_t n2 = hi0 ^ c[3] ^ k[1];
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

static inline uint64_t ps_philox_u64(uint64_t seed, uint64_t pos, uint32_t draw) {
  uint32_t c[4] = {(uint32_t)pos, (uint32_t)(pos >> 32), draw, 0u};
  uint32_t k[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  for (int r = 0; r < 10; ++r) {
    ps_philox_round(c, k);
    k[0] += 0x9E3779B9u;
    k[1] += 0xBB67AE85u;
  }
  return (uint64_t)c[0] | ((uint64_t)c[1] << 32);
}

static inline int64_t ps_philox_randint(uint64_t seed, uint64_t pos, uint32_t draw,
                                        int64_t low, int64_t high) {
  return low + (int64_t)(ps_philox_u64(seed, pos, draw) % (uint64_t)(high - low));
}
#endif
