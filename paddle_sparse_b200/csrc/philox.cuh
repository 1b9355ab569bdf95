// Philox4x32-10 (Salmon et al., SC'11) — the counter-based generator behind sample_adj.
// Draw u of seed-row position p under `seed` is philox(key = seed, counter = (p_lo, p_hi, u, 0))
// -> x0 | x1 << 32, reduced to [low, high) by low + r % (high - low): the "fixed seed mapping"
// that stands in for the reference's per-draw paddle::experimental::randint
// (csrc/cpu/utils.h:22-34), which no device code can reproduce.
#pragma once
#include <stdint.h>

namespace ps {

__host__ __device__ __forceinline__ uint64_t philox_u64(uint64_t seed, uint64_t pos, uint32_t draw) {
  uint32_t c0 = (uint32_t)pos, c1 = (uint32_t)(pos >> 32), c2 = draw, c3 = 0u;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c0 = n0;
    c1 = (uint32_t)p1;
    c2 = n2;
    c3 = (uint32_t)p0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return (uint64_t)c0 | ((uint64_t)c1 << 32);
}

__host__ __device__ __forceinline__ int64_t philox_randint(uint64_t seed, uint64_t pos, uint32_t draw, int64_t low,
                                                           int64_t high) {
  return low + (int64_t)(philox_u64(seed, pos, draw) % (uint64_t)(high - low));
}

}  // namespace ps
