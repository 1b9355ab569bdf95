// Deterministic synthetic COO generator used by bench.py and the full-size
// tests: the same counter hash is restated in NumPy (bench/tests) so host and device
// inputs are identical.
#include "common.cuh"

namespace ps {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

__global__ void __launch_bounds__(256)
synth_coo_kernel(uint64_t seed, int64_t first, int64_t n, uint64_t M, uint64_t N, int64_t *__restrict__ row,
                 int64_t *__restrict__ col, float *__restrict__ value) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t c = (uint64_t)(first + i);
  const uint64_t h = splitmix64(seed + c);
  row[i] = (int64_t)((h >> 32) % M);
  col[i] = (int64_t)((h & 0xffffffffull) % N);
  if (value) {
    const uint64_t g = splitmix64((seed ^ 0x9e3779b97f4a7c15ull) + c);
    value[i] = (float)(g >> 40) * (1.0f / 16777216.0f);
  }
}

// Skewed variant: R-MAT with (a, b, c, d) = (0.57, 0.19, 0.19, 0.05) on the same counter stream.
// Entry c descends `levels` quadrant choices; level l uses the high (l even) or low (l odd) 32 bits of
// splitmix64(seed + 16 * c + l / 2) against the integer thresholds floor(0.57 * 2^32), floor(0.76 * 2^32),
// floor(0.95 * 2^32); the accumulated row / col bit strings are reduced mod M / mod N.
constexpr uint32_t RMAT_A = 2448131358u, RMAT_AB = 3264175144u, RMAT_ABC = 4080218931u;

__global__ void __launch_bounds__(256)
synth_rmat_kernel(uint64_t seed, int64_t first, int64_t n, uint64_t M, uint64_t N, int levels,
                  int64_t *__restrict__ row, int64_t *__restrict__ col, float *__restrict__ value) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t c = (uint64_t)(first + i);
  uint64_t r = 0, q = 0, h = 0;
  for (int l = 0; l < levels; ++l) {
    if ((l & 1) == 0) h = splitmix64(seed + 16ull * c + (uint64_t)(l >> 1));
    const uint32_t u = (l & 1) ? (uint32_t)(h & 0xffffffffull) : (uint32_t)(h >> 32);
    const uint64_t rb = u >= RMAT_AB ? 1ull : 0ull;
    const uint64_t cb = (u >= RMAT_A && u < RMAT_AB) || u >= RMAT_ABC ? 1ull : 0ull;
    r = (r << 1) | rb;
    q = (q << 1) | cb;
  }
  row[i] = (int64_t)(r % M);
  col[i] = (int64_t)(q % N);
  if (value) {
    const uint64_t g = splitmix64((seed ^ 0x9e3779b97f4a7c15ull) + c);
    value[i] = (float)(g >> 40) * (1.0f / 16777216.0f);
  }
}

}  // namespace ps

using namespace ps;

extern "C" int ps_synth_rmat(uint64_t seed, int64_t first, int64_t n, int64_t M, int64_t N, int64_t *out_row,
                             int64_t *out_col, float *out_value, void *stream) {
  PS_CHECK(n >= 0 && M > 0 && N > 0, "synth_rmat: bad sizes");
  PS_CHECK(M <= (int64_t(1) << 32) && N <= (int64_t(1) << 32), "synth_rmat: sizes above 2^32");
  if (n == 0) return PS_OK;
  const int levels = bits_for(M > N ? M : N);
  synth_rmat_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
      seed, first, n, (uint64_t)M, (uint64_t)N, levels, out_row, out_col, out_value);
  PS_LAUNCH_OK();
  return PS_OK;
}

extern "C" int ps_synth_coo(uint64_t seed, int64_t first, int64_t n, int64_t M, int64_t N, int64_t *out_row,
                            int64_t *out_col, float *out_value, void *stream) {
  PS_CHECK(n >= 0 && M > 0 && N > 0, "synth_coo: bad sizes");
  if (n == 0) return PS_OK;
  synth_coo_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(seed, first, n, (uint64_t)M,
                                                                                 (uint64_t)N, out_row, out_col, out_value);
  PS_LAUNCH_OK();
  return PS_OK;
}
