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

}  // namespace ps

using namespace ps;

extern "C" int ps_synth_coo(uint64_t seed, int64_t first, int64_t n, int64_t M, int64_t N, int64_t *out_row,
                            int64_t *out_col, float *out_value, void *stream) {
  PS_CHECK(n >= 0 && M > 0 && N > 0, "synth_coo: bad sizes");
  if (n == 0) return PS_OK;
  synth_coo_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(seed, first, n, (uint64_t)M,
                                                                                 (uint64_t)N, out_row, out_col, out_value);
  PS_LAUNCH_OK();
  return PS_OK;
}
