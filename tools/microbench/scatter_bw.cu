// Memory-system ceiling of a radix pass's access pattern on B200: stream-read (key u64, payload u32), write them
// as R interleaved output streams in runs of L elements (R*L = tile).  No ranking, no shared memory: the time of
// this kernel is what the DRAM/L2 path allows a perfectly balanced pass with tile size R*L.
// nvcc -arch=sm_100a -O3 -o scatter_bw scatter_bw.cu && ./scatter_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k(const uint64_t *__restrict__ kin, const uint32_t *__restrict__ pin,
                                         uint64_t *__restrict__ kout, uint32_t *__restrict__ pout, long long n, int R,
                                         int L, int shift) {
  const long long tile_elems = (long long)R * L;
  const long long tiles = n / tile_elems;
  const long long stream_len = tiles * L;  // elements per output stream
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long base = t * tile_elems;
    for (int j = threadIdx.x; j < tile_elems; j += 256) {
      const int d = j / L, r = j - d * L;
      // shift: misalign the run starts by (d % 4) elements like real data (runs start anywhere)
      const long long dst = (long long)d * stream_len + t * L + r + (shift ? (d & 3) : 0);
      kout[dst] = kin[base + j];
      pout[dst] = pin[base + j];
    }
  }
}

int main() {
  const long long n = 61859140 / 12288 * 12288;
  uint64_t *kin, *kout;
  uint32_t *pin, *pout;
  cudaMalloc(&kin, n * 8 + 64);
  cudaMalloc(&kout, n * 8 + 4096);
  cudaMalloc(&pin, n * 4 + 64);
  cudaMalloc(&pout, n * 4 + 4096);
  cudaMemset(kin, 1, n * 8);
  cudaMemset(pin, 1, n * 4);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  const int cfg[][2] = {{1, 3072}, {256, 12}, {256, 24}, {256, 48}, {256, 96}, {512, 12}, {512, 24}, {2048, 6}, {64, 48}, {16, 192}};
  for (int shift = 0; shift < 2; ++shift)
    for (auto &c : cfg) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        k<<<148 * 8, 256>>>(kin, pin, kout, pout, n, c[0], c[1], shift);
        cudaEventRecord(b);
        cudaDeviceSynchronize();
      }
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      printf("streams %4d run %4d (tile %5d) shift %d: %.3f ms  %.2f TB/s  err %d\n", c[0], c[1], c[0] * c[1], shift, ms,
             n * 24.0 / ms / 1e9, (int)cudaGetLastError());
    }
  return 0;
}
