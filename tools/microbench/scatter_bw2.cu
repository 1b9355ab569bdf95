// Memory-system ceiling of a partition pass's write pattern on B200, round 2: which (bins, run length, tuple
// width, tile order) the HBM / L2 path sustains when NO ranking work is done.  Reads a tuple stream (key of KB bytes
// + NP 32-bit payload words, structure of arrays), writes it as R interleaved output streams in runs of L tuples
// (R * L = tile).  order 0: tiles dealt round-robin over the CTAs (what an atomic ticket gives); order 1: every CTA
// owns a contiguous range of tiles (its runs of one bin are adjacent in memory and written one after the other).
// Run starts are misaligned by a per-bin hash like real data unless align = 1 (the write-combined case: every run
// starts on a 128-byte line).
//   nvcc -arch=sm_100a -O3 -o scatter_bw2 scatter_bw2.cu && ./scatter_bw2
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <typename KeyT, int NP>
__global__ void __launch_bounds__(256) k(const KeyT *__restrict__ kin, const uint32_t *__restrict__ pin, KeyT *__restrict__ kout,
                                         uint32_t *__restrict__ pout, long long n, int R, int L, int order, int align,
                                         long long pstride) {
  const long long tile_elems = (long long)R * L;
  const long long tiles = n / tile_elems;
  const long long stream_len = tiles * L + 64;  // elements per output stream (+ slack for the misalignment)
  const long long per_cta = (tiles + gridDim.x - 1) / gridDim.x;
  for (long long q = 0;; ++q) {
    const long long t = order ? (long long)blockIdx.x * per_cta + q : (long long)blockIdx.x + q * gridDim.x;
    if (order ? (q >= per_cta || t >= tiles) : (t >= tiles)) break;
    const long long base = t * tile_elems;
    for (int j = threadIdx.x; j < tile_elems; j += 256) {
      const int d = j / L, r = j - d * L;
      const int mis = align ? 0 : (int)((d * 2654435761u) >> 27) % 13;
      const long long dst = (long long)d * stream_len + t * L + r + mis;
      kout[dst] = kin[base + j];
#pragma unroll
      for (int p = 0; p < NP; ++p) pout[p * pstride + dst] = pin[p * pstride + base + j];
    }
  }
}

template <typename KeyT, int NP>
void sweep(const char *name) {
  const long long n0 = 61859140;
  const size_t slack = 1 << 22;
  KeyT *kin, *kout;
  uint32_t *pin, *pout;
  const long long pstride = n0 + slack;
  cudaMalloc(&kin, (n0 + slack) * sizeof(KeyT));
  cudaMalloc(&kout, (n0 + slack) * sizeof(KeyT));
  cudaMalloc(&pin, (size_t)(NP ? NP : 1) * pstride * 4);
  cudaMalloc(&pout, (size_t)(NP ? NP : 1) * pstride * 4);
  cudaMemset(kin, 1, n0 * sizeof(KeyT));
  cudaMemset(pin, 1, (size_t)(NP ? NP : 1) * pstride * 4);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  const int cfg[][2] = {{1, 8192},  {64, 128},  {64, 256},  {128, 96},  {128, 128}, {256, 16},  {256, 32}, {256, 48},
                        {256, 64},  {256, 96},  {512, 24},  {512, 32},  {512, 48},  {1024, 12}, {1024, 16}, {1024, 24},
                        {2048, 4},  {2048, 6},  {2048, 8},  {2048, 12}, {4096, 2},  {4096, 3},  {4096, 4},  {4096, 6}};
  const double bytes = 2.0 * (sizeof(KeyT) + 4.0 * NP);
  for (int order = 0; order < 2; ++order)
    for (int align = 0; align < 2; ++align) {
      for (auto &c : cfg) {
        const long long tile = (long long)c[0] * c[1];
        if (align && (c[1] % 32)) continue;  // the aligned case only makes sense for whole-line runs
        const long long n = n0 / tile * tile;
        if (n + (long long)c[0] * 64 > n0 + (long long)slack) continue;
        float ms = 0;
        for (int rep = 0; rep < 3; ++rep) {
          cudaEventRecord(a);
          k<KeyT, NP><<<148 * 8, 256>>>(kin, pin, kout, pout, n, c[0], c[1], order, align, pstride);
          cudaEventRecord(b);
          cudaDeviceSynchronize();
          float m;
          cudaEventElapsedTime(&m, a, b);
          if (rep == 0 || m < ms) ms = m;
        }
        printf("%s order %d align %d bins %4d run %4d (tile %5lld): %.3f ms  %.2f TB/s  err %d\n", name, order, align, c[0],
               c[1], tile, ms, n * bytes / ms / 1e9, (int)cudaGetLastError());
        fflush(stdout);
      }
    }
  cudaFree(kin); cudaFree(kout); cudaFree(pin); cudaFree(pout);
}

int main() {
  sweep<uint64_t, 1>("u64+1w");
  sweep<uint32_t, 1>("u32+1w");
  sweep<uint32_t, 3>("u32+3w");
  sweep<uint32_t, 0>("u32   ");
  return 0;
}
