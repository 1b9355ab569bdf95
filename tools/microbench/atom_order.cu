// Does a shared-memory atomicAdd issued by one warp hand out its return values in ascending LANE order when
// several lanes hit the same address?  (A radix pass that ranks with ATOMS needs exactly that for stability.)
// nvcc -arch=sm_100a -O3 -o atom_order atom_order.cu && ./atom_order
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

__global__ void __launch_bounds__(256) k(int iters, int nvals, int stride, unsigned long long *bad, unsigned long long *coll) {
  __shared__ uint32_t cnt[8][256];
  const unsigned tid = threadIdx.x, warp = tid >> 5;
  for (int j = tid; j < 8 * 256; j += 256) (&cnt[0][0])[j] = 0;
  __syncthreads();
  uint32_t x = (blockIdx.x * 256 + tid) * 2654435761u + 12345u;
  const unsigned lt = lanemask_lt();
  unsigned long long nb = 0, nc = 0;
  for (int it = 0; it < iters; ++it) {
    x = x * 1664525u + 1013904223u;
    const unsigned d = (((x >> 11) % (unsigned)nvals) * (unsigned)stride) & 255u;
    const uint32_t before = cnt[warp][d];
    __syncwarp();
    const unsigned m = __match_any_sync(0xffffffffu, d);
    const uint32_t got = atomicAdd(&cnt[warp][d], 1u);
    __syncwarp();
    if (got != before + __popc(m & lt)) ++nb;
    if (__popc(m) > 1) ++nc;
  }
  atomicAdd(bad, nb);
  atomicAdd(coll, nc);
}
int main() {
  unsigned long long *d, h[2];
  cudaMalloc(&d, 16);
  const int cfg[][2] = {{256, 1}, {64, 1}, {16, 1}, {4, 1}, {1, 1}, {8, 32}, {4, 32}, {16, 16}, {200, 1}, {32, 8}};
  for (auto &c : cfg) {
    cudaMemset(d, 0, 16);
    k<<<148 * 8, 256>>>(20000, c[0], c[1], d, d + 1);
    cudaDeviceSynchronize();
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("values %3d stride %2d: %llu lane-instances checked, %llu in collisions, %llu out of lane order (err %d)\n", c[0], c[1],
           148ull * 8 * 256 * 20000, h[1], h[0], (int)cudaGetLastError());
  }
  return 0;
}
