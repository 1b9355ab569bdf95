// Micro-benchmark (run on the GPU box): per-SM cost of the warp-level "which lanes hold my digit" primitives
// that a radix-sort pass can rank with.  nvcc -arch=sm_100a -O3 -o rank_bench rank_bench.cu && ./rank_bench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

template <int MODE>
__global__ void __launch_bounds__(256) k(const uint32_t *in, uint32_t *out, int iters, long long *cycles) {
  __shared__ uint32_t cnt[8][256];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < 8 * 256; j += 256) (&cnt[0][0])[j] = 0;
  __syncthreads();
  uint32_t x = in[blockIdx.x * 256 + tid];
  uint32_t acc = 0;
  const unsigned lt = lanemask_lt();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    x = x * 1664525u + 1013904223u;
    const unsigned d = (x >> 13) & 255u;
    if (MODE == 0) {  // match.any
      const unsigned m = __match_any_sync(0xffffffffu, d);
      acc += __popc(m & lt) + __popc(m);
    } else if (MODE == 1) {  // 8 ballots
      unsigned m = 0xffffffffu;
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        const unsigned bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        m &= ((d >> b) & 1u) ? bal : ~bal;
      }
      acc += __popc(m & lt) + __popc(m);
    } else if (MODE == 2) {  // shared-memory atomic (returns the rank)
      acc += atomicAdd(&cnt[warp][d], 1u);
    } else if (MODE == 3) {  // match + LDS/STS counter chain (what the pass kernel does)
      const unsigned m = __match_any_sync(0xffffffffu, d);
      const uint32_t before = cnt[warp][d];
      __syncwarp();
      if ((m & lt) == 0) cnt[warp][d] = before + __popc(m);
      __syncwarp();
      acc += before + __popc(m & lt);
    } else if (MODE == 4) {  // ballots + LDS/STS counter chain
      unsigned m = 0xffffffffu;
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        const unsigned bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        m &= ((d >> b) & 1u) ? bal : ~bal;
      }
      const uint32_t before = cnt[warp][d];
      __syncwarp();
      if ((m & lt) == 0) cnt[warp][d] = before + __popc(m);
      __syncwarp();
      acc += before + __popc(m & lt);
    } else if (MODE == 5) {  // plain random LDS + STS (no ranking): the shared-memory floor
      const uint32_t before = cnt[warp][d];
      cnt[warp][(d + 1) & 255] = before + 1;
      acc += before;
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * 256 + tid] = acc;
  if (tid == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int blocks_per_sm) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * blocks_per_sm, iters = 4096;
  uint32_t *in, *out;
  long long *cyc;
  cudaMalloc(&in, blocks * 256 * 4);
  cudaMalloc(&out, blocks * 256 * 4);
  cudaMalloc(&cyc, blocks * 8);
  cudaMemset(in, 1, blocks * 256 * 4);
  uint32_t *h = new uint32_t[blocks * 256];
  for (int i = 0; i < blocks * 256; ++i) h[i] = i * 2654435761u;
  cudaMemcpy(in, h, blocks * 256 * 4, cudaMemcpyHostToDevice);
  k<MODE><<<blocks, 256>>>(in, out, iters, cyc);
  cudaDeviceSynchronize();
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  cudaEventRecord(a);
  k<MODE><<<blocks, 256>>>(in, out, iters, cyc);
  cudaEventRecord(b);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  long long *hc = new long long[blocks];
  cudaMemcpy(hc, cyc, blocks * 8, cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < blocks; ++i) avg += hc[i];
  avg /= blocks;
  // warp-items per SM = blocks_per_sm * 8 warps * iters; SM cycles ~ avg
  printf("%-34s %d CTAs/SM: %.1f SM-cycles per warp-item (%.3f ms, err=%d)\n", name, blocks_per_sm,
         avg / (blocks_per_sm * 8.0 * iters), ms, (int)cudaGetLastError());
  cudaFree(in); cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int bps : {4, 8}) {
    run<0>("match.any + 2 popc", bps);
    run<1>("8 ballots + 2 popc", bps);
    run<2>("smem atomicAdd (rank)", bps);
    run<3>("match + LDS/STS chain", bps);
    run<4>("8 ballots + LDS/STS chain", bps);
    run<5>("random LDS + STS only", bps);
  }
  return 0;
}
