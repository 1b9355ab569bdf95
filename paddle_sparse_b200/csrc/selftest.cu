// Device self-test of the one hardware behaviour the sorts' STABILITY rests on: when several lanes of one warp
// instruction atomicAdd the same shared-memory word, ATOMS hands out the return values in ascending lane order
// (radix_sort.cuh phase 2, local_sort.cuh).  The CUDA programming model does not promise that order; sm_100a
// delivers it (6e10 lane-instances in round 1, none out of order).  ps_selftest_atom_order() re-checks it on the
// device it runs on, with both counter layouts the kernels use (one 32-bit counter per digit; two 16-bit counters
// packed in a word) under every collision pattern from "all lanes on one counter" to "no collision"; smoke() and a
// -m gpu test call it.  There is no second ranking mode in this build: a device that failed the test must not run
// the sorts at all (they would be unstable), which is what the callers of the test enforce.
#include <stdlib.h>

#include "common.cuh"

namespace ps {

__global__ void __launch_bounds__(256)
atom_order_kernel(int iters, int nvals, int stride, int packed, unsigned long long *out /*[checked, collided, bad]*/) {
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
    const unsigned m = __match_any_sync(0xffffffffu, d);
    uint32_t before, got;
    if (packed) {  // two 16-bit counters per word: lanes of DIFFERENT digits share a word too
      const unsigned sh = (d & 1u) * 16u;
      before = (cnt[warp][d >> 1] >> sh) & 0xffffu;
      __syncwarp();
      got = (atomicAdd(&cnt[warp][d >> 1], 1u << sh) >> sh) & 0xffffu;
      __syncwarp();
    } else {
      before = cnt[warp][d];
      __syncwarp();
      got = atomicAdd(&cnt[warp][d], 1u);
      __syncwarp();
    }
    if (got != before + (uint32_t)__popc(m & lt)) ++nb;
    if (__popc(m) > 1) ++nc;
    if (packed && (it & 1023) == 1023) {  // keep the 16-bit fields from overflowing (32 lanes x 1024 < 65536)
      __syncwarp();
      for (int q = tid & 31; q < 128; q += 32) cnt[warp][q] = 0;
      __syncwarp();
    }
  }
  atomicAdd(&out[0], (unsigned long long)iters);  // per thread: lane-instances
  atomicAdd(&out[1], nc);
  atomicAdd(&out[2], nb);
}

int rank_mode() { return 0; }  // 0 = one shared-memory atomic per item: the only ranking this build has

}  // namespace ps

using namespace ps;

extern "C" {

int ps_rank_mode(void) { return rank_mode(); }

int ps_selftest_atom_order(int64_t *host_out /*[3]: lane-instances checked, in collisions, out of lane order*/,
                           void *stream) {
  PS_CHECK(host_out != nullptr, "selftest_atom_order: host_out is NULL");
  cudaStream_t s = (cudaStream_t)stream;
  Scratch scratch(s);
  PS_SCRATCH(d, scratch, unsigned long long, 3);
  PS_CUDA(cudaMemsetAsync(d, 0, 3 * sizeof(unsigned long long), s));
  const int cfg[][2] = {{256, 1}, {64, 1}, {16, 1}, {4, 1}, {1, 1}, {8, 32}, {4, 32}, {16, 16}, {200, 1}, {32, 8}};
  const int blocks = sm_count() * 4;
  for (int packed = 0; packed < 2; ++packed)
    for (auto &c : cfg) {
      atom_order_kernel<<<blocks, 256, 0, s>>>(2048, c[0], c[1], packed, d);
      PS_LAUNCH_OK();
    }
  unsigned long long h[3];
  PS_TRY(read_back(h, d, sizeof(h), s));
  host_out[0] = (int64_t)h[0];
  host_out[1] = (int64_t)h[1];
  host_out[2] = (int64_t)h[2];
  return PS_OK;
}

}  // extern "C"
