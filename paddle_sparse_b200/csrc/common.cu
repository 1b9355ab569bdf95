// Host-side plumbing shared by every entry point: error string, scratch pool, read-back.
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace ps {

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

cudaError_t Scratch::alloc_(void **p, size_t bytes) {
  static std::once_flag once[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64) {
    std::call_once(once[dev], [dev]() {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        uint64_t keep = UINT64_MAX;  // retain freed blocks: no OS round trip per call
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      }
    });
  }
  return cudaMallocAsync(p, bytes, s_);
}

int read_back(void *host_dst, const void *dev_src, size_t bytes, cudaStream_t s) {
  static thread_local void *pinned = nullptr;
  static thread_local size_t cap = 0;
  if (bytes > cap) {
    if (pinned) cudaFreeHost(pinned);
    pinned = nullptr;
    cap = 0;
    size_t want = bytes < 256 ? 256 : bytes;
    PS_CUDA(cudaHostAlloc(&pinned, want, cudaHostAllocDefault));
    cap = want;
  }
  PS_CUDA(cudaMemcpyAsync(pinned, dev_src, bytes, cudaMemcpyDeviceToHost, s));
  PS_CUDA(cudaStreamSynchronize(s));
  memcpy(host_dst, pinned, bytes);
  return PS_OK;
}

}  // namespace ps

extern "C" {

// csrc/version.cpp:8-33: the reference returns CUDA_VERSION of its build (or -1 without CUDA)
int64_t ps_cuda_version(void) { return (int64_t)CUDA_VERSION; }

const char *ps_last_error(void) { return ps::g_err; }

}  // extern "C"
