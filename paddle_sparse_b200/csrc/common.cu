// Host-side plumbing shared by every entry point: error string, scratch pool, read-back.
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <set>
#include <utility>

#include "common.cuh"

namespace ps {

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- live kernel timing (see common.cuh)
struct ProfPair {
  cudaEvent_t a, b;
};
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mu;
static std::vector<ProfPair> g_prof[PROF_COUNT];
static const char *const g_prof_names[PROF_COUNT] = {"rs_hist_kernel",  "rs_pass_kernel",   "dedup_kernel",
                                                     "ind2ptr_kernel",  "ptr2ind_kernel",   "invert_perm_kernel",
                                                     "transpose_gather_kernel", "segment_reduce_kernel",
                                                     "gather_rows_kernel", "ls2_sort_dedup_kernel",
                                                     "ls_bucket_hist_kernel + cl_bucket_hist_kernel",
                                                     "cl_local_kernel"};

ProfScope::ProfScope(int id, cudaStream_t s) : id_(id), s_(s), slot_(nullptr) {
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  ProfPair *p = new ProfPair;
  if (cudaEventCreate(&p->a) != cudaSuccess || cudaEventCreate(&p->b) != cudaSuccess) {
    delete p;
    return;
  }
  cudaEventRecord(p->a, s_);
  slot_ = p;
}
ProfScope::~ProfScope() {
  if (!slot_) return;
  ProfPair *p = static_cast<ProfPair *>(slot_);
  cudaEventRecord(p->b, s_);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof[id_].push_back(*p);
  delete p;
}

int rs_variant() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_RS_VARIANT");
    v = e ? atoi(e) : 0;
  }
  return v;
}

int hybrid_variant() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_HYBRID");
    v = e ? atoi(e) : 1;
  }
  return v;
}

int ls_variant() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_LS_VARIANT");
    v = e ? atoi(e) : 2;
  }
  return v;
}

int csc_variant() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("PS_CSC_HYBRID");
    v = e ? atoi(e) : 1;
  }
  return v;
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

bool func_needs_config(const void *kernel) {
  static std::mutex mu;
  static std::set<std::pair<int, const void *>> seen;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return true;
  std::lock_guard<std::mutex> lk(mu);
  return seen.insert(std::make_pair(dev, kernel)).second;
}

static void retain_pool_blocks() {
  static std::once_flag once[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
  std::call_once(once[dev], [dev]() {
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
      uint64_t keep = UINT64_MAX;  // retain freed blocks: no OS round trip per call
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  });
}

cudaError_t Scratch::alloc_(void **p, size_t bytes) {
  retain_pool_blocks();
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

int64_t ps_launch_count(void) { return (int64_t)ps::g_launches.load(); }

// ---- the shim-owned device allocator: stream-ordered blocks for outputs the DLPack boundary
// (paddle_sparse_ops.py) hands to a caller that brought no allocator of its own
int ps_alloc(int64_t bytes, void **out_ptr, void *stream) {
  using namespace ps;
  PS_CHECK(bytes >= 0 && out_ptr, "alloc: bad arguments");
  retain_pool_blocks();
  PS_CUDA(cudaMallocAsync(out_ptr, (size_t)(bytes > 0 ? bytes : 1), (cudaStream_t)stream));
  return PS_OK;
}
int ps_free(void *ptr, void *stream) {
  using namespace ps;
  if (ptr) PS_CUDA(cudaFreeAsync(ptr, (cudaStream_t)stream));
  return PS_OK;
}
int ps_memcpy(void *dst, const void *src, int64_t bytes, int kind, void *stream) {
  using namespace ps;
  PS_CHECK(bytes >= 0 && kind >= 1 && kind <= 3, "memcpy: bad arguments");
  if (bytes == 0) return PS_OK;
  const cudaMemcpyKind k = kind == 1 ? cudaMemcpyHostToDevice : kind == 2 ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
  PS_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, k, (cudaStream_t)stream));
  if (kind != 3) PS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return PS_OK;
}
int ps_stream_sync(void *stream) {
  using namespace ps;
  PS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return PS_OK;
}
int ps_current_device(void) {
  int dev = -1;
  return cudaGetDevice(&dev) == cudaSuccess ? dev : -1;
}
int ps_set_device(int dev) {
  using namespace ps;
  PS_CUDA(cudaSetDevice(dev));
  return PS_OK;
}

void ps_prof_enable(int on) { ps::g_prof_on.store(on != 0); }

int ps_prof_kinds(void) { return (int)ps::PROF_COUNT; }

const char *ps_prof_name(int id) { return (id >= 0 && id < ps::PROF_COUNT) ? ps::g_prof_names[id] : ""; }

// Sum of the recorded durations of kernel family `id` since the last read (synchronises on the
// recorded events, then forgets them).
int ps_prof_read(int id, double *total_ms, int64_t *launches) {
  using namespace ps;
  PS_CHECK(id >= 0 && id < PROF_COUNT && total_ms && launches, "prof_read: bad arguments");
  std::vector<ProfPair> v;
  {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    v.swap(g_prof[id]);
  }
  double tot = 0;
  for (auto &p : v) {
    float ms = 0;
    PS_CUDA(cudaEventSynchronize(p.b));
    PS_CUDA(cudaEventElapsedTime(&ms, p.a, p.b));
    tot += ms;
    cudaEventDestroy(p.a);
    cudaEventDestroy(p.b);
  }
  *total_ms = tot;
  *launches = (int64_t)v.size();
  return PS_OK;
}

}  // extern "C"
