// TEST INFRASTRUCTURE ONLY (oracle).
//
// extern "C" doorway onto the reference's own CUDA functions, compiled UNMODIFIED for sm_100a from
// /root/reference/csrc/cuda/convert_cuda.cu against oracle/stub_gpu/paddle/extension.h (oracle/Makefile,
// target `refcuda`).  Only tests/ may load the resulting oracle/_ref/libps_ref_cuda.so: it is the
// "reference GPU baseline" of SURVEY.md 8(d) - the reference's kernels recompiled for B200 - used to check the
// product's ind2ptr / ptr2ind bit for bit on the device and to time them side by side.
//
//   refcuda_ind2ptr -> ind2ptr_cuda  csrc/cuda/convert_cuda.cu:23-40 (kernel :6-21)
//   refcuda_ptr2ind -> ptr2ind_cuda  csrc/cuda/convert_cuda.cu:56-68 (kernel :42-54)
//
// All pointers are device pointers owned by the caller; `out` is the buffer paddle::empty hands back.
#include <cstdint>
#include <string>

#include "cuda/convert_cuda.h"

namespace {
thread_local std::string g_err;
}

extern "C" {

const char *refcuda_last_error() { return g_err.c_str(); }

int refcuda_ind2ptr(const int64_t *ind, int64_t numel, int64_t M, int64_t *out, void *stream) {
  try {
    cudaStream_t s = (cudaStream_t)stream;
    paddle::ps_stub_next_output() = {out, M + 1, s};
    paddle::Tensor t(const_cast<int64_t *>(ind), numel, s);
    paddle::Tensor r = ind2ptr_cuda(t, M);
    if (r.data<int64_t>() != out) throw std::runtime_error("unexpected output buffer");
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e));
    return 0;
  } catch (const std::exception &ex) {
    g_err = ex.what();
    return 1;
  }
}

int refcuda_ptr2ind(const int64_t *ptr, int64_t numel, int64_t E, int64_t *out, void *stream) {
  try {
    cudaStream_t s = (cudaStream_t)stream;
    paddle::ps_stub_next_output() = {out, E > 0 ? E : 1, s};
    paddle::Tensor t(const_cast<int64_t *>(ptr), numel, s);
    paddle::Tensor r = ptr2ind_cuda(t, E);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e));
    (void)r;
    return 0;
  } catch (const std::exception &ex) {
    g_err = ex.what();
    return 1;
  }
}

}  // extern "C"
