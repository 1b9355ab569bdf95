// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product path.
//
// Device-side stand-in for <paddle/extension.h>, written for this repo so that the reference's own CUDA
// source csrc/cuda/convert_cuda.cu compiles UNMODIFIED for sm_100a from where it lies under /root/reference
// (oracle/Makefile, target `refcuda`).  It supplies exactly what that file touches:
//   paddle::Tensor {numel, data<T>, dtype, place, stream, is_gpu}, paddle::empty / paddle::zeros, PD_CHECK.
// Tensors are borrowed views over device memory the caller owns; paddle::empty / zeros hand out the output
// buffer the shim registered beforehand (ps_stub_next_output), so nothing is allocated inside a timed region.
// Nothing here is derived from Paddle's implementation.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <sstream>
#include <stdexcept>
#include <vector>

namespace paddle {

enum class DataType { INT64 = 0 };
struct Place {
  int kind = 1;  // 1 = device
};

class Tensor {
 public:
  Tensor() = default;
  Tensor(int64_t *borrowed, int64_t n, cudaStream_t s) : n_(n), ptr_(borrowed), stream_(s) {}
  int64_t numel() const { return n_; }
  template <typename T>
  T *data() const {
    static_assert(sizeof(T) == sizeof(int64_t), "oracle stub only carries int64");
    return reinterpret_cast<T *>(ptr_);
  }
  DataType dtype() const { return DataType::INT64; }
  Place place() const { return Place{}; }
  cudaStream_t stream() const { return stream_; }
  bool is_cpu() const { return false; }
  bool is_gpu() const { return true; }

 private:
  int64_t n_ = 0;
  int64_t *ptr_ = nullptr;
  cudaStream_t stream_ = nullptr;
};

struct StubOutput {
  int64_t *ptr = nullptr;
  int64_t capacity = 0;
  cudaStream_t stream = nullptr;
};
inline StubOutput &ps_stub_next_output() {
  static thread_local StubOutput o;
  return o;
}
inline Tensor empty(const std::vector<int64_t> &shape, DataType, const Place &) {
  StubOutput &o = ps_stub_next_output();
  const int64_t n = shape.empty() ? 1 : shape[0];
  if (o.ptr == nullptr || n > o.capacity) throw std::runtime_error("oracle gpu stub: no output buffer registered");
  return Tensor(o.ptr, n, o.stream);
}
inline Tensor zeros(const std::vector<int64_t> &shape, DataType dt, const Place &p) {
  Tensor t = empty(shape, dt, p);
  cudaMemsetAsync(t.data<int64_t>(), 0, sizeof(int64_t) * (size_t)t.numel(), t.stream());
  return t;
}

}  // namespace paddle

#define PD_THROW(...)                              \
  do {                                             \
    std::ostringstream pd_throw_os__;              \
    pd_throw_os__ << __VA_ARGS__;                  \
    throw std::runtime_error(pd_throw_os__.str()); \
  } while (0)

#define PD_CHECK(cond, ...)             \
  do {                                  \
    if (!(cond)) PD_THROW(__VA_ARGS__); \
  } while (0)
