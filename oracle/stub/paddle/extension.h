// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product path.
//
// Minimal stand-in for <paddle/extension.h>, written for this repo so that the
// reference's own CPU sources (csrc/cpu/convert_cpu.cpp, csrc/cpu/sample_cpu.cpp)
// compile UNMODIFIED from where they lie under /root/reference.  It supplies
// exactly the symbols those two files touch (census: SURVEY.md section 2.2):
//   paddle::Tensor {numel, data<T>, dtype, place, is_cpu, is_gpu, shape}
//   paddle::empty / paddle::zeros, paddle::DataType, paddle::Place
//   paddle::experimental::randint  -> an INJECTED generator (ps_stub_set_rng)
//   PD_CHECK / PD_THROW
// Nothing here is derived from Paddle's implementation.
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace paddle {

enum class DataType { INT64 = 0 };

struct Place {
  int kind = 0;  // 0 = host
};
inline Place CPUPlace() { return Place{}; }

class Tensor {
 public:
  Tensor() = default;
  // Owning tensor of n int64 elements.
  explicit Tensor(int64_t n, bool zero = false)
      : n_(n), own_(new int64_t[n > 0 ? n : 1]), ptr_(own_.get()) {
    if (zero) std::memset(ptr_, 0, sizeof(int64_t) * (n > 0 ? n : 1));
  }
  // Borrowed view over caller memory (never freed here).
  Tensor(int64_t *borrowed, int64_t n) : n_(n), ptr_(borrowed) {}

  int64_t numel() const { return n_; }
  template <typename T>
  T *data() const {
    static_assert(sizeof(T) == sizeof(int64_t), "oracle stub only carries int64");
    return reinterpret_cast<T *>(ptr_);
  }
  DataType dtype() const { return DataType::INT64; }
  Place place() const { return Place{}; }
  bool is_cpu() const { return true; }
  bool is_gpu() const { return false; }
  std::vector<int64_t> shape() const { return {n_}; }

 private:
  int64_t n_ = 0;
  std::shared_ptr<int64_t[]> own_;
  int64_t *ptr_ = nullptr;
};

inline Tensor empty(const std::vector<int64_t> &shape, DataType, const Place &) {
  return Tensor(shape.empty() ? 1 : shape[0], false);
}
inline Tensor zeros(const std::vector<int64_t> &shape, DataType, const Place &) {
  return Tensor(shape.empty() ? 1 : shape[0], true);
}

namespace experimental {
// Injected generator: returns one integer uniform on [low, high).
using stub_rng_fn = int64_t (*)(int64_t low, int64_t high, void *state);
struct StubRng {
  stub_rng_fn fn = nullptr;
  void *state = nullptr;
};
inline StubRng &stub_rng() {
  static thread_local StubRng r;
  return r;
}
inline Tensor randint(int64_t low, int64_t high, const std::vector<int64_t> &,
                      DataType, const Place &) {
  StubRng &r = stub_rng();
  if (r.fn == nullptr) throw std::runtime_error("oracle stub: no RNG injected");
  Tensor t(1);
  t.data<int64_t>()[0] = r.fn(low, high, r.state);
  return t;
}
}  // namespace experimental

}  // namespace paddle

#define PD_THROW(...)                                   \
  do {                                                  \
    std::ostringstream pd_throw_os__;                   \
    pd_throw_os__ << __VA_ARGS__;                       \
    throw std::runtime_error(pd_throw_os__.str());      \
  } while (0)

#define PD_CHECK(cond, ...)                             \
  do {                                                  \
    if (!(cond)) PD_THROW(__VA_ARGS__);                 \
  } while (0)
