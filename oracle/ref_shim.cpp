// TEST INFRASTRUCTURE ONLY (oracle).
//
// extern "C" doorway onto the reference's own CPU functions, which are compiled
// UNMODIFIED from /root/reference/csrc/cpu/{convert_cpu,sample_cpu}.cpp against
// oracle/stub/paddle/extension.h (see oracle/Makefile).  Only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the resulting
// oracle/_ref/libps_ref_cpu.so.
//
//   ref_ind2ptr      -> ind2ptr_cpu     csrc/cpu/convert_cpu.cpp:6-30
//   ref_ptr2ind      -> ptr2ind_cpu     csrc/cpu/convert_cpu.cpp:32-48
//   ref_sample_adj   -> sample_adj_cpu  csrc/cpu/sample_cpu.cpp:10-148
//
// RNG: the reference draws every random number through
// paddle::experimental::randint (csrc/cpu/utils.h:22-34).  The stub routes that
// call to the callback below, which replays the Philox mapping of
// oracle/philox.h: draw u of seed-row position p.  The (p, u) schedule is known
// in advance because the reference visits seed rows in order and makes exactly
// `num_neighbors` draws for each row that samples (sample_cpu.cpp:75, 102).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "cpu/convert_cpu.h"
#include "cpu/sample_cpu.h"
#include "philox.h"

namespace {
thread_local std::string g_err;

struct DrawSchedule {
  uint64_t seed;
  std::vector<int64_t> row_pos;  // seed-row position of every draw, in call order
  std::vector<uint32_t> draw;    // draw index within that row
  size_t cursor = 0;
};

int64_t philox_cb(int64_t low, int64_t high, void *state) {
  auto *s = static_cast<DrawSchedule *>(state);
  if (s->cursor >= s->row_pos.size())
    throw std::runtime_error("oracle shim: more draws than scheduled");
  const size_t c = s->cursor++;
  return ps_philox_randint(s->seed, (uint64_t)s->row_pos[c], s->draw[c], low, high);
}

struct SampleResult {
  std::vector<int64_t> rowptr, col, n_id, e_id;
};
}  // namespace

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

int ref_ind2ptr(const int64_t *ind, int64_t numel, int64_t M, int64_t *out) {
  try {
    paddle::Tensor t(const_cast<int64_t *>(ind), numel);
    paddle::Tensor r = ind2ptr_cpu(t, M);
    std::memcpy(out, r.data<int64_t>(), sizeof(int64_t) * (size_t)(M + 1));
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

int ref_ptr2ind(const int64_t *ptr, int64_t M, int64_t E, int64_t *out) {
  try {
    paddle::Tensor t(const_cast<int64_t *>(ptr), M + 1);
    paddle::Tensor r = ptr2ind_cpu(t, E);
    if (E > 0) std::memcpy(out, r.data<int64_t>(), sizeof(int64_t) * (size_t)E);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

// Returns an opaque handle (nullptr on error); sizes[0] = E, sizes[1] = |n_id|.
void *ref_sample_adj(const int64_t *rowptr, int64_t M, const int64_t *col, int64_t nnz,
                     const int64_t *idx, int64_t S, int64_t num_neighbors, int replace,
                     uint64_t seed, int64_t *sizes) {
  try {
    DrawSchedule sched;
    sched.seed = seed;
    if (num_neighbors >= 0) {
      for (int64_t p = 0; p < S; ++p) {
        const int64_t deg = rowptr[idx[p] + 1] - rowptr[idx[p]];
        const bool draws = replace ? (deg > 0) : (deg > num_neighbors);
        if (draws)
          for (int64_t u = 0; u < num_neighbors; ++u) {
            sched.row_pos.push_back(p);
            sched.draw.push_back((uint32_t)u);
          }
      }
    }
    auto &rng = paddle::experimental::stub_rng();
    rng.fn = philox_cb;
    rng.state = &sched;

    paddle::Tensor t_rowptr(const_cast<int64_t *>(rowptr), M + 1);
    paddle::Tensor t_col(const_cast<int64_t *>(col), nnz);
    paddle::Tensor t_idx(const_cast<int64_t *>(idx), S);
    auto ret = sample_adj_cpu(t_rowptr, t_col, t_idx, num_neighbors, replace != 0);
    rng.fn = nullptr;
    rng.state = nullptr;

    auto *res = new SampleResult();
    auto grab = [](const paddle::Tensor &t, std::vector<int64_t> &v) {
      v.assign(t.data<int64_t>(), t.data<int64_t>() + t.numel());
    };
    grab(std::get<0>(ret), res->rowptr);
    const int64_t E = res->rowptr.empty() ? 0 : res->rowptr.back();
    // paddle::empty({0}) in the stub still reports numel 0, so the copies are exact.
    grab(std::get<1>(ret), res->col);
    grab(std::get<2>(ret), res->n_id);
    grab(std::get<3>(ret), res->e_id);
    res->col.resize((size_t)E);
    res->e_id.resize((size_t)E);
    sizes[0] = E;
    sizes[1] = (int64_t)res->n_id.size();
    return res;
  } catch (const std::exception &e) {
    g_err = e.what();
    return nullptr;
  }
}

void ref_sample_adj_fetch(void *h, int64_t *rowptr, int64_t *col, int64_t *n_id,
                          int64_t *e_id) {
  auto *r = static_cast<SampleResult *>(h);
  auto put = [](const std::vector<int64_t> &v, int64_t *dst) {
    if (!v.empty()) std::memcpy(dst, v.data(), sizeof(int64_t) * v.size());
  };
  put(r->rowptr, rowptr);
  put(r->col, col);
  put(r->n_id, n_id);
  put(r->e_id, e_id);
}

void ref_sample_adj_free(void *h) { delete static_cast<SampleResult *>(h); }

}  // extern "C"
