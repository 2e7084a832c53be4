// ctx.h — the opaque context behind include/h2ok_hist.h and the helpers every translation unit of the C ABI shares.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "engine.h"
#include "kernels.h"

namespace h2ok {

// thread-local message of the last failed call (h2ok_last_error)
char* err_buf();
constexpr size_t kErrBufLen = 512;

inline int32_t fail(int32_t code, const char* fmt, const char* a = "", const char* b = "") {
  snprintf(err_buf(), kErrBufLen, fmt, a, b);
  return code;
}

#define CK(call)                                                                                       \
  do {                                                                                                 \
    cudaError_t _e = (call);                                                                           \
    if (_e != cudaSuccess) return ::h2ok::fail(H2OK_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(_e)); \
  } while (0)

#define REQUIRE(cond, msg)                                                        \
  do {                                                                            \
    if (!(cond)) return ::h2ok::fail(H2OK_ERR_ARG, "%s (%s)", msg, #cond);        \
  } while (0)

template <typename T>
struct DevBuf {  // grow-only device buffer
  T* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, n * sizeof(T));
    if (e == cudaSuccess) cap = n;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

struct TreeLoop;  // tree_driver.cu: the per-level loop inside the library

}  // namespace h2ok

struct h2ok_ctx {
  int device = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // scratch of next_level (a8) and find_splits (a7)
  h2ok::DevBuf<h2ok_col_desc> tmp_descs;
  h2ok::DevBuf<int32_t> tmp_valid;
  h2ok::DevBuf<h2ok::PairCand> cand;
  // scratch of build_level: packed row records + range-error flag
  h2ok::DevBuf<uint2> rec;
  h2ok::DevBuf<int32_t> rec_err;
  // resident data of the host-buffer API
  h2ok::DevBuf<float> X, y, w;
  h2ok::DevBuf<int32_t> nid, leaf, wq, wyq, wyyq;
  h2ok::DevBuf<double> absmax;
  h2ok::DevBuf<uint32_t> cmin, cmax;
  int64_t rows = 0, ld = 0;
  int cols = 0;
  bool have_data = false, have_targets = false, weighted = false;
  h2ok_quant q{};
  // level state
  h2ok::DevBuf<h2ok_col_desc> descs, child_descs;
  h2ok::DevBuf<int64_t> acc;
  h2ok::DevBuf<uint32_t> mn, mx;
  h2ok::DevBuf<double> hist;
  h2ok::DevBuf<float> min2, maxin;
  h2ok::DevBuf<h2ok_split> splits;
  h2ok::DevBuf<int32_t> child_ids, n_children;
  int n_nodes = 0, max_bins = 0, gid_base = 0;
  bool level_built = false, splits_found = false, children_made = false;
  bool timed = false;      // ev0/ev1 bracket the last hist_build launches
  h2ok::Engine eng;        // resident tree engine (h2ok_hist_tree_*)
  int eng_level_bins = 0;  // h2ok_hist_tree_level_bins hint for the next tree_build_level_dev (0: none)
  bool eng_ready = false;
  h2ok_stats stats{};
  h2ok::TreeLoop* loop = nullptr;  // h2ok_hist_loop_* state
  void* comm = nullptr;            // ncclComm_t once h2ok_hist_comm_init has run
  int comm_rank = 0, comm_ranks = 1;
};

namespace h2ok {
void loop_free(h2ok_ctx* c);  // tree_driver.cu
void comm_free(h2ok_ctx* c);
}  // namespace h2ok
