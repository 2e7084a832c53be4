// capi.cu — the extern "C" boundary declared in include/h2ok_hist.h.  No torch types, plain pointers and sizes.
// There is no CPU fallback: every entry point drives the sm_100a kernels or fails with a status.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <thread>
#include <vector>

#include "common.cuh"
#include "ctx.h"

using namespace h2ok;

static thread_local char g_err[h2ok::kErrBufLen] = "";
char* h2ok::err_buf() { return g_err; }

// ---- host copies of the a1 rules (DHistogram.<init> / find_maxEx) for root descriptors -------------------
static double h_ulp(double x) {
  x = fabs(x);
  if (isinf(x)) return x;
  if (x == 1.7976931348623157e308) return x - nextafter(x, 0.0);
  return nextafter(x, INFINITY) - x;
}
static double h_find_maxex(double maxin, int is_int) {
  double ulp = h_ulp(maxin);
  if (is_int > 0 && 1 > ulp) ulp = 1;
  double res = maxin + ulp;
  return isinf(res) ? maxin : res;
}
static int h_make_desc(double mn, double maxex, int is_int, int nbins, h2ok_col_desc* out) {
  out->min = mn;
  out->maxex = maxex;
  out->is_int = is_int;
  out->nbin = 0;
  out->step = 0.0;
  if (!(maxex > mn)) return 0;
  int xbins = nbins;
  double step;
  if (is_int > 0 && maxex - mn <= (double)xbins) {
    xbins = (int)((long long)maxex - (long long)mn);
    step = 1.0;
  } else {
    volatile double range = maxex - mn;  // keep the two roundings separate (no contraction on the host either)
    step = (double)xbins / range;
    if (!(step > 0.0) || isinf(step)) return 0;
  }
  if (xbins < 1) return 0;
  out->step = step;
  out->nbin = xbins;
  return 1;
}
static double pow2_scale(double absmax, int qbits) {
  if (!(absmax > 0.0) || isinf(absmax)) return 1.0;
  int ex;
  frexp(absmax, &ex);
  return ldexp(1.0, qbits - ex);
}

extern "C" {

const char* h2ok_last_error(void) { return g_err; }
int32_t h2ok_hist_abi_version(void) { return 2; }  // 2: h2ok_split carries equal + cat_bits

int32_t h2ok_hist_create(int32_t device, h2ok_ctx** out) {
  REQUIRE(out != nullptr, "null out pointer");
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (device < 0 || device >= n) return fail(H2OK_ERR_ARG, "%s", "device index out of range");
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail(H2OK_ERR_UNSUPPORTED, "%s", "this library is built for sm_100a (B200) only");
  h2ok_ctx* c = new h2ok_ctx();
  c->device = device;
  cudaError_t e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev1);
  if (e != cudaSuccess) {  // nothing half-made is left behind
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
    return fail(H2OK_ERR_CUDA, "h2ok_hist_create: %s", cudaGetErrorString(e));
  }
  c->stream = c->own_stream;
  *out = c;
  return H2OK_OK;
}

int32_t h2ok_hist_destroy(h2ok_ctx* c) {
  if (!c) return H2OK_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  loop_free(c);
  comm_free(c);
  engine_free(c->eng);
  c->tmp_descs.release(); c->tmp_valid.release(); c->cand.release(); c->rec.release(); c->rec_err.release(); c->X.release(); c->y.release(); c->w.release();
  c->nid.release(); c->leaf.release(); c->wq.release(); c->wyq.release(); c->wyyq.release(); c->absmax.release();
  c->cmin.release(); c->cmax.release(); c->descs.release(); c->child_descs.release(); c->acc.release();
  c->mn.release(); c->mx.release(); c->hist.release(); c->min2.release(); c->maxin.release(); c->splits.release();
  c->child_ids.release(); c->n_children.release();
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  cudaStreamDestroy(c->own_stream);
  delete c;
  return H2OK_OK;
}

int32_t h2ok_hist_set_stream(h2ok_ctx* c, void* stream, int32_t use_own) {
  REQUIRE(c != nullptr, "null context");
  // a NULL handle is a real stream (CUDA's legacy default stream, which is what torch uses by default)
  c->stream = use_own ? c->own_stream : (cudaStream_t)stream;
  return H2OK_OK;
}

int32_t h2ok_hist_sync(h2ok_ctx* c) {
  REQUIRE(c != nullptr, "null context");
  CK(cudaStreamSynchronize(c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_copy_to_host(h2ok_ctx* c, const void* d_src, void* host_dst, size_t bytes) {
  REQUIRE(c && (bytes == 0 || (d_src && host_dst)), "null pointer");
  if (bytes == 0) return H2OK_OK;
  CK(cudaSetDevice(c->device));
  CK(cudaMemcpyAsync(host_dst, d_src, bytes, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return H2OK_OK;
}

// ---------------------------------------------------------------------------------------- device-resident
int32_t h2ok_hist_synth_fill_dev(h2ok_ctx* c, int32_t shape, uint64_t seed, int64_t row0, int64_t rows, int32_t cols,
                                 int64_t ld, int32_t na_permille, int32_t weight_mode, float* d_X, float* d_y, float* d_w) {
  REQUIRE(c != nullptr, "null context");
  REQUIRE(rows >= 0 && cols > 0 && ld >= rows, "bad shape");
  REQUIRE(shape == H2OK_SHAPE_HIGGS28 || shape == H2OK_SHAPE_WIDE64 || shape == H2OK_SHAPE_SPARSE, "unknown shape id");
  if (rows == 0) return H2OK_OK;
  CK(launch_synth(shape, seed, row0, rows, cols, ld, na_permille, weight_mode, d_X, d_y, d_w, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_col_ranges_dev(h2ok_ctx* c, const float* d_X, int64_t ld, int64_t rows, int32_t cols,
                                 uint32_t* d_min_enc, uint32_t* d_max_enc) {
  REQUIRE(c && d_X && d_min_enc && d_max_enc, "null pointer");
  REQUIRE(rows >= 0 && cols > 0 && ld >= rows, "bad shape");
  CK(launch_col_ranges(d_X, ld, rows, cols, d_min_enc, d_max_enc, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_target_absmax_dev(h2ok_ctx* c, const float* d_y, const float* d_w, int64_t rows, double* d_absmax) {
  REQUIRE(c && d_y && d_absmax, "null pointer");
  CK(launch_target_absmax(d_y, d_w, rows, d_absmax, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_quantize_rows_dev(h2ok_ctx* c, const float* d_y, const float* d_w, int64_t rows, const h2ok_quant* q,
                                    int32_t* d_wq, int32_t* d_wyq, int32_t* d_wyyq) {
  REQUIRE(c && d_y && q && d_wyq && d_wyyq, "null pointer");
  REQUIRE(!q->weighted || (d_w && d_wq), "weighted quantisation needs w and wq");
  if (rows == 0) return H2OK_OK;
  CK(launch_quantize_rows(d_y, d_w, rows, *q, d_wq, d_wyq, d_wyyq, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_bin_dev(h2ok_ctx* c, const float* d_x, int64_t n, const h2ok_col_desc* desc, int32_t* d_bins) {
  REQUIRE(c && d_x && desc && d_bins, "null pointer");
  if (n == 0) return H2OK_OK;
  CK(launch_bin(d_x, n, *desc, d_bins, c->stream));
  return H2OK_OK;
}

static int32_t check_level_shape(int32_t n_nodes, int32_t cols, int32_t max_bins) {
  if (n_nodes < 1 || n_nodes > H2OK_MAX_LEVEL_NODES) return fail(H2OK_ERR_ARG, "%s", "n_nodes out of range");
  if (cols < 1) return fail(H2OK_ERR_ARG, "%s", "cols < 1");
  if (max_bins < 1 || max_bins > H2OK_MAX_BINS) return fail(H2OK_ERR_UNSUPPORTED, "%s", "max_bins outside [1, H2OK_MAX_BINS]");
  return H2OK_OK;
}

int32_t h2ok_hist_build_level_dev(h2ok_ctx* c, const float* d_X, int64_t ld, int64_t rows, int32_t cols,
                                  const int32_t* d_nid, const int32_t* d_wq, const int32_t* d_wyq, const int32_t* d_wyyq,
                                  int32_t n_nodes, const h2ok_col_desc* d_descs, int32_t max_bins, int32_t clear,
                                  int64_t* d_acc, uint32_t* d_min2_enc, uint32_t* d_maxin_enc) {
  REQUIRE(c && d_X && d_nid && d_wyq && d_wyyq && d_descs && d_acc && d_min2_enc && d_maxin_enc, "null pointer");
  int32_t rc = check_level_shape(n_nodes, cols, max_bins);
  if (rc) return rc;
  REQUIRE(rows >= 0 && ld >= rows && ld % 4 == 0, "ld must be >= rows and a multiple of 4");
  REQUIRE(((uintptr_t)d_X % 16) == 0 && ((uintptr_t)d_nid % 16) == 0 && ((uintptr_t)d_wyq % 16) == 0 &&
              ((uintptr_t)d_wyyq % 16) == 0 && ((uintptr_t)d_wq % 16) == 0,
          "row arrays must be 16-byte aligned");
  const int nb1 = max_bins + 1;
  const int64_t ncell = (int64_t)n_nodes * cols;
  if (clear) {
    CK(cudaMemsetAsync(d_acc, 0, sizeof(int64_t) * (size_t)(ncell * nb1 * 3), c->stream));
    CK(launch_init_enc(d_min2_enc, d_maxin_enc, ncell, c->stream));
  }
  HistPlan plan = plan_hist(n_nodes, cols, rows, nb1);
  c->stats.passes = plan.passes;
  c->stats.copies = plan.copies;
  c->stats.launches = plan.passes + 1 + (clear ? 1 : 0);  // hist passes + pack_rows + init_enc
  c->stats.algo_bytes = (double)rows * cols * 4.0 + (double)rows * (4.0 + 4.0 + 4.0 + (d_wq ? 4.0 : 0.0)) +
                        (double)ncell * nb1 * 3 * 8.0;
  if (rows == 0) return H2OK_OK;
  const int64_t padded = (rows + 3) / 4 * 4;
  CK(c->rec.ensure((size_t)padded));
  if (!c->rec_err.p) {
    CK(c->rec_err.ensure(1));
    CK(cudaMemsetAsync(c->rec_err.p, 0, sizeof(int32_t), c->stream));
  }
  HistArgs a{};
  a.X = d_X; a.ld = ld; a.rows = rows; a.cols = cols; a.rec = c->rec.p; a.wq = d_wq;
  a.descs = d_descs; a.nb1 = nb1; a.acc = d_acc; a.min2_enc = d_min2_enc; a.maxin_enc = d_maxin_enc;
  CK(cudaEventRecord(c->ev0, c->stream));  // device time of pack + hist_build launches (read by last_level_stats)
  CK(launch_pack_rows(d_nid, d_wq, d_wyq, d_wyyq, rows, padded, n_nodes, c->rec.p, c->rec_err.p, c->stream));
  CK(launch_hist_build(a, plan, n_nodes, d_wq != nullptr, c->stream));
  CK(cudaEventRecord(c->ev1, c->stream));
  c->timed = true;
  return H2OK_OK;
}

int32_t h2ok_hist_build_level_passes(int32_t n_nodes, int32_t cols, int64_t rows, int32_t max_bins) {
  if (n_nodes < 1 || cols < 1 || max_bins < 1 || max_bins > H2OK_MAX_BINS) return 0;
  return plan_hist(n_nodes, cols, rows, max_bins + 1).passes;
}

int32_t h2ok_hist_finalize_level_dev(h2ok_ctx* c, const int64_t* d_acc, const uint32_t* d_min2_enc,
                                     const uint32_t* d_maxin_enc, int32_t n_nodes, int32_t cols, int32_t max_bins,
                                     const h2ok_quant* q, double* d_hist, float* d_min2, float* d_maxin) {
  REQUIRE(c && q, "null pointer");
  REQUIRE((d_hist == nullptr) == (d_acc == nullptr), "acc and hist go together");
  REQUIRE((d_min2 == nullptr) == (d_min2_enc == nullptr) && (d_min2 == nullptr) == (d_maxin == nullptr) &&
              (d_min2 == nullptr) == (d_maxin_enc == nullptr),
          "encoded and decoded ranges go together");
  int32_t rc = check_level_shape(n_nodes, cols, max_bins);
  if (rc) return rc;
  CK(launch_finalize(d_acc, d_min2_enc, d_maxin_enc, (int64_t)n_nodes * cols, max_bins + 1, *q, d_hist, d_min2, d_maxin,
                     c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_find_splits_dev(h2ok_ctx* c, const double* d_hist, const h2ok_col_desc* d_descs, int32_t n_nodes,
                                  int32_t cols, int32_t max_bins, double min_rows, double min_split_improvement,
                                  h2ok_split* d_splits) {
  REQUIRE(c && d_hist && d_descs && d_splits, "null pointer");
  int32_t rc = check_level_shape(n_nodes, cols, max_bins);
  if (rc) return rc;
  CK(c->cand.ensure((size_t)n_nodes * cols));
  CK(launch_find_splits(d_hist, d_descs, n_nodes, cols, max_bins + 1, min_rows, min_split_improvement, c->cand.p, d_splits,
                        c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_next_level_dev(h2ok_ctx* c, const h2ok_split* d_splits, const h2ok_col_desc* d_descs,
                                 const float* d_min2, const float* d_maxin, int32_t n_nodes, int32_t cols,
                                 const h2ok_params* params, int32_t depth, h2ok_col_desc* d_child_descs,
                                 int32_t* d_child_ids, int32_t* d_n_children) {
  (void)depth;
  REQUIRE(c && d_splits && d_descs && d_min2 && d_maxin && params && d_child_descs && d_child_ids && d_n_children,
          "null pointer");
  int32_t rc = check_level_shape(n_nodes, cols, params->nbins);
  if (rc) return rc;
  CK(c->tmp_descs.ensure((size_t)2 * n_nodes * cols));
  CK(c->tmp_valid.ensure((size_t)2 * n_nodes));
  CK(launch_next_level(d_splits, d_descs, d_min2, d_maxin, n_nodes, cols, *params, c->tmp_descs.p, c->tmp_valid.p,
                       d_child_descs, d_child_ids, d_n_children, c->stream));
  return H2OK_OK;
}

// ---------------------------------------------------------------------------------------- a9: tree-level caller
int32_t h2ok_hist_gbm_residual_dev(h2ok_ctx* c, int32_t dist, const float* d_label, const float* d_f, const float* d_w,
                                   int64_t rows, double den_scale, float* d_res, int32_t* d_denq) {
  REQUIRE(c && d_label && d_f && d_res, "null pointer");
  REQUIRE(dist == H2OK_DIST_GAUSSIAN || dist == H2OK_DIST_BERNOULLI, "unknown distribution");
  REQUIRE(rows >= 0 && den_scale > 0, "bad rows / den_scale");
  if (rows == 0) return H2OK_OK;
  CK(launch_gbm_residual(dist, d_label, d_f, d_w, rows, den_scale, d_res, d_denq, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_gbm_residual_multi_dev(h2ok_ctx* c, int32_t n_classes, const float* d_label, const float* d_f, int64_t ld,
                                         const float* d_w, int64_t rows, double den_scale, float* d_res, int32_t* d_denq) {
  REQUIRE(c && d_label && d_f && d_res && d_denq, "null pointer");
  REQUIRE(n_classes >= 2 && n_classes <= 1024, "n_classes outside [2, 1024]");
  REQUIRE(rows >= 0 && ld >= rows && den_scale > 0, "bad rows / ld / den_scale");
  if (rows == 0) return H2OK_OK;
  CK(launch_gbm_residual_multi(n_classes, d_label, d_f, ld, d_w, rows, den_scale, d_res, d_denq, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_gbm_gamma_pass_dev(h2ok_ctx* c, const int32_t* d_leaf, const int32_t* d_numq, const int32_t* d_denq,
                                     int64_t rows, int32_t n_codes, int64_t* d_num_acc, int64_t* d_den_acc) {
  REQUIRE(c && d_leaf && d_numq && d_num_acc && d_den_acc, "null pointer");
  REQUIRE(rows >= 0 && n_codes > 0, "bad rows / n_codes");
  if (rows == 0) {  // a rank with an empty shard still has to produce zero sums
    CK(cudaMemsetAsync(d_num_acc, 0, sizeof(int64_t) * (size_t)n_codes, c->stream));
    CK(cudaMemsetAsync(d_den_acc, 0, sizeof(int64_t) * (size_t)n_codes, c->stream));
    return H2OK_OK;
  }
  CK(launch_gbm_gamma_pass(d_leaf, d_numq, d_denq, rows, n_codes, d_num_acc, d_den_acc, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_gbm_add_gamma_dev(h2ok_ctx* c, const int32_t* d_leaf, int64_t rows, const double* d_gamma,
                                    int32_t n_codes, double learn_rate, float* d_f) {
  REQUIRE(c && d_leaf && d_gamma && d_f, "null pointer");
  REQUIRE(rows >= 0 && n_codes > 0, "bad rows / n_codes");
  if (rows == 0) return H2OK_OK;
  CK(launch_gbm_add_gamma(d_leaf, rows, d_gamma, n_codes, learn_rate, d_f, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_synth_fill_cols_dev(h2ok_ctx* c, int32_t shape, uint64_t seed, int64_t row0, int64_t rows, int32_t col0,
                                      int32_t ncols, int64_t ld, int32_t na_permille, float* d_X) {
  REQUIRE(c && d_X, "null pointer");
  REQUIRE(rows >= 0 && col0 >= 0 && ncols > 0 && ld >= rows, "bad shape");
  REQUIRE(shape == H2OK_SHAPE_HIGGS28 || shape == H2OK_SHAPE_WIDE64 || shape == H2OK_SHAPE_SPARSE, "unknown shape id");
  if (rows == 0) return H2OK_OK;
  CK(launch_synth_cols(shape, seed, row0, rows, col0, ncols, ld, na_permille, d_X, c->stream));
  return H2OK_OK;
}

// ---------------------------------------------------------------------------------------- resident tree engine
int32_t h2ok_hist_tree_create(h2ok_ctx* c, int64_t rows, int32_t cols, int32_t flags) {
  REQUIRE(c != nullptr, "null context");
  REQUIRE(rows > 0 && cols > 0, "empty shard");
  REQUIRE((flags & ~(H2OK_TREE_WEIGHTED | H2OK_TREE_SPARSE | H2OK_TREE_U8)) == 0, "unknown flag");
  if ((flags & H2OK_TREE_WEIGHTED) && (flags & (H2OK_TREE_SPARSE | H2OK_TREE_U8)))
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "sparse and u8 storage take unweighted trees only");
  if ((flags & H2OK_TREE_SPARSE) && (flags & H2OK_TREE_U8))
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "one storage format per shard");
  CK(cudaSetDevice(c->device));
  c->eng_ready = false;
  CK(cudaStreamSynchronize(c->stream));
  loop_free(c);
  CK(engine_alloc(c->eng, rows, cols, (flags & H2OK_TREE_WEIGHTED) != 0, (flags & H2OK_TREE_SPARSE) != 0,
                  (flags & H2OK_TREE_U8) != 0));
  c->eng_ready = true;
  return H2OK_OK;
}

int32_t h2ok_hist_tree_destroy(h2ok_ctx* c) {
  REQUIRE(c != nullptr, "null context");
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  loop_free(c);
  engine_free(c->eng);
  c->eng_ready = false;
  return H2OK_OK;
}

int32_t h2ok_hist_tree_load_cols_dev(h2ok_ctx* c, int32_t col0, int32_t ncols, const float* d_X, int64_t ld) {
  REQUIRE(c && d_X, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_load_cols_dev before tree_create");
  REQUIRE(col0 >= 0 && ncols > 0 && col0 + ncols <= c->eng.cols && ld >= c->eng.rows, "column block outside the shard");
  CK(engine_load_cols(c->eng, col0, ncols, d_X, ld, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_tree_load_bins_dev(h2ok_ctx* c, int32_t col0, int32_t ncols, const uint8_t* d_bins, int64_t ld) {
  REQUIRE(c && d_bins, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_load_bins_dev before tree_create");
  if (!c->eng.u8) return fail(H2OK_ERR_STATE, "%s", "tree_load_bins_dev needs a shard created with H2OK_TREE_U8");
  REQUIRE(col0 >= 0 && ncols > 0 && col0 + ncols <= c->eng.cols && ld >= c->eng.rows, "column block outside the shard");
  CK(engine_load_bins(c->eng, col0, ncols, d_bins, ld, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_prebin_dev(h2ok_ctx* c, const float* d_X, int64_t ld, int64_t rows, int32_t ncols, const h2ok_col_desc* descs,
                             uint8_t* d_bins, int64_t ld_bins) {
  REQUIRE(c && d_X && descs && d_bins, "null pointer");
  REQUIRE(rows >= 0 && ncols > 0 && ld >= rows && ld_bins >= rows, "bad shape");
  for (int32_t j = 0; j < ncols; j++)
    if (descs[j].nbin > 255) return fail(H2OK_ERR_UNSUPPORTED, "%s", "prebin: a byte holds at most 255 bins (255 is the NA code)");
  if (rows == 0) return H2OK_OK;
  CK(cudaSetDevice(c->device));
  CK(c->tmp_descs.ensure((size_t)ncols));
  CK(cudaMemcpyAsync(c->tmp_descs.p, descs, sizeof(h2ok_col_desc) * (size_t)ncols, cudaMemcpyHostToDevice, c->stream));
  CK(launch_prebin(d_X, ld, rows, ncols, c->tmp_descs.p, d_bins, ld_bins, c->stream));
  CK(cudaStreamSynchronize(c->stream));  // descs is the caller's host memory
  return H2OK_OK;
}

int32_t h2ok_hist_tree_begin_dev(h2ok_ctx* c, const int32_t* d_wq, const int32_t* d_wyq, const int32_t* d_wyyq,
                                 const h2ok_quant* q) {
  REQUIRE(c && d_wyq && d_wyyq, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_begin_dev before tree_create");
  REQUIRE(!c->eng.weighted || d_wq, "weighted engine needs d_wq");
  CK(cudaMemsetAsync(c->eng.err, 0, sizeof(int32_t), c->stream));  // range_errors count per tree
  CK(engine_begin_tree(c->eng, d_wq, d_wyq, d_wyyq, q, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_tree_build_level_dev(h2ok_ctx* c, int32_t n_nodes, const h2ok_col_desc* d_descs, int32_t max_bins,
                                       int32_t clear, int64_t* d_acc, uint32_t* d_min2_enc, uint32_t* d_maxin_enc) {
  REQUIRE(c && d_descs && d_acc && d_min2_enc && d_maxin_enc, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_build_level_dev before tree_create");
  int32_t rc = check_level_shape(n_nodes, c->eng.cols, max_bins);
  if (rc) return rc;
  const int nb1 = max_bins + 1;
  const int level_nb1 = c->eng_level_bins > 0 && c->eng_level_bins < max_bins ? c->eng_level_bins + 1 : nb1;
  c->eng_level_bins = 0;  // the hint covers one level
  if (level_nb1 > engine_max_nb1(c->eng.cols))
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "tree engine: 16 columns x (bins+1) of the level must fit 225 KB (bins <= 1199)");
  const int64_t ncell = (int64_t)n_nodes * c->eng.cols;
  if (clear) {
    CK(cudaMemsetAsync(d_acc, 0, sizeof(int64_t) * (size_t)(ncell * nb1 * 3), c->stream));
    CK(launch_init_enc(d_min2_enc, d_maxin_enc, ncell, c->stream));
  }
  c->stats.passes = 1;
  c->stats.copies = 1;
  c->stats.launches = 1 + (clear ? 1 : 0);  // hist (+ init_enc)
  c->stats.algo_bytes = (double)c->eng.rows * c->eng.cols * 4.0 + (double)c->eng.rows * 16.0 + (double)ncell * nb1 * 3 * 8.0;
  CK(cudaEventRecord(c->ev0, c->stream));
  CK(engine_build_level(c->eng, n_nodes, d_descs, nb1, level_nb1, d_acc, d_min2_enc, d_maxin_enc, c->stream));
  CK(cudaEventRecord(c->ev1, c->stream));
  c->timed = true;
  return H2OK_OK;
}

int32_t h2ok_hist_tree_level_bins(h2ok_ctx* c, int32_t level_bins) {
  REQUIRE(c != nullptr, "null context");
  REQUIRE(level_bins >= 0 && level_bins <= H2OK_MAX_BINS, "level_bins outside [0, H2OK_MAX_BINS]");
  c->eng_level_bins = level_bins;
  return H2OK_OK;
}

int32_t h2ok_hist_tree_apply_splits_dev(h2ok_ctx* c, const h2ok_split* d_splits, const int32_t* d_child_ids,
                                        int32_t gid_base, int32_t n_nodes, int32_t last_level) {
  REQUIRE(c && d_splits && d_child_ids, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_apply_splits_dev before tree_create");
  REQUIRE(n_nodes >= 1 && n_nodes <= H2OK_MAX_LEVEL_NODES, "n_nodes out of range");
  CK(engine_apply_splits(c->eng, d_splits, d_child_ids, gid_base, n_nodes, last_level == 0, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_tree_nid_dev(h2ok_ctx* c, int32_t* d_nid_out) {
  REQUIRE(c && d_nid_out, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_nid_dev before tree_create");
  CK(engine_nid(c->eng, d_nid_out, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_tree_leaf_dev(h2ok_ctx* c, int32_t* d_leaf_out) {
  REQUIRE(c && d_leaf_out, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_leaf_dev before tree_create");
  CK(cudaMemcpyAsync(d_leaf_out, c->eng.leaf, sizeof(int32_t) * (size_t)c->eng.rows, cudaMemcpyDeviceToDevice, c->stream));
  return H2OK_OK;
}

// Allocation-free a8 for callers that may not allocate device memory while running (see include/h2ok_hist.h).
// Same kernels as h2ok_hist_next_level_dev; the scratch is the library's fixed device array instead of ctx buffers.
int32_t h2ok_hist_next_level_noalloc_dev(h2ok_ctx* c, const h2ok_split* d_splits, const h2ok_col_desc* d_descs,
                                         const float* d_min2, const float* d_maxin, int32_t n_nodes, int32_t cols,
                                         const h2ok_params* params, int32_t depth, h2ok_col_desc* d_child_descs,
                                         int32_t* d_child_ids, int32_t* d_n_children) {
  (void)depth;
  REQUIRE(c && d_splits && d_descs && d_min2 && d_maxin && params && d_child_descs && d_child_ids && d_n_children,
          "null pointer");
  int32_t rc = check_level_shape(n_nodes, cols, params->nbins);
  if (rc) return rc;
  if ((int64_t)2 * n_nodes * cols > H2OK_NOALLOC_MAX_SLOT_COLS)
    return fail(H2OK_ERR_ARG, "%s", "next_level_noalloc: 2*n_nodes*cols exceeds H2OK_NOALLOC_MAX_SLOT_COLS (16384); "
                                    "use h2ok_hist_next_level_dev for larger levels");
  if (2 * n_nodes > H2OK_NOALLOC_MAX_SLOTS)
    return fail(H2OK_ERR_ARG, "%s", "next_level_noalloc: 2*n_nodes exceeds H2OK_NOALLOC_MAX_SLOTS (2048); "
                                    "use h2ok_hist_next_level_dev for larger levels");
  CK(launch_next_level_noalloc(d_splits, d_descs, d_min2, d_maxin, n_nodes, cols, *params, d_child_descs, d_child_ids,
                               d_n_children, c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_apply_splits_dev(h2ok_ctx* c, const float* d_X, int64_t ld, int64_t rows, const h2ok_split* d_splits,
                                   const int32_t* d_child_ids, int32_t gid_base, int32_t n_nodes, int32_t* d_nid,
                                   int32_t* d_leaf) {
  REQUIRE(c && d_X && d_splits && d_child_ids && d_nid, "null pointer");
  REQUIRE(n_nodes >= 1 && rows >= 0, "bad shape");
  if (rows == 0) return H2OK_OK;
  CK(launch_apply_splits(d_X, ld, rows, d_splits, d_child_ids, gid_base, n_nodes, d_nid, d_leaf, c->stream));
  return H2OK_OK;
}

// ---------------------------------------------------------------------------------------- host-buffer API
// The host-buffer entry points run on the resident tree engine (row_engine.cu): the columns pass through a small
// staging buffer into the engine's row-major copy, so the context holds the data once.
// Host columns -> one column-major device block, through pinned staging: kUploadThreads worker threads each copy chunks
// of the caller's (pageable) columns into their own pinned buffer and send them off on their own stream, so the host
// copy of one chunk runs beside the bus transfer of the others.  (One cudaMemcpyAsync per column from pageable memory
// moved 10 GB/s: the driver stages through a single pinned buffer on one thread.)
static constexpr int kUploadThreads = 4;
static constexpr size_t kUploadChunk = (size_t)8 << 20;  // floats per chunk (32 MB)

struct Uploader {  // lives for one h2ok_hist_set_data_host call: the pinned buffers are made once, not per block
  struct Worker {
    float* pin = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t free_ev = nullptr;
    cudaError_t err = cudaSuccess;
  } w[kUploadThreads];
  cudaError_t init() {
    cudaError_t err = cudaSuccess;
    for (auto& x : w) {
      if (err == cudaSuccess) err = cudaMallocHost((void**)&x.pin, kUploadChunk * sizeof(float));
      if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&x.stream, cudaStreamNonBlocking);
      if (err == cudaSuccess) err = cudaEventCreateWithFlags(&x.free_ev, cudaEventDisableTiming);
    }
    return err;
  }
  ~Uploader() {
    for (auto& x : w) {
      if (x.free_ev) cudaEventDestroy(x.free_ev);
      if (x.stream) cudaStreamDestroy(x.stream);
      if (x.pin) cudaFreeHost(x.pin);
    }
  }
  cudaError_t columns(int device, float* d_block, int64_t ld, const float* const* cols_host, int32_t ncols, int64_t rows) {
    const int64_t chunks_per_col = (rows + (int64_t)kUploadChunk - 1) / (int64_t)kUploadChunk;
    const int64_t items = chunks_per_col * ncols;
    std::vector<std::thread> pool;
    for (int t = 0; t < kUploadThreads; t++)
      pool.emplace_back([&, t] {
        Worker& me = w[t];
        me.err = cudaSetDevice(device);
        for (int64_t i = t; i < items && me.err == cudaSuccess; i += kUploadThreads) {
          const int64_t j = i / chunks_per_col, off = (i % chunks_per_col) * (int64_t)kUploadChunk;
          const size_t n = (size_t)(rows - off < (int64_t)kUploadChunk ? rows - off : (int64_t)kUploadChunk);
          me.err = cudaEventSynchronize(me.free_ev);  // the previous transfer out of this buffer has finished
          if (me.err != cudaSuccess) break;
          memcpy(me.pin, cols_host[j] + off, n * sizeof(float));
          me.err = cudaMemcpyAsync(d_block + j * ld + off, me.pin, n * sizeof(float), cudaMemcpyHostToDevice, me.stream);
          if (me.err == cudaSuccess) me.err = cudaEventRecord(me.free_ev, me.stream);
        }
        if (me.err == cudaSuccess) me.err = cudaStreamSynchronize(me.stream);
      });
    for (auto& th : pool) th.join();
    for (auto& x : w)
      if (x.err != cudaSuccess) return x.err;
    return cudaSuccess;
  }
};

int32_t h2ok_hist_set_data_host(h2ok_ctx* c, const float* const* cols_host, int64_t rows, int32_t cols) {
  REQUIRE(c && cols_host, "null pointer");
  REQUIRE(rows > 0 && cols > 0, "empty data");
  for (int32_t j = 0; j < cols; j++) REQUIRE(cols_host[j] != nullptr, "null column");
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  c->have_data = c->have_targets = c->eng_ready = false;
  CK(engine_alloc(c->eng, rows, cols, false));
  c->eng_ready = true;
  const int64_t ld = (rows + 63) / 64 * 64;
  int64_t per = ((int64_t)1 << 30) / (4 * ld);  // columns per staging block (<= 1 GB)
  per = per < 1 ? 1 : (per > cols ? cols : per);
  CK(c->X.ensure((size_t)ld * per));
  CK(c->cmin.ensure(cols)); CK(c->cmax.ensure(cols));
  Uploader up;
  CK(up.init());
  for (int32_t c0 = 0; c0 < cols; c0 += (int32_t)per) {
    const int32_t nc = cols - c0 < per ? cols - c0 : (int32_t)per;
    CK(cudaStreamSynchronize(c->stream));  // the block is free again (the previous columns have been taken in)
    CK(up.columns(c->device, c->X.p, ld, cols_host + c0, nc, rows));
    CK(launch_col_ranges(c->X.p, ld, rows, nc, c->cmin.p + c0, c->cmax.p + c0, c->stream));
    CK(engine_load_cols(c->eng, c0, nc, c->X.p, ld, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  c->X.release();
  c->rows = rows;
  c->ld = ld;
  c->cols = cols;
  c->have_data = true;
  c->level_built = c->splits_found = c->children_made = false;
  return H2OK_OK;
}

int32_t h2ok_hist_set_targets_host(h2ok_ctx* c, const float* y_host, const float* w_host) {
  REQUIRE(c && y_host, "null pointer");
  if (!c->have_data) return fail(H2OK_ERR_STATE, "%s", "set_targets_host before set_data_host");
  CK(cudaSetDevice(c->device));
  const int64_t n = c->rows;
  const size_t pad = (size_t)(n + 3) / 4 * 4;
  CK(c->y.ensure(pad)); CK(c->wyq.ensure(pad)); CK(c->wyyq.ensure(pad));
  CK(c->absmax.ensure(3));
  c->weighted = w_host != nullptr;
  if (c->weighted) { CK(c->w.ensure(pad)); CK(c->wq.ensure(pad)); }
  CK(engine_set_weighted(c->eng, c->weighted));
  CK(cudaMemcpyAsync(c->y.p, y_host, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  if (c->weighted) CK(cudaMemcpyAsync(c->w.p, w_host, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  CK(launch_target_absmax(c->y.p, c->weighted ? c->w.p : nullptr, n, c->absmax.p, c->stream));
  double m[3];
  CK(cudaMemcpyAsync(m, c->absmax.p, sizeof(m), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  h2ok_hist_choose_quant(m, c->weighted ? 1 : 0, 25, &c->q);
  CK(launch_quantize_rows(c->y.p, c->weighted ? c->w.p : nullptr, n, c->q, c->weighted ? c->wq.p : nullptr, c->wyq.p,
                          c->wyyq.p, c->stream));
  CK(engine_begin_tree(c->eng, c->weighted ? c->wq.p : nullptr, c->wyq.p, c->wyyq.p, &c->q, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->gid_base = 0;
  c->have_targets = true;
  c->level_built = c->splits_found = c->children_made = false;
  return H2OK_OK;
}

int32_t h2ok_hist_root_descs_from_ranges(const float* col_min, const float* col_max, int32_t cols,
                                         const h2ok_params* params, const int32_t* is_int, h2ok_col_desc* descs_out) {
  REQUIRE(col_min && col_max && params && descs_out && cols > 0, "null pointer");
  for (int j = 0; j < cols; j++) {
    float lo = col_min[j] + 0.0f, hi = col_max[j] + 0.0f;
    int ii = is_int ? is_int[j] : 0;
    if (ii < 0 || ii > 2) return fail(H2OK_ERR_ARG, "%s", "is_int: 0 real, 1 integer, 2 categorical");
    if (ii == 2 && lo <= hi && (lo < 0.0f || hi >= (float)(32 * H2OK_CAT_WORDS)))
      return fail(H2OK_ERR_UNSUPPORTED, "%s", "categorical columns hold the codes 0 .. 1023 (a split's bit set covers those)");
    h2ok_col_desc* o = &descs_out[j];
    memset(o, 0, sizeof(*o));
    o->is_int = ii;
    if (!(lo < hi) || isinf(lo) || isinf(hi)) continue;  // DHistogram.initialHist: constant / all-NA columns
    h_make_desc((double)lo, h_find_maxex((double)hi, ii), ii, params->nbins_top_level, o);
  }
  return H2OK_OK;
}

int32_t h2ok_hist_choose_quant(const double* m, int32_t weighted, int32_t qbits, h2ok_quant* q) {
  REQUIRE(m && q, "null pointer");
  REQUIRE(qbits >= 8 && qbits <= 25, "qbits outside [8,25] (the packed row records hold 27-bit targets)");
  q->weighted = weighted ? 1 : 0;
  q->qbits = qbits;
  q->scale_w = weighted ? pow2_scale(m[0], qbits) : 1.0;
  q->scale_wy = pow2_scale(m[1], qbits);
  q->scale_wyy = pow2_scale(m[2], qbits);
  return H2OK_OK;
}

int32_t h2ok_hist_root_descs(h2ok_ctx* c, const h2ok_params* params, const int32_t* is_int, h2ok_col_desc* descs_out) {
  REQUIRE(c && params && descs_out, "null pointer");
  if (!c->have_data) return fail(H2OK_ERR_STATE, "%s", "root_descs before set_data_host");
  CK(cudaSetDevice(c->device));
  std::vector<uint32_t> mn(c->cols), mx(c->cols);  // the column ranges were taken while the data went through staging
  CK(cudaMemcpyAsync(mn.data(), c->cmin.p, sizeof(uint32_t) * c->cols, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(mx.data(), c->cmax.p, sizeof(uint32_t) * c->cols, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  std::vector<float> lo(c->cols), hi(c->cols);
  for (int j = 0; j < c->cols; j++) {
    uint32_t lo_b = dec_f32_bits(mn[j]), hi_b = dec_f32_bits(mx[j]);
    memcpy(&lo[j], &lo_b, 4);
    memcpy(&hi[j], &hi_b, 4);
  }
  return h2ok_hist_root_descs_from_ranges(lo.data(), hi.data(), c->cols, params, is_int, descs_out);
}

int32_t h2ok_hist_build_level(h2ok_ctx* c, int32_t n_nodes, const h2ok_col_desc* descs, int32_t max_bins,
                              double* hist_out, float* min2_out, float* maxin_out) {
  REQUIRE(c && descs, "null pointer");
  if (!c->have_targets) return fail(H2OK_ERR_STATE, "%s", "build_level before set_targets_host");
  int32_t rc = check_level_shape(n_nodes, c->cols, max_bins);
  if (rc) return rc;
  CK(cudaSetDevice(c->device));
  const int nb1 = max_bins + 1;
  const size_t ncell = (size_t)n_nodes * c->cols;
  CK(c->descs.ensure(ncell)); CK(c->acc.ensure(ncell * nb1 * 3)); CK(c->hist.ensure(ncell * nb1 * 3));
  CK(c->mn.ensure(ncell)); CK(c->mx.ensure(ncell)); CK(c->min2.ensure(ncell)); CK(c->maxin.ensure(ncell));
  CK(cudaMemcpyAsync(c->descs.p, descs, sizeof(h2ok_col_desc) * ncell, cudaMemcpyHostToDevice, c->stream));
  int32_t level_bins = 1;  // the descriptors are on the host: size the shared histogram for this level's widest one
  for (size_t i = 0; i < ncell; i++) level_bins = descs[i].nbin > level_bins ? descs[i].nbin : level_bins;
  c->eng_level_bins = level_bins < max_bins ? level_bins : 0;
  rc = h2ok_hist_tree_build_level_dev(c, n_nodes, c->descs.p, max_bins, 1, c->acc.p, c->mn.p, c->mx.p);
  if (rc) return rc;
  rc = h2ok_hist_finalize_level_dev(c, c->acc.p, c->mn.p, c->mx.p, n_nodes, c->cols, max_bins, &c->q, c->hist.p,
                                    c->min2.p, c->maxin.p);
  if (rc) return rc;
  c->stats.launches += 2;
  if (hist_out) CK(cudaMemcpyAsync(hist_out, c->hist.p, sizeof(double) * ncell * nb1 * 3, cudaMemcpyDeviceToHost, c->stream));
  if (min2_out) CK(cudaMemcpyAsync(min2_out, c->min2.p, sizeof(float) * ncell, cudaMemcpyDeviceToHost, c->stream));
  if (maxin_out) CK(cudaMemcpyAsync(maxin_out, c->maxin.p, sizeof(float) * ncell, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->n_nodes = n_nodes;
  c->max_bins = max_bins;
  c->level_built = true;
  c->splits_found = c->children_made = false;
  return H2OK_OK;
}

int32_t h2ok_hist_find_splits(h2ok_ctx* c, double min_rows, double min_split_improvement, h2ok_split* splits_out) {
  REQUIRE(c != nullptr, "null context");
  if (!c->level_built) return fail(H2OK_ERR_STATE, "%s", "find_splits before build_level");
  CK(cudaSetDevice(c->device));
  CK(c->splits.ensure(c->n_nodes));
  int32_t rc = h2ok_hist_find_splits_dev(c, c->hist.p, c->descs.p, c->n_nodes, c->cols, c->max_bins, min_rows,
                                         min_split_improvement, c->splits.p);
  if (rc) return rc;
  if (splits_out) CK(cudaMemcpyAsync(splits_out, c->splits.p, sizeof(h2ok_split) * c->n_nodes, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->splits_found = true;
  c->children_made = false;
  return H2OK_OK;
}

int32_t h2ok_hist_next_level(h2ok_ctx* c, const h2ok_params* params, int32_t depth, h2ok_col_desc* child_descs_out,
                             int32_t* child_ids_out, int32_t* n_children_out) {
  REQUIRE(c && params && n_children_out, "null pointer");
  if (!c->splits_found) return fail(H2OK_ERR_STATE, "%s", "next_level before find_splits");
  CK(cudaSetDevice(c->device));
  const size_t slots = (size_t)2 * c->n_nodes;
  CK(c->child_descs.ensure(slots * c->cols)); CK(c->child_ids.ensure(slots)); CK(c->n_children.ensure(1));
  int32_t rc = h2ok_hist_next_level_dev(c, c->splits.p, c->descs.p, c->min2.p, c->maxin.p, c->n_nodes, c->cols, params,
                                        depth, c->child_descs.p, c->child_ids.p, c->n_children.p);
  if (rc) return rc;
  int32_t n = 0;
  CK(cudaMemcpyAsync(&n, c->n_children.p, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  if (child_ids_out) CK(cudaMemcpyAsync(child_ids_out, c->child_ids.p, sizeof(int32_t) * slots, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (child_descs_out && n > 0)
    CK(cudaMemcpy(child_descs_out, c->child_descs.p, sizeof(h2ok_col_desc) * (size_t)n * c->cols, cudaMemcpyDeviceToHost));
  *n_children_out = n;
  c->children_made = true;
  return H2OK_OK;
}

int32_t h2ok_hist_apply_splits(h2ok_ctx* c) {
  REQUIRE(c != nullptr, "null context");
  if (!c->splits_found) return fail(H2OK_ERR_STATE, "%s", "apply_splits before find_splits");
  CK(cudaSetDevice(c->device));
  if (!c->children_made) {  // last level of the tree: every side becomes a leaf
    CK(c->child_ids.ensure((size_t)2 * c->n_nodes));
    CK(cudaMemsetAsync(c->child_ids.p, 0xFF, sizeof(int32_t) * 2 * (size_t)c->n_nodes, c->stream));
  }
  int32_t rc = h2ok_hist_tree_apply_splits_dev(c, c->splits.p, c->child_ids.p, c->gid_base, c->n_nodes,
                                               c->children_made ? 0 : 1);
  if (rc) return rc;
  CK(cudaStreamSynchronize(c->stream));
  c->gid_base += c->n_nodes;
  c->level_built = c->splits_found = c->children_made = false;
  return H2OK_OK;
}

int32_t h2ok_hist_get_nid_host(h2ok_ctx* c, int32_t* nid_out, int32_t* leaf_out) {
  REQUIRE(c != nullptr, "null context");
  if (!c->have_targets) return fail(H2OK_ERR_STATE, "%s", "get_nid_host before set_targets_host");
  CK(cudaSetDevice(c->device));
  if (nid_out) {
    CK(c->nid.ensure((size_t)c->rows));
    CK(engine_nid(c->eng, c->nid.p, c->stream));
    CK(cudaMemcpyAsync(nid_out, c->nid.p, sizeof(int32_t) * (size_t)c->rows, cudaMemcpyDeviceToHost, c->stream));
  }
  if (leaf_out) CK(cudaMemcpyAsync(leaf_out, c->eng.leaf, sizeof(int32_t) * (size_t)c->rows, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return H2OK_OK;
}

int32_t h2ok_hist_get_quant(h2ok_ctx* c, h2ok_quant* out) {
  REQUIRE(c && out, "null pointer");
  if (!c->have_targets) return fail(H2OK_ERR_STATE, "%s", "get_quant before set_targets_host");
  *out = c->q;
  return H2OK_OK;
}

int32_t h2ok_hist_last_level_stats(h2ok_ctx* c, h2ok_stats* out) {
  REQUIRE(c && out, "null pointer");
  if (c->timed) {  // waits only for the histogram kernels, which callers have normally synchronised past already
    CK(cudaEventSynchronize(c->ev1));
    CK(cudaEventElapsedTime(&c->stats.build_ms, c->ev0, c->ev1));
    c->timed = false;
    if (c->rec_err.p) {
      int32_t e = 0;
      CK(cudaMemcpyAsync(&e, c->rec_err.p, sizeof(e), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      c->stats.range_errors = e;
    }
    if (c->eng_ready && c->eng.err) {  // (node, column) pairs wider than a tree_level_bins hint allowed
      int32_t e = 0;
      CK(cudaMemcpyAsync(&e, c->eng.err, sizeof(e), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      c->stats.range_errors += e;
    }
  }
  *out = c->stats;
  return H2OK_OK;
}

}  // extern "C"
