/*
 * hist_oracle.h — CPU ORACLE for the GBM histogram path.  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker or as the timed CPU baseline.  Nothing under h2o-kubernetes_b200/
 * may import, link or call it.
 *
 * PARITY UNPINNED: the reference tree (/root/reference, a Rust Kubernetes CLI) holds no implementation,
 * test, golden vector or fixture for this path; the arithmetic lives in H2O-3 (ai.h2o:h2o-algos, package
 * hex.tree), which the reference only names as a container image with tag `latest`
 * (src/k8s/templates.rs:29-30, src/k8s/mod.rs:115) — no pinned version, no jar, no JVM on this machine.
 * Every function here restates the H2O-3 unit named in its comment AS RECALLED from H2O-3's public source
 * (h2o-algos/src/main/java/hex/tree/{DHistogram,DTree,ScoreBuildHistogram}.java, 3.30-era, contemporary with
 * the reference); where a detail matters for bit-exactness this file is the normative definition.
 */
#ifndef H2OK_HIST_ORACLE_H
#define H2OK_HIST_ORACLE_H

#include "../include/h2ok_hist.h"

#ifdef __cplusplus
extern "C" {
#endif

/* generator; identical integer/fp32 recipe is evaluated on the device */
uint64_t h2ok_oracle_splitmix64(uint64_t z);
float h2ok_oracle_gen_x(int32_t shape, uint64_t seed, int64_t row, int32_t col, int32_t na_permille);
float h2ok_oracle_gen_y(uint64_t seed, int64_t row, int32_t cols);
float h2ok_oracle_gen_w(uint64_t seed, int64_t row, int32_t weight_mode);
void h2ok_oracle_synth_fill(int32_t shape, uint64_t seed, int64_t row0, int64_t rows, int32_t cols, int64_t ld,
                            int32_t na_permille, int32_t weight_mode, float* X, float* y, float* w);

/* a1 */
int32_t h2ok_oracle_make_desc(double min, double maxex, int32_t is_int, int32_t nbins, h2ok_col_desc* out);
double h2ok_oracle_find_maxex(double maxin, int32_t is_int);
void h2ok_oracle_col_ranges(const float* X, int64_t ld, int64_t rows, int32_t cols, float* mn, float* mx);
int32_t h2ok_oracle_root_descs(const float* X, int64_t ld, int64_t rows, int32_t cols, const h2ok_params* p,
                               const int32_t* is_int, h2ok_col_desc* out);
/* a2 */
int32_t h2ok_oracle_bin(float x, const h2ok_col_desc* d);
void h2ok_oracle_bin_many(const float* x, int64_t n, const h2ok_col_desc* d, int32_t* bins);

/* quantisation shared definition */
void h2ok_oracle_target_absmax(const float* y, const float* w, int64_t rows, double* absmax3);
void h2ok_oracle_choose_quant(const double* absmax3, int32_t weighted, int32_t qbits, h2ok_quant* q);
void h2ok_oracle_quantize_rows(const float* y, const float* w, int64_t rows, const h2ok_quant* q, int32_t* wq,
                               int32_t* wyq, int32_t* wyyq);

/* a3/a4: fp64 accumulation of the unquantised fp32 inputs (north_star's oracle).  abs_wy (may be NULL) receives
 * sum |w*y| per bin, [n_nodes][cols][max_bins+1], the denominator of the relative tolerance. threads<=1: scalar. */
void h2ok_oracle_build_level(const float* X, int64_t ld, int64_t rows, int32_t cols, const float* y, const float* w,
                             const int32_t* nid, int32_t n_nodes, const h2ok_col_desc* descs, int32_t max_bins,
                             double* hist, double* abs_wy, float* min2, float* maxin, int32_t threads);
/* same walk over the fixed-point targets: every channel is an exact integer → bit-exact check of the kernels */
void h2ok_oracle_build_level_q(const float* X, int64_t ld, int64_t rows, int32_t cols, const int32_t* nid,
                               const int32_t* wq, const int32_t* wyq, const int32_t* wyyq, int32_t n_nodes,
                               const h2ok_col_desc* descs, int32_t max_bins, int64_t* acc, float* min2, float* maxin);
void h2ok_oracle_finalize(const int64_t* acc, int64_t n, int32_t max_bins, const h2ok_quant* q, double* hist);

/* a7 */
void h2ok_oracle_find_splits(const double* hist, const h2ok_col_desc* descs, int32_t n_nodes, int32_t cols,
                             int32_t max_bins, double min_rows, double min_split_improvement, h2ok_split* splits);
/* a8 */
void h2ok_oracle_next_level(const h2ok_split* splits, const h2ok_col_desc* descs, const float* min2,
                            const float* maxin, int32_t n_nodes, int32_t cols, const h2ok_params* p, int32_t depth,
                            h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children);
/* a6 */
void h2ok_oracle_apply_splits(const float* X, int64_t ld, int64_t rows, const h2ok_split* splits,
                              const int32_t* child_ids, int32_t gid_base, int32_t n_nodes, int32_t* nid, int32_t* leaf);

/* a9 */
void h2ok_oracle_gbm_residual(int32_t dist, const float* label, const float* f, const float* w, int64_t rows,
                              double den_scale, float* res, int32_t* denq);
void h2ok_oracle_gbm_residual_multi(int32_t K, const float* label, const float* f, int64_t ld, const float* w, int64_t rows,
                                    double den_scale, float* res, int32_t* denq);
void h2ok_oracle_gbm_gamma_pass(const int32_t* leaf, const int32_t* numq, const int32_t* denq, int64_t rows,
                                int32_t n_codes, int64_t* num_acc, int64_t* den_acc);
void h2ok_oracle_gbm_add_gamma(const int32_t* leaf, int64_t rows, const double* gamma, int32_t n_codes, double learn_rate,
                               float* f);

int32_t h2ok_oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
