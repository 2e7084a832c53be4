// kernels.h — host-side launch interface between capi.cu and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/h2ok_hist.h"

namespace h2ok {

constexpr int kHistThreads = 1024;

struct HistArgs {
  const float* X;
  int64_t ld, rows;
  int cols;
  const uint2* rec;    // packed row records (pack_rows_kernel): node + wyq + wyyq, 8 B per row
  const int32_t* wq;   // NULL when unweighted
  const h2ok_col_desc* descs;  // [n_nodes][cols]
  int nb1;                     // max_bins + 1
  int node_base, node_cnt;     // node group handled by this launch
  int copies;                  // privatised sub-histograms per (node,bin), power of two <= 32
  int64_t chunk_rows;          // rows per CTA, multiple of 4*kHistThreads
  int64_t* acc;                // [n_nodes][cols][nb1][3]
  uint32_t* min2_enc;          // [n_nodes][cols]
  uint32_t* maxin_enc;
};

struct HistPlan {
  int group = 1, passes = 1, copies = 1, chunks = 1;
  int64_t chunk_rows = 0;
  size_t smem = 0;
};

HistPlan plan_hist(int n_nodes, int cols, int64_t rows, int nb1);
size_t hist_smem_bytes(int node_cnt, int nb1, int copies);

cudaError_t launch_synth(int shape, uint64_t seed, int64_t row0, int64_t rows, int cols, int64_t ld, int na_permille,
                         int weight_mode, float* X, float* y, float* w, cudaStream_t s);
cudaError_t launch_synth_cols(int shape, uint64_t seed, int64_t row0, int64_t rows, int col0, int ncols, int64_t ld,
                              int na_permille, float* X, cudaStream_t s);
cudaError_t launch_init_enc(uint32_t* mn, uint32_t* mx, int64_t n, cudaStream_t s);
cudaError_t launch_col_ranges(const float* X, int64_t ld, int64_t rows, int cols, uint32_t* mn, uint32_t* mx, cudaStream_t s);
cudaError_t launch_target_absmax(const float* y, const float* w, int64_t rows, double* out, cudaStream_t s);
cudaError_t launch_quantize_rows(const float* y, const float* w, int64_t rows, const h2ok_quant& q, int32_t* wq,
                                 int32_t* wyq, int32_t* wyyq, cudaStream_t s);
// the same with the scales in device memory (no host round trip between absmax and the quantised targets)
cudaError_t launch_choose_quant_dev(const double* absmax3, int weighted, int qbits, h2ok_quant* q_dev, cudaStream_t s);
cudaError_t launch_quantize_rows_dev(const float* y, const float* w, int64_t rows, bool weighted, const h2ok_quant* q_dev,
                                     int32_t* wq, int32_t* wyq, int32_t* wyyq, cudaStream_t s);
cudaError_t launch_finalize_dev(const int64_t* acc, int64_t n_node_cols, int nb1, const h2ok_quant* q_dev, double* hist,
                                cudaStream_t s);
cudaError_t launch_bin(const float* x, int64_t n, const h2ok_col_desc& d, int32_t* out, cudaStream_t s);
cudaError_t launch_pack_rows(const int32_t* nid, const int32_t* wq, const int32_t* wyq, const int32_t* wyyq, int64_t rows,
                             int64_t padded, int n_nodes, uint2* rec, int32_t* err, cudaStream_t s);
cudaError_t launch_hist_build(const HistArgs& base, const HistPlan& plan, int n_nodes, bool weighted, cudaStream_t s);
cudaError_t launch_finalize(const int64_t* acc, const uint32_t* mn, const uint32_t* mx, int64_t n_node_cols, int nb1,
                            const h2ok_quant& q, double* hist, float* min2, float* maxin, cudaStream_t s);

// a7: the best split point of one (node, column) pair (DTree.findBestSplitPoint + Split.splat), before bestCol
struct PairCand {
  double key;  // se_left + se_right
  double seL, seR, nL, nR, pL, pR, se_before;
  int32_t bin, na, found;  // found: 0 none, 1 ranged split, 2 categorical split (bins sorted by mean response)
  float splat;
};
// pairs [pair0, pair1) of the flat [n_nodes * cols] list, from the fp64 histogram (hist != NULL) or straight from the
// int64 accumulators with q's scales (hist == NULL); cand is indexed by pair
// (q_dev != NULL: q's scales are read from device memory instead)
// (node_class != NULL: a forest — q_dev is an array and node_class[node] picks the scales of the node's class tree)
cudaError_t launch_split_points(const double* hist, const int64_t* acc, const h2ok_quant* q, const h2ok_quant* q_dev,
                                const h2ok_col_desc* descs,
                                int64_t pair0, int64_t pair1, int nb1, double min_rows, double msi, PairCand* cand,
                                cudaStream_t s, const unsigned char* node_class = nullptr, int cols = 0,
                                bool categorical = false);
cudaError_t launch_best_col(const PairCand* cand, int n_nodes, int cols, h2ok_split* splits, cudaStream_t s);
// the bit sets of the level's categorical splits (found == 2 candidates that won their node), computed from the
// histogram pairs [pair0, pair1) this rank holds: into the split records (bits == NULL) or into bits[n_nodes][32]
// (zeros for nodes whose winning pair lies elsewhere), which launch_set_cat_bits copies into the records
cudaError_t launch_cat_bits(const double* hist, const int64_t* acc, const h2ok_quant* q, const h2ok_quant* q_dev,
                            const h2ok_col_desc* descs, int n_nodes, int cols, int64_t pair0, int64_t pair1, int nb1,
                            h2ok_split* splits, uint32_t* bits, cudaStream_t s, const unsigned char* node_class = nullptr);
cudaError_t launch_set_cat_bits(const uint32_t* bits, int n_nodes, h2ok_split* splits, cudaStream_t s);
// both steps over every pair; cand: [n_nodes * cols] scratch
cudaError_t launch_find_splits(const double* hist, const h2ok_col_desc* descs, int n_nodes, int cols, int nb1,
                               double min_rows, double msi, PairCand* cand, h2ok_split* splits, cudaStream_t s);
// tmp_descs: [2*n_nodes][cols] scratch; tmp_valid: [2*n_nodes] scratch
cudaError_t launch_next_level(const h2ok_split* splits, const h2ok_col_desc* descs, const float* min2, const float* maxin,
                              int n_nodes, int cols, const h2ok_params& p, h2ok_col_desc* tmp_descs, int32_t* tmp_valid,
                              h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children, cudaStream_t s);
// the same from the encoded observed ranges (what the histogram pass and the cross-rank reduce produce)
cudaError_t launch_next_level_enc(const h2ok_split* splits, const h2ok_col_desc* descs, const uint32_t* min2_enc,
                                  const uint32_t* maxin_enc, int n_nodes, int cols, const h2ok_params& p,
                                  h2ok_col_desc* tmp_descs, int32_t* tmp_valid, h2ok_col_desc* child_descs,
                                  int32_t* child_ids, int32_t* n_children, cudaStream_t s);
cudaError_t launch_next_level_noalloc(const h2ok_split* splits, const h2ok_col_desc* descs, const float* min2,
                                      const float* maxin, int n_nodes, int cols, const h2ok_params& p,
                                      h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children,
                                      cudaStream_t s);
cudaError_t launch_apply_splits(const float* X, int64_t ld, int64_t rows, const h2ok_split* splits,
                                const int32_t* child_ids, int gid_base, int n_nodes, int32_t* nid, int32_t* leaf,
                                cudaStream_t s);

cudaError_t launch_gbm_residual(int dist, const float* label, const float* f, const float* w, int64_t rows,
                                double den_scale, float* res, int32_t* denq, cudaStream_t s);
cudaError_t launch_gbm_residual_multi(int K, const float* label, const float* f, int64_t ld, const float* w, int64_t rows,
                                      double den_scale, float* res, int32_t* denq, cudaStream_t s);
cudaError_t launch_gbm_gamma_pass(const int32_t* leaf, const int32_t* numq, const int32_t* denq, int64_t rows, int n_codes,
                                  int64_t* num_acc, int64_t* den_acc, cudaStream_t s);
cudaError_t launch_gbm_add_gamma(const int32_t* leaf, int64_t rows, const double* gamma, int n_codes, double learn_rate,
                                 float* f, cudaStream_t s);

}  // namespace h2ok
