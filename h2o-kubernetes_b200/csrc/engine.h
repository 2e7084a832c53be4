// engine.h — resident row-wise tree engine (SURVEY.md §8 a3/a4/a6 at depth).
//
// The stateless histogram kernel (hist_kernels.cu) walks one column of the rows in their original order, so at depth
// d a CTA sees rows of all 2^d nodes: the per-row descriptor comes from shared memory and 64 nodes x 257 bins leave
// room for a single histogram copy — its floor is ~23 shared-memory wavefronts per 32 cells.
//
// The engine turns the problem by 90 degrees.  X is kept ROW-major (one row = cols consecutive floats, 256 B at 64
// columns) and never moves.  What is regrouped per level is a 12-byte row record {row, wyq, wyyq} (+ wq on weighted
// shards): the records are kept grouped by node (H2O does the same per chunk: ScoreBuildHistogram2 buckets row
// indices by nid before calling updateHisto).  A histogram CTA owns a contiguous run of records, i.e. the rows of
// ONE node at a time:
//   * a warp takes one row at a time; lane L owns columns {2L, 2L+1} of a 64-column block, so the row is one
//     coalesced 256-byte load whose every sector is used — the gather costs no extra HBM traffic;
//   * the node's descriptors for the lane's columns sit in registers: no per-row table lookups at any depth;
//   * the shared histogram is laid out [column slot][bin][channel][lane]: a lane only ever touches bank == lane,
//     so every shared atomic is conflict-free at every depth and for every data distribution, with ONE copy;
//   * the row record is the same for all lanes: broadcast loads, amortised over 64 cells;
//   * the rows of a warp's next batch are pulled into L2 while the current batch is accumulated (one prefetch per
//     row in which every lane names its own sector), so row loads are L2 hits and no register holds data in flight.
// Integer accumulation makes the result independent of the row order, so the engine's histograms are bit-identical
// to the stateless path's.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/h2ok_hist.h"

namespace h2ok {

constexpr int kEngMaxNodes = 1024;  // >= H2OK_MAX_LEVEL_NODES + 1
constexpr int kEngMaxClasses = 16;  // class trees a forest builds in lockstep

struct Engine {
  int64_t rows = 0;  // rows of the shard
  int64_t ldt = 0;   // floats per row of XT (cols rounded up to a multiple of 8: rows start on 32-byte sectors)
  int64_t ld8 = 0;   // bytes per column of code8 (rows rounded up to 128)
  int cols = 0;
  int sms = 0;         // SMs of the device: the histogram pass runs one CTA per SM and column block
  int route_ctas = 0;  // CTAs of the routing / regrouping kernels (each owns one contiguous range of rows)
  bool weighted = false;
  float* XT = nullptr;       // [rows][ldt] row-major copy of the shard (dense storage)
  // u8 storage instead (row_engine.cu "u8 storage"): pre-binned integer codes, one byte per cell, 255 = NA
  bool u8 = false;
  unsigned char* XB = nullptr;  // [rows][ldb] row-major codes (code8 holds the same codes column-major)
  int64_t ldb = 0;
  int32_t* load_err = nullptr;  // != 0: a cell loaded as fp32 was not a code (device)
  bool load_checked = true, load_failed = false;  // host copy of it, read once after such a load
  // sparse storage instead (row_engine.cu "sparse storage"): per column the list of its nonzero cells {row, value}
  bool sparse = false;
  const uint2** csc_ptr = nullptr;   // [cols] device table of the columns' lists
  int64_t* csc_len = nullptr;        // [cols] their lengths
  std::vector<void*> csc_blocks;     // the allocations behind them (one per loaded block of columns)
  int64_t nnz = 0;                   // nonzero cells of the shard
  unsigned long long* node_tot = nullptr;  // [kEngMaxNodes][3] rows / sum wyq / sum wyyq of every node of the level
  uint2* rec = nullptr;              // [rows][K] {node, wyq, wyyq} of every row packed in 8 bytes, rebuilt per level
  // a forest (sparse storage): K class trees in lockstep.  nid16 / leaf are then [K][rows], the targets [K][ld_targets],
  // q_dev [K], and every node of the flat list knows its class tree
  int K = 1, K_alloc = 1;
  int64_t ld_targets = 0;            // != 0: a forest is being built
  unsigned char* node_class[2] = {nullptr, nullptr};
  uint8_t* code8 = nullptr;  // [cols][ld8] order-preserving 8-bit code of every cell (255 = NaN), for routing
  uint32_t* thr = nullptr;   // [cols][256] the codes' thresholds (encoded floats, ascending; unused = 0xFFFFFFFF)
  uint32_t* cmin = nullptr;  // [cols] encoded column minimum / maximum of the shard (NaN skipped), taken while loading
  uint32_t* cmax = nullptr;
  // row records grouped by node, double buffered: the row, its fixed-point targets {wyq, wyyq}, its weight
  int32_t* gidx[2] = {nullptr, nullptr};
  uint2* gval[2] = {nullptr, nullptr};
  int32_t* gw[2] = {nullptr, nullptr};  // weighted shards only
  int32_t* seg[2] = {nullptr, nullptr}; // [n_nodes + 1] first record of every node of the level; seg[n_nodes] = live rows
  int cur = 0;
  int n_nodes = 0;                      // nodes of the current level (0: no tree in progress)
  uint16_t* nid16 = nullptr;            // [rows] node of every row in the current level, 0xFFFF = in a leaf
  int32_t* cta_cnt = nullptr;           // [kEngMaxNodes][route_ctas] rows per (child, routing CTA), then scanned
  int32_t* leaf = nullptr;              // [rows] leaf code of every row (original order), -1 while still in the tree
  int32_t* mnode[2] = {nullptr, nullptr};  // [kEngMaxNodes][2] centring offsets {M1, M2} of every node (row_engine.cu)
  // the current tree's targets in original row order (the caller's arrays: they must stay valid until the tree is done)
  const int32_t* wq_rows = nullptr;
  const int32_t* wyq_rows = nullptr;
  const int32_t* wyyq_rows = nullptr;
  double scale_wy = 0, scale_wyy = 0;   // fixed-point scales of the tree's targets (0: no centring)
  const h2ok_quant* q_dev = nullptr;    // the same scales in device memory, when the tree was started without a host copy
  int32_t* na_flag = nullptr;           // != 0 once a NaN cell has been seen while loading (device)
  bool na_known = false, has_na = true; // host copy of it, read once per data set: NaN-free shards skip the NA test
  int32_t* err = nullptr;               // (node, column) pairs that had more bins than a level's shared histogram held
};

cudaError_t engine_alloc(Engine& e, int64_t rows, int cols, bool weighted, bool sparse = false, bool u8 = false);
// u8 storage: columns [col0, col0 + ncols) of a column-major block of codes (one byte per cell)
cudaError_t engine_load_bins(Engine& e, int col0, int ncols, const unsigned char* B, int64_t ld, cudaStream_t s);
// u8 storage: did every fp32 cell handed to engine_load_cols hold a code?  (one blocking read after such a load)
cudaError_t engine_check_load(Engine& e, cudaStream_t s, bool* bad);
// a2 as a pre-binning pass: codes of a column-major fp32 block against one (device) descriptor per column; NaN and
// untracked columns give 255
cudaError_t launch_prebin(const float* X, int64_t ld, int64_t rows, int ncols, const h2ok_col_desc* d_descs, unsigned char* B,
                          int64_t ldb, cudaStream_t s);
void engine_free(Engine& e);
// switch the shard between unweighted and weighted trees (allocates the weight arrays the first time)
cudaError_t engine_set_weighted(Engine& e, bool weighted);
// copy columns [col0, col0 + ncols) of a column-major block (leading dimension ld) into the row-major copy
cudaError_t engine_load_cols(Engine& e, int col0, int ncols, const float* X, int64_t ld, cudaStream_t s);
// new tree: every row at the root, records in row order.  q (may be NULL) gives the targets' fixed-point scales, from
// which the children's centring offsets are derived; without it every node keeps the root's offsets.  The target
// arrays are read again by every level's regrouping: they must stay valid and unchanged until the tree is finished.
// a forest of K class trees on the sparse storage (wyq / wyyq: [K][ld] targets, q_dev: [K] scales on the device)
cudaError_t engine_set_forest(Engine& e, int K);
cudaError_t engine_begin_forest(Engine& e, int K, const int32_t* wyq, const int32_t* wyyq, int64_t ld, const h2ok_quant* q_dev,
                                cudaStream_t s);
// (q_dev: q in device memory instead — a tree then starts without a host round trip; q may be NULL with it)
cudaError_t engine_begin_tree(Engine& e, const int32_t* wq_rows, const int32_t* wyq_rows, const int32_t* wyyq_rows,
                              const h2ok_quant* q, cudaStream_t s, const h2ok_quant* q_dev = nullptr);
// one level's histogram pass.  nb1 = bins + 1 of the accumulators' layout; level_nb1 (0: same as nb1) = bins + 1 of
// the widest histogram of THIS level, which sizes the shared histogram (H2O halves nbins_top_level per level)
cudaError_t engine_build_level(Engine& e, int n_nodes, const h2ok_col_desc* descs, int nb1, int level_nb1, int64_t* acc,
                               uint32_t* min2_enc, uint32_t* maxin_enc, cudaStream_t s);
// the largest bins + 1 a level may use (16 columns per block must fit the shared memory)
int engine_max_nb1(int cols);
// route rows with the level's splits, retire rows into leaves, regroup the records by child
// (regroup == false: the tree's last level — rows only get their leaves, no records are laid out)
cudaError_t engine_apply_splits(Engine& e, const h2ok_split* splits, const int32_t* child_ids, int gid_base, int n_nodes,
                                bool regroup, cudaStream_t s);
// row -> node map of the current level in original row order (tests / inspection)
cudaError_t engine_nid(Engine& e, int32_t* nid_rows, cudaStream_t s);

}  // namespace h2ok
