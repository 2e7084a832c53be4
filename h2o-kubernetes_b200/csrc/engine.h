// engine.h — resident, node-grouped tree engine (SURVEY.md §8 a3/a4/a6 at depth).
//
// The stateless histogram kernel (hist_kernels.cu) takes rows in their original order, so at depth d a CTA sees
// rows of all 2^d nodes: the per-row descriptor comes from shared memory and 64 nodes x 257 bins leave room for a
// single histogram copy, i.e. 3-4-way bank conflicts — its floor is ~23 shared-memory wavefronts per 32 cells.
// The engine instead keeps the shard physically grouped by node.  H2O does the same per chunk (ScoreBuildHistogram2
// buckets row indices by nid before calling updateHisto); here the data itself is regrouped, because a 4-byte gather
// would waste 7/8 of every 32-byte sector.  The regrouping write is fused into the histogram pass: level d's pass
// reads the layout grouped by level d-1 (each segment holds the two children of one parent), accumulates with both
// descriptors in registers and 32 private histogram copies per child, and stores every x at its position in the
// layout grouped by level d, which the next pass reads.  Integer accumulation makes the result independent of the
// row order, so the engine's histograms are bit-identical to the stateless path's.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/h2ok_hist.h"

namespace h2ok {

constexpr int kEngTile = 4096;      // positions per partition tile
constexpr int kEngSegAlign = 128;   // every segment starts at a multiple of 128 positions (one warp's block)
constexpr int kEngMaxGroups = 2050; // children of one level + the dead group

struct Engine {
  int64_t rows = 0;  // rows of the shard
  int64_t cap = 0;   // positions per layout buffer (rows + alignment padding)
  int cols = 0;
  bool weighted = false;
  // layout buffers, double buffered: positions hold rows grouped by node; padding positions have rowid -1
  float* X[2] = {nullptr, nullptr};       // [cols][cap]
  uint2* rec[2] = {nullptr, nullptr};     // packed (node, wyq, wyyq) per position
  int32_t* wq[2] = {nullptr, nullptr};    // weights per position (weighted only)
  int32_t* rowid[2] = {nullptr, nullptr}; // original row of each position
  int32_t* pnode[2] = {nullptr, nullptr}; // node of each position at the current level (-1: in a leaf / padding)
  int32_t* seg[2] = {nullptr, nullptr};   // per layout: [0..kEngMaxGroups) lo, [kEngMaxGroups..2*) hi, then used extent
  int32_t* seg_child = nullptr;           // [2*kEngMaxGroups] nodes (at the level being built) held by each segment
  int cur = 0;                            // which buffer holds the current layout
  int n_seg = 1;                          // segments of the current layout (the dead zone comes after them)
  bool dest_valid = false;                // apply() has prepared the next layout; the next pass must write it
  // per-position scratch
  int32_t* newnode = nullptr;
  int32_t* dest = nullptr;
  int32_t* tile_counts = nullptr;  // [tiles][groups]
  int64_t tile_cap = 0;
  // per-row (original order) arrays
  int32_t* leaf = nullptr;
  // work items of the histogram pass
  int4* items = nullptr;  // {segment, first position, end position, 0}
  int32_t* n_items = nullptr;
  int max_items = 0;
  int64_t chunk_rows = 0;
};

cudaError_t engine_alloc(Engine& e, int64_t rows, int cols, bool weighted);
void engine_free(Engine& e);
// the identity layout: position p = row p
cudaError_t engine_reset_layout(Engine& e, cudaStream_t s);
// new tree: gather the quantised targets into layout order, every row back at the root
cudaError_t engine_begin_tree(Engine& e, const int32_t* wq_rows, const int32_t* wyq_rows, const int32_t* wyyq_rows,
                              cudaStream_t s);
// one level's histogram pass (+ the fused regrouping write when apply() prepared one)
cudaError_t engine_build_level(Engine& e, int n_nodes, const h2ok_col_desc* descs, int nb1, int64_t* acc,
                               uint32_t* min2_enc, uint32_t* maxin_enc, cudaStream_t s);
// route rows with the level's splits, retire rows into leaves, prepare the layout grouped by the children
cudaError_t engine_apply_splits(Engine& e, const h2ok_split* splits, const int32_t* child_ids, int gid_base, int n_nodes,
                                bool last_level, cudaStream_t s);
// row -> node map of the current level in original row order (tests / inspection)
cudaError_t engine_nid(Engine& e, int32_t* nid_rows, cudaStream_t s);

}  // namespace h2ok
