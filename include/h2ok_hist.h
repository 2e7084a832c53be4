/*
 * h2ok_hist.h — C ABI of the B200-native GBM histogram path ("hist" layer).
 *
 * The reference (isgasho/h2o-kubernetes, crate `h2ok`) has NO FFI, plugin or operator interface
 * and no numeric code: the GBM it stands up runs inside the unvendored H2O-3 jar started by
 * src/k8s/templates.rs:29-30 (`java ... -jar /opt/h2oai/h2o-3/h2o.jar`, image pinned only as
 * `latest` at src/k8s/mod.rs:115).  This header therefore DEFINES the boundary; each entry point
 * names the H2O-3 unit it stands in for (package hex.tree, recalled from H2O-3's public source,
 * not present under /root/reference — see DESIGN.md "parity unpinned") and the reference line
 * through which that unit is reached today.
 *
 * Conventions inherited from the reference:
 *  - fallible operations return a status instead of aborting (Result<..> everywhere, e.g.
 *    src/k8s/mod.rs:85-93, src/main.rs:54-59): every function returns int32_t, 0 == H2OK_OK;
 *    the message is fetched with h2ok_last_error() (thread-local).
 *  - callers are single-threaded and blocking (fresh Runtime + block_on per op,
 *    src/k8s/mod.rs:86,130,167): a context is used from one thread; the *_dev entry points
 *    enqueue on the context's stream and return, the host-buffer entry points are complete on return.
 *  - no hidden state beyond the opaque context (state in the reference is an explicit file,
 *    src/main.rs:72-101).
 *
 * Data layout: X is column-major fp32, column c at X + c*ld (ld >= rows, ld % 4 == 0, base
 * 16-byte aligned) — one contiguous vector per column like an H2O Vec/Chunk.  Row-indexed arrays
 * (y, w, nid, quantised targets) have `rows` entries.  nid[r] is the index of the row's node in
 * the current level's active-node list, or a negative value if the row already sits in a leaf.
 *
 * Histogram layout: [n_nodes][cols][max_bins+1][3], channels {w, wY, wYY}; slot nbin of a
 * (node,col) is the NA bucket (DHistogram._vals, hex/tree/DHistogram.java), slots above nbin unused.
 */
#ifndef H2OK_HIST_H
#define H2OK_HIST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define H2OK_OK 0
#define H2OK_ERR_ARG 1         /* bad argument (null, size, alignment) */
#define H2OK_ERR_CUDA 2        /* a CUDA call failed; message has the CUDA error string */
#define H2OK_ERR_STATE 3       /* call sequence violated (e.g. build_level before set_data) */
#define H2OK_ERR_UNSUPPORTED 4 /* > H2OK_MAX_BINS bins, category codes past 1023, ... */

#define H2OK_MAX_BINS 1024
#define H2OK_MAX_LEVEL_NODES 1023 /* node indices travel in 10 bits of the packed row records */
/* limits of h2ok_hist_next_level_noalloc_dev: 2*n_nodes*cols and 2*n_nodes may not exceed these */
#define H2OK_NOALLOC_MAX_SLOT_COLS 16384
#define H2OK_NOALLOC_MAX_SLOTS 2048

/* NA direction of a split (DHistogram.NASplitDir). */
#define H2OK_NA_NONE 0
#define H2OK_NA_VS_REST 1 /* non-NA left, NA right */
#define H2OK_NA_LEFT 2
#define H2OK_NA_RIGHT 3

/* Distributions of the tree-level caller (a9). */
#define H2OK_DIST_GAUSSIAN 0
#define H2OK_DIST_BERNOULLI 1
#define H2OK_DIST_MULTINOMIAL 2

/* How h2ok_hist_loop_level combines the ranks' histograms (a5, MRTask.reduce). */
#define H2OK_REDUCE_SCATTER 0 /* reduce-scatter over (node, column) pairs + sharded split search + all-gather of candidates */
#define H2OK_REDUCE_ALL 1     /* plain all-reduce: every rank holds the whole reduced histogram */
#define H2OK_COMM_ID_BYTES 128 /* size of the rendezvous id of h2ok_hist_comm_unique_id (an ncclUniqueId) */

/* Flags of h2ok_hist_tree_create. */
#define H2OK_TREE_WEIGHTED 1 /* trees with per-row weights (8 more bytes per row) */
#define H2OK_TREE_SPARSE 2   /* store only the nonzero cells: for matrices that are mostly exact zeros (see below) */
#define H2OK_TREE_U8 4       /* pre-binned input: one byte per cell, an integer code 0..254, 255 = NA (see below) */

/* Synthetic shapes of SURVEY.md §8(d). */
#define H2OK_SHAPE_HIGGS28 2 /* config 2: 28 HIGGS-shaped columns */
#define H2OK_SHAPE_WIDE64 3  /* config 3: the 28-column recipe tiled to any width with affine offsets */
#define H2OK_SHAPE_SPARSE 4  /* config 4: 90 % exact zeros, else u */

typedef struct h2ok_ctx h2ok_ctx;
typedef struct h2ok_group h2ok_group; /* several devices driven by one host thread (h2ok_hist_group_*) */

/* a1 — one DHistogram's constructor fields for one (node, column).  nbin == 0 means the column is
 * not tracked for this node (constant / all-NA there) and contributes nothing. */
typedef struct h2ok_col_desc {
  double min;     /* inclusive lower bound (_min) */
  double maxex;   /* exclusive upper bound (_maxEx) */
  double step;    /* _step: nbin/(maxex-min), or 1.0 for integer columns with few values */
  int32_t nbin;   /* _nbin <= max_bins */
  int32_t is_int; /* _isInt: 0 real, 1 integer, 2 categorical (integer codes 0 .. 1023, at most nbins of them) */
} h2ok_col_desc;

/* a7 — DTree.Split as chosen by UndecidedNode.bestCol.  col < 0: the node does not split.
 * equal == 0: a ranged split, x >= splat goes right.  equal == 2: a categorical split (DTree.Split._equal 2 / 3 with its
 * IcedBitSet): the column is categorical (h2ok_col_desc.is_int == 2), `bin` is the cut position in the order of the
 * bins sorted by mean response, splat is unused, and a row whose category code is c goes right when bit c of cat_bits
 * is set (word c / 32, bit c % 32).  NaN follows na_dir either way. */
#define H2OK_CAT_WORDS 32 /* category codes 0 .. 1023 */
typedef struct h2ok_split {
  int32_t col;
  int32_t bin;    /* bins [0,bin) left, [bin,nbin) right */
  int32_t na_dir; /* H2OK_NA_* */
  float splat;    /* row routing threshold: x >= splat goes right (Split.splat) */
  double se_before, se_left, se_right;
  double n_left, n_right;       /* weighted row counts */
  double pred_left, pred_right; /* wY/w of each side */
  int32_t equal;
  int32_t pad;
  uint32_t cat_bits[H2OK_CAT_WORDS];
} h2ok_split;

/* Fixed-point scales for the integer accumulators.  All three are powers of two so that the
 * conversion back to double is exact.  weighted == 0: w is absent (== 1) and channel 0 counts rows. */
typedef struct h2ok_quant {
  double scale_w, scale_wy, scale_wyy;
  int32_t weighted;
  int32_t qbits; /* |quantised value| <= 2^qbits; at most 25 */
} h2ok_quant;

typedef struct h2ok_params {
  int32_t nbins;           /* 256 in BASELINE configs 2/3, 64 in config 4 */
  int32_t nbins_top_level; /* H2O halves this per level down to nbins; set == nbins to keep it fixed */
  int32_t max_depth;
  int32_t pad;
  double min_rows;              /* 10 */
  double min_split_improvement; /* 1e-5 */
} h2ok_params;

typedef struct h2ok_stats {
  float build_ms;   /* device time of the last build_level's kernels */
  int32_t launches; /* kernels launched by the last build_level */
  int32_t passes;   /* node-group passes over X */
  int32_t copies;   /* privatised sub-histograms per (node,col) */
  int32_t range_errors; /* != 0: some quantised target was outside the packed range (|wyq| < 2^26, wyyq, wq < 2^27) */
  int32_t pad;
  double algo_bytes;
} h2ok_stats;

/* Host buffers that h2ok_hist_loop_level fills for the level it ran; any pointer may be NULL (not wanted), and the
 * whole struct may be NULL.  Shapes as in the per-step entry points.  With several ranks under H2OK_REDUCE_SCATTER a
 * rank's `hist` receives only the (node, column) pairs it reduced — pairs [done_pair0, done_pair1) of h2ok_loop_view,
 * the rest of the buffer is left untouched — so that the reduced histogram reaches host memory once, in parallel over
 * the ranks' links; everything else is complete on every rank. */
typedef struct h2ok_level_out {
  double* hist;               /* [n_nodes][cols][max_bins+1][3] */
  float* min2;                /* [n_nodes][cols] */
  float* maxin;               /* [n_nodes][cols] */
  h2ok_split* splits;         /* [n_nodes] */
  h2ok_col_desc* child_descs; /* [n_children][cols] */
  int32_t* child_ids;         /* [2*n_nodes] */
} h2ok_level_out;

/* Where the loop stands, and device pointers to the level that h2ok_hist_loop_level finished last (valid until the
 * next loop_level / loop_begin_tree call; tests and callers that keep everything on the device read them). */
typedef struct h2ok_loop_view {
  int32_t depth, n_nodes, gid_base; /* the level the NEXT loop_level call will run; n_nodes == 0: tree finished */
  int32_t max_nodes, max_bins, cols, rank, n_ranks, reduce_mode;
  int32_t done_depth;               /* -1: no level has run since loop_begin_tree */
  int32_t done_nodes, done_children, done_gid_base;
  int32_t n_trees; /* class trees built in lockstep (h2ok_hist_loop_create_forest); 1: one tree.  With n_trees = K > 1:
                    * d_leaf is [K][rows], d_wyq / d_wyyq are [K][rows rounded up to 4], the node list of a level is flat
                    * (class tree 0's nodes, then tree 1's ...) and leaf codes use the flat tree-wide node ids */
  int64_t done_pair0, done_pair1;   /* (node, column) pairs whose reduced sums this rank holds in d_done_acc */
  h2ok_quant quant;                 /* fixed-point scales of the current tree's targets */
  const h2ok_col_desc* d_descs;     /* [n_nodes][cols] descriptors of the next level */
  const int32_t *d_wq, *d_wyq, *d_wyyq; /* the tree's quantised targets, original row order (d_wq NULL-able) */
  const int32_t* d_leaf;            /* [rows] leaf codes, -1 while a row is still in the tree */
  const h2ok_col_desc* d_done_descs;
  const int64_t* d_done_acc;        /* [done_nodes][cols][max_bins+1][3] int64 fixed-point sums */
  const uint32_t *d_done_min2_enc, *d_done_maxin_enc;
  const h2ok_split* d_done_splits;
  const int32_t* d_done_child_ids;
} h2ok_loop_view;

/* ---- lifecycle ------------------------------------------------------------------------------ */
int32_t h2ok_hist_create(int32_t device, h2ok_ctx** out);
int32_t h2ok_hist_destroy(h2ok_ctx* ctx);
const char* h2ok_last_error(void);
int32_t h2ok_hist_abi_version(void);
/* use_own == 0: run every call on `stream` (a cudaStream_t; NULL is CUDA's legacy default stream, a valid choice).
 * use_own != 0: go back to the context's own non-blocking stream (the state after create). */
int32_t h2ok_hist_set_stream(h2ok_ctx* ctx, void* stream, int32_t use_own);
int32_t h2ok_hist_sync(h2ok_ctx* ctx);
/* Blocking device -> host copy ordered after everything enqueued on the context's stream (how a caller without its own
 * CUDA runtime reads the device pointers of h2ok_loop_view). */
int32_t h2ok_hist_copy_to_host(h2ok_ctx* ctx, const void* d_src, void* host_dst, size_t bytes);

/* ---- device-resident entry points (pointers are device pointers; asynchronous) -------------- */

/* Counter-based generator of SURVEY.md §8(d); fills rows [row0,row0+rows) of the global matrix.
 * d_w may be NULL.  na_permille: per-mille of cells set to NaN.  weight_mode: 0 w=1, 1 w=0.5+u with 2% zeros. */
int32_t h2ok_hist_synth_fill_dev(h2ok_ctx* ctx, int32_t shape, uint64_t seed, int64_t row0, int64_t rows,
                                 int32_t cols, int64_t ld, int32_t na_permille, int32_t weight_mode,
                                 float* d_X, float* d_y, float* d_w);
/* The same generator for columns [col0, col0 + ncols) only (column c at d_X + (c - col0) * ld): lets a caller stage a
 * wide matrix block by block (h2ok_hist_tree_load_cols_dev). */
int32_t h2ok_hist_synth_fill_cols_dev(h2ok_ctx* ctx, int32_t shape, uint64_t seed, int64_t row0, int64_t rows,
                                      int32_t col0, int32_t ncols, int64_t ld, int32_t na_permille, float* d_X);

/* Root ranges (DHistogram.initialHist reads Vec.min()/max()): finite-or-inf min and max per column, NaN skipped.
 * Output is the order-preserving uint32 encoding (see h2ok_hist_decode_f32) so ranks can be combined with integer min/max. */
int32_t h2ok_hist_col_ranges_dev(h2ok_ctx* ctx, const float* d_X, int64_t ld, int64_t rows, int32_t cols,
                                 uint32_t* d_min_enc, uint32_t* d_max_enc);

/* max |w|, |w*y|, |w*y*y| over rows → d_absmax[3] (doubles); used to choose h2ok_quant scales. */
int32_t h2ok_hist_target_absmax_dev(h2ok_ctx* ctx, const float* d_y, const float* d_w, int64_t rows, double* d_absmax);

/* Per-row fixed-point targets: wq = rint(w*scale_w) (absent when !weighted), wyq = rint(w*y*scale_wy),
 * wyyq = rint(w*y*y*scale_wyy), all computed in fp64, round-half-even. */
int32_t h2ok_hist_quantize_rows_dev(h2ok_ctx* ctx, const float* d_y, const float* d_w, int64_t rows,
                                    const h2ok_quant* q, int32_t* d_wq, int32_t* d_wyq, int32_t* d_wyyq);

/* a2 — DHistogram.bin for n values against one descriptor (test hook for bit-exact bin indices). */
int32_t h2ok_hist_bin_dev(h2ok_ctx* ctx, const float* d_x, int64_t n, const h2ok_col_desc* desc, int32_t* d_bins);

/* a3/a4 — ScoreBuildHistogram local phase: accumulate every active row into its (node,col) histogram.
 * d_acc: int64 [n_nodes][cols][max_bins+1][3] fixed-point sums, zeroed here when clear != 0.
 * d_min2_enc/d_maxin_enc: uint32 [n_nodes][cols] encoded observed min / max (DHistogram._min2/_maxIn), reset when clear != 0.
 * d_descs: device array [n_nodes][cols]. d_wq may be NULL (unweighted).  Targets must satisfy |wyq| < 2^26 and
 * 0 <= wyyq, wq < 2^27 (true for h2ok_hist_choose_quant's scales); violations are counted in h2ok_stats.range_errors
 * and the offending rows are left out. */
int32_t h2ok_hist_build_level_dev(h2ok_ctx* ctx, const float* d_X, int64_t ld, int64_t rows, int32_t cols,
                                  const int32_t* d_nid, const int32_t* d_wq, const int32_t* d_wyq,
                                  const int32_t* d_wyyq, int32_t n_nodes, const h2ok_col_desc* d_descs,
                                  int32_t max_bins, int32_t clear, int64_t* d_acc, uint32_t* d_min2_enc,
                                  uint32_t* d_maxin_enc);

/* Passes over X (= histogram kernel launches) h2ok_hist_build_level_dev makes for a level of this shape: the nodes are
 * taken in groups whose histograms and descriptor tables fit one CTA's 227 KB of shared memory.  Host arithmetic only;
 * returns 0 for shapes build_level_dev rejects.  (Not a status code.) */
int32_t h2ok_hist_build_level_passes(int32_t n_nodes, int32_t cols, int64_t rows, int32_t max_bins);

/* a5 tail — after the cross-rank reduce (MRTask.reduce / DHistogram.add, done by the caller on d_acc with an
 * integer sum and on the encoded ranges with integer min/max): fixed-point → double, encoded → float. */
int32_t h2ok_hist_finalize_level_dev(h2ok_ctx* ctx, const int64_t* d_acc, const uint32_t* d_min2_enc,
                                     const uint32_t* d_maxin_enc, int32_t n_nodes, int32_t cols, int32_t max_bins,
                                     const h2ok_quant* q, double* d_hist, float* d_min2, float* d_maxin);

/* a7 — DTree.findBestSplitPoint per (node,col) + UndecidedNode.bestCol per node. */
int32_t h2ok_hist_find_splits_dev(h2ok_ctx* ctx, const double* d_hist, const h2ok_col_desc* d_descs, int32_t n_nodes,
                                  int32_t cols, int32_t max_bins, double min_rows, double min_split_improvement,
                                  h2ok_split* d_splits);

/* a8 — DTree.Split.nextLevelHistos for both children of every split node.  d_child_ids[2*i+way] receives the
 * child's index in the next level's active list or -1; d_child_descs is [2*n_nodes][cols] (compacted, first
 * *d_n_children entries valid).  depth = level of the parents. */
int32_t h2ok_hist_next_level_dev(h2ok_ctx* ctx, const h2ok_split* d_splits, const h2ok_col_desc* d_descs,
                                 const float* d_min2, const float* d_maxin, int32_t n_nodes, int32_t cols,
                                 const h2ok_params* params, int32_t depth, h2ok_col_desc* d_child_descs,
                                 int32_t* d_child_ids, int32_t* d_n_children);
/* (depth is reserved: the children's bin count follows from the parent's descriptor, max(parent nbin >> 1, nbins).) */

/* a8 without any allocation — same inputs, outputs and results as h2ok_hist_next_level_dev.  No counterpart in the
 * reference (nor in H2O-3): it exists for hosts that cannot allocate device memory while they run — a process whose
 * GPU memory is wholly owned by a pool allocator, or a loop that cannot tolerate the device-wide synchronisation of
 * cudaMalloc/cudaFree.  h2ok_hist_next_level_dev grows a scratch buffer inside the context on demand; this variant
 * keeps that scratch in a fixed 520 KiB device array that is part of the library image (resident on every GPU that
 * loads the library until the process exits) and calls no cudaMalloc, cudaFree or other allocating routine.
 * Limits: 2*n_nodes*cols <= H2OK_NOALLOC_MAX_SLOT_COLS and 2*n_nodes <= H2OK_NOALLOC_MAX_SLOTS, otherwise
 * H2OK_ERR_ARG with a message naming the limit (128 nodes x 64 columns fits exactly).
 * The first call on each device looks up the array's device address (with that device current) and caches it; make
 * that first call before starting any stream capture.  All calls on a device share the one array, which keeps the
 * previous call's contents (the flags are cleared at the start of each call): two calls that overlap on the GPU —
 * different streams or contexts — race on it, so serialise them. */
int32_t h2ok_hist_next_level_noalloc_dev(h2ok_ctx* ctx, const h2ok_split* d_splits, const h2ok_col_desc* d_descs,
                                         const float* d_min2, const float* d_maxin, int32_t n_nodes, int32_t cols,
                                         const h2ok_params* params, int32_t depth, h2ok_col_desc* d_child_descs,
                                         int32_t* d_child_ids, int32_t* d_n_children);

/* a6 — ScoreBuildHistogram.score_decide / DecidedNode.getChildNodeID: move each active row to its child.
 * A row that retires gets nid -1 and d_leaf[r] = 3*(gid_base+node) + k, k = 0 node did not split, 1 / 2 the left /
 * right side has no child (gid_base: tree-wide id of this level's node 0; d_leaf may be NULL). */
int32_t h2ok_hist_apply_splits_dev(h2ok_ctx* ctx, const float* d_X, int64_t ld, int64_t rows, const h2ok_split* d_splits,
                                   const int32_t* d_child_ids, int32_t gid_base, int32_t n_nodes, int32_t* d_nid,
                                   int32_t* d_leaf);

/* ---- resident tree engine (device-resident; the fast path at depth) ---------------------------------------------
 * Same results as build_level_dev / apply_splits_dev, bit for bit.  The shard lives inside the context as a ROW-major
 * copy of X that never moves; per level only a 12-byte row record {row, wyq, wyyq} (+ wq) is kept grouped by node (what
 * ScoreBuildHistogram2 does with row indices per chunk, hex/tree/ScoreBuildHistogram2.java).  A histogram CTA then
 * works on one node at a time: a warp takes whole rows (one coalesced load, every sector used), each lane owns its
 * columns' descriptors in registers and its own shared-memory bank, so there are no descriptor lookups and no
 * bank conflicts at any depth.
 * Call order: tree_create -> tree_load_cols_dev for every block of columns (in any order; the caller's column-major
 * block can be freed or reused afterwards) -> per tree: tree_begin_dev -> per level: tree_build_level_dev, [reduce /
 * finalize / find_splits / next_level as usual], tree_apply_splits_dev.  Device memory: rows * (4 * cols rounded up
 * to 8 + cols + 30) bytes, + 8 per row for weighted shards.  Limits: rows < 2^31 - 4096, cols < 32768; every bin count up to
 * H2OK_MAX_BINS runs (64 columns per block up to 299 bins + NA, 32 up to 599, 16 beyond).
 *
 * H2OK_TREE_SPARSE (an input format beyond dense fp32; H2O keeps mostly-zero columns in sparse
 * chunks too): the shard is stored column by column as the list of its nonzero cells {row, value} — 8 bytes per
 * nonzero instead of 4 bytes per cell — and the histogram pass streams those lists, every lane on a nonzero cell, with
 * the row's node and targets looked up by row; what the zero cells contribute follows exactly from the node totals
 * afterwards.  No row records are kept or regrouped.  Same results bit for bit, any density (it pays below ~25 %
 * nonzeros).  Unweighted trees only.  Algorithmic bytes per level: nnz * 8 + rows * 10 (row -> node map and targets,
 * once) + histogram.
 *
 * H2OK_TREE_U8 (pre-binned input, the other format of SURVEY.md §8(f) row 4): every cell is an integer code 0..254,
 * 255 = NA — bins computed once (h2ok_hist_prebin_dev against the root descriptors, i.e. global bins as in H2O's
 * histogram types with fixed split points) or supplied by the caller.  The shard keeps one byte per cell (64 B per
 * row at 64 columns) and the trees are built on the CODES as integer columns (is_int == 1 for every column, nbins
 * >= 255): DHistogram's integer rule then gives step-1 bins at every level, so the histogram pass bins with one
 * integer subtraction — no fp64 — and routing compares the byte itself.  Results equal the standard path run on the
 * code matrix, bit for bit.  Unweighted trees.  Algorithmic bytes per level: rows * cols * 1 + rows * 8 + histogram.
 * Load with h2ok_hist_tree_load_bins_dev (or tree_load_cols_dev with fp32 cells that hold the codes). */
int32_t h2ok_hist_tree_create(h2ok_ctx* ctx, int64_t rows, int32_t cols, int32_t flags /*H2OK_TREE_**/);
int32_t h2ok_hist_tree_destroy(h2ok_ctx* ctx);
/* columns [col0, col0 + ncols) from a column-major device block (column c at d_X + (c - col0) * ld), rows in
 * original order */
int32_t h2ok_hist_tree_load_cols_dev(h2ok_ctx* ctx, int32_t col0, int32_t ncols, const float* d_X, int64_t ld);
/* H2OK_TREE_U8 shards: columns [col0, col0 + ncols) from a column-major device block of codes (column c at
 * d_bins + (c - col0) * ld, one byte per cell). */
int32_t h2ok_hist_tree_load_bins_dev(h2ok_ctx* ctx, int32_t col0, int32_t ncols, const uint8_t* d_bins, int64_t ld);
/* a2 as a pre-binning pass (DHistogram.bin once per cell): d_bins[c * ld_bins + r] = bin of d_X[c * ld + r] against
 * descs[c] (HOST array, one descriptor per column, nbin <= 255), 255 for NaN and for untracked columns.  Synchronous. */
int32_t h2ok_hist_prebin_dev(h2ok_ctx* ctx, const float* d_X, int64_t ld, int64_t rows, int32_t ncols,
                             const h2ok_col_desc* descs, uint8_t* d_bins, int64_t ld_bins);
/* new tree: per-row fixed-point targets in ORIGINAL row order (d_wq NULL when unweighted); all rows back at the root.
 * The target arrays are read again by every tree_apply_splits_dev of the tree: keep them valid and unchanged until
 * the tree is finished.
 * q = the h2ok_quant the targets were quantised with (host pointer; may be NULL): the engine stores every node's
 * targets relative to an offset near the node's mean, which it derives from the parent's split record and these
 * scales, so that its 32-bit histogram words rarely wrap.  Results do not depend on it. */
int32_t h2ok_hist_tree_begin_dev(h2ok_ctx* ctx, const int32_t* d_wq, const int32_t* d_wyq, const int32_t* d_wyyq,
                                 const h2ok_quant* q);
/* a3/a4 for the current level; outputs as h2ok_hist_build_level_dev.  n_nodes must be the number of children the
 * previous tree_apply_splits_dev produced (1 after tree_begin_dev). */
int32_t h2ok_hist_tree_build_level_dev(h2ok_ctx* ctx, int32_t n_nodes, const h2ok_col_desc* d_descs, int32_t max_bins,
                                       int32_t clear, int64_t* d_acc, uint32_t* d_min2_enc, uint32_t* d_maxin_enc);
/* Optional hint for the NEXT tree_build_level_dev only: no node of that level has more than level_bins bins (H2O
 * halves nbins_top_level per level: level d uses max(nbins, nbins_top_level >> d)).  The shared histogram is then
 * sized for the level instead of for max_bins, which keeps 64 columns per block at 256 bins and lets a 1024-bin root
 * level run (16 columns per block).  A descriptor with more bins than the hint is not accumulated and counted in
 * h2ok_stats.range_errors. */
int32_t h2ok_hist_tree_level_bins(h2ok_ctx* ctx, int32_t level_bins);
/* a6 for the current level + regrouping of the records by child.  Child ids must be numbered in slot order
 * (2 * node + way), which is what next_level_dev produces.  last_level != 0: the tree ends here — rows get their leaf
 * codes and no records are laid out for a next level (pass 0 whenever another tree_build_level_dev follows). */
int32_t h2ok_hist_tree_apply_splits_dev(h2ok_ctx* ctx, const h2ok_split* d_splits, const int32_t* d_child_ids,
                                        int32_t gid_base, int32_t n_nodes, int32_t last_level);
/* row -> node of the current level / leaf codes, in original row order */
int32_t h2ok_hist_tree_nid_dev(h2ok_ctx* ctx, int32_t* d_nid_out);
int32_t h2ok_hist_tree_leaf_dev(h2ok_ctx* ctx, int32_t* d_leaf_out);

/* The same generator straight into the engine: fills every column of the shard created by tree_create with rows
 * [row0, row0 + rows) of the global synthetic matrix (no caller-side staging; synchronous). */
int32_t h2ok_hist_tree_load_synth(h2ok_ctx* ctx, int32_t shape, uint64_t seed, int64_t row0, int32_t na_permille);

/* ---- the level loop inside the library (SharedTree.buildLayer, hex/tree/SharedTree.java) ------------------------
 * One call per tree level: ScoreBuildHistogram2's local phase on the resident engine, MRTask.reduce over the ranks
 * (NCCL, called from inside the library), DTree.findBestSplitPoint + bestCol, Split.nextLevelHistos, score_decide and
 * the regrouping of the row records — enqueued back to back without a host round trip; the call returns when the
 * level's small results (and the host buffers asked for) have arrived, while routing is still running.
 * Call order: create -> [comm_init] -> tree_create + data (tree_load_cols_dev / tree_load_synth) -> loop_create ->
 * per tree: loop_begin_tree_* -> loop_level until *n_children == 0 or max_depth levels have run.
 * Results are bit-identical to the per-step entry points and to any number of ranks (integer sums).
 *
 * Ranks.  One context per process and GPU: rank 0 calls h2ok_hist_comm_unique_id, hands the H2OK_COMM_ID_BYTES bytes
 * to the other processes by any means, and every rank calls h2ok_hist_comm_init (collective, blocking) BEFORE
 * loop_create.  Rows are sharded (each rank loads its own rows); descriptors, splits and child ids are replicated.
 * The reference's counterpart is the H2O cloud the StatefulSet forms (src/k8s/templates.rs:18-32, port 54321): pods
 * there, GPUs of one node here.  A single-threaded caller that owns several GPUs uses h2ok_hist_group_* below. */
int32_t h2ok_hist_comm_unique_id(void* id_out /*[H2OK_COMM_ID_BYTES]*/);
int32_t h2ok_hist_comm_init(h2ok_ctx* ctx, const void* id /*[H2OK_COMM_ID_BYTES]*/, int32_t rank, int32_t n_ranks);
/* Root descriptors from the column ranges of all ranks (DHistogram.initialHist), level buffers for
 * min(2^(max_depth-1), H2OK_MAX_LEVEL_NODES) nodes.  Collective when n_ranks > 1.  is_int: [cols] or NULL. */
int32_t h2ok_hist_loop_create(h2ok_ctx* ctx, const h2ok_params* params, const int32_t* is_int);
/* A forest: n_trees (<= 16) trees on the same rows built in lockstep — the K class trees of one multinomial iteration
 * (GBM.buildNextKTrees builds K trees per iteration; BASELINE config 4 has K = 10).  H2OK_TREE_SPARSE shards only: the
 * column lists are streamed ONCE per level for all the trees (every nonzero cell serves the K trees, whose row records
 * sit side by side), which is SURVEY.md §8(d)'s "fused level".  Every tree has its own targets and fixed-point scales
 * and equals the tree loop_begin_tree_* would have built alone, bit for bit; the level's node list is flat (tree 0's
 * nodes, then tree 1's, ...), n_trees * 2^(max_depth-1) must not exceed H2OK_MAX_LEVEL_NODES.  h2ok_level_out.hist is
 * not delivered for a forest. */
int32_t h2ok_hist_loop_create_forest(h2ok_ctx* ctx, const h2ok_params* params, const int32_t* is_int, int32_t n_trees);
/* New forest: d_y = the n_trees target vectors, [n_trees][ld] (device; read by work enqueued on the context's stream). */
int32_t h2ok_hist_loop_begin_forest_dev(h2ok_ctx* ctx, const float* d_y, int64_t ld);
/* The fixed-point scales of the current forest's trees (out: [n_trees], host); blocks until they have arrived. */
int32_t h2ok_hist_loop_forest_quants(h2ok_ctx* ctx, h2ok_quant* out);
int32_t h2ok_hist_loop_destroy(h2ok_ctx* ctx);
int32_t h2ok_hist_loop_root_descs(h2ok_ctx* ctx, h2ok_col_desc* descs_out /*[cols], host*/);
int32_t h2ok_hist_loop_set_reduce(h2ok_ctx* ctx, int32_t mode /*H2OK_REDUCE_**/);
/* checksum != 0: keep the running result checksums of h2ok_hist_loop_checksum (one extra read of every level's
 * accumulators); time_phases != 0: bracket the phases of every level with CUDA events (h2ok_hist_loop_phase_stats). */
int32_t h2ok_hist_loop_set_options(h2ok_ctx* ctx, int32_t checksum, int32_t time_phases);
/* New tree on targets y (and weights w, NULL: unweighted) in original row order: common fixed-point scales over all
 * ranks (chosen on the device: no host round trip), quantised targets, every row back at the root.  Asynchronous.
 * _dev: device pointers, read by work enqueued on the context's stream (order later writes to them after it);
 * _host: the library copies them in; _synth: the generator's y / w for rows from row0. */
int32_t h2ok_hist_loop_begin_tree_dev(h2ok_ctx* ctx, const float* d_y, const float* d_w);
int32_t h2ok_hist_loop_begin_tree_host(h2ok_ctx* ctx, const float* y_host, const float* w_host);
int32_t h2ok_hist_loop_begin_tree_synth(h2ok_ctx* ctx, uint64_t seed, int64_t row0, int32_t weight_mode);
/* One level.  out may be NULL.  *n_nodes_done: nodes of the level that ran; *n_children: nodes of the next level
 * (0: the tree is finished; every row then has its leaf code). */
int32_t h2ok_hist_loop_level(h2ok_ctx* ctx, const h2ok_level_out* out, int32_t* n_nodes_done, int32_t* n_children);
int32_t h2ok_hist_loop_view(h2ok_ctx* ctx, h2ok_loop_view* view_out);
/* Device milliseconds of the timed levels since the last reset, added up per tree depth d < max_depth:
 * count_out[d] levels of nodes_out[d] nodes, ms_out[5*d + k] with k = 0 the histogram kernel alone, 1 clear + histogram
 * pass, 2 reduce over ranks, 3 split search + candidate gather + children, 4 routing + regrouping.  The events are read
 * back many levels after they were recorded, so timing never holds the host up; this call synchronises the stream.
 * nodes_out may be NULL; max_depth 0 with reset != 0 just clears the sums. */
int32_t h2ok_hist_loop_phase_stats(h2ok_ctx* ctx, int32_t reset, int32_t max_depth, int32_t* count_out, int32_t* nodes_out,
                                   float* ms_out);
/* 64-bit checksums of the current tree so far, the same for any number of ranks: out[0] every level's reduced integer
 * accumulators, out[1] the splits {col, bin, na_dir, splat} of every node, out[2] the leaf code of every row.  Each is
 * sum(splitmix64(global index ^ salt) * value) mod 2^64 (tests/oracle_lib.py has the host restatement).  row0 = global
 * index of this rank's first row.  Collective when n_ranks > 1. */
int32_t h2ok_hist_loop_checksum(h2ok_ctx* ctx, int64_t row0, uint64_t* out /*[3]*/);

/* ---- a group of devices behind one host thread -------------------------------------------------------------------
 * What a caller shaped like the reference needs to use all GPUs of a node: `h2ok` is single-threaded and blocking
 * (src/main.rs:18-37; one Runtime + block_on per operation, src/k8s/mod.rs:85-93), so the library drives the devices.
 * A group owns one context per device and one NCCL communicator over them (ncclCommInitAll); rows are sharded in
 * contiguous blocks, member i holding rows [i*R/n, (i+1)*R/n).  Every call is the per-context call of the same name run
 * on all members, the NCCL phases grouped.  Results are the single-device results, bit for bit. */
int32_t h2ok_hist_group_create(const int32_t* device_ids, int32_t n_dev, h2ok_group** out);
int32_t h2ok_hist_group_destroy(h2ok_group* group);
int32_t h2ok_hist_group_size(h2ok_group* group); /* number of members (not a status code) */
/* Member i's context (for the per-context calls: tree_create + tree_load_cols_dev with caller-made device blocks,
 * loop_view, copy_to_host, last_level_stats ...), the global index of its first row and its row count. */
int32_t h2ok_hist_group_member(h2ok_group* group, int32_t i, h2ok_ctx** ctx_out, int64_t* row0_out, int64_t* rows_out);
/* Data: the generator's rows_total x cols matrix, every member its own block / the caller's host columns (all rows). */
int32_t h2ok_hist_group_load_synth(h2ok_group* group, int32_t shape, uint64_t seed, int64_t rows_total, int32_t cols,
                                   int32_t na_permille);
int32_t h2ok_hist_group_set_data_host(h2ok_group* group, const float* const* cols_host, int64_t rows_total, int32_t cols);
int32_t h2ok_hist_group_loop_create(h2ok_group* group, const h2ok_params* params, const int32_t* is_int);
int32_t h2ok_hist_group_set_options(h2ok_group* group, int32_t reduce_mode, int32_t checksum, int32_t time_phases);
int32_t h2ok_hist_group_begin_tree_synth(h2ok_group* group, uint64_t seed, int32_t weight_mode);
/* y_host / w_host: all rows_total targets in original row order (w_host NULL: unweighted) */
int32_t h2ok_hist_group_begin_tree_host(h2ok_group* group, const float* y_host, const float* w_host);
/* One level on every member.  `out` is ONE set of host buffers: under H2OK_REDUCE_SCATTER each member writes the
 * (node, column) pairs it reduced into the same `hist`, which is therefore complete on return and travels over all the
 * devices' links at once; ranges, splits and children are written by member 0. */
int32_t h2ok_hist_group_level(h2ok_group* group, const h2ok_level_out* out, int32_t* n_nodes_done, int32_t* n_children);
int32_t h2ok_hist_group_checksum(h2ok_group* group, uint64_t* out /*[3]*/);
int32_t h2ok_hist_group_hist_ms(h2ok_group* group, float* ms_out); /* last level's histogram pass, slowest member */
int32_t h2ok_hist_group_sync(h2ok_group* group);

/* ---- a9: the tree-level caller (GBM.buildNextKTrees) ---------------------------------------------------------
 * ComputePredAndRes: d_res[r] = label - mu(f) (gaussian mu = f, bernoulli mu = 1/(1+exp(-f))), and, when d_denq is
 * given, the fixed-point GammaPass denominator rint(w * mu * (1 - mu) * den_scale) (bernoulli) or rint(w * den_scale)
 * / 1 (gaussian).  d_w may be NULL. */
int32_t h2ok_hist_gbm_residual_dev(h2ok_ctx* ctx, int32_t dist, const float* d_label, const float* d_f, const float* d_w,
                                   int64_t rows, double den_scale, float* d_res, int32_t* d_denq);
/* GammaPass: integer sums of d_numq and d_denq (NULL: row count) per leaf code (d_leaf as written by apply_splits);
 * d_num_acc / d_den_acc: int64 [n_codes], overwritten.  Combine across ranks with an integer sum. */
/* ComputePredAndRes for K-class (multinomial) GBM, all classes in one pass over f: p = softmax(f[0..K)[r]) in fp32
 * (max-subtracted, expf, summed in class order), d_res[k*ld + r] = (label[r] == k) - p_k, and
 * d_denq[k*ld + r] = rint(w * |res| * (1 - |res|) * den_scale), the GammaPass denominator.  d_label holds class
 * ids 0..K-1 as floats; d_f, d_res, d_denq are [K][ld].  The class-k tree is then built on d_res + k*ld and its
 * leaf values are (K-1)/K * sum(res) / sum(|res|(1-|res|)) (GBM.java GammaPass, m1class). */
int32_t h2ok_hist_gbm_residual_multi_dev(h2ok_ctx* ctx, int32_t n_classes, const float* d_label, const float* d_f,
                                         int64_t ld, const float* d_w, int64_t rows, double den_scale, float* d_res,
                                         int32_t* d_denq);
int32_t h2ok_hist_gbm_gamma_pass_dev(h2ok_ctx* ctx, const int32_t* d_leaf, const int32_t* d_numq, const int32_t* d_denq,
                                     int64_t rows, int32_t n_codes, int64_t* d_num_acc, int64_t* d_den_acc);
/* AddTreeContributions: f[r] += (float)(learn_rate * d_gamma[d_leaf[r]]). */
int32_t h2ok_hist_gbm_add_gamma_dev(h2ok_ctx* ctx, const int32_t* d_leaf, int64_t rows, const double* d_gamma,
                                    int32_t n_codes, double learn_rate, float* d_f);

/* a1 — DHistogram.initialHist from column ranges that the caller has already combined across ranks (plain host
 * arithmetic, no device work): constant / all-NA / infinite-range columns get nbin 0. */
int32_t h2ok_hist_root_descs_from_ranges(const float* col_min, const float* col_max, int32_t cols,
                                         const h2ok_params* params, const int32_t* is_int, h2ok_col_desc* descs_out);
/* Power-of-two fixed-point scales from max|w|, max|w*y|, max|w*y*y| (host arithmetic).
 * Limit: the split search and finalize turn the integer sums into doubles, which is exact (and therefore equal to the
 * oracle's sequential sums, bit for bit) while |sum| < 2^53 quanta, i.e. while (rows over ALL ranks) * 2^qbits < 2^53:
 * 2^28 = 268M rows at the default qbits = 25.  Beyond that ranks still agree with each other (the sums are integers)
 * but equality with a sequential fp64 walk is no longer guaranteed; lower qbits for larger jobs. */
int32_t h2ok_hist_choose_quant(const double* absmax3, int32_t weighted, int32_t qbits, h2ok_quant* out);

/* ---- host-buffer entry points (what an H2O-side caller binds; synchronous) ------------------ */

/* Library copies the columns (cols_host[c] points at `rows` floats). */
int32_t h2ok_hist_set_data_host(h2ok_ctx* ctx, const float* const* cols_host, int64_t rows, int32_t cols);
/* New tree: copy y (and w, may be NULL), choose scales, quantise, put every row in node 0. */
int32_t h2ok_hist_set_targets_host(h2ok_ctx* ctx, const float* y_host, const float* w_host);
/* Root descriptors from the resident data (DHistogram.initialHist). */
int32_t h2ok_hist_root_descs(h2ok_ctx* ctx, const h2ok_params* params, const int32_t* is_int /*[cols] or NULL*/,
                             h2ok_col_desc* descs_out /*[cols]*/);
/* One level: histogram + observed ranges to host buffers. */
int32_t h2ok_hist_build_level(h2ok_ctx* ctx, int32_t n_nodes, const h2ok_col_desc* descs, int32_t max_bins,
                              double* hist_out, float* min2_out, float* maxin_out);
/* Split search on the level just built; splits to host. */
int32_t h2ok_hist_find_splits(h2ok_ctx* ctx, double min_rows, double min_split_improvement, h2ok_split* splits_out);
/* Children of the level just searched: descriptors + child ids to host; *n_children_out active nodes next. */
int32_t h2ok_hist_next_level(h2ok_ctx* ctx, const h2ok_params* params, int32_t depth, h2ok_col_desc* child_descs_out,
                             int32_t* child_ids_out, int32_t* n_children_out);
/* Route rows with the splits found (device-resident copies are used; nothing crosses the bus). */
int32_t h2ok_hist_apply_splits(h2ok_ctx* ctx);
/* Copy the row→node map / leaf assignment back (tests). */
int32_t h2ok_hist_get_nid_host(h2ok_ctx* ctx, int32_t* nid_out, int32_t* leaf_out);
int32_t h2ok_hist_get_quant(h2ok_ctx* ctx, h2ok_quant* out);
int32_t h2ok_hist_last_level_stats(h2ok_ctx* ctx, h2ok_stats* out);

#ifdef __cplusplus
}
#endif
#endif /* H2OK_HIST_H */
