// tree_driver.cu — SharedTree.buildLayer inside the library (h2ok_hist_loop_*, h2ok_hist_comm_*, h2ok_hist_group_*).
//
// One call per tree level enqueues, without a host round trip in between:
//   clear -> histogram pass (row_engine.cu) -> [reduce over ranks] -> split points per (node, column) -> [all-gather of
//   the candidates] -> bestCol -> children's descriptors -> routing + regrouping of the row records
// and returns once the level's (small) results are on the host; the routing kernels are still running then and the
// next level's pass queues up behind them.
//
// Across ranks (rows are sharded) the exchange is the "preferred form": the int64 accumulators are
// reduce-scattered over the flat (node, column) list, every rank searches the split points of the pairs it received,
// the 80-byte candidates are all-gathered and every rank picks the same bestCol from identical bytes.  The histogram
// crosses the wire once (half an all-reduce) and the split search is spread over the ranks.  Integer sums make the
// result independent of the number of ranks, bit for bit.  H2OK_REDUCE_ALL keeps the plain all-reduce (every rank then
// holds the whole reduced histogram).
//
// NCCL is looked up at run time (dlopen of libnccl.so.2, the copy already loaded into the process if there is one), so
// single-GPU callers need no NCCL at all.
#include <dlfcn.h>
#include <math.h>
#include <nccl.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "common.cuh"
#include "ctx.h"

namespace h2ok {

// ------------------------------------------------------------------------------------------------ NCCL, late bound
namespace {

struct NcclApi {
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclReduceScatter) ReduceScatter = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  bool ok = false;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the process's own copy first (e.g. torch's)
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) return;
#define H2OK_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(h, "nccl" #name))
    H2OK_SYM(GetUniqueId); H2OK_SYM(CommInitRank); H2OK_SYM(CommInitAll); H2OK_SYM(CommDestroy); H2OK_SYM(AllReduce);
    H2OK_SYM(ReduceScatter); H2OK_SYM(AllGather); H2OK_SYM(GroupStart); H2OK_SYM(GroupEnd); H2OK_SYM(GetErrorString);
#undef H2OK_SYM
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommInitAll && api.CommDestroy && api.AllReduce &&
             api.ReduceScatter && api.AllGather && api.GroupStart && api.GroupEnd && api.GetErrorString;
  });
  return api.ok ? &api : nullptr;
}

#define NK(call)                                                                                              \
  do {                                                                                                        \
    ncclResult_t _r = (call);                                                                                 \
    if (_r != ncclSuccess) return fail(H2OK_ERR_CUDA, "%s: %s", #call, nccl_api()->GetErrorString(_r));       \
  } while (0)

// ---------------------------------------------------------------------------------------------- result checksum
// Order-independent 64-bit sums: every element is multiplied by a hash of its GLOBAL index and added up mod 2^64, so
// ranks can hash their own part and the parts add up to the same value whatever the sharding.
__device__ __forceinline__ unsigned long long warp_add_u64(unsigned long long v) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__global__ void hash_i64_kernel(const int64_t* __restrict__ p, int64_t n, uint64_t idx0, uint64_t salt,
                                unsigned long long* __restrict__ out) {
  unsigned long long s = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    s += splitmix64((idx0 + (uint64_t)i) ^ salt) * (unsigned long long)p[i];
  s = warp_add_u64(s);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}
__global__ void hash_i32_kernel(const int32_t* __restrict__ p, int64_t n, uint64_t idx0, uint64_t salt,
                                unsigned long long* __restrict__ out) {
  unsigned long long s = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    s += splitmix64((idx0 + (uint64_t)i) ^ salt) * (unsigned long long)(long long)p[i];
  s = warp_add_u64(s);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}
// the tree's shape: {col, bin, na_dir, bits of splat} of every node, indexed by the tree-wide node id
__global__ void hash_splits_kernel(const h2ok_split* __restrict__ sp, int n, int gid_base, uint64_t salt,
                                   unsigned long long* __restrict__ out) {
  unsigned long long s = 0;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) {
    const h2ok_split t = sp[g];
    const uint64_t b = (uint64_t)(gid_base + g) * 4;
    s += splitmix64((b + 0) ^ salt) * (unsigned long long)(long long)t.col;
    s += splitmix64((b + 1) ^ salt) * (unsigned long long)(long long)t.bin;
    s += splitmix64((b + 2) ^ salt) * (unsigned long long)(long long)t.na_dir;
    s += splitmix64((b + 3) ^ salt) * (unsigned long long)__float_as_uint(t.splat);
    if (t.col >= 0 && t.equal == 2)  // a categorical split: its bit set, word by word
      for (int w = 0; w < H2OK_CAT_WORDS; w++)
        s += splitmix64((b + 1) ^ salt ^ ((uint64_t)(w + 1) << 32)) * (unsigned long long)t.cat_bits[w];
  }
  s = warp_add_u64(s);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

constexpr uint64_t kSaltAcc = 0xACC0ull, kSaltSplit = 0x5911ull, kSaltLeaf = 0x1EAFull;

inline unsigned grid_for(int64_t n, int t, int cap) {
  int64_t g = (n + t - 1) / t;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

// ------------------------------------------------------------------------------------------------ loop state
struct TreeLoop {
  h2ok_params params{};
  std::vector<int32_t> is_int;
  std::vector<h2ok_col_desc> root;
  int cols = 0, max_nodes = 0, max_bins = 0, nb1 = 0;
  int forest = 1;  // class trees built in lockstep (h2ok_hist_loop_create_forest); 1: one tree
  int reduce_mode = H2OK_REDUCE_SCATTER;
  bool has_cat = false;  // some column is categorical: the split search sorts its bins and splits carry bit sets
  // level buffers (device)
  h2ok_col_desc* descs[2] = {nullptr, nullptr};  // [max_nodes][cols], this level / the next
  // accumulators and ranges are double buffered by level parity: the NEXT level's histogram pass is enqueued before
  // the call returns (see enqueue_hist), while the level that just ran must stay readable (h2ok_hist_loop_view)
  int64_t* acc2[2] = {nullptr, nullptr};         // [pairs padded to a multiple of the ranks][nb1][3]
  uint32_t* rng2[2] = {nullptr, nullptr};        // [2][max_nodes * cols] encoded observed min | max
  int64_t* acc = nullptr;                        // the current level's
  uint32_t* rng = nullptr;
  bool pre = false;                              // the current level's histogram pass is already in the stream

  int64_t* done_acc = nullptr;                   // the level h2ok_hist_loop_view shows
  uint32_t* done_rng = nullptr;
  PairCand* cand = nullptr;                      // [padded pairs]
  h2ok_split* splits = nullptr;                  // [max_nodes]
  int32_t* child_ids = nullptr;                  // [2 * max_nodes]
  int32_t* n_children = nullptr;                 // [1]
  h2ok_col_desc* d_root = nullptr;               // [cols] the root descriptors, kept on the device
  h2ok_col_desc* tmp_descs = nullptr;            // [2 * max_nodes][cols] scratch of next_level
  int32_t* tmp_valid = nullptr;                  // [2 * max_nodes]
  double* hist = nullptr;                        // fp64 copy of acc, made only when a caller asks for it
  float* rngf = nullptr;                         // [2][max_nodes * cols] decoded ranges, likewise
  unsigned long long* sums = nullptr;            // [3] checksums {acc, splits, leaf}
  unsigned long long* sums_tmp = nullptr;        // [3] this rank's share on its way through the all-reduce
  size_t pairs_cap = 0, acc_pairs = 0;
  int32_t* h_nch = nullptr;                      // pinned
  cudaStream_t out_stream = nullptr;             // device -> host copies of a level's results
  cudaEvent_t ev_ready = nullptr, ev_out = nullptr;
  // phase timing (loop_set_options): a ring of event sets, read back sixteen levels later — by then they have long
  // completed, so timing never makes the host wait for a level — and added up per tree depth
  static constexpr int kRing = 16, kMaxDepth = 64;
  struct EvSet {
    cudaEvent_t a = nullptr, h0 = nullptr, h1 = nullptr, b = nullptr, c = nullptr, d = nullptr, e = nullptr;
    int depth = 0;
    bool used = false;      // recorded (at least its first half) and not yet read back
    bool complete = false;  // ... up to the end of routing: a pass enqueued ahead is only half a level
  } ring[kRing];
  int ring_pos = 0;
  EvSet* cur_set = nullptr;
  EvSet* pre_set = nullptr;  // timing events of a histogram pass that was enqueued ahead (pre)
  double ph_sum[kMaxDepth][5] = {};  // per depth: histogram kernel alone, clear + histogram, reduce, search + children, routing
  int ph_cnt[kMaxDepth] = {};
  int ph_nodes[kMaxDepth] = {};
  // targets of the current tree (device, owned)
  DevBuf<float> y, w;
  DevBuf<int32_t> wq, wyq, wyyq;
  DevBuf<double> absmax;
  h2ok_quant* q_dev = nullptr;  // the tree's fixed-point scales, chosen on the device
  h2ok_quant* h_q = nullptr;    // pinned host copy (arrives with the first level's results)
  const float* tree_y = nullptr;  // the current tree's targets between the phases of loop_begin_tree
  const float* tree_w = nullptr;
  int64_t tree_ld = 0;
  bool have_tree = false, checksum = false;
  // tree state
  int cur = 0, n_nodes = 0, depth = 0, gid_base = 0;
  // the level just finished (what h2ok_hist_loop_view shows)
  int done_nodes = 0, done_children = 0, done_depth = -1, done_gid_base = 0, done_cur = 0;
  bool phases_timed = false;
  // pending level (between the phases of one h2ok_hist_loop_level)
  int64_t pairs = 0, chunk = 0, p0 = 0, p1 = 0;
  bool last = false;
  const h2ok_level_out* out = nullptr;
};

void comm_free(h2ok_ctx* c) {
  if (c->comm) {
    NcclApi* n = nccl_api();
    if (n) n->CommDestroy((ncclComm_t)c->comm);
    c->comm = nullptr;
  }
  c->comm_rank = 0;
  c->comm_ranks = 1;
}

void loop_free(h2ok_ctx* c) {
  TreeLoop* L = c->loop;
  if (!L) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (L->out_stream) cudaStreamSynchronize(L->out_stream);
  for (int k = 0; k < 2; k++) cudaFree(L->descs[k]);
  for (int k = 0; k < 2; k++) {
    cudaFree(L->acc2[k]);
    cudaFree(L->rng2[k]);
  }
  cudaFree(L->cand); cudaFree(L->d_root); cudaFree(L->splits); cudaFree(L->child_ids);
  cudaFree(L->n_children); cudaFree(L->tmp_descs); cudaFree(L->tmp_valid); cudaFree(L->hist); cudaFree(L->rngf);
  cudaFree(L->sums); cudaFree(L->sums_tmp); cudaFree(L->q_dev);
  if (L->h_nch) cudaFreeHost(L->h_nch);
  if (L->h_q) cudaFreeHost(L->h_q);
  if (L->out_stream) cudaStreamDestroy(L->out_stream);
  for (cudaEvent_t e : {L->ev_ready, L->ev_out})
    if (e) cudaEventDestroy(e);
  for (auto& st : L->ring)
    for (cudaEvent_t e : {st.a, st.h0, st.h1, st.b, st.c, st.d, st.e})
      if (e) cudaEventDestroy(e);
  L->y.release(); L->w.release(); L->wq.release(); L->wyq.release(); L->wyyq.release(); L->absmax.release();
  delete L;
  c->loop = nullptr;
}

namespace {

int32_t need_loop(h2ok_ctx* c, const char* who) {
  if (!c) return fail(H2OK_ERR_ARG, "%s: null context", who);
  if (!c->loop) return fail(H2OK_ERR_STATE, "%s before h2ok_hist_loop_create", who);
  return H2OK_OK;
}

// one timed level's device times into the per-depth sums (its events completed long ago, or the caller has synchronised)
int32_t drain_set(TreeLoop& L, TreeLoop::EvSet& st) {
  float ms[5];
  CK(cudaEventSynchronize(st.e));
  CK(cudaEventElapsedTime(&ms[0], st.h0, st.h1));
  CK(cudaEventElapsedTime(&ms[1], st.a, st.b));
  CK(cudaEventElapsedTime(&ms[2], st.b, st.c));
  CK(cudaEventElapsedTime(&ms[3], st.c, st.d));
  CK(cudaEventElapsedTime(&ms[4], st.d, st.e));
  if (st.depth < TreeLoop::kMaxDepth) {
    for (int k = 0; k < 5; k++) L.ph_sum[st.depth][k] += ms[k];
    L.ph_cnt[st.depth] += 1;
  }
  st.used = false;
  return H2OK_OK;
}

// ---- the phases of one level; a group of contexts (one host thread, several devices) runs each phase for every
//      member before the next, with the NCCL calls of a phase inside one ncclGroupStart / ncclGroupEnd ---------------
// Clear + histogram pass of level `depth` with at most n_bound nodes (descriptors in descs[cur]).  The pass takes the
// exact node count from the engine's own node table on the device, so it can be enqueued for the NEXT level before the
// host knows how many children there are: level_finish does that right behind routing, and the GPU never waits for
// the host's round trip between two levels (on an eighth of the rows that round trip was 0.1 ms of a 1.4 ms level).
int32_t enqueue_hist(h2ok_ctx* c, int n_bound, int depth, int cur) {
  TreeLoop& L = *c->loop;
  const int R = c->comm_ranks;
  const size_t cell = (size_t)L.nb1 * 3;
  const int64_t pairs = (int64_t)n_bound * L.cols, chunk = (pairs + R - 1) / R;
  int64_t* acc = L.acc2[depth & 1];
  uint32_t* rng = L.rng2[depth & 1];
  L.cur_set = nullptr;
  if (L.phases_timed) {
    TreeLoop::EvSet& st = L.ring[L.ring_pos];
    L.ring_pos = (L.ring_pos + 1) % TreeLoop::kRing;
    if (st.used && st.complete) {
      const int32_t rc = drain_set(L, st);
      if (rc) return rc;
    }
    st.depth = depth;
    st.used = true;
    st.complete = false;
    L.cur_set = &st;
    CK(cudaEventRecord(st.a, c->stream));
  }
  CK(cudaMemsetAsync(acc, 0, sizeof(int64_t) * cell * (size_t)(chunk * R), c->stream));
  CK(launch_init_enc(rng, rng + L.pairs_cap, pairs, c->stream));
  // H2O halves nbins_top_level per level (Split.nextLevelHistos: max(parent bins >> 1, nbins)): the shared histogram
  // is sized for the level's widest one
  int level_bins = L.params.nbins_top_level >> (depth < 30 ? depth : 30);
  if (level_bins < L.params.nbins) level_bins = L.params.nbins;
  const int level_nb1 = level_bins < L.max_bins ? level_bins + 1 : L.nb1;
  CK(cudaEventRecord(c->ev0, c->stream));
  if (L.cur_set) CK(cudaEventRecord(L.cur_set->h0, c->stream));
  CK(engine_build_level(c->eng, n_bound, L.descs[cur], L.nb1, level_nb1, acc, rng, rng + L.pairs_cap, c->stream));
  CK(cudaEventRecord(c->ev1, c->stream));
  c->timed = true;
  if (L.cur_set) {
    CK(cudaEventRecord(L.cur_set->h1, c->stream));
    CK(cudaEventRecord(L.cur_set->b, c->stream));
  }
  return H2OK_OK;
}

int32_t level_local(h2ok_ctx* c, const h2ok_level_out* out) {
  TreeLoop& L = *c->loop;
  if (!L.have_tree) return fail(H2OK_ERR_STATE, "%s", "loop_level before loop_begin_tree");
  if (L.n_nodes < 1 || L.depth >= L.params.max_depth) return fail(H2OK_ERR_STATE, "%s", "loop_level: the tree is finished");
  CK(cudaSetDevice(c->device));
  bool bad_load = false;
  CK(engine_check_load(c->eng, c->stream, &bad_load));
  if (bad_load) return fail(H2OK_ERR_ARG, "%s", "u8 storage: a cell loaded from fp32 was not an integer code in [0, 254] or NaN");
  const int n = L.n_nodes, R = c->comm_ranks;
  L.out = out;
  L.pairs = (int64_t)n * L.cols;
  L.chunk = (L.pairs + R - 1) / R;
  const bool scatter = R > 1 && L.reduce_mode == H2OK_REDUCE_SCATTER;
  L.p0 = scatter ? L.chunk * c->comm_rank : 0;
  L.p1 = scatter ? (L.p0 + L.chunk < L.pairs ? L.p0 + L.chunk : L.pairs) : L.pairs;
  if (L.p0 > L.pairs) L.p0 = L.pairs;
  L.last = L.depth + 1 >= L.params.max_depth;
  L.acc = L.acc2[L.depth & 1];
  L.rng = L.rng2[L.depth & 1];
  if (!L.pre) {  // (otherwise the pass was enqueued behind the previous level's routing)
    const int32_t rc = enqueue_hist(c, n, L.depth, L.cur);
    if (rc) return rc;
  }
  L.pre = false;
  if (L.depth < TreeLoop::kMaxDepth) L.ph_nodes[L.depth] = n;
  const size_t cell = (size_t)L.nb1 * 3;
  c->stats.passes = 1;
  c->stats.copies = 1;
  c->stats.launches = 3;
  c->stats.algo_bytes = (double)c->eng.rows * c->eng.cols * 4.0 + (double)c->eng.rows * (c->eng.weighted ? 12.0 : 8.0) +
                        (double)L.pairs * cell * 8.0;
  return H2OK_OK;
}

// MRTask.reduce / DHistogram.add across ranks
int32_t level_reduce(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  if (c->comm_ranks <= 1) return H2OK_OK;
  NcclApi* n = nccl_api();
  ncclComm_t comm = (ncclComm_t)c->comm;
  const size_t cell = (size_t)L.nb1 * 3;
  NK(n->GroupStart());  // one launch for the three collectives
  if (L.reduce_mode == H2OK_REDUCE_SCATTER)
    NK(n->ReduceScatter(L.acc, L.acc + (size_t)L.chunk * c->comm_rank * cell, (size_t)L.chunk * cell, ncclInt64, ncclSum, comm,
                        c->stream));
  else
    NK(n->AllReduce(L.acc, L.acc, (size_t)L.pairs * cell, ncclInt64, ncclSum, comm, c->stream));
  NK(n->AllReduce(L.rng, L.rng, (size_t)L.pairs, ncclUint32, ncclMin, comm, c->stream));
  NK(n->AllReduce(L.rng + L.pairs_cap, L.rng + L.pairs_cap, (size_t)L.pairs, ncclUint32, ncclMax, comm, c->stream));
  NK(n->GroupEnd());
  return H2OK_OK;
}

// DTree.findBestSplitPoint on this rank's pairs, straight from the integer accumulators
int32_t level_search(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  if (L.cur_set) CK(cudaEventRecord(L.cur_set->c, c->stream));
  CK(launch_split_points(nullptr, L.acc, nullptr, L.q_dev, L.descs[L.cur], L.p0, L.p1, L.nb1, L.params.min_rows,
                         L.params.min_split_improvement, L.cand, c->stream,
                         L.forest > 1 ? c->eng.node_class[c->eng.cur] : nullptr, L.cols, L.has_cat));
  if (L.checksum && L.p1 > L.p0) {
    const int64_t cell = (int64_t)L.nb1 * 3, cnt = (L.p1 - L.p0) * cell;
    hash_i64_kernel<<<grid_for(cnt, 256, 1184), 256, 0, c->stream>>>(L.acc + L.p0 * cell, cnt, (uint64_t)(L.p0 * cell),
                                                                    splitmix64(kSaltAcc + (uint64_t)L.depth), L.sums);
    CK(cudaGetLastError());
  }
  return H2OK_OK;
}

int32_t level_gather(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  if (c->comm_ranks <= 1 || L.reduce_mode != H2OK_REDUCE_SCATTER) return H2OK_OK;
  NcclApi* n = nccl_api();
  const size_t bytes = (size_t)L.chunk * sizeof(PairCand);
  NK(n->AllGather(reinterpret_cast<char*>(L.cand) + bytes * c->comm_rank, L.cand, bytes, ncclInt8, (ncclComm_t)c->comm,
                  c->stream));
  return H2OK_OK;
}

// bestCol, the children's descriptors, the results' way to the host, then routing + regrouping
int32_t level_finish(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  const int n = L.n_nodes, nx = 1 - L.cur;
  const h2ok_level_out* out = L.out;
  if (out && out->hist && L.forest > 1)
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "a forest level does not deliver the fp64 histogram (read the accumulators: loop_view)");
  CK(launch_best_col(L.cand, n, L.cols, L.splits, c->stream));
  if (L.has_cat)  // (reduce mode ALL: every rank holds every pair)
    CK(launch_cat_bits(nullptr, L.acc, nullptr, L.q_dev, L.descs[L.cur], n, L.cols, 0, (int64_t)n * L.cols, L.nb1, L.splits,
                       nullptr, c->stream, L.forest > 1 ? c->eng.node_class[c->eng.cur] : nullptr));
  if (L.last) {
    CK(cudaMemsetAsync(L.child_ids, 0xFF, sizeof(int32_t) * 2 * (size_t)n, c->stream));
    CK(cudaMemsetAsync(L.n_children, 0, sizeof(int32_t), c->stream));
  } else {
    CK(launch_next_level_enc(L.splits, L.descs[L.cur], L.rng, L.rng + L.pairs_cap, n, L.cols, L.params, L.tmp_descs,
                             L.tmp_valid, L.descs[nx], L.child_ids, L.n_children, c->stream));
  }
  if (L.checksum) {
    hash_splits_kernel<<<grid_for(n, 128, 8), 128, 0, c->stream>>>(L.splits, n, L.gid_base, kSaltSplit, L.sums + 1);
    CK(cudaGetLastError());
  }
  if (out && out->hist && !L.hist) CK(cudaMalloc((void**)&L.hist, sizeof(double) * (size_t)L.nb1 * 3 * L.acc_pairs));
  if (out && (out->min2 || out->maxin) && !L.rngf) CK(cudaMalloc((void**)&L.rngf, sizeof(float) * 2 * L.pairs_cap));
  if (out && out->hist && L.p1 > L.p0) {
    const int64_t cell = (int64_t)L.nb1 * 3;
    CK(launch_finalize_dev(L.acc + L.p0 * cell, L.p1 - L.p0, L.nb1, L.q_dev, L.hist + L.p0 * cell, c->stream));
  }
  if (out && (out->min2 || out->maxin))
    CK(launch_finalize(nullptr, L.rng, L.rng + L.pairs_cap, L.pairs, L.nb1, h2ok_quant{}, nullptr, L.rngf, L.rngf + L.pairs_cap,
                       c->stream));
  if (L.cur_set) CK(cudaEventRecord(L.cur_set->d, c->stream));
  // results to the host on their own stream, so that routing does not wait for the copies
  CK(cudaEventRecord(L.ev_ready, c->stream));
  CK(cudaStreamWaitEvent(L.out_stream, L.ev_ready, 0));
  CK(cudaMemcpyAsync(L.h_nch, L.n_children, sizeof(int32_t), cudaMemcpyDeviceToHost, L.out_stream));
  if (out) {
    const size_t cell = (size_t)L.nb1 * 3;
    if (out->hist && L.p1 > L.p0)
      CK(cudaMemcpyAsync(out->hist + L.p0 * cell, L.hist + L.p0 * cell, sizeof(double) * cell * (size_t)(L.p1 - L.p0),
                         cudaMemcpyDeviceToHost, L.out_stream));
    if (out->min2) CK(cudaMemcpyAsync(out->min2, L.rngf, sizeof(float) * (size_t)L.pairs, cudaMemcpyDeviceToHost, L.out_stream));
    if (out->maxin)
      CK(cudaMemcpyAsync(out->maxin, L.rngf + L.pairs_cap, sizeof(float) * (size_t)L.pairs, cudaMemcpyDeviceToHost, L.out_stream));
    if (out->splits) CK(cudaMemcpyAsync(out->splits, L.splits, sizeof(h2ok_split) * (size_t)n, cudaMemcpyDeviceToHost, L.out_stream));
    if (out->child_ids)
      CK(cudaMemcpyAsync(out->child_ids, L.child_ids, sizeof(int32_t) * 2 * (size_t)n, cudaMemcpyDeviceToHost, L.out_stream));
  }
  CK(cudaEventRecord(L.ev_out, L.out_stream));
  CK(engine_apply_splits(c->eng, L.splits, L.child_ids, L.gid_base, n, !L.last, c->stream));
  if (L.cur_set) {
    CK(cudaEventRecord(L.cur_set->e, c->stream));
    L.cur_set->complete = true;
  }
  // the next level's histogram pass goes in right behind routing, with the node count bounded by twice this level's
  // (the pass reads the exact count from the engine's node table).  Not for sparse storage, whose pass sizes its
  // tables from the count.
  if (!L.last && !c->eng.sparse) {
    const int n_up = 2 * n < L.max_nodes ? 2 * n : L.max_nodes;
    TreeLoop::EvSet* mine = L.cur_set;
    const int32_t rc = enqueue_hist(c, n_up, L.depth + 1, nx);
    if (rc) return rc;
    L.pre_set = L.cur_set;
    L.cur_set = mine;
    L.pre = true;
  }
  c->stats.launches += L.last ? 6 : 9;
  return H2OK_OK;
}

int32_t level_wait(h2ok_ctx* c, int32_t* n_done, int32_t* n_children) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  CK(cudaEventSynchronize(L.ev_out));
  const int nch = L.last ? 0 : *L.h_nch;
  if (nch < 0 || nch > L.max_nodes) {
    L.have_tree = false;
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "loop_level: the next level has more nodes than max_depth allows for");
  }
  if (L.out && L.out->child_descs && nch > 0) {
    CK(cudaMemcpyAsync(L.out->child_descs, L.descs[1 - L.cur], sizeof(h2ok_col_desc) * (size_t)nch * L.cols,
                       cudaMemcpyDeviceToHost, L.out_stream));
    CK(cudaStreamSynchronize(L.out_stream));
  }
  L.done_acc = L.acc;
  L.done_rng = L.rng;
  if (L.pre && nch == 0) {  // the tree ended early: the pass enqueued ahead found no node and did nothing
    if (L.pre_set) L.pre_set->used = false;
    L.pre = false;
  }
  L.cur_set = L.pre ? L.pre_set : nullptr;
  L.done_nodes = L.n_nodes;
  L.done_children = nch;
  L.done_depth = L.depth;
  L.done_gid_base = L.gid_base;
  L.done_cur = L.cur;
  L.gid_base += L.n_nodes;
  L.depth += 1;
  L.cur = 1 - L.cur;
  L.n_nodes = nch;
  L.out = nullptr;
  if (n_done) *n_done = L.done_nodes;
  if (n_children) *n_children = nch;
  return H2OK_OK;
}

// a tree's start in three phases (a group runs each phase for all its members, the NCCL one inside a group call)
int32_t begin_tree_local(h2ok_ctx* c, const float* d_y, const float* d_w) {
  TreeLoop& L = *c->loop;
  if (L.forest > 1) return fail(H2OK_ERR_STATE, "%s", "this loop builds a forest: start its trees with loop_begin_forest_dev");
  CK(cudaSetDevice(c->device));
  const int64_t n = c->eng.rows;
  const size_t pad = (size_t)(n + 3) / 4 * 4;
  const bool weighted = d_w != nullptr;
  if (weighted && (c->eng.sparse || c->eng.u8))
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "sparse and u8 storage take unweighted trees only");
  if (weighted != c->eng.weighted) CK(engine_set_weighted(c->eng, weighted));
  CK(L.wyq.ensure(pad)); CK(L.wyyq.ensure(pad)); CK(L.absmax.ensure(3));
  if (weighted) CK(L.wq.ensure(pad));
  L.tree_y = d_y;
  L.tree_w = d_w;
  CK(launch_target_absmax(d_y, d_w, n, L.absmax.p, c->stream));
  return H2OK_OK;
}
int32_t begin_tree_reduce(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  if (c->comm_ranks > 1)
    NK(nccl_api()->AllReduce(L.absmax.p, L.absmax.p, 3, ncclFloat64, ncclMax, (ncclComm_t)c->comm, c->stream));
  return H2OK_OK;
}
int32_t begin_tree_finish(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  const int64_t n = c->eng.rows;
  const bool weighted = L.tree_w != nullptr;
  // no host round trip: the scales are chosen and used on the device; the host copy travels with the first level's results
  CK(launch_choose_quant_dev(L.absmax.p, weighted ? 1 : 0, 25, L.q_dev, c->stream));
  CK(cudaMemcpyAsync(L.h_q, L.q_dev, sizeof(h2ok_quant), cudaMemcpyDeviceToHost, c->stream));
  CK(launch_quantize_rows_dev(L.tree_y, L.tree_w, n, weighted, L.q_dev, weighted ? L.wq.p : nullptr, L.wyq.p, L.wyyq.p, c->stream));
  CK(cudaMemsetAsync(c->eng.err, 0, sizeof(int32_t), c->stream));
  CK(engine_begin_tree(c->eng, weighted ? L.wq.p : nullptr, L.wyq.p, L.wyyq.p, nullptr, c->stream, L.q_dev));
  // (device to device: a copy from pageable host memory would make the host wait for the stream to drain first)
  CK(cudaMemcpyAsync(L.descs[0], L.d_root, sizeof(h2ok_col_desc) * L.cols, cudaMemcpyDeviceToDevice, c->stream));
  CK(cudaMemsetAsync(L.sums, 0, 3 * sizeof(unsigned long long), c->stream));
  if (L.pre && L.pre_set) L.pre_set->used = false;  // a tree abandoned half way: its pass enqueued ahead is void
  L.pre = false;
  L.cur = 0;
  L.n_nodes = 1;
  L.depth = 0;
  L.gid_base = 0;
  L.done_depth = -1;
  L.have_tree = true;
  return H2OK_OK;
}
int32_t begin_tree_common(h2ok_ctx* c, const float* d_y, const float* d_w) {
  int32_t rc = begin_tree_local(c, d_y, d_w);
  if (rc) return rc;
  if ((rc = begin_tree_reduce(c))) return rc;
  return begin_tree_finish(c);
}

// K class trees in lockstep (sparse storage): d_y holds the K target vectors [K][ld_y]
int32_t begin_forest_local(h2ok_ctx* c, const float* d_y, int64_t ld_y) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  if (!c->eng.sparse) return fail(H2OK_ERR_UNSUPPORTED, "%s", "a forest needs a shard created with H2OK_TREE_SPARSE");
  const int K = L.forest;
  const int64_t n = c->eng.rows;
  const size_t pad = (size_t)(n + 3) / 4 * 4;
  CK(L.wyq.ensure(pad * K)); CK(L.wyyq.ensure(pad * K)); CK(L.absmax.ensure(3 * (size_t)K));
  L.tree_y = d_y;
  L.tree_w = nullptr;
  L.tree_ld = ld_y;
  for (int k = 0; k < K; k++) CK(launch_target_absmax(d_y + (int64_t)k * ld_y, nullptr, n, L.absmax.p + 3 * k, c->stream));
  return H2OK_OK;
}
int32_t begin_forest_reduce(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  if (c->comm_ranks > 1)
    NK(nccl_api()->AllReduce(L.absmax.p, L.absmax.p, 3 * (size_t)L.forest, ncclFloat64, ncclMax, (ncclComm_t)c->comm, c->stream));
  return H2OK_OK;
}
int32_t begin_forest_finish(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  const int K = L.forest;
  const int64_t n = c->eng.rows;
  const size_t pad = (size_t)(n + 3) / 4 * 4;
  for (int k = 0; k < K; k++) {
    CK(launch_choose_quant_dev(L.absmax.p + 3 * k, 0, 25, L.q_dev + k, c->stream));
    CK(launch_quantize_rows_dev(L.tree_y + (int64_t)k * L.tree_ld, nullptr, n, false, L.q_dev + k, nullptr, L.wyq.p + pad * k,
                                L.wyyq.p + pad * k, c->stream));
    CK(cudaMemcpyAsync(L.descs[0] + (size_t)k * L.cols, L.d_root, sizeof(h2ok_col_desc) * L.cols, cudaMemcpyDeviceToDevice,
                       c->stream));
  }
  CK(cudaMemcpyAsync(L.h_q, L.q_dev, sizeof(h2ok_quant) * (size_t)K, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemsetAsync(c->eng.err, 0, sizeof(int32_t), c->stream));
  CK(engine_begin_forest(c->eng, K, L.wyq.p, L.wyyq.p, (int64_t)pad, L.q_dev, c->stream));
  CK(cudaMemsetAsync(L.sums, 0, 3 * sizeof(unsigned long long), c->stream));
  L.pre = false;
  L.cur = 0;
  L.n_nodes = K;
  L.depth = 0;
  L.gid_base = 0;
  L.done_depth = -1;
  L.have_tree = true;
  return H2OK_OK;
}

// the checksum query in three phases
int32_t checksum_local(h2ok_ctx* c, int64_t row0) {
  TreeLoop& L = *c->loop;
  if (!L.checksum) return fail(H2OK_ERR_STATE, "%s", "loop_checksum: switch the checksum on with loop_set_options first");
  CK(cudaSetDevice(c->device));
  CK(cudaMemsetAsync(L.sums + 2, 0, sizeof(unsigned long long), c->stream));
  for (int k = 0; k < L.forest; k++)  // (a forest: every class tree's leaves, each under its own salt)
    hash_i32_kernel<<<grid_for(c->eng.rows, 256, 1184), 256, 0, c->stream>>>(c->eng.leaf + (int64_t)k * c->eng.rows, c->eng.rows,
                                                                            (uint64_t)row0, kSaltLeaf + 0x100ull * k, L.sums + 2);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(L.sums_tmp, L.sums, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToDevice, c->stream));
  if (c->comm_ranks > 1 && c->comm_rank != 0) {
    // every rank hashed the same splits: rank 0's count.  Under H2OK_REDUCE_ALL the same holds for the accumulators;
    // under H2OK_REDUCE_SCATTER a rank hashed only the pairs it reduced
    CK(cudaMemsetAsync(L.sums_tmp + 1, 0, sizeof(unsigned long long), c->stream));
    if (L.reduce_mode != H2OK_REDUCE_SCATTER) CK(cudaMemsetAsync(L.sums_tmp, 0, sizeof(unsigned long long), c->stream));
  }
  return H2OK_OK;
}
int32_t checksum_reduce(h2ok_ctx* c) {
  TreeLoop& L = *c->loop;
  if (c->comm_ranks > 1)
    NK(nccl_api()->AllReduce(L.sums_tmp, L.sums_tmp, 3, ncclUint64, ncclSum, (ncclComm_t)c->comm, c->stream));
  return H2OK_OK;
}
int32_t checksum_fetch(h2ok_ctx* c, uint64_t* out3) {
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  unsigned long long h[3];
  CK(cudaMemcpyAsync(h, L.sums_tmp, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  for (int k = 0; k < 3; k++) out3[k] = h[k];
  return H2OK_OK;
}

}  // namespace
}  // namespace h2ok

using namespace h2ok;

extern "C" {

// ------------------------------------------------------------------------------------------------ communicator
int32_t h2ok_hist_comm_unique_id(void* id_out) {
  REQUIRE(id_out != nullptr, "null pointer");
  NcclApi* n = nccl_api();
  if (!n) return fail(H2OK_ERR_UNSUPPORTED, "%s", "libnccl.so.2 could not be loaded");
  static_assert(sizeof(ncclUniqueId) == H2OK_COMM_ID_BYTES, "H2OK_COMM_ID_BYTES must be NCCL's unique id size");
  ncclUniqueId id;
  NK(n->GetUniqueId(&id));
  memcpy(id_out, &id, sizeof(id));
  return H2OK_OK;
}

int32_t h2ok_hist_comm_init(h2ok_ctx* c, const void* id_bytes, int32_t rank, int32_t n_ranks) {
  REQUIRE(c && id_bytes, "null pointer");
  REQUIRE(n_ranks >= 1 && rank >= 0 && rank < n_ranks, "rank outside [0, n_ranks)");
  if (c->loop) return fail(H2OK_ERR_STATE, "%s", "comm_init after loop_create: join the communicator first");
  comm_free(c);
  if (n_ranks == 1) return H2OK_OK;
  NcclApi* n = nccl_api();
  if (!n) return fail(H2OK_ERR_UNSUPPORTED, "%s", "libnccl.so.2 could not be loaded");
  CK(cudaSetDevice(c->device));
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof(id));
  ncclComm_t comm = nullptr;
  NK(n->CommInitRank(&comm, n_ranks, id, rank));
  c->comm = comm;
  c->comm_rank = rank;
  c->comm_ranks = n_ranks;
  return H2OK_OK;
}

// ------------------------------------------------------------------------------------------------ data without a host
int32_t h2ok_hist_tree_load_synth(h2ok_ctx* c, int32_t shape, uint64_t seed, int64_t row0, int32_t na_permille) {
  REQUIRE(c != nullptr, "null context");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "tree_load_synth before tree_create");
  REQUIRE(shape == H2OK_SHAPE_HIGGS28 || shape == H2OK_SHAPE_WIDE64 || shape == H2OK_SHAPE_SPARSE, "unknown shape id");
  CK(cudaSetDevice(c->device));
  const int64_t rows = c->eng.rows, ld = (rows + 63) / 64 * 64;
  const int cols = c->eng.cols;
  int64_t per = ((int64_t)512 << 20) / (4 * ld);  // columns per staging block (<= 512 MB)
  per = per < 1 ? 1 : (per > cols ? cols : per);
  CK(c->X.ensure((size_t)ld * per));
  for (int c0 = 0; c0 < cols; c0 += (int)per) {
    const int nc = cols - c0 < per ? cols - c0 : (int)per;
    CK(launch_synth_cols(shape, seed, row0, rows, c0, nc, ld, na_permille, c->X.p, c->stream));
    CK(engine_load_cols(c->eng, c0, nc, c->X.p, ld, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  c->X.release();
  return H2OK_OK;
}

// ------------------------------------------------------------------------------------------------ the loop
static int32_t loop_alloc(h2ok_ctx* c, const h2ok_params* params, const int32_t* is_int, int forest = 1) {
  REQUIRE(c && params, "null pointer");
  if (!c->eng_ready) return fail(H2OK_ERR_STATE, "%s", "loop_create before tree_create / the data is loaded");
  REQUIRE(params->max_depth >= 1 && params->nbins >= 1 && params->nbins_top_level >= 1, "bad params");
  const int max_bins = params->nbins > params->nbins_top_level ? params->nbins : params->nbins_top_level;
  if (max_bins > H2OK_MAX_BINS) return fail(H2OK_ERR_UNSUPPORTED, "%s", "bins outside [1, H2OK_MAX_BINS]");
  const int cols = c->eng.cols;
  bool has_cat = false;
  for (int j = 0; is_int && j < cols; j++) {
    if (is_int[j] < 0 || is_int[j] > 2) return fail(H2OK_ERR_ARG, "%s", "is_int: 0 real, 1 integer, 2 categorical");
    has_cat = has_cat || is_int[j] == 2;
  }
  if (c->eng.u8) {
    // the cells are integer codes: every column is an integer column whose range never exceeds nbins, so that
    // DHistogram's integer rule gives step-1 bins at every level (what the byte-wide histogram pass assumes)
    if (params->nbins < 255 || params->nbins_top_level < 255)
      return fail(H2OK_ERR_UNSUPPORTED, "%s", "u8 storage needs nbins >= 255 (one bin per code at every level)");
    for (int j = 0; j < cols; j++)
      if (!is_int || is_int[j] < 1)
        return fail(H2OK_ERR_ARG, "%s", "u8 storage: every column is an integer or categorical column (is_int 1 or 2)");
  }
  CK(cudaSetDevice(c->device));
  loop_free(c);
  TreeLoop* L = new TreeLoop();
  c->loop = L;
  L->params = *params;
  L->cols = cols;
  L->max_bins = max_bins;
  L->nb1 = max_bins + 1;
  if (is_int) L->is_int.assign(is_int, is_int + cols);
  L->has_cat = has_cat;
  // a categorical split's bit set is worked out from the winning column's reduced histogram, which under
  // H2OK_REDUCE_SCATTER only one rank holds: such loops reduce with an all-reduce, every rank then has every pair
  if (has_cat) L->reduce_mode = H2OK_REDUCE_ALL;
  int64_t mn = (int64_t)1 << (params->max_depth - 1 < 10 ? params->max_depth - 1 : 10);
  L->forest = forest;
  if (forest > 1) {
    mn *= forest;  // the flat node list of a level holds every class tree's nodes
    if (mn > H2OK_MAX_LEVEL_NODES) {
      loop_free(c);
      return fail(H2OK_ERR_UNSUPPORTED, "%s", "forest: n_trees * 2^(max_depth-1) must not exceed H2OK_MAX_LEVEL_NODES");
    }
  }
  L->max_nodes = (int)(mn < H2OK_MAX_LEVEL_NODES ? mn : H2OK_MAX_LEVEL_NODES);
  const int R = c->comm_ranks;
  L->pairs_cap = (size_t)L->max_nodes * cols;
  L->acc_pairs = (L->pairs_cap + R - 1) / R * R;
  const size_t cell = (size_t)L->nb1 * 3;
  cudaError_t e = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(p, bytes);
  };
  for (int k = 0; k < 2; k++) A((void**)&L->descs[k], sizeof(h2ok_col_desc) * L->pairs_cap);
  for (int k = 0; k < 2; k++) {
    A((void**)&L->acc2[k], sizeof(int64_t) * cell * L->acc_pairs);
    A((void**)&L->rng2[k], sizeof(uint32_t) * 2 * L->pairs_cap);
  }
  A((void**)&L->cand, sizeof(PairCand) * L->acc_pairs);
  A((void**)&L->d_root, sizeof(h2ok_col_desc) * (size_t)cols);
  A((void**)&L->splits, sizeof(h2ok_split) * (size_t)L->max_nodes);
  A((void**)&L->child_ids, sizeof(int32_t) * 2 * (size_t)L->max_nodes);
  A((void**)&L->n_children, sizeof(int32_t));
  A((void**)&L->tmp_descs, sizeof(h2ok_col_desc) * 2 * L->pairs_cap);
  A((void**)&L->tmp_valid, sizeof(int32_t) * 2 * (size_t)L->max_nodes);
  A((void**)&L->sums, 3 * sizeof(unsigned long long));
  A((void**)&L->sums_tmp, 3 * sizeof(unsigned long long));
  A((void**)&L->q_dev, sizeof(h2ok_quant) * (size_t)forest);
  if (e == cudaSuccess) e = cudaMallocHost((void**)&L->h_nch, sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMallocHost((void**)&L->h_q, sizeof(h2ok_quant) * (size_t)forest);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&L->out_stream, cudaStreamNonBlocking);
  for (cudaEvent_t* ev : {&L->ev_ready, &L->ev_out})
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
  for (auto& st : L->ring)
    for (cudaEvent_t* ev : {&st.a, &st.h0, &st.h1, &st.b, &st.c, &st.d, &st.e})
      if (e == cudaSuccess) e = cudaEventCreate(ev);
  if (e != cudaSuccess) {
    loop_free(c);
    return fail(H2OK_ERR_CUDA, "loop_create: %s", cudaGetErrorString(e));
  }
  return H2OK_OK;
}

// DHistogram.initialHist reads the column ranges of ALL ranks (each took its own while the data was loaded)
static int32_t loop_reduce_ranges(h2ok_ctx* c) {
  if (c->comm_ranks <= 1) return H2OK_OK;
  NcclApi* n = nccl_api();
  const size_t cols = (size_t)c->eng.cols;
  NK(n->AllReduce(c->eng.cmin, c->eng.cmin, cols, ncclUint32, ncclMin, (ncclComm_t)c->comm, c->stream));
  NK(n->AllReduce(c->eng.cmax, c->eng.cmax, cols, ncclUint32, ncclMax, (ncclComm_t)c->comm, c->stream));
  return H2OK_OK;
}

static int32_t loop_make_root(h2ok_ctx* c) {
  TreeLoop* L = c->loop;
  const int cols = L->cols;
  CK(cudaSetDevice(c->device));
  std::vector<uint32_t> emn(cols), emx(cols);
  CK(cudaMemcpyAsync(emn.data(), c->eng.cmin, sizeof(uint32_t) * cols, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(emx.data(), c->eng.cmax, sizeof(uint32_t) * cols, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  std::vector<float> lo(cols), hi(cols);
  for (int j = 0; j < cols; j++) {
    const uint32_t a = dec_f32_bits(emn[j]), b = dec_f32_bits(emx[j]);
    memcpy(&lo[j], &a, 4);
    memcpy(&hi[j], &b, 4);
  }
  L->root.resize(cols);
  const int32_t rc = h2ok_hist_root_descs_from_ranges(lo.data(), hi.data(), cols, &L->params,
                                                      L->is_int.empty() ? nullptr : L->is_int.data(), L->root.data());
  if (rc) return rc;
  CK(cudaMemcpy(L->d_root, L->root.data(), sizeof(h2ok_col_desc) * cols, cudaMemcpyHostToDevice));
  return H2OK_OK;
}

int32_t h2ok_hist_loop_create(h2ok_ctx* c, const h2ok_params* params, const int32_t* is_int) {
  int32_t rc = loop_alloc(c, params, is_int);
  if (rc) return rc;
  if ((rc = loop_reduce_ranges(c))) return rc;
  return loop_make_root(c);
}

int32_t h2ok_hist_loop_create_forest(h2ok_ctx* c, const h2ok_params* params, const int32_t* is_int, int32_t n_trees) {
  REQUIRE(c != nullptr, "null context");
  REQUIRE(n_trees >= 1 && n_trees <= kEngMaxClasses, "n_trees outside [1, 16]");
  if (n_trees > 1 && !(c->eng_ready && c->eng.sparse))
    return fail(H2OK_ERR_UNSUPPORTED, "%s", "a forest needs a shard created with H2OK_TREE_SPARSE");
  int32_t rc = loop_alloc(c, params, is_int, n_trees);
  if (rc) return rc;
  if ((rc = loop_reduce_ranges(c))) return rc;
  return loop_make_root(c);
}

int32_t h2ok_hist_loop_begin_forest_dev(h2ok_ctx* c, const float* d_y, int64_t ld) {
  int32_t rc = need_loop(c, "loop_begin_forest_dev");
  if (rc) return rc;
  REQUIRE(d_y != nullptr && ld >= c->eng.rows, "bad arguments");
  if (c->loop->forest == 1) return begin_tree_common(c, d_y, nullptr);
  if ((rc = begin_forest_local(c, d_y, ld))) return rc;
  if ((rc = begin_forest_reduce(c))) return rc;
  return begin_forest_finish(c);
}

int32_t h2ok_hist_loop_forest_quants(h2ok_ctx* c, h2ok_quant* out) {
  int32_t rc = need_loop(c, "loop_forest_quants");
  if (rc) return rc;
  REQUIRE(out != nullptr, "null pointer");
  if (!c->loop->have_tree) return fail(H2OK_ERR_STATE, "%s", "loop_forest_quants before a tree was begun");
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));  // the host copy of the scales travels behind the tree's set-up
  memcpy(out, c->loop->h_q, sizeof(h2ok_quant) * (size_t)c->loop->forest);
  return H2OK_OK;
}

int32_t h2ok_hist_loop_destroy(h2ok_ctx* c) {
  REQUIRE(c != nullptr, "null context");
  loop_free(c);
  return H2OK_OK;
}

int32_t h2ok_hist_loop_root_descs(h2ok_ctx* c, h2ok_col_desc* descs_out) {
  int32_t rc = need_loop(c, "loop_root_descs");
  if (rc) return rc;
  REQUIRE(descs_out != nullptr, "null pointer");
  memcpy(descs_out, c->loop->root.data(), sizeof(h2ok_col_desc) * c->loop->cols);
  return H2OK_OK;
}

int32_t h2ok_hist_loop_set_reduce(h2ok_ctx* c, int32_t mode) {
  int32_t rc = need_loop(c, "loop_set_reduce");
  if (rc) return rc;
  REQUIRE(mode == H2OK_REDUCE_SCATTER || mode == H2OK_REDUCE_ALL, "unknown reduce mode");
  c->loop->reduce_mode = c->loop->has_cat ? H2OK_REDUCE_ALL : mode;  // (categorical columns: see loop_alloc)
  return H2OK_OK;
}

int32_t h2ok_hist_loop_set_options(h2ok_ctx* c, int32_t checksum, int32_t time_phases) {
  int32_t rc = need_loop(c, "loop_set_options");
  if (rc) return rc;
  c->loop->checksum = checksum != 0;
  c->loop->phases_timed = time_phases != 0;
  return H2OK_OK;
}

int32_t h2ok_hist_loop_begin_tree_dev(h2ok_ctx* c, const float* d_y, const float* d_w) {
  int32_t rc = need_loop(c, "loop_begin_tree_dev");
  if (rc) return rc;
  REQUIRE(d_y != nullptr, "null pointer");
  CK(cudaSetDevice(c->device));
  return begin_tree_common(c, d_y, d_w);
}

int32_t h2ok_hist_loop_begin_tree_host(h2ok_ctx* c, const float* y_host, const float* w_host) {
  int32_t rc = need_loop(c, "loop_begin_tree_host");
  if (rc) return rc;
  REQUIRE(y_host != nullptr, "null pointer");
  CK(cudaSetDevice(c->device));
  TreeLoop& L = *c->loop;
  const size_t n = (size_t)c->eng.rows, pad = (n + 3) / 4 * 4;
  CK(L.y.ensure(pad));
  CK(cudaMemcpyAsync(L.y.p, y_host, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
  if (w_host) {
    CK(L.w.ensure(pad));
    CK(cudaMemcpyAsync(L.w.p, w_host, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
  }
  return begin_tree_common(c, L.y.p, w_host ? L.w.p : nullptr);
}

int32_t h2ok_hist_loop_begin_tree_synth(h2ok_ctx* c, uint64_t seed, int64_t row0, int32_t weight_mode) {
  int32_t rc = need_loop(c, "loop_begin_tree_synth");
  if (rc) return rc;
  CK(cudaSetDevice(c->device));
  TreeLoop& L = *c->loop;
  const size_t n = (size_t)c->eng.rows, pad = (n + 3) / 4 * 4;
  CK(L.y.ensure(pad));
  if (weight_mode) CK(L.w.ensure(pad));
  CK(launch_synth(H2OK_SHAPE_HIGGS28, seed, row0, (int64_t)n, c->eng.cols, (int64_t)pad, 0, weight_mode, nullptr, L.y.p,
                  weight_mode ? L.w.p : nullptr, c->stream));
  return begin_tree_common(c, L.y.p, weight_mode ? L.w.p : nullptr);
}

int32_t h2ok_hist_loop_level(h2ok_ctx* c, const h2ok_level_out* out, int32_t* n_nodes_done, int32_t* n_children) {
  int32_t rc = need_loop(c, "loop_level");
  if (rc) return rc;
  if ((rc = level_local(c, out))) return rc;
  if ((rc = level_reduce(c))) return rc;
  if ((rc = level_search(c))) return rc;
  if ((rc = level_gather(c))) return rc;
  if ((rc = level_finish(c))) return rc;
  return level_wait(c, n_nodes_done, n_children);
}

int32_t h2ok_hist_loop_view(h2ok_ctx* c, h2ok_loop_view* v) {
  int32_t rc = need_loop(c, "loop_view");
  if (rc) return rc;
  REQUIRE(v != nullptr, "null pointer");
  TreeLoop& L = *c->loop;
  memset(v, 0, sizeof(*v));
  v->depth = L.depth;
  v->n_nodes = L.n_nodes;
  v->gid_base = L.gid_base;
  v->max_nodes = L.max_nodes;
  v->max_bins = L.max_bins;
  v->cols = L.cols;
  v->rank = c->comm_rank;
  v->n_ranks = c->comm_ranks;
  v->reduce_mode = L.reduce_mode;
  v->n_trees = L.forest;
  CK(cudaStreamSynchronize(c->stream));  // the host copy of the scales is in flight until the tree's first level is done
  v->quant = *L.h_q;
  v->d_descs = L.descs[L.cur];
  v->d_wq = L.wq.p;
  v->d_wyq = L.wyq.p;
  v->d_wyyq = L.wyyq.p;
  v->d_leaf = c->eng.leaf;
  if (L.done_depth >= 0) {
    const int R = c->comm_ranks;
    const int64_t pairs = (int64_t)L.done_nodes * L.cols, chunk = (pairs + R - 1) / R;
    const bool scatter = R > 1 && L.reduce_mode == H2OK_REDUCE_SCATTER;
    v->done_depth = L.done_depth;
    v->done_nodes = L.done_nodes;
    v->done_children = L.done_children;
    v->done_gid_base = L.done_gid_base;
    v->done_pair0 = scatter ? (chunk * c->comm_rank < pairs ? chunk * c->comm_rank : pairs) : 0;
    v->done_pair1 = scatter ? (v->done_pair0 + chunk < pairs ? v->done_pair0 + chunk : pairs) : pairs;
    v->d_done_descs = L.descs[L.done_cur];
    v->d_done_acc = L.done_acc;
    v->d_done_min2_enc = L.done_rng;
    v->d_done_maxin_enc = L.done_rng + L.pairs_cap;
    v->d_done_splits = L.splits;
    v->d_done_child_ids = L.child_ids;
  } else {
    v->done_depth = -1;
  }
  return H2OK_OK;
}

int32_t h2ok_hist_loop_phase_stats(h2ok_ctx* c, int32_t reset, int32_t max_depth, int32_t* count_out, int32_t* nodes_out,
                                   float* ms_out) {
  int32_t rc = need_loop(c, "loop_phase_stats");
  if (rc) return rc;
  REQUIRE(max_depth >= 0 && max_depth <= TreeLoop::kMaxDepth, "max_depth outside [0, 64]");
  REQUIRE(max_depth == 0 || (count_out && ms_out), "null pointer");
  TreeLoop& L = *c->loop;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  for (auto& st : L.ring)
    if (st.used && st.complete && (rc = drain_set(L, st))) return rc;
  for (int d = 0; d < max_depth; d++) {
    count_out[d] = L.ph_cnt[d];
    if (nodes_out) nodes_out[d] = L.ph_nodes[d];
    for (int k = 0; k < 5; k++) ms_out[5 * d + k] = (float)L.ph_sum[d][k];
  }
  if (reset) {
    memset(L.ph_sum, 0, sizeof(L.ph_sum));
    memset(L.ph_cnt, 0, sizeof(L.ph_cnt));
  }
  return H2OK_OK;
}

int32_t h2ok_hist_loop_checksum(h2ok_ctx* c, int64_t row0, uint64_t* out3) {
  int32_t rc = need_loop(c, "loop_checksum");
  if (rc) return rc;
  REQUIRE(out3 != nullptr && row0 >= 0, "bad arguments");
  if ((rc = checksum_local(c, row0))) return rc;
  if ((rc = checksum_reduce(c))) return rc;
  return checksum_fetch(c, out3);
}

// ------------------------------------------------------------------------------------------------ a group of devices
// One host thread, several GPUs (what a caller shaped like the reference's main() needs: src/main.rs:18-37 is single
// threaded and blocking).  A group owns one context per device and one NCCL communicator over them; every group call
// runs the phases of the per-context call for all members, with the phases that talk to NCCL inside one
// ncclGroupStart / ncclGroupEnd (several devices driven from one thread must be grouped).
struct h2ok_group {
  std::vector<h2ok_ctx*> m;
  std::vector<int64_t> row0;  // global index of every member's first row
  int64_t rows_total = 0;
};

#define EACH(g, expr)                          \
  do {                                         \
    for (h2ok_ctx * c : (g)->m) {              \
      int32_t _rc = (expr);                    \
      if (_rc) return _rc;                     \
    }                                          \
  } while (0)
#define EACH_NCCL(g, expr)                                  \
  do {                                                      \
    if ((g)->m.size() > 1) NK(nccl_api()->GroupStart());    \
    int32_t _first = H2OK_OK;                               \
    for (h2ok_ctx * c : (g)->m) {                           \
      int32_t _rc = (expr);                                 \
      if (_rc && !_first) _first = _rc;                     \
    }                                                       \
    if ((g)->m.size() > 1) NK(nccl_api()->GroupEnd());      \
    if (_first) return _first;                              \
  } while (0)

int32_t h2ok_hist_group_create(const int32_t* device_ids, int32_t n_dev, h2ok_group** out) {
  REQUIRE(device_ids && out, "null pointer");
  REQUIRE(n_dev >= 1 && n_dev <= 64, "n_dev outside [1, 64]");
  for (int i = 0; i < n_dev; i++)
    for (int j = 0; j < i; j++) REQUIRE(device_ids[i] != device_ids[j], "a device appears twice");
  h2ok_group* g = new h2ok_group();
  int32_t rc = H2OK_OK;
  for (int i = 0; i < n_dev && !rc; i++) {
    h2ok_ctx* c = nullptr;
    rc = h2ok_hist_create(device_ids[i], &c);
    if (!rc) g->m.push_back(c);
  }
  if (!rc && n_dev > 1) {
    NcclApi* n = nccl_api();
    if (!n) rc = fail(H2OK_ERR_UNSUPPORTED, "%s", "libnccl.so.2 could not be loaded");
    else {
      std::vector<ncclComm_t> comms(n_dev);
      std::vector<int> devs(device_ids, device_ids + n_dev);
      ncclResult_t r = n->CommInitAll(comms.data(), n_dev, devs.data());
      if (r != ncclSuccess) rc = fail(H2OK_ERR_CUDA, "ncclCommInitAll: %s", n->GetErrorString(r));
      else
        for (int i = 0; i < n_dev; i++) {
          g->m[i]->comm = comms[i];
          g->m[i]->comm_rank = i;
          g->m[i]->comm_ranks = n_dev;
        }
    }
  }
  if (rc) {
    for (h2ok_ctx* c : g->m) h2ok_hist_destroy(c);
    delete g;
    return rc;
  }
  g->row0.assign(n_dev, 0);
  *out = g;
  return H2OK_OK;
}

int32_t h2ok_hist_group_destroy(h2ok_group* g) {
  if (!g) return H2OK_OK;
  for (h2ok_ctx* c : g->m) h2ok_hist_destroy(c);
  delete g;
  return H2OK_OK;
}

int32_t h2ok_hist_group_size(h2ok_group* g) { return g ? (int32_t)g->m.size() : 0; }

int32_t h2ok_hist_group_member(h2ok_group* g, int32_t i, h2ok_ctx** out, int64_t* row0, int64_t* rows) {
  REQUIRE(g && out, "null pointer");
  REQUIRE(i >= 0 && i < (int32_t)g->m.size(), "member index out of range");
  *out = g->m[i];
  if (row0) *row0 = g->row0[i];
  if (rows) *rows = g->m[i]->eng_ready ? g->m[i]->eng.rows : 0;
  return H2OK_OK;
}

// contiguous row blocks: member i holds rows [i * R / n, (i + 1) * R / n) of every column
int32_t h2ok_hist_group_load_synth(h2ok_group* g, int32_t shape, uint64_t seed, int64_t rows_total, int32_t cols,
                                   int32_t na_permille) {
  REQUIRE(g != nullptr, "null group");
  const int64_t n = (int64_t)g->m.size();
  REQUIRE(rows_total >= n && cols > 0, "fewer rows than devices");
  g->rows_total = rows_total;
  for (int64_t i = 0; i < n; i++) {
    g->row0[i] = rows_total * i / n;
    const int64_t rows = rows_total * (i + 1) / n - g->row0[i];
    int32_t rc = h2ok_hist_tree_create(g->m[i], rows, cols, 0);
    if (rc) return rc;
  }
  // every device generates its own rows; the members work at the same time (the calls below only enqueue)
  for (int64_t i = 0; i < n; i++) {
    int32_t rc = h2ok_hist_tree_load_synth(g->m[i], shape, seed, g->row0[i], na_permille);
    if (rc) return rc;
  }
  return H2OK_OK;
}

// The caller's columns (host memory, all rows): every member copies its own row block.
int32_t h2ok_hist_group_set_data_host(h2ok_group* g, const float* const* cols_host, int64_t rows_total, int32_t cols) {
  REQUIRE(g && cols_host, "null pointer");
  const int64_t n = (int64_t)g->m.size();
  REQUIRE(rows_total >= n && cols > 0, "fewer rows than devices");
  g->rows_total = rows_total;
  std::vector<const float*> ptrs(cols);
  for (int64_t i = 0; i < n; i++) {
    g->row0[i] = rows_total * i / n;
    const int64_t rows = rows_total * (i + 1) / n - g->row0[i];
    for (int j = 0; j < cols; j++) ptrs[j] = cols_host[j] + g->row0[i];
    int32_t rc = h2ok_hist_set_data_host(g->m[i], ptrs.data(), rows, cols);
    if (rc) return rc;
  }
  return H2OK_OK;
}

int32_t h2ok_hist_group_loop_create(h2ok_group* g, const h2ok_params* params, const int32_t* is_int) {
  REQUIRE(g && params, "null pointer");
  EACH(g, loop_alloc(c, params, is_int));
  EACH_NCCL(g, loop_reduce_ranges(c));
  EACH(g, loop_make_root(c));
  return H2OK_OK;
}

int32_t h2ok_hist_group_set_options(h2ok_group* g, int32_t reduce_mode, int32_t checksum, int32_t time_phases) {
  REQUIRE(g != nullptr, "null group");
  EACH(g, h2ok_hist_loop_set_reduce(c, reduce_mode));
  EACH(g, h2ok_hist_loop_set_options(c, checksum, time_phases));
  return H2OK_OK;
}

int32_t h2ok_hist_group_begin_tree_synth(h2ok_group* g, uint64_t seed, int32_t weight_mode) {
  REQUIRE(g != nullptr, "null group");
  for (size_t i = 0; i < g->m.size(); i++) {
    h2ok_ctx* c = g->m[i];
    int32_t rc = need_loop(c, "group_begin_tree_synth");
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    TreeLoop& L = *c->loop;
    const size_t n = (size_t)c->eng.rows, pad = (n + 3) / 4 * 4;
    CK(L.y.ensure(pad));
    if (weight_mode) CK(L.w.ensure(pad));
    CK(launch_synth(H2OK_SHAPE_HIGGS28, seed, g->row0[i], (int64_t)n, c->eng.cols, (int64_t)pad, 0, weight_mode, nullptr, L.y.p,
                    weight_mode ? L.w.p : nullptr, c->stream));
    if ((rc = begin_tree_local(c, L.y.p, weight_mode ? L.w.p : nullptr))) return rc;
  }
  EACH_NCCL(g, begin_tree_reduce(c));
  EACH(g, begin_tree_finish(c));
  return H2OK_OK;
}

int32_t h2ok_hist_group_begin_tree_host(h2ok_group* g, const float* y_host, const float* w_host) {
  REQUIRE(g && y_host, "null pointer");
  for (size_t i = 0; i < g->m.size(); i++) {
    h2ok_ctx* c = g->m[i];
    int32_t rc = need_loop(c, "group_begin_tree_host");
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    TreeLoop& L = *c->loop;
    const size_t n = (size_t)c->eng.rows, pad = (n + 3) / 4 * 4;
    CK(L.y.ensure(pad));
    CK(cudaMemcpyAsync(L.y.p, y_host + g->row0[i], sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
    if (w_host) {
      CK(L.w.ensure(pad));
      CK(cudaMemcpyAsync(L.w.p, w_host + g->row0[i], sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
    }
    if ((rc = begin_tree_local(c, L.y.p, w_host ? L.w.p : nullptr))) return rc;
  }
  EACH_NCCL(g, begin_tree_reduce(c));
  EACH(g, begin_tree_finish(c));
  return H2OK_OK;
}

// One level on every member.  `out` is ONE set of host buffers: under H2OK_REDUCE_SCATTER every member writes the
// (node, column) pairs it reduced into the same `hist`, so the buffer is complete on return and arrives over all the
// devices' links at once; the replicated results (ranges, splits, children) are written by member 0.
int32_t h2ok_hist_group_level(h2ok_group* g, const h2ok_level_out* out, int32_t* n_nodes_done, int32_t* n_children) {
  REQUIRE(g != nullptr, "null group");
  for (h2ok_ctx* c : g->m) {
    int32_t rc = need_loop(c, "group_level");
    if (rc) return rc;
  }
  h2ok_level_out rest{};  // members other than the first deliver their share of the histogram only
  if (out) rest.hist = out->hist;
  const bool all = g->m[0]->loop->reduce_mode == H2OK_REDUCE_ALL;
  for (size_t i = 0; i < g->m.size(); i++) {
    int32_t rc = level_local(g->m[i], i == 0 ? out : (out && !all ? &rest : nullptr));
    if (rc) return rc;
  }
  EACH_NCCL(g, level_reduce(c));
  EACH(g, level_search(c));
  EACH_NCCL(g, level_gather(c));
  EACH(g, level_finish(c));
  int32_t done = 0, nch = 0;
  for (size_t i = 0; i < g->m.size(); i++) {
    int32_t d = 0, n = 0;
    int32_t rc = level_wait(g->m[i], &d, &n);
    if (rc) return rc;
    if (i == 0) {
      done = d;
      nch = n;
    } else if (d != done || n != nch) {
      return fail(H2OK_ERR_STATE, "%s", "group_level: the members disagree on the level (not the same data set-up on every device?)");
    }
  }
  if (n_nodes_done) *n_nodes_done = done;
  if (n_children) *n_children = nch;
  return H2OK_OK;
}

int32_t h2ok_hist_group_checksum(h2ok_group* g, uint64_t* out3) {
  REQUIRE(g && out3, "null pointer");
  for (size_t i = 0; i < g->m.size(); i++) {
    int32_t rc = need_loop(g->m[i], "group_checksum");
    if (rc) return rc;
    if ((rc = checksum_local(g->m[i], g->row0[i]))) return rc;
  }
  EACH_NCCL(g, checksum_reduce(c));
  return checksum_fetch(g->m[0], out3);
}

// device milliseconds of the last level's histogram pass, the slowest member's
int32_t h2ok_hist_group_hist_ms(h2ok_group* g, float* ms_out) {
  REQUIRE(g && ms_out, "null pointer");
  float worst = 0;
  for (h2ok_ctx* c : g->m) {
    h2ok_stats st;
    int32_t rc = h2ok_hist_last_level_stats(c, &st);
    if (rc) return rc;
    worst = st.build_ms > worst ? st.build_ms : worst;
  }
  *ms_out = worst;
  return H2OK_OK;
}

int32_t h2ok_hist_group_sync(h2ok_group* g) {
  REQUIRE(g != nullptr, "null group");
  EACH(g, h2ok_hist_sync(c));
  return H2OK_OK;
}

}  // extern "C"
