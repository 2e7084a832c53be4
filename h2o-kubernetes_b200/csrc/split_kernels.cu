// split_kernels.cu — sm_100a kernels for SURVEY.md §8 rows a6 (score_decide), a7 (findBestSplitPoint + bestCol)
// and a8 (nextLevelHistos).  Arithmetic mirrors oracle/hist_oracle.c operation for operation with explicit
// round-to-nearest intrinsics (no FMA contraction), so results are bit-identical to the oracle.
//
// a7 exactness note: the oracle's running prefix/suffix sums are sequential fp64 additions.  Histograms produced
// by this library are exact multiples of a power of two with |sum| < 2^53 quanta, so every partial sum is exact and
// the warp-shuffle scan below returns the same bits as the sequential loop (and suffix == total - prefix).
#include <math.h>

#include "common.cuh"
#include "kernels.h"

namespace h2ok {

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kSplitWarps = 8;

struct Cand {  // running best of DTree.findBestSplitPoint
  double key;  // se_left + se_right
  double seL, seR, nL, nR, pL, pR;
  int dist;   // |b - nbins/2|
  int order;  // position in the sequential scan: 2*b + option; -1 = initial state
  int bin;
  int na;
};

__device__ __forceinline__ bool cand_better(double key, int dist, int order, const Cand& b) {
  // sequential rule: strictly smaller error, or equal error and strictly closer to the middle; on full ties the
  // earlier candidate stays
  if (key < b.key) return true;
  if (key == b.key) {
    if (dist < b.dist) return true;
    if (dist == b.dist && order < b.order) return true;
  }
  return false;
}

__device__ __forceinline__ double se_of(double wyy, double wy, double w) {
  double se = __dsub_rn(wyy, __ddiv_rn(__dmul_rn(wy, wy), w));
  return se < 0 ? 0 : se;  // NaN stays NaN, like the Java/C comparison
}

__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o; o >>= 1) v = __dadd_rn(v, __shfl_xor_sync(kFull, v, o));
  return v;
}

__device__ __forceinline__ double warp_incl_scan(double v, int lane) {
  for (int o = 1; o < 32; o <<= 1) {
    double t = __shfl_up_sync(kFull, v, o);
    if (lane >= o) v = __dadd_rn(v, t);
  }
  return v;
}

// Where a (node, column)'s bins come from: the fp64 histogram, or the int64 fixed-point accumulators converted on the
// fly with finalize's rule (one exact multiplication by a power of two), which skips materialising the histogram.
struct ValsF64 {
  const double* __restrict__ p;
  __device__ __forceinline__ double operator()(int b, int ch) const { return p[3 * b + ch]; }
  __device__ __forceinline__ void scales_from(const h2ok_quant*) {}
};
struct ValsAcc {
  const int64_t* __restrict__ p;
  double inv[3];
  __device__ __forceinline__ double operator()(int b, int ch) const { return __dmul_rn((double)p[3 * b + ch], inv[ch]); }
  __device__ __forceinline__ void scales_from(const h2ok_quant* q) {  // powers of two: exact reciprocals
    inv[0] = 1.0 / q->scale_w;
    inv[1] = 1.0 / q->scale_wy;
    inv[2] = 1.0 / q->scale_wyy;
  }
};

// A categorical column's bins in the order of their mean response (DTree.findBestSplitPoint sorts them and then scans
// the sorted copy like any ordered column); the NA bucket keeps its place behind the last bin.
template <class V>
struct ValsPerm {
  V v;
  const unsigned short* idx;
  int nb;
  __device__ __forceinline__ double operator()(int b, int ch) const { return v(b < nb ? (int)idx[b] : b, ch); }
};

// Which columns are sorted: categorical, one category per bin, four categories or more (with three the ordered scan
// already tries two of the three partitions; H2O-3's own threshold as recalled — oracle/hist_oracle.c cat_sorted()).
__device__ __forceinline__ bool cat_sorted(const h2ok_col_desc& d) { return d.is_int == 2 && d.step == 1.0 && d.nbin >= 4; }

// idx[rank] = bin, ranks by (mean response, bin) ascending — a stable sort, done by counting: bin i's rank is the
// number of bins that come before it.  An empty bin's mean is -DBL_MAX (they come first).
template <class V>
__device__ void warp_sort_bins(const V& vals, int nb, int lane, double* avg, unsigned short* idx) {
  for (int b = lane; b < nb; b += 32) {
    const double w = vals(b, 0);
    avg[b] = w == 0 ? -1.7976931348623157e308 : __ddiv_rn(vals(b, 1), w);
  }
  __syncwarp();
  for (int i = lane; i < nb; i += 32) {
    const double a = avg[i];
    int r = 0;
    for (int j = 0; j < nb; j++) {
      const double aj = avg[j];
      r += (aj < a || (aj == a && j < i)) ? 1 : 0;
    }
    idx[r < nb ? r : nb - 1] = (unsigned short)i;
  }
  __syncwarp();
}

// The IcedBitSet of a categorical split: the categories of the sorted positions [bin, nb) go right.  A category the
// node has not seen (empty bin) goes where most of the seen categories went, left on a tie (as recalled).
template <class V>
__device__ void warp_cat_bits(const V& vals, const h2ok_col_desc& d, const unsigned short* idx, int bin, int lane,
                              unsigned* bits /* [H2OK_CAT_WORDS], shared */) {
  bits[lane] = 0;
  __syncwarp();
  int nl = 0, nr = 0;
  for (int i = lane; i < d.nbin; i += 32)
    if (vals((int)idx[i], 0) != 0) {
      if (i >= bin) nr++;
      else nl++;
    }
  nl = __reduce_add_sync(kFull, nl);
  nr = __reduce_add_sync(kFull, nr);
  const bool empties_left = nl >= nr;
  const int off = (int)d.min;
  for (int i = lane; i < d.nbin; i += 32) {
    const int b = (int)idx[i];
    const bool right = vals(b, 0) != 0 ? i >= bin : !empties_left;
    const int c = b + off;
    if (right && c >= 0 && c < 32 * H2OK_CAT_WORDS) atomicOr(&bits[c >> 5], 1u << (c & 31));
  }
  __syncwarp();
}

// One warp evaluates one (node, column).  Returns found != 0 in every lane and the winning candidate in `out`.
template <class V>
__device__ int warp_best_split_point(const V& vals, int nbins, double min_rows, double msi, int lane,
                                     Cand& out, double& se_before_out) {
  if (nbins <= 1) return 0;
  double W = 0, WY = 0, WYY = 0;
  for (int b = lane; b < nbins; b += 32) {
    W = __dadd_rn(W, vals(b, 0));
    WY = __dadd_rn(WY, vals(b, 1));
    WYY = __dadd_rn(WYY, vals(b, 2));
  }
  W = warp_sum(W);
  WY = warp_sum(WY);
  WYY = warp_sum(WYY);
  const double wNA = vals(nbins, 0), wYNA = vals(nbins, 1), wYYNA = vals(nbins, 2);
  const double tot = __dadd_rn(W, wNA);
  if (tot < __dmul_rn(2.0, min_rows)) return 0;
  const double sy = __dadd_rn(WY, wYNA);
  const double var = __dsub_rn(__dmul_rn(__dadd_rn(WYY, wYYNA), tot), __dmul_rn(sy, sy));
  if ((float)var == 0.0f) return 0;

  Cand best;
  best.key = __dadd_rn(1.7976931348623157e308, 1.7976931348623157e308);  // DBL_MAX + DBL_MAX
  best.seL = best.seR = 1.7976931348623157e308;
  best.nL = best.nR = best.pL = best.pR = 0;
  const int mid = nbins >> 1;
  best.dist = mid;  // |0 - mid|
  best.order = -1;
  best.bin = 0;
  best.na = H2OK_NA_NONE;
  const double seNonNA = se_of(WYY, WY, W);
  double seBefore = seNonNA;
  if (wNA >= min_rows) {
    double sall = __dadd_rn(WY, wYNA);
    double seAll = __dsub_rn(__dadd_rn(WYY, wYYNA), __ddiv_rn(__dmul_rn(sall, sall), __dadd_rn(W, wNA)));
    double seNA = se_of(wYYNA, wYNA, wNA);
    best.seL = seNonNA;
    best.seR = seNA;
    best.key = __dadd_rn(seNonNA, seNA);
    best.na = H2OK_NA_VS_REST;
    seBefore = seAll;
    best.nL = W;
    best.pL = WY;
    best.nR = wNA;
    best.pR = wYNA;
  }
  const Cand init = best;
  const bool init_nan = !(init.key == init.key);

  double cw = 0, cwy = 0, cwyy = 0;  // carries of the row-wise scan
  const int nrows = (nbins + 31) / 32;
  for (int k = 0; k < nrows; k++) {
    const int b = 32 * k + lane;
    double v0 = 0, v1 = 0, v2 = 0;
    if (b < nbins) {
      v0 = vals(b, 0);
      v1 = vals(b, 1);
      v2 = vals(b, 2);
    }
    double i0 = warp_incl_scan(v0, lane), i1 = warp_incl_scan(v1, lane), i2 = warp_incl_scan(v2, lane);
    const double wlo = __dadd_rn(cw, __dsub_rn(i0, v0));  // exclusive prefix: bins [0,b)
    const double wYlo = __dadd_rn(cwy, __dsub_rn(i1, v1));
    const double wYYlo = __dadd_rn(cwyy, __dsub_rn(i2, v2));
    cw = __dadd_rn(cw, __shfl_sync(kFull, i0, 31));
    cwy = __dadd_rn(cwy, __shfl_sync(kFull, i1, 31));
    cwyy = __dadd_rn(cwyy, __shfl_sync(kFull, i2, 31));
    if (b < 1 || b > nbins - 1 || v0 == 0) continue;
    const double whi = __dsub_rn(W, wlo), wYhi = __dsub_rn(WY, wYlo), wYYhi = __dsub_rn(WYY, wYYlo);
    if (__dadd_rn(wlo, wNA) < min_rows) continue;
    if (__dadd_rn(whi, wNA) < min_rows) continue;  // H2O breaks here; whi only shrinks, so skipping is the same
    const int dist = abs(b - mid);
    if (wNA == 0) {
      double selo = se_of(wYYlo, wYlo, wlo), sehi = se_of(wYYhi, wYhi, whi);
      double key = __dadd_rn(selo, sehi);
      if (cand_better(key, dist, 2 * b, best) && wlo >= min_rows && whi >= min_rows) {
        best.key = key; best.seL = selo; best.seR = sehi; best.dist = dist; best.order = 2 * b; best.bin = b;
        best.nL = wlo; best.nR = whi; best.pL = wYlo; best.pR = wYhi; best.na = H2OK_NA_NONE;
      }
    } else {
      {  // NAs to the left
        double nl = __dadd_rn(wlo, wNA);
        double selo = se_of(__dadd_rn(wYYlo, wYYNA), __dadd_rn(wYlo, wYNA), nl), sehi = se_of(wYYhi, wYhi, whi);
        double key = __dadd_rn(selo, sehi);
        if (cand_better(key, dist, 2 * b, best) && nl >= min_rows && whi >= min_rows) {
          best.key = key; best.seL = selo; best.seR = sehi; best.dist = dist; best.order = 2 * b; best.bin = b;
          best.nL = nl; best.nR = whi; best.pL = __dadd_rn(wYlo, wYNA); best.pR = wYhi; best.na = H2OK_NA_LEFT;
        }
      }
      {  // NAs to the right
        double nr = __dadd_rn(whi, wNA);
        double selo = se_of(wYYlo, wYlo, wlo), sehi = se_of(__dadd_rn(wYYhi, wYYNA), __dadd_rn(wYhi, wYNA), nr);
        double key = __dadd_rn(selo, sehi);
        if (cand_better(key, dist, 2 * b + 1, best) && wlo >= min_rows && nr >= min_rows) {
          best.key = key; best.seL = selo; best.seR = sehi; best.dist = dist; best.order = 2 * b + 1; best.bin = b;
          best.nL = wlo; best.nR = nr; best.pL = wYlo; best.pR = __dadd_rn(wYhi, wYNA); best.na = H2OK_NA_RIGHT;
        }
      }
    }
  }
  // cross-lane lexicographic minimum over (key, dist, order); lanes still holding the initial state carry order -1
  // and the initial key, so the initial state wins every tie exactly as it does in the sequential scan
  int winner = lane;
  {
    double k = best.key;
    int d = best.dist, o = best.order, wl = lane;
    for (int off = 16; off; off >>= 1) {
      double k2 = __shfl_xor_sync(kFull, k, off);
      int d2 = __shfl_xor_sync(kFull, d, off), o2 = __shfl_xor_sync(kFull, o, off), w2 = __shfl_xor_sync(kFull, wl, off);
      bool take = (k2 < k) || (k2 == k && (d2 < d || (d2 == d && (o2 < o || (o2 == o && w2 < wl)))));
      if (take) { k = k2; d = d2; o = o2; wl = w2; }
    }
    winner = wl;
  }
  if (init_nan) best = init;  // nothing compares better than NaN in the sequential scan
  else {
    Cand w;
    w.key = __shfl_sync(kFull, best.key, winner);
    w.seL = __shfl_sync(kFull, best.seL, winner);
    w.seR = __shfl_sync(kFull, best.seR, winner);
    w.nL = __shfl_sync(kFull, best.nL, winner);
    w.nR = __shfl_sync(kFull, best.nR, winner);
    w.pL = __shfl_sync(kFull, best.pL, winner);
    w.pR = __shfl_sync(kFull, best.pR, winner);
    w.dist = __shfl_sync(kFull, best.dist, winner);
    w.order = __shfl_sync(kFull, best.order, winner);
    w.bin = __shfl_sync(kFull, best.bin, winner);
    w.na = __shfl_sync(kFull, best.na, winner);
    best = w;
  }
  if (best.bin == 0 && best.na == H2OK_NA_NONE) return 0;
  if (!(__dadd_rn(best.seL, best.seR) < __dmul_rn(seBefore, __dsub_rn(1.0, msi)))) return 0;
  if (best.na == H2OK_NA_NONE) best.na = best.nL > best.nR ? H2OK_NA_LEFT : H2OK_NA_RIGHT;
  best.pL = __ddiv_rn(best.pL, best.nL);
  best.pR = __ddiv_rn(best.pR, best.nR);
  out = best;
  se_before_out = seBefore;
  return 1;
}

// DTree.Split.splat for the chosen (node, col, bin): warp-parallel search of the neighbouring non-empty bins.
template <class V>
__device__ float warp_splat(const V& vals, const h2ok_col_desc& d, int bin, int na, int lane) {
  if (na == H2OK_NA_VS_REST) return -1.0f;
  int x = -1, n = d.nbin;
  for (int b = lane; b < d.nbin; b += 32) {
    if (vals(b, 0) != 0) {
      if (b < bin) x = max(x, b);
      else n = min(n, b);
    }
  }
  x = __reduce_max_sync(kFull, x);
  n = __reduce_min_sync(kFull, n);
  double lo = __dadd_rn(d.min, __ddiv_rn((double)(x + 1), d.step));
  double hi = __dadd_rn(d.min, __ddiv_rn((double)n, d.step));
  if (d.is_int > 0) lo = d.step == 1 ? __dsub_rn(lo, 1.0) : floor(lo);
  if (d.is_int > 0) hi = d.step == 1 ? hi : ceil(hi);
  return (float)__ddiv_rn(__dadd_rn(lo, hi), 2.0);
}

// a7 in two steps, so that the (node, column) pairs can be spread over all SMs — and over ranks (SURVEY.md §8(e)
// "preferred form": every rank searches the pairs whose histograms the reduce-scatter left with it):
//   split_points_kernel  one warp per pair of [pair0, pair1): DTree.findBestSplitPoint + Split.splat -> PairCand
//   best_col_kernel      one warp per node: UndecidedNode.bestCol over the node's PairCands -> h2ok_split
// n_nodes may live on the device (n_dev != NULL): grids are then sized by an upper bound and surplus warps leave.
template <class V>
__global__ void __launch_bounds__(kSplitWarps * 32) split_points_kernel(V base, const h2ok_quant* __restrict__ q_dev,
                                                                        const unsigned char* __restrict__ node_class, int cols,
                                                                        const h2ok_col_desc* __restrict__ descs,
                                                                        int64_t pair0, int64_t pair1, int nb1,
                                                                        double min_rows, double msi,
                                                                        PairCand* __restrict__ out, int cat_smem) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t pair = pair0 + (int64_t)blockIdx.x * kSplitWarps + warp;
  if (pair >= pair1) return;
  const h2ok_col_desc d = descs[pair];
  PairCand pc;
  pc.found = 0;
  pc.bin = pc.na = 0;
  pc.splat = 0.0f;
  pc.key = pc.seL = pc.seR = pc.nL = pc.nR = pc.pL = pc.pR = pc.se_before = 0;
  if (d.nbin > 1) {
    V vals = base;
    // (a forest: the node's class tree has its own scales)
    if (q_dev) vals.scales_from(node_class ? q_dev + node_class[pair / cols] : q_dev);
    vals.p += pair * nb1 * 3;
    Cand cand;
    double before;
    if (cat_smem && cat_sorted(d)) {
      extern __shared__ __align__(16) unsigned char sort_smem[];
      double* avg = reinterpret_cast<double*>(sort_smem) + (size_t)warp * nb1;
      unsigned short* idx = reinterpret_cast<unsigned short*>(reinterpret_cast<double*>(sort_smem) + (size_t)kSplitWarps * nb1) +
                            (size_t)warp * nb1;
      warp_sort_bins(vals, d.nbin, lane, avg, idx);
      const ValsPerm<V> pv{vals, idx, d.nbin};
      if (warp_best_split_point(pv, d.nbin, min_rows, msi, lane, cand, before)) {
        pc.found = cand.na == H2OK_NA_VS_REST ? 1 : 2;  // (all non-NA rows on one side: no bit set is needed)
        pc.key = cand.key;
        pc.seL = cand.seL; pc.seR = cand.seR; pc.nL = cand.nL; pc.nR = cand.nR; pc.pL = cand.pL; pc.pR = cand.pR;
        pc.se_before = before;
        pc.bin = cand.bin;
        pc.na = cand.na;
        pc.splat = cand.na == H2OK_NA_VS_REST ? -1.0f : 0.0f;
      }
    } else if (warp_best_split_point(vals, d.nbin, min_rows, msi, lane, cand, before)) {
      pc.found = 1;
      pc.key = cand.key;
      pc.seL = cand.seL; pc.seR = cand.seR; pc.nL = cand.nL; pc.nR = cand.nR; pc.pL = cand.pL; pc.pR = cand.pR;
      pc.se_before = before;
      pc.bin = cand.bin;
      pc.na = cand.na;
      pc.splat = warp_splat(vals, d, cand.bin, cand.na, lane);
    }
  }
  if (lane == 0) out[pair] = pc;
}

__global__ void __launch_bounds__(kSplitWarps * 32) best_col_kernel(const PairCand* __restrict__ cand, int n_nodes, int cols,
                                                                   h2ok_split* __restrict__ splits) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.x * kSplitWarps + warp;
  if (g >= n_nodes) return;
  // bestCol: strictly smaller se wins, lowest column on ties
  double bk = 0;
  int bc = 0x7fffffff;
  for (int c = lane; c < cols; c += 32) {
    const PairCand& p = cand[(int64_t)g * cols + c];
    if (p.found && (bc == 0x7fffffff || p.key < bk)) {
      bk = p.key;
      bc = c;
    }
  }
  for (int off = 16; off; off >>= 1) {
    const double k2 = __shfl_xor_sync(kFull, bk, off);
    const int c2 = __shfl_xor_sync(kFull, bc, off);
    if (c2 != 0x7fffffff && (bc == 0x7fffffff || k2 < bk || (k2 == bk && c2 < bc))) {
      bk = k2;
      bc = c2;
    }
  }
  if (lane != 0) return;
  h2ok_split out;
  out.col = -1;
  out.bin = 0;
  out.na_dir = 0;
  out.splat = 0.0f;
  out.se_before = out.se_left = out.se_right = out.n_left = out.n_right = out.pred_left = out.pred_right = 0;
  out.equal = out.pad = 0;
  for (int w = 0; w < H2OK_CAT_WORDS; w++) out.cat_bits[w] = 0;  // (filled in by cat_bits_kernel)
  if (bc != 0x7fffffff) {
    const PairCand p = cand[(int64_t)g * cols + bc];
    out.col = bc;
    out.bin = p.bin;
    out.na_dir = p.na;
    out.splat = p.splat;
    out.se_before = p.se_before;
    out.se_left = p.seL;
    out.se_right = p.seR;
    out.n_left = p.nL;
    out.n_right = p.nR;
    out.pred_left = p.pL;
    out.pred_right = p.pR;
    out.equal = p.found == 2 ? 2 : 0;
  }
  splits[g] = out;
}

// The bit sets of the level's categorical splits, one warp per node: the winning column's bins are sorted again and the
// set is written into the split record — or, when this rank holds only the pairs [pair0, pair1) of the reduced
// histogram, into bits[node] (zeros for the nodes whose winner another rank holds; the ranks' arrays are then summed).
template <class V>
__global__ void __launch_bounds__(kSplitWarps * 32) cat_bits_kernel(V base, const h2ok_quant* __restrict__ q_dev,
                                                                    const unsigned char* __restrict__ node_class, int cols,
                                                                    const h2ok_col_desc* __restrict__ descs, int n_nodes,
                                                                    int64_t pair0, int64_t pair1, int nb1,
                                                                    h2ok_split* __restrict__ splits, uint32_t* __restrict__ bits) {
  extern __shared__ __align__(16) unsigned char sort_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.x * kSplitWarps + warp;
  if (g >= n_nodes) return;
  double* avg = reinterpret_cast<double*>(sort_smem) + (size_t)warp * nb1;
  unsigned short* idx = reinterpret_cast<unsigned short*>(reinterpret_cast<double*>(sort_smem) + (size_t)kSplitWarps * nb1) +
                        (size_t)warp * nb1;
  unsigned* wbits = reinterpret_cast<unsigned*>(reinterpret_cast<unsigned short*>(reinterpret_cast<double*>(sort_smem) +
                                                                                   (size_t)kSplitWarps * nb1) +
                                                (size_t)kSplitWarps * (nb1 + (nb1 & 1))) + warp * H2OK_CAT_WORDS;
  const int col = splits[g].col, bin = splits[g].bin;
  const bool cat = col >= 0 && splits[g].equal == 2;
  const int64_t pair = (int64_t)g * cols + (cat ? col : 0);
  unsigned word = 0;
  if (cat && pair >= pair0 && pair < pair1) {
    const h2ok_col_desc d = descs[pair];
    V vals = base;
    if (q_dev) vals.scales_from(node_class ? q_dev + node_class[g] : q_dev);
    vals.p += pair * nb1 * 3;
    warp_sort_bins(vals, d.nbin, lane, avg, idx);
    warp_cat_bits(vals, d, idx, bin, lane, wbits);
    word = wbits[lane];
  }
  if (bits) bits[(int64_t)g * H2OK_CAT_WORDS + lane] = word;
  else if (cat) splits[g].cat_bits[lane] = word;
}

__global__ void set_cat_bits_kernel(const uint32_t* __restrict__ bits, int n_nodes, h2ok_split* __restrict__ splits) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes * H2OK_CAT_WORDS) return;
  const int g = i / H2OK_CAT_WORDS;
  if (splits[g].col >= 0 && splits[g].equal == 2) splits[g].cat_bits[i % H2OK_CAT_WORDS] = bits[i];
}

// ---- a8 ---------------------------------------------------------------------------------------------------
__device__ inline double ulp_d(double x) {  // java.lang.Math.ulp(double)
  x = fabs(x);
  if (isinf(x)) return x;
  if (x == 1.7976931348623157e308) return __dsub_rn(x, nextafter(x, 0.0));
  return __dsub_rn(nextafter(x, (double)INFINITY), x);
}

__device__ inline double find_maxex(double maxin, int is_int) {  // DHistogram.find_maxEx
  double ulp = ulp_d(maxin);
  if (is_int > 0 && 1 > ulp) ulp = 1;
  double res = __dadd_rn(maxin, ulp);
  return isinf(res) ? maxin : res;
}

__device__ inline int make_desc(double mn, double maxex, int is_int, int nbins, h2ok_col_desc& out) {  // DHistogram.<init>
  out.min = mn;
  out.maxex = maxex;
  out.is_int = is_int;
  out.nbin = 0;
  out.step = 0.0;
  if (!(maxex > mn)) return 0;
  int xbins = nbins;
  double step;
  if (is_int > 0 && __dsub_rn(maxex, mn) <= (double)xbins) {
    xbins = (int)((long long)maxex - (long long)mn);
    step = 1.0;
  } else {
    step = __ddiv_rn((double)xbins, __dsub_rn(maxex, mn));
    if (!(step > 0.0) || isinf(step)) return 0;
  }
  if (xbins < 1) return 0;
  out.step = step;
  out.nbin = xbins;
  return 1;
}

__device__ inline bool equals_within_one_small_ulp(double a, double b) {
  if (isinf(a) || (isinf(b) && (a < b || b < a))) return false;
  double ua = ulp_d(a), ub = ulp_d(b);
  double su = ua < ub ? ua : ub;
  return fabs(__dsub_rn(a, b)) <= su;
}

// ENC: min2 / maxin are the order-preserving uint32 encodings the histogram pass (and the cross-rank reduce) work on,
// decoded here with finalize's rule instead of in a kernel of their own
template <bool ENC>
__global__ void next_level_descs_kernel(const h2ok_split* __restrict__ splits, const h2ok_col_desc* __restrict__ descs,
                                        const void* __restrict__ min2_v, const void* __restrict__ maxin_v, int n_nodes,
                                        int cols, h2ok_params p, h2ok_col_desc* __restrict__ tmp,
                                        int32_t* __restrict__ valid) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)2 * n_nodes * cols) return;
  const int j = (int)(i % cols);
  const int slot = (int)(i / cols);
  const int g = slot >> 1, way = slot & 1;
  h2ok_col_desc o;
  o.min = o.maxex = o.step = 0;
  o.nbin = 0;
  const h2ok_split s = splits[g];
  const h2ok_col_desc h = descs[(int64_t)g * cols + j];
  o.is_int = h.is_int;
  bool ok = s.col >= 0;
  if (ok) {
    double n = way == 0 ? s.n_left : s.n_right;
    double se = way == 0 ? s.se_left : s.se_right;
    ok = !(n < p.min_rows) && !(se <= 1e-30) && h.nbin > 0;
  }
  if (ok) {
    int adj_nbins = (h.nbin >> 1) > p.nbins ? (h.nbin >> 1) : p.nbins;
    double mn, mxin;
    if (ENC) {
      mn = (double)__fadd_rn(__uint_as_float(dec_f32_bits(static_cast<const uint32_t*>(min2_v)[(int64_t)g * cols + j])), 0.0f);
      mxin = (double)__fadd_rn(__uint_as_float(dec_f32_bits(static_cast<const uint32_t*>(maxin_v)[(int64_t)g * cols + j])), 0.0f);
    } else {
      mn = (double)static_cast<const float*>(min2_v)[(int64_t)g * cols + j];
      mxin = (double)static_cast<const float*>(maxin_v)[(int64_t)g * cols + j];
    }
    ok = !(mxin == mn) && (mn <= mxin);
    if (ok) {
      double maxex = find_maxex(mxin, h.is_int);
      if (s.col == j) {
        if (s.na_dir == H2OK_NA_VS_REST) {
          if (way == 1) ok = false;
        } else if (s.equal >= 2) {
          // a set of categories on each side: nothing is known about either side's range (Split.nextLevelHistos case 2 / 3)
        } else {
          double split = (double)s.splat;
          if (h.is_int > 0) split = (double)(float)ceil(split);
          if (way == 0) maxex = split;
          else mn = split;
        }
      }
      if (ok) {
        if (mn > maxex) ok = false;
        else if (equals_within_one_small_ulp(mn, maxex)) ok = false;
        else if (isinf(__ddiv_rn((double)adj_nbins, __dsub_rn(maxex, mn)))) ok = false;
        else if (h.is_int > 0 && !(__dadd_rn(mn, 1.0) < maxex)) ok = false;
        if (ok) {
          h2ok_col_desc t;
          if (make_desc(mn, maxex, h.is_int, adj_nbins, t)) {
            o = t;
            atomicOr(&valid[slot], 1);
          }
        }
      }
    }
  }
  tmp[i] = o;
}

__global__ void __launch_bounds__(1024) next_level_compact_kernel(const h2ok_col_desc* __restrict__ tmp,
                                                                  const int32_t* __restrict__ valid, int n_nodes, int cols,
                                                                  h2ok_col_desc* __restrict__ child_descs,
                                                                  int32_t* __restrict__ child_ids,
                                                                  int32_t* __restrict__ n_children) {
  __shared__ int s_part[1024];
  __shared__ int s_ids[2 * H2OK_MAX_LEVEL_NODES];
  const int slots = 2 * n_nodes;
  const int per = (slots + 1023) / 1024;
  const int lo = threadIdx.x * per, hi = min(slots, lo + per);
  int cnt = 0;
  for (int sl = lo; sl < hi; sl++) cnt += valid[sl] ? 1 : 0;
  s_part[threadIdx.x] = cnt;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {  // Hillis-Steele inclusive scan of the per-thread counts
    int t = threadIdx.x >= o ? s_part[threadIdx.x - o] : 0;
    __syncthreads();
    s_part[threadIdx.x] += t;
    __syncthreads();
  }
  int run = s_part[threadIdx.x] - cnt;
  for (int sl = lo; sl < hi; sl++) {
    int id = valid[sl] ? run++ : -1;
    s_ids[sl] = id;
    child_ids[sl] = id;
  }
  if (threadIdx.x == 1023) *n_children = s_part[1023];
  __syncthreads();
  const int64_t total = (int64_t)slots * cols;
  for (int64_t i = threadIdx.x; i < total; i += 1024) {
    int sl = (int)(i / cols), j = (int)(i % cols);
    int id = s_ids[sl];
    if (id >= 0) child_descs[(int64_t)id * cols + j] = tmp[i];
  }
}

// ---- a6 ---------------------------------------------------------------------------------------------------
__global__ void apply_splits_kernel(const float* __restrict__ X, int64_t ld, int64_t rows,
                                    const h2ok_split* __restrict__ splits, const int32_t* __restrict__ child_ids,
                                    int gid_base, int n_nodes, int32_t* __restrict__ nid, int32_t* __restrict__ leaf) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  int g = nid[r];
  if (g < 0 || g >= n_nodes) return;
  const int col = splits[g].col;
  if (col < 0) {
    nid[r] = -1;
    if (leaf) leaf[r] = (gid_base + g) * 3;
    return;
  }
  const int na = splits[g].na_dir;
  const float splat = splits[g].splat;
  const float d = X[(int64_t)col * ld + r];
  int way;
  if (d != d) way = (na == H2OK_NA_LEFT) ? 0 : 1;
  else if (na == H2OK_NA_VS_REST) way = 0;
  else if (splits[g].equal >= 2) way = cat_goes_right(splits[g].cat_bits, d, na);
  else way = d >= splat ? 1 : 0;
  const int ch = child_ids[2 * g + way];
  nid[r] = ch;
  if (ch < 0 && leaf) leaf[r] = (gid_base + g) * 3 + 1 + way;
}

}  // namespace

// shared memory of the categorical sort: per warp nb1 means, nb1 bin indices (and a bit set in cat_bits_kernel)
static size_t cat_smem_bytes(int nb1) {
  return (size_t)kSplitWarps * ((size_t)nb1 * 8 + (size_t)(nb1 + (nb1 & 1)) * 2 + H2OK_CAT_WORDS * 4);
}

cudaError_t launch_split_points(const double* hist, const int64_t* acc, const h2ok_quant* q, const h2ok_quant* q_dev,
                                const h2ok_col_desc* descs, int64_t pair0, int64_t pair1, int nb1, double min_rows,
                                double msi, PairCand* cand, cudaStream_t s, const unsigned char* node_class, int cols,
                                bool categorical) {
  if (pair1 <= pair0) return cudaSuccess;
  const unsigned grid = (unsigned)((pair1 - pair0 + kSplitWarps - 1) / kSplitWarps);
  const size_t smem = categorical ? cat_smem_bytes(nb1) : 0;
  if (hist) {
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(split_points_kernel<ValsF64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    split_points_kernel<ValsF64><<<grid, kSplitWarps * 32, smem, s>>>(ValsF64{hist}, nullptr, nullptr, 1, descs, pair0, pair1, nb1,
                                                                      min_rows, msi, cand, categorical ? 1 : 0);
  } else {
    ValsAcc v{acc, {1.0, 1.0, 1.0}};
    if (!q_dev) {
      v.inv[0] = 1.0 / q->scale_w;
      v.inv[1] = 1.0 / q->scale_wy;
      v.inv[2] = 1.0 / q->scale_wyy;
    }
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(split_points_kernel<ValsAcc>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    split_points_kernel<ValsAcc><<<grid, kSplitWarps * 32, smem, s>>>(v, q_dev, node_class, cols > 0 ? cols : 1, descs, pair0,
                                                                      pair1, nb1, min_rows, msi, cand, categorical ? 1 : 0);
  }
  return cudaGetLastError();
}

cudaError_t launch_cat_bits(const double* hist, const int64_t* acc, const h2ok_quant* q, const h2ok_quant* q_dev,
                            const h2ok_col_desc* descs, int n_nodes, int cols, int64_t pair0, int64_t pair1, int nb1,
                            h2ok_split* splits, uint32_t* bits, cudaStream_t s, const unsigned char* node_class) {
  const unsigned grid = (unsigned)((n_nodes + kSplitWarps - 1) / kSplitWarps);
  const size_t smem = cat_smem_bytes(nb1);
  if (hist) {
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(cat_bits_kernel<ValsF64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    cat_bits_kernel<ValsF64><<<grid, kSplitWarps * 32, smem, s>>>(ValsF64{hist}, nullptr, nullptr, cols, descs, n_nodes, pair0,
                                                                  pair1, nb1, splits, bits);
  } else {
    ValsAcc v{acc, {1.0, 1.0, 1.0}};
    if (!q_dev) {
      v.inv[0] = 1.0 / q->scale_w;
      v.inv[1] = 1.0 / q->scale_wy;
      v.inv[2] = 1.0 / q->scale_wyy;
    }
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(cat_bits_kernel<ValsAcc>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    cat_bits_kernel<ValsAcc><<<grid, kSplitWarps * 32, smem, s>>>(v, q_dev, node_class, cols, descs, n_nodes, pair0, pair1, nb1,
                                                                  splits, bits);
  }
  return cudaGetLastError();
}

cudaError_t launch_set_cat_bits(const uint32_t* bits, int n_nodes, h2ok_split* splits, cudaStream_t s) {
  const int n = n_nodes * H2OK_CAT_WORDS;
  set_cat_bits_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(bits, n_nodes, splits);
  return cudaGetLastError();
}

cudaError_t launch_best_col(const PairCand* cand, int n_nodes, int cols, h2ok_split* splits, cudaStream_t s) {
  best_col_kernel<<<(unsigned)((n_nodes + kSplitWarps - 1) / kSplitWarps), kSplitWarps * 32, 0, s>>>(cand, n_nodes, cols, splits);
  return cudaGetLastError();
}

// (the caller's descriptors live on the device: the categorical sort's shared memory is always asked for here)
cudaError_t launch_find_splits(const double* hist, const h2ok_col_desc* descs, int n_nodes, int cols, int nb1,
                               double min_rows, double msi, PairCand* cand, h2ok_split* splits, cudaStream_t s) {
  const int64_t pairs = (int64_t)n_nodes * cols;
  cudaError_t e = launch_split_points(hist, nullptr, nullptr, nullptr, descs, 0, pairs, nb1, min_rows, msi, cand, s, nullptr, 0,
                                      true);
  if (e != cudaSuccess) return e;
  e = launch_best_col(cand, n_nodes, cols, splits, s);
  if (e != cudaSuccess) return e;
  return launch_cat_bits(hist, nullptr, nullptr, nullptr, descs, n_nodes, cols, 0, pairs, nb1, splits, nullptr, s, nullptr);
}

static cudaError_t next_level_any(bool enc, const h2ok_split* splits, const h2ok_col_desc* descs, const void* min2,
                                  const void* maxin, int n_nodes, int cols, const h2ok_params& p, h2ok_col_desc* tmp_descs,
                                  int32_t* tmp_valid, h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children,
                                  cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(tmp_valid, 0, sizeof(int32_t) * 2 * (size_t)n_nodes, s);
  if (e != cudaSuccess) return e;
  int64_t n = (int64_t)2 * n_nodes * cols;
  if (enc)
    next_level_descs_kernel<true><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(splits, descs, min2, maxin, n_nodes, cols, p,
                                                                              tmp_descs, tmp_valid);
  else
    next_level_descs_kernel<false><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(splits, descs, min2, maxin, n_nodes, cols, p,
                                                                               tmp_descs, tmp_valid);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  next_level_compact_kernel<<<1, 1024, 0, s>>>(tmp_descs, tmp_valid, n_nodes, cols, child_descs, child_ids, n_children);
  return cudaGetLastError();
}

cudaError_t launch_next_level(const h2ok_split* splits, const h2ok_col_desc* descs, const float* min2, const float* maxin,
                              int n_nodes, int cols, const h2ok_params& p, h2ok_col_desc* tmp_descs, int32_t* tmp_valid,
                              h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children, cudaStream_t s) {
  return next_level_any(false, splits, descs, min2, maxin, n_nodes, cols, p, tmp_descs, tmp_valid, child_descs, child_ids,
                        n_children, s);
}

cudaError_t launch_next_level_enc(const h2ok_split* splits, const h2ok_col_desc* descs, const uint32_t* min2_enc,
                                  const uint32_t* maxin_enc, int n_nodes, int cols, const h2ok_params& p,
                                  h2ok_col_desc* tmp_descs, int32_t* tmp_valid, h2ok_col_desc* child_descs,
                                  int32_t* child_ids, int32_t* n_children, cudaStream_t s) {
  return next_level_any(true, splits, descs, min2_enc, maxin_enc, n_nodes, cols, p, tmp_descs, tmp_valid, child_descs,
                        child_ids, n_children, s);
}

// ---- allocation-free variant of a8 ---------------------------------------------------------------------------
// Some hosts of this library must never cause a device allocation while they run: an H2O node (or any embedding
// process) that has handed all of the GPU's memory to a pool allocator cannot let a raw cudaMalloc compete with
// that pool, and a service loop with a latency budget cannot accept the device-wide synchronisation cudaMalloc /
// cudaFree imply.  For them the scratch of next_level (candidate child descriptors + one flag per child slot) lives
// in this fixed array instead of a grow-on-demand buffer.  It costs kNoAllocScratchBytes (520 KiB) on every GPU that
// loads the library, until the process exits.
constexpr size_t kNoAllocDescBytes = (size_t)H2OK_NOALLOC_MAX_SLOT_COLS * sizeof(h2ok_col_desc);
constexpr size_t kNoAllocScratchBytes = kNoAllocDescBytes + (size_t)H2OK_NOALLOC_MAX_SLOTS * sizeof(int32_t);
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// The array's name is only a host-side handle; the device address is looked up once per device (with that device
// current) and cached.  The first call on a device therefore does a symbol lookup and should be made before any
// stream capture begins.
cudaError_t launch_next_level_noalloc(const h2ok_split* splits, const h2ok_col_desc* descs, const float* min2,
                                      const float* maxin, int n_nodes, int cols, const h2ok_params& p,
                                      h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children,
                                      cudaStream_t s) {
  constexpr int kMaxDevices = 64;
  static unsigned char* dev_ptr[kMaxDevices] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  if (dev_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    dev_ptr[dev] = static_cast<unsigned char*>(ptr);
  }
  h2ok_col_desc* tmp_descs = reinterpret_cast<h2ok_col_desc*>(dev_ptr[dev]);
  int32_t* tmp_valid = reinterpret_cast<int32_t*>(dev_ptr[dev] + kNoAllocDescBytes);
  // the array still holds the previous call's flags: launch_next_level clears the ones this call uses
  return launch_next_level(splits, descs, min2, maxin, n_nodes, cols, p, tmp_descs, tmp_valid, child_descs, child_ids,
                           n_children, s);
}

cudaError_t launch_apply_splits(const float* X, int64_t ld, int64_t rows, const h2ok_split* splits,
                                const int32_t* child_ids, int gid_base, int n_nodes, int32_t* nid, int32_t* leaf,
                                cudaStream_t s) {
  apply_splits_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(X, ld, rows, splits, child_ids, gid_base, n_nodes,
                                                                    nid, leaf);
  return cudaGetLastError();
}

}  // namespace h2ok
