// row_engine.cu — the resident row-wise tree engine (see engine.h for the design).
//
// Kernels:
//   eng_load_kernel      column-major block -> row-major XT (tiled transpose), once per data set.
//   eng_begin_kernel     new tree: records in row order, every row at the root.
//   eng_hist_kernel      one level's histogram pass.  CTA = (contiguous run of records, block of 32*CPL columns); a
//                        warp takes whole rows, lane L owns CPL (1 or 2) adjacent columns; descriptors and observed
//                        ranges in registers; shared histogram [slot][bin][channel][lane] -> bank == lane, no
//                        conflicts, one copy; the next batch's rows are prefetched into L2.
//   eng_thresholds_kernel / eng_code_kernel   once per data set: 8-bit order-preserving code of every cell, column-major.
//   eng_route_kernel     a6 in original row order: row -> child (code compare, exact float compare on code ties),
//                        leaves recorded, rows per (child, CTA) counted.
//   eng_scan_kernel      first record of every (child, CTA), next level's node table and centring offsets.
//   eng_scatter_kernel   next level's records {row, targets - child offset} written to their child's run.
#include <atomic>
#include <vector>

#include "common.cuh"
#include "engine.h"
#include "kernels.h"

namespace h2ok {

namespace {

inline unsigned nblk(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

constexpr unsigned kNoNidConst = 0xFFFFu;  // row -> node map: the row sits in a leaf
constexpr int kHistSmemCap = 227 * 1024 - 2048;  // dynamic shared memory left beside the kernel's static arrays (<= 1.2 KB)
constexpr int kSegNodes = kEngMaxNodes + 1;       // seg[kSegNodes] = number of nodes the table describes
constexpr int kSegInts = kEngMaxNodes + 2;

// ------------------------------------------------------------------------------------------------ data / tree set-up
__global__ void __launch_bounds__(256) eng_load_kernel(const float* __restrict__ X, int64_t ld, int64_t rows, int col0,
                                                       int ncols, float* __restrict__ XT, int64_t ldt) {
  __shared__ float t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int c = c0 + ty + i * 8;
    const int64_t r = r0 + tx;
    if (c < ncols && r < rows) t[ty + i * 8][tx] = X[(int64_t)c * ld + r];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int64_t r = r0 + ty + i * 8;
    const int c = c0 + tx;
    if (r < rows && c < ncols) XT[r * ldt + col0 + c] = t[tx][ty + i * 8];
  }
}

// Centring offsets.  The record arrays hold (target - M[node]) for an offset M near the node's mean target, so that the
// 32-bit histogram words wander around their starting point instead of marching through the range (a wrap costs a
// trip to the global accumulator).  Any M gives exact results: the flush adds count * M back.  The root's M is the
// mean over the shard's first rows; a child's M comes from its parent's split record (mean y and mean y^2 of the
// side, scaled like the targets).  Weighted shards keep M = 0 — their count word holds the weight sum, not the rows.
__global__ void __launch_bounds__(1024) eng_root_mean_kernel(int64_t rows, const int32_t* __restrict__ wyq_rows,
                                                             const int32_t* __restrict__ wyyq_rows, bool weighted,
                                                             int32_t* __restrict__ mnode) {
  __shared__ int s1[32], s2[32];
  const int64_t cnt = rows < 65536 ? rows : 65536;
  int t1 = 0, t2 = 0;
  if (!weighted)
    for (int64_t i = threadIdx.x; i < cnt; i += 1024) {  // >> 16 keeps the sums in 32 bits
      t1 += wyq_rows[i] >> 16;
      t2 += wyyq_rows[i] >> 16;
    }
  t1 = __reduce_add_sync(0xffffffffu, t1);
  t2 = __reduce_add_sync(0xffffffffu, t2);
  if ((threadIdx.x & 31) == 0) {
    s1[threadIdx.x >> 5] = t1;
    s2[threadIdx.x >> 5] = t2;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    t1 = __reduce_add_sync(0xffffffffu, s1[threadIdx.x]);
    t2 = __reduce_add_sync(0xffffffffu, s2[threadIdx.x]);
    if (threadIdx.x == 0) {
      mnode[0] = (int32_t)((long long)t1 * 65536 / cnt);
      mnode[1] = (int32_t)((long long)t2 * 65536 / cnt);
    }
  }
}

__global__ void eng_begin_kernel(int64_t rows, const int32_t* __restrict__ wq_rows, const int32_t* __restrict__ wyq_rows,
                                 const int32_t* __restrict__ wyyq_rows, const int32_t* __restrict__ mnode,
                                 int32_t* __restrict__ gidx, uint2* __restrict__ gval, int32_t* __restrict__ gw,
                                 int32_t* __restrict__ leaf, uint16_t* __restrict__ nid16, int32_t* __restrict__ seg,
                                 int root = 0) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    seg[0] = 0;
    seg[1] = (int32_t)rows;
    seg[kSegNodes] = 1;
  }
  if (p >= rows) return;
  leaf[p] = -1;
  nid16[p] = (uint16_t)root;  // (a forest: class tree k starts at node k of the flat list)
  if (!gidx) return;  // sparse storage keeps no row records: the row -> node map is all it needs
  // zero-weight rows still travel down the tree (a6); the histogram pass skips them by weight
  gidx[p] = (int32_t)p;
  // a row whose quantised weight is 0 takes no part, whatever its (independently scaled) targets round to — the
  // histogram pass sends it to the NA bin with weight 0, so its targets must be 0 too (weighted shards keep M == 0)
  const bool dead = gw && wq_rows[p] == 0;
  gval[p] = dead ? make_uint2(0u, 0u) : make_uint2((unsigned)(wyq_rows[p] - mnode[0]), (unsigned)(wyyq_rows[p] - mnode[1]));
  if (gw) gw[p] = wq_rows[p];
}

// first g in [0, n) with table[g + 1] > v (table non-decreasing, table[n] > v)
__device__ __forceinline__ int seg_of(const int32_t* __restrict__ table, int n, int64_t v) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (table[mid + 1] > v) hi = mid;
    else lo = mid + 1;
  }
  return lo;
}

// ------------------------------------------------------------------------------------------------ histogram pass
struct EngHistArgs {
  const float* XT;
  const unsigned char* XB;  // u8 storage instead of XT: one byte per cell (an integer code, 255 = NA), row pitch ldt4
  int64_t ldt;
  unsigned ldt4;  // row pitch in bytes
  int cols;
  const int32_t* gidx;  // row of every record
  const uint2* gval;    // {wyq - M1[node], wyyq - M2[node]}
  const int32_t* gw;    // wq (weighted only)
  const int32_t* mnode; // centring offsets {M1, M2} of every node of the level
  const int32_t* seg;
  int n_nodes;
  const h2ok_col_desc* descs;
  int nb1;  // bins + 1 of the global accumulators (row stride of acc / hist)
  int nbs;  // bins + 1 of the shared histogram: the most any node of THIS level uses (<= nb1)
  int64_t* acc;
  uint32_t* min2_enc;
  uint32_t* maxin_enc;
  int32_t* err;  // counts (node, column) pairs with more bins than the level's shared histogram holds
};

// 32-bit shared-window addressing: keeps the per-cell address math to one IMAD (no generic->shared conversion)
__device__ __forceinline__ void sh_inc(unsigned addr) { asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr) : "memory"); }
__device__ __forceinline__ unsigned sh_add(unsigned addr, unsigned v) {
  unsigned o;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(addr), "r"(v) : "memory");
  return o;
}

// base + idx * pitch: with both factors 32-bit this is one IMAD.WIDE.U32 with the base as its 64-bit addend (ptxas
// splits an explicit mad.wide.u32 into three instructions)
__device__ __forceinline__ const char* row_addr(const char* base, unsigned idx, unsigned pitch) {
  return base + (unsigned long long)idx * pitch;
}
// hb + b * pitch in one IMAD (the compiler splits it when b comes out of a select)
__device__ __forceinline__ unsigned cell_addr(unsigned hb, unsigned b, unsigned pitch) {
  unsigned r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(b), "r"(pitch), "r"(hb));
  return r;
}

// cold: a 32-bit shared word wrapped; the carry / borrow goes straight to the 64-bit global accumulator
__device__ __noinline__ void eng_carry(int64_t* acc, const h2ok_col_desc* __restrict__ descs, int node, int cols, int col,
                                       int nb1, int b, int d0, int d1, int d2) {
  if (col >= cols || descs[(int64_t)node * cols + col].nbin <= 0) return;  // untracked: its words are never flushed
  const int64_t cell = ((int64_t)node * cols + col) * nb1 + b;
  if (d0) carry_to_global(acc, cell, 0, d0);
  if (d1) carry_to_global(acc, cell, 1, d1);
  if (d2) carry_to_global(acc, cell, 2, d2);
}

template <int CPL, bool U8 = false> struct XVec;
template <> struct XVec<1, false> { using type = float; };
template <> struct XVec<2, false> { using type = float2; };
template <> struct XVec<1, true> { using type = unsigned char; };
template <> struct XVec<2, true> { using type = unsigned short; };

__device__ __forceinline__ float xget(const float& v, int) { return v; }
__device__ __forceinline__ float xget(const float2& v, int k) { return k ? v.y : v.x; }
__device__ __forceinline__ int xcode(const unsigned char& v, int) { return (int)v; }
__device__ __forceinline__ int xcode(const unsigned short& v, int k) { return k ? (int)(v >> 8) : (int)(v & 0xFFu); }


// What a lane keeps in registers while it walks the rows of one node: the descriptors and running extremes of its CPL
// columns.  A (node, column) that is not tracked gets a harmless 1-bin descriptor — its rows land in words that the
// flush skips — so the hot loop has no per-column test.
template <int CPL>
struct EngLane {
  double mn[CPL], st[CPL];
  int nb[CPL];
  float fmn[CPL], fmx[CPL];
  int imn[CPL], icmn[CPL], icmx[CPL];  // u8 storage: integer lower bound of the bins, running extremes of the codes
};

// A signed value v (|v| < 2^31) was added to a word that held `old`; did the 32-bit word wrap?  Bit 31 of the result
// is set when it did: the word crossed between its halves (old and new differ in bit 31) in the direction that cannot
// be reached without leaving the range (old's bit 31 differs from v's sign).
__device__ __forceinline__ unsigned wrap_bit(unsigned old, int v) {
  const unsigned n = old + (unsigned)v;
  return (old ^ n) & (old ^ (unsigned)v);
}
// __byte_perm selectors that put byte 3 of the second operand into byte 0 / 1 / 2 / 3 of the first
__device__ constexpr unsigned kPark[4] = {0x3217u, 0x3270u, 0x3710u, 0x7210u};

// U rows x CPL cells in three phases, so that the independent cells overlap their latencies instead of running one
// behind the other: (1) eng_bins: every cell's bin and shared address; (2) every cell's atomics; (3) ONE test for
// "some 32-bit word wrapped" over flag bytes parked in mask registers, which the cold path reads.  wq == 0 rows (weighted only) are turned into NaNs
// that add zeros: the NA bin gets +0 and the running extremes ignore NaN — which is what skipping the row means.
// v1 / v2 are the row's targets minus the run's centring offsets (see the kernel).
// LPR = lanes per row (32, or 16 for levels whose histograms are too wide for 32 columns per block): a bin of one
// column slot is 3 channels x LPR lanes x 4 bytes.
template <int CPL, bool WEIGHTED, int U, int LPR, bool NA, bool U8 = false>
__device__ __forceinline__ void eng_bins(EngLane<CPL>& c, const unsigned (&hb)[CPL],
                                         const typename XVec<CPL, U8>::type (&xv)[U], const int (&w)[U],
                                         unsigned (&pa)[U * CPL]) {
  if constexpr (U8) {
    // u8 storage: the cell IS an integer code and every descriptor has step 1 (DHistogram's integer rule: the range
    // of a code column never exceeds nbins >= 255), so the bin is code - min, clamped: no fp64, no conversions.
#pragma unroll
    for (int j = 0; j < U; j++) {
#pragma unroll
      for (int k = 0; k < CPL; k++) {
        const int code = xcode(xv[j], k);
        int b = min(max(code - c.imn[k], 0), c.nb[k] - 1);
        if (NA) {
          const bool na = code == 255;
          b = na ? c.nb[k] : b;
          c.icmn[k] = na ? c.icmn[k] : min(c.icmn[k], code);
          c.icmx[k] = na ? c.icmx[k] : max(c.icmx[k], code);
        } else {
          c.icmn[k] = min(c.icmn[k], code);
          c.icmx[k] = max(c.icmx[k], code);
        }
        pa[j * CPL + k] = cell_addr(hb[k], (unsigned)b, 12u * LPR);
      }
    }
  } else {
  float xs[U][CPL];
#pragma unroll
  for (int j = 0; j < U; j++) {
#pragma unroll
    for (int k = 0; k < CPL; k++) {
      float x = xget(xv[j], k);
      if (WEIGHTED) x = w[j] != 0 ? x : __int_as_float(0x7FC00000);
      xs[j][k] = x;
      // NA == false: the shard holds no NaN at all (census taken while the data was loaded), the NA test is dropped
      const int b = NA ? bin_fast(x, c.mn[k], c.st[k], c.nb[k]) : bin_fast_no_na(x, c.mn[k], c.st[k], c.nb[k]);
      pa[j * CPL + k] = cell_addr(hb[k], (unsigned)b, 12u * LPR);
    }
  }
  // running extremes (NaN-ignoring and exact), two rows at a time: min(min(m, a), b) is one three-input FMNMX3
#pragma unroll
  for (int k = 0; k < CPL; k++) {
#pragma unroll
    for (int j = 0; j + 1 < U; j += 2) {
      c.fmn[k] = fminf(fminf(c.fmn[k], xs[j][k]), xs[j + 1][k]);
      c.fmx[k] = fmaxf(fmaxf(c.fmx[k], xs[j][k]), xs[j + 1][k]);
    }
    if (U & 1) {
      c.fmn[k] = fminf(c.fmn[k], xs[U - 1][k]);
      c.fmx[k] = fmaxf(c.fmx[k], xs[U - 1][k]);
    }
  }
  }
}

template <int CPL, bool WEIGHTED, int U, int LPR>
__device__ __forceinline__ void eng_accum(const EngHistArgs& a, const unsigned (&hb)[CPL], int node, int col_lane,
                                          const unsigned (&pa)[U * CPL], const int (&v1)[U], const int (&v2)[U],
                                          const int (&w)[U]) {
  unsigned o1[U * CPL], o2[U * CPL], o0[WEIGHTED ? U * CPL : 1];
#pragma unroll
  for (int j = 0; j < U; j++) {
#pragma unroll
    for (int k = 0; k < CPL; k++) {
      const unsigned p = pa[j * CPL + k];
      if (WEIGHTED) o0[j * CPL + k] = sh_add(p, (unsigned)w[j]);
      else sh_inc(p);
      o1[j * CPL + k] = sh_add(p + 4u * LPR, (unsigned)v1[j]);
      o2[j * CPL + k] = sh_add(p + 8u * LPR, (unsigned)v2[j]);
    }
  }
  // Wrap flags without predicates: bit 31 of (old ^ new) & (old ^ v) says that the add wrapped the word (see
  // wrap_bit).  The top byte of every flag word is parked in a byte of a mask register with one PRMT, four flags per
  // register; one test of the masks decides whether the cold path runs, and the cold path reads the parked bytes.
  constexpr int PER = WEIGHTED ? 3 : 2, NF = PER * U * CPL, NM = (NF + 3) / 4;
  unsigned m[NM];
#pragma unroll
  for (int i = 0; i < NM; i++) m[i] = 0;
#pragma unroll
  for (int i = 0; i < U * CPL; i++) {
    const int j = i / CPL;
    const int f0 = PER * i, f1 = PER * i + 1, f2 = PER * i + 2;
    m[f0 >> 2] = __byte_perm(m[f0 >> 2], wrap_bit(o1[i], v1[j]), kPark[f0 & 3]);
    m[f1 >> 2] = __byte_perm(m[f1 >> 2], wrap_bit(o2[i], v2[j]), kPark[f1 & 3]);
    if (WEIGHTED) m[f2 >> 2] = __byte_perm(m[f2 >> 2], wrap_bit(o0[WEIGHTED ? i : 0], w[j]), kPark[f2 & 3]);
  }
  unsigned any = 0;
#pragma unroll
  for (int i = 0; i < NM; i++) any |= m[i];
  if (any & 0x80808080u) {
#pragma unroll
    for (int i = 0; i < U * CPL; i++) {
      const int j = i / CPL, k = i - j * CPL;
      const int f0 = PER * i, f1 = PER * i + 1, f2 = PER * i + 2;
      const bool e1 = (m[f0 >> 2] >> (8 * (f0 & 3) + 7)) & 1u, e2 = (m[f1 >> 2] >> (8 * (f1 & 3) + 7)) & 1u;
      const bool ev = WEIGHTED ? (m[(WEIGHTED ? f2 : 0) >> 2] >> (8 * (f2 & 3) + 7)) & 1u : false;
      if (ev | e1 | e2) {
        const int b = (int)((pa[i] - hb[k]) / (12u * LPR));
        eng_carry(a.acc, a.descs, node, a.cols, col_lane + k, a.nb1, b, ev ? 1 : 0, e1 ? (v1[j] < 0 ? -1 : 1) : 0,
                  e2 ? (v2[j] < 0 ? -1 : 1) : 0);
      }
    }
  }
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// 1024 threads leave 64 registers per thread, and a single spilled register in the row loop costs ~30% (with 227 KB
// of shared memory carved out, local memory has almost no L1 behind it): rows per batch are chosen so that every
// instantiation compiles without spills — eight cells per lane in flight, four where the weights need registers too.
template <int CPL, bool WEIGHTED>
struct EngCfg {
  static constexpr int threads = 1024;
  static constexpr int U = WEIGHTED ? 4 / CPL : (CPL == 1 ? 6 : 4);  // 6, 4 (unweighted) or 4, 2 (weighted): even
};

template <int CPL, bool WEIGHTED, int LPR, bool NA, bool U8 = false>
__global__ void __launch_bounds__(EngCfg<CPL, WEIGHTED>::threads, 1) eng_hist_kernel(const __grid_constant__ EngHistArgs a) {
  static_assert(NA || !WEIGHTED, "zero-weight rows are turned into NaNs: weighted shards need the NA test");
  static_assert(!U8 || (!WEIGHTED && LPR == 32), "u8 storage: unweighted trees, 32 lanes per row");
  using V = typename XVec<CPL, U8>::type;
  constexpr int THREADS = EngCfg<CPL, WEIGHTED>::threads, NW = THREADS / 32;
  constexpr int U = LPR == 16 && !WEIGHTED ? 4 : EngCfg<CPL, WEIGHTED>::U;  // (the 16-lane layout spills at eight rows)
  extern __shared__ __align__(16) unsigned h[];
  __shared__ unsigned s_mn[CPL * 32], s_mx[CPL * 32];
  __shared__ unsigned char s_tracked[CPL * 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nb1 = a.nb1, nbs = a.nbs;
  const int triples = CPL * nbs * LPR;
  const bool on = LPR == 32 || lane < LPR;  // with 16 lanes per row the upper half of every warp only loads and prefetches
  const int col_lane = (int)blockIdx.y * LPR * CPL + (lane & (LPR - 1)) * CPL;
  const int n_nodes = min(a.n_nodes, a.seg[kSegNodes]);  // the table's own count bounds every read of it
  if (n_nodes < 1) return;
  const int64_t live = a.seg[n_nodes];
  int p0 = (int)(live * (int64_t)blockIdx.x / gridDim.x);
  const int p1 = (int)(live * (int64_t)(blockIdx.x + 1) / gridDim.x);
  if (p0 >= p1) return;
  int node = seg_of(a.seg, n_nodes, p0);

  for (int i = tid; i < triples * 3; i += THREADS) h[i] = ((i / LPR) % 3 == 0) ? 0u : kWyBias;

  unsigned hb[CPL];  // shared address of (slot k, bin 0, channel 0, this lane)
#pragma unroll
  for (int k = 0; k < CPL; k++) {
    hb[k] = (unsigned)__cvta_generic_to_shared(h) + (lane & (LPR - 1)) * 4 + (unsigned)(k * nbs) * (12u * LPR);
    asm volatile("" : "+r"(hb[k]));  // keep it in a register: recomputing it costs seven instructions per cell
  }
  // lanes past the last column read the row's last CPL floats instead (never accumulated: their columns are untracked)
  const unsigned ldt4 = a.ldt4;  // row pitch in bytes: one IMAD.WIDE per row address
  const char* xrow = U8 ? reinterpret_cast<const char*>(a.XB + min(col_lane, (int)ldt4 - CPL))
                        : reinterpret_cast<const char*>(a.XT + min(col_lane, (int)a.ldt - CPL));
  asm volatile("" : "+l"(xrow));  // kept in a register pair: rebuilt per batch it turns every row address into three instructions
  const int32_t* __restrict__ gidx = a.gidx;
  const uint2* __restrict__ gval = a.gval;
  const int32_t* __restrict__ gw = a.gw;
  EngLane<CPL> c;

  while (p0 < p1) {
    const int nend = a.seg[node + 1];
    const int send = nend < p1 ? nend : p1;
    if (send <= p0) {  // a node without rows on this shard
      node++;
      continue;
    }
    bool tracked[CPL];
#pragma unroll
    for (int k = 0; k < CPL; k++) {
      const int col = col_lane + k;
      tracked[k] = false;
      c.nb[k] = 1;
      c.mn[k] = c.st[k] = 0.0;
      if (on && col < a.cols) {
        const h2ok_col_desc d = a.descs[(int64_t)node * a.cols + col];
        if (d.nbin >= nbs) atomicAdd(a.err, 1);  // the caller's level_bins hint was too small: counted, not accumulated
        else if (U8 && d.nbin > 0 && !(d.step == 1.0 && d.min == (double)(int)d.min)) atomicAdd(a.err, 1);  // not a step-1 integer descriptor
        else if (d.nbin > 0) {
          tracked[k] = true;
          c.nb[k] = d.nbin;
          c.mn[k] = d.min;
          c.st[k] = d.step;
          c.imn[k] = (int)d.min;
        }
      }
      if (!tracked[k]) c.imn[k] = 0;
      c.icmn[k] = 0x7fffffff;
      c.icmx[k] = -1;
      c.fmn[k] = __int_as_float(0x7f800000);
      c.fmx[k] = __int_as_float(0xff800000);
    }
    if (warp == 0) {
#pragma unroll
      for (int k = 0; k < CPL; k++) {
        s_mn[k * 32 + lane] = kEncPosInf;
        s_mx[k * 32 + lane] = kEncNegInf;
        s_tracked[k * 32 + lane] = tracked[k] ? 1 : 0;
      }
    }
    __syncthreads();  // histogram words and range slots are ready

    // Every warp takes U consecutive records at a time; the warps of the CTA interleave.  The row numbers of a batch
    // are fetched one batch ahead, and as soon as they are there (after the bin phase) the rows themselves are pulled
    // into L2 with one prefetch per row (each lane names its own sector), so that the next batch's row loads are L2
    // hits instead of DRAM round trips — without holding a single register for data in flight.
    // Batches start on multiples of U records, so the (warp-uniform) record loads are 128-bit: one load brings four
    // row numbers or two {wyq, wyyq} pairs.  The records before the first aligned batch and after the last whole one
    // go one row at a time.
    {
      constexpr int stride = NW * U;
      auto one_row = [&](int q) {
        int v1[1], v2[1], w[1];
        V x[1];
        x[0] = __ldg(reinterpret_cast<const V*>(row_addr(xrow, (unsigned)__ldg(gidx + q), ldt4)));
        const uint2 r = __ldg(gval + q);
        v1[0] = (int)r.x;
        v2[0] = (int)r.y;
        w[0] = WEIGHTED ? __ldg(gw + q) : 1;
        unsigned pa[CPL];
        if (on) {
          eng_bins<CPL, WEIGHTED, 1, LPR, NA, U8>(c, hb, x, w, pa);
          eng_accum<CPL, WEIGHTED, 1, LPR>(a, hb, node, col_lane, pa, v1, v2, w);
        }
      };
      auto load_idx = [&](int q, int (&idx)[U]) {
        if (U % 4 == 0) {
#pragma unroll
          for (int j = 0; j < U; j += 4) {
            const int4 t = __ldg(reinterpret_cast<const int4*>(gidx + q + j));
            idx[j] = t.x; idx[j + 1 < U ? j + 1 : j] = t.y; idx[j + 2 < U ? j + 2 : j] = t.z; idx[j + 3 < U ? j + 3 : j] = t.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < U; j += 2) {
            const int2 t = __ldg(reinterpret_cast<const int2*>(gidx + q + j));
            idx[j] = t.x; idx[j + 1 < U ? j + 1 : j] = t.y;
          }
        }
      };
      const int pal = (p0 + U - 1) / U * U;  // first aligned record of the run
      if (warp == NW - 1)
        for (int q = p0; q < pal && q < send; q++) one_row(q);
      int p = pal + warp * U;
      int nx[U];
      if (p + U <= send) load_idx(p, nx);
      for (; p + U <= send; p += stride) {
        int v1[U], v2[U], w[U];
        V x[U];
#pragma unroll
        for (int j = 0; j < U; j++) x[j] = __ldg(reinterpret_cast<const V*>(row_addr(xrow, (unsigned)nx[j], ldt4)));
        const bool more = p + stride + U <= send;
        if (more) load_idx(p + stride, nx);
#pragma unroll
        for (int j = 0; j < U; j += 2) {
          const uint4 r = __ldg(reinterpret_cast<const uint4*>(gval + p + j));
          v1[j] = (int)r.x;
          v2[j] = (int)r.y;
          v1[j + 1 < U ? j + 1 : j] = (int)r.z;
          v2[j + 1 < U ? j + 1 : j] = (int)r.w;
        }
#pragma unroll
        for (int j = 0; j < U; j++) w[j] = WEIGHTED ? __ldg(gw + p + j) : 1;
        unsigned pa[U * CPL];
        if (on) eng_bins<CPL, WEIGHTED, U, LPR, NA, U8>(c, hb, x, w, pa);
        // The two warp barriers are there for ptxas, not for the threads: without them it starts a cell's atomics as
        // soon as that cell's bin is known and threads the prefetches through the atomics, i.e. it undoes the phases
        // (all bins, then the prefetches, then all atomics back to back).  Memory operations do not move across a
        // barrier; the arithmetic still does.  100M x 64: 6.2-6.3 ms per level against 6.3-6.65 (same box, same run).
        __syncwarp();
        if (more) {
#pragma unroll
          for (int j = 0; j < U; j++) prefetch_l2(row_addr(xrow, (unsigned)nx[j], ldt4));
        }
        __syncwarp();
        if (on) eng_accum<CPL, WEIGHTED, U, LPR>(a, hb, node, col_lane, pa, v1, v2, w);
      }
      for (; p < send; p++) one_row(p);  // the run's last, partial batch
    }

    // observed ranges: registers -> shared (integer min / max of the order-preserving encoding) -> global
#pragma unroll
    for (int k = 0; k < CPL; k++) {
      if (U8) {  // codes are small non-negative integers: exact as floats
        c.fmn[k] = c.icmx[k] < 0 ? __int_as_float(0x7f800000) : (float)c.icmn[k];
        c.fmx[k] = c.icmx[k] < 0 ? __int_as_float(0xff800000) : (float)c.icmx[k];
      }
      if (tracked[k]) {
        atomicMin(&s_mn[k * 32 + lane], enc_f32_bits(__float_as_uint(c.fmn[k])));
        atomicMax(&s_mx[k * 32 + lane], enc_f32_bits(__float_as_uint(c.fmx[k])));
      }
    }
    __syncthreads();
    if (tid < CPL * 32 && s_tracked[tid]) {
      const int col = (int)blockIdx.y * LPR * CPL + (tid & 31) * CPL + (tid >> 5);
      const unsigned e0 = s_mn[tid], e1 = s_mx[tid];
      if (e0 != kEncPosInf) atomicMin(&a.min2_enc[(int64_t)node * a.cols + col], e0);
      if (e1 != kEncNegInf) atomicMax(&a.maxin_enc[(int64_t)node * a.cols + col], e1);
    }
    // flush the touched (slot, bin, lane) triples to the 64-bit global accumulators, undo the centring with the
    // triple's own count, and re-arm the words for the next node
    const int m1 = a.mnode[2 * node], m2 = a.mnode[2 * node + 1];
    for (int t = tid; t < triples; t += THREADS) {
      const int kb = t / LPR, ln = t % LPR;
      unsigned* w0 = h + (kb * 3) * LPR + ln;
      const unsigned c0 = w0[0], u1 = w0[LPR], u2 = w0[2 * LPR];
      if (c0 | (u1 ^ kWyBias) | (u2 ^ kWyBias)) {
        w0[0] = 0u;
        w0[LPR] = kWyBias;
        w0[2 * LPR] = kWyBias;
        const int k = kb / nbs, b = kb - k * nbs;
        if (s_tracked[k * 32 + ln]) {
          const int col = (int)blockIdx.y * LPR * CPL + ln * CPL + k;
          unsigned long long* g = reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)node * a.cols + col) * nb1 + b) * 3;
          const long long s1 = (long long)u1 - (long long)kWyBias + (long long)c0 * m1;
          const long long s2 = (long long)u2 - (long long)kWyBias + (long long)c0 * m2;
          if (c0) atomicAdd(g, (unsigned long long)c0);
          if (s1) atomicAdd(g + 1, (unsigned long long)s1);
          if (s2) atomicAdd(g + 2, (unsigned long long)s2);
        }
      }
    }
    __syncthreads();  // s_tracked, the offsets and the range slots are rewritten at the top of the next run
    p0 = send;
    node++;
  }
}

// 1024-thread exclusive scan; returns the exclusive prefix, *total gets the block sum (valid for every thread)
__device__ __forceinline__ int block_exscan_1024(int v, int* s_warp, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = s_warp[lane];
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    s_warp[lane] = winc - w;
    if (lane == 31) s_warp[32] = winc;
  }
  __syncthreads();
  const int ex = s_warp[warp] + inc - v;
  *total = s_warp[32];
  __syncthreads();  // s_warp may be reused
  return ex;
}

// ------------------------------------------------------------------------------------------------ u8 storage
// Pre-binned input: every cell is an integer code 0..254 (255 = NA), one byte.  The shard keeps
// the codes twice, row-major for the histogram pass (64 B per row at 64 columns) and column-major for routing, which
// compares the code itself with the threshold — exact, no float to fall back to.
__global__ void __launch_bounds__(256) eng_load_u8_kernel(const unsigned char* __restrict__ B, int64_t ld, int64_t rows,
                                                          int col0, int ncols, unsigned char* __restrict__ XB, int64_t ldb,
                                                          unsigned char* __restrict__ code8, int64_t ld8,
                                                          int* __restrict__ na_flag) {
  __shared__ unsigned char t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int c = c0 + ty + i * 8;
    const int64_t r = r0 + tx;
    if (c < ncols && r < rows) {
      const unsigned char v = B[(int64_t)c * ld + r];
      t[ty + i * 8][tx] = v;
      code8[(int64_t)(col0 + c) * ld8 + r] = v;
      if (v == 255) *na_flag = 1;
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int64_t r = r0 + ty + i * 8;
    const int c = c0 + tx;
    if (r < rows && c < ncols) XB[r * ldb + col0 + c] = t[tx][ty + i * 8];
  }
}

// column ranges of a block of codes, as encoded floats (what DHistogram.initialHist reads)
__global__ void __launch_bounds__(256) eng_u8_ranges_kernel(const unsigned char* __restrict__ B, int64_t ld, int64_t rows,
                                                            uint32_t* __restrict__ mn, uint32_t* __restrict__ mx) {
  const unsigned char* b = B + (int64_t)blockIdx.y * ld;
  int lo = 256, hi = -1;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int v = b[r];
    if (v != 255) {
      lo = min(lo, v);
      hi = max(hi, v);
    }
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  if ((threadIdx.x & 31) == 0 && hi >= 0) {
    atomicMin(&mn[blockIdx.y], enc_f32_bits(__float_as_uint((float)lo)));
    atomicMax(&mx[blockIdx.y], enc_f32_bits(__float_as_uint((float)hi)));
  }
}

// fp32 cells that already hold codes (0..254, NaN = NA) -> bytes; anything else is flagged
__global__ void eng_f32_codes_kernel(const float* __restrict__ X, int64_t n, unsigned char* __restrict__ B, int* __restrict__ bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = X[i];
  if (x != x) B[i] = 255;
  else if (x >= 0.0f && x <= 254.0f && x == (float)(int)x) B[i] = (unsigned char)(int)x;
  else {
    B[i] = 255;
    *bad = 1;
  }
}

// a2 as a pre-binning pass: the code of every cell against its column's (root) descriptor
__global__ void __launch_bounds__(256) eng_prebin_kernel(const float* __restrict__ X, int64_t ld, int64_t rows,
                                                         const h2ok_col_desc* __restrict__ descs, unsigned char* __restrict__ B,
                                                         int64_t ldb) {
  const h2ok_col_desc d = descs[blockIdx.y];
  const int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (r >= rows) return;
  const float x = X[(int64_t)blockIdx.y * ld + r];
  int b = 255;
  if (x == x && d.nbin > 0) b = bin_of(x, d.min, d.step, d.nbin);
  B[(int64_t)blockIdx.y * ldb + r] = (unsigned char)b;
}

// ------------------------------------------------------------------------------------------------ sparse storage
// For matrices that are mostly exact zeros (BASELINE config 4: 90 % zeros; H2O itself keeps such columns in sparse
// chunks, CXIChunk / CX0Chunk) the shard is stored column by column as the list of its NONZERO cells {row, value},
// rows ascending (a NaN is a nonzero, -0.0 a zero): 8 bytes per nonzero instead of 4 bytes per cell.
//
// The histogram pass then works like H2O's own (one column at a time, every node of the level at once) but on the
// nonzeros only: a CTA streams a stretch of one column's list, looks the row's node up in the row -> node map (2 bytes,
// L2-resident) and its targets in the row-order target arrays, bins against the node's descriptor for this column
// (a shared table indexed by node) and adds into a shared histogram [copy][node][bin][channel].  Every lane works on
// a nonzero cell: a warp instruction updates 32 cells where the row-wise layouts keep 3 of 32 lanes busy at 10 %
// density.  No row records are regrouped for such a shard — routing alone keeps the row -> node map current.
// The zeros are accounted for exactly afterwards (eng_csc_zero_kernel): a node's totals over all its rows minus what
// the column's nonzeros added is what its zero cells contribute; it goes to the bin 0.0 falls into, and a (node,
// column) with a zero cell has 0.0 inside its observed range.  Integer sums make this exact.
constexpr int kCscChunk = 65536;  // rows per compaction CTA

// nonzeros of every (column, chunk of kCscChunk rows)
__global__ void __launch_bounds__(1024) eng_csc_count_kernel(const float* __restrict__ X, int64_t ld, int64_t rows, int n_chunks,
                                                             unsigned* __restrict__ cnt) {
  __shared__ int s_warp[32];
  const float* x = X + (int64_t)blockIdx.y * ld;
  const int64_t r0 = (int64_t)blockIdx.x * kCscChunk;
  const int64_t r1 = r0 + kCscChunk < rows ? r0 + kCscChunk : rows;
  int n = 0;
  for (int64_t r = r0 + threadIdx.x; r < r1; r += 1024) n += !(x[r] == 0.0f);
  n = __reduce_add_sync(0xffffffffu, n);
  if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x < 32) {
    n = __reduce_add_sync(0xffffffffu, s_warp[threadIdx.x]);
    if (threadIdx.x == 0) cnt[(int64_t)blockIdx.y * n_chunks + blockIdx.x] = (unsigned)n;
  }
}

// stable compaction of one (column, chunk): {row, value bits} pairs, rows ascending
__global__ void __launch_bounds__(1024) eng_csc_fill_kernel(const float* __restrict__ X, int64_t ld, int64_t rows, int n_chunks,
                                                            const int64_t* __restrict__ start, uint2* __restrict__ out) {
  __shared__ int s_warp[33];
  const float* x = X + (int64_t)blockIdx.y * ld;
  const int64_t r0 = (int64_t)blockIdx.x * kCscChunk;
  const int64_t r1 = r0 + kCscChunk < rows ? r0 + kCscChunk : rows;
  int64_t base = start[(int64_t)blockIdx.y * n_chunks + blockIdx.x];
  const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  for (int64_t rb = r0; rb < r1; rb += 1024) {
    const int64_t r = rb + threadIdx.x;
    const float v = r < r1 ? x[r] : 0.0f;
    const bool nz = !(v == 0.0f);
    const unsigned bal = __ballot_sync(0xffffffffu, nz);
    int tot;
    const int ex = block_exscan_1024(lane == 0 ? __popc(bal) : 0, s_warp, &tot);
    const int wbase = __shfl_sync(0xffffffffu, ex, 0);
    if (nz) out[base + wbase + __popc(bal & lt)] = make_uint2((unsigned)r, __float_as_uint(v));
    base += tot;
  }
}

// one cell of the sparse shard (routing's exact comparison on a code tie): binary search of the row in its column
__device__ __forceinline__ float csc_get(const uint2* const* __restrict__ cptr, const int64_t* __restrict__ clen, int64_t r,
                                         int col) {
  const uint2* p = cptr[col];
  int64_t lo = 0, hi = clen[col];
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if ((int64_t)p[mid].x < r) lo = mid + 1;
    else hi = mid;
  }
  if (lo < clen[col] && (int64_t)p[lo].x == r) return __uint_as_float(p[lo].y);
  return 0.0f;
}

// per node: rows, sum of wyq, sum of wyyq over ALL its rows (what the zero cells are worked out from).  Rows of one warp
// that share a node are added up with shuffles first; the CTA's partial sums live in shared memory as 32-bit words
// (the only native shared atomic): a group sum s (|s| < 2^31) goes in as s >> 16 and s & 0xFFFF, and the words are
// flushed to the 64-bit global totals every 32768 rows, before either half can overflow.  (64-bit shared atomics are
// CAS loops: with 32 nodes they made this pass cost 8 ms per level.)
// The same pass packs {node, wyq, wyyq} of every row into ONE 8-byte record (common.cuh rec_pack): the histogram pass
// then fetches a row's node and targets with a single gather instead of three — its gathers are sector-granular L2
// traffic, 168 GB per level on config 4 with three arrays against 64 GB with one.
__global__ void __launch_bounds__(512) eng_node_totals_kernel(int64_t rows, const uint16_t* __restrict__ nid16,
                                                              const int32_t* __restrict__ wyq, const int32_t* __restrict__ wyyq,
                                                              int n_nodes, unsigned long long* __restrict__ tot,
                                                              uint2* __restrict__ rec, int K, int k) {
  extern __shared__ int s_t[];  // [n_nodes][5]: rows, hi/lo of sum wyq, hi/lo of sum wyyq
  const unsigned lane = threadIdx.x & 31u;
  const int64_t per = ((rows + gridDim.x - 1) / gridDim.x + 511) / 512 * 512;
  const int64_t r0 = (int64_t)blockIdx.x * per, r1 = r0 + per < rows ? r0 + per : rows;
  for (int64_t rs = r0; rs < r1; rs += 32768) {
    for (int i = threadIdx.x; i < 5 * n_nodes; i += 512) s_t[i] = 0;
    __syncthreads();
    const int64_t re = rs + 32768 < r1 ? rs + 32768 : r1;
    for (int64_t rb = rs; rb < re; rb += 512) {
      const int64_t r = rb + threadIdx.x;
      const unsigned nd = r < re ? nid16[r] : kNoNidConst;
      const int a = nd != kNoNidConst ? wyq[r] : 0, b = nd != kNoNidConst ? wyyq[r] : 0;
      // (a forest of K class trees keeps the K records of a row next to each other: rec[row][K])
      if (r < re) rec[r * K + k] = rec_pack(a, b, nd != kNoNidConst && (int)nd < n_nodes ? nd : kNoNode);
      unsigned todo = __ballot_sync(0xffffffffu, nd != kNoNidConst);
      while (todo) {
        const int src = __ffs(todo) - 1;
        const unsigned key = __shfl_sync(0xffffffffu, nd, src);
        const unsigned grp = __ballot_sync(0xffffffffu, nd == key);
        const int sa = __reduce_add_sync(0xffffffffu, nd == key ? a : 0), sb = __reduce_add_sync(0xffffffffu, nd == key ? b : 0);
        if ((int)lane == src && (int)key < n_nodes) {
          int* t = s_t + 5 * key;
          atomicAdd(t, __popc(grp));
          atomicAdd(t + 1, sa >> 16);
          atomicAdd(t + 2, sa & 0xFFFF);
          atomicAdd(t + 3, sb >> 16);
          atomicAdd(t + 4, sb & 0xFFFF);
        }
        todo &= ~grp;
      }
    }
    __syncthreads();
    for (int g = threadIdx.x; g < n_nodes; g += 512) {
      const int* t = s_t + 5 * g;
      if (t[0]) {
        atomicAdd(&tot[3 * g], (unsigned long long)t[0]);
        atomicAdd(&tot[3 * g + 1], (unsigned long long)((long long)t[1] * 65536 + (long long)t[2]));
        atomicAdd(&tot[3 * g + 2], (unsigned long long)((long long)t[3] * 65536 + (long long)t[4]));
      }
    }
    __syncthreads();
  }
}

// What the pass needs of one node for one column: two 16-byte shared-memory loads per cell.  The node minimum is a
// float whenever the descriptor came from this library (column ranges, min2 / maxIn and split points are all floats);
// a caller's descriptor with a minimum that is not sets kCscWide and the exact double is read from a side array.
struct CscNode {
  double st;
  float mn;
  unsigned nbw;  // nbin (0: (node, column) not tracked) | kCscWide
};
struct CscNodeB {
  int m1, m2;           // centring offsets of the node
  unsigned rmin, rmax;  // observed range so far (order-preserving codes)
};
constexpr unsigned kCscWide = 1u << 31;

struct CscHistArgs {
  const uint2* const* cptr;  // [cols] nonzeros of every column
  const int64_t* clen;       // [cols]
  int cols, chunks;
  int node_base;  // this launch covers nodes [node_base, node_base + n_nodes)
  const uint2* rec;  // [rows][K] {node, wyq, wyyq} packed (rec_pack), the K class trees of a forest side by side
  int K;
  const int32_t* mnode;
  int n_nodes;
  const h2ok_col_desc* descs;
  int nb1, nbs, copies;
  int64_t* acc;
  uint32_t* min2_enc;
  uint32_t* maxin_enc;
  int32_t* err;
};

// 512 threads: the pass is a chain of dependent loads per cell (list entry, then the row's record), so throughput
// comes from independent chains in flight — one cell per thread at a time ran at 18.6 ms per level on config 4,
// latency-bound; four at a time 8.1.  Two CTAs share an SM whenever their histograms fit.
constexpr int kCscThreads = 512;

// The three parts the two passes below share: (1) node tables and zeroed histogram copies in shared memory, (2) one cell
// of one node added in, (3) the copies added up and flushed.
struct CscSmem {
  CscNode* tab;
  CscNodeB* tab2;
  double* mn_wide;
  unsigned* h;
};
constexpr size_t kCscNodeBytes = sizeof(CscNode) + sizeof(CscNodeB) + 8;  // two table entries, the exact minimum

__device__ __forceinline__ CscSmem csc_setup(const CscHistArgs& a, unsigned char* smem_raw, int col, int tid) {
  const int n = a.n_nodes, nbs = a.nbs, nbase = a.node_base;
  CscSmem m;
  m.tab = reinterpret_cast<CscNode*>(smem_raw);
  m.tab2 = reinterpret_cast<CscNodeB*>(m.tab + n);
  m.mn_wide = reinterpret_cast<double*>(m.tab2 + n);
  m.h = reinterpret_cast<unsigned*>(m.mn_wide + n);  // [copy][node][bin][channel]
  for (int g = tid; g < n; g += kCscThreads) {
    const h2ok_col_desc d = a.descs[(int64_t)(nbase + g) * a.cols + col];
    CscNode t;
    t.mn = (float)d.min;
    t.st = d.step;
    unsigned nb = d.nbin > 0 ? (unsigned)d.nbin : 0u;
    if (d.nbin >= nbs) {
      if (blockIdx.x == 0) atomicAdd(a.err, 1);
      nb = 0;
    }
    t.nbw = nb | ((double)t.mn != d.min ? kCscWide : 0u);
    m.tab[g] = t;
    CscNodeB t2;
    t2.m1 = a.mnode[2 * (nbase + g)];
    t2.m2 = a.mnode[2 * (nbase + g) + 1];
    t2.rmin = kEncPosInf;
    t2.rmax = kEncNegInf;
    m.tab2[g] = t2;
    m.mn_wide[g] = d.min;
  }
  const int words = a.copies * n * nbs * 3;
  for (int i = tid; i < words; i += kCscThreads) m.h[i] = (i % 3 == 0) ? 0u : kWyBias;
  __syncthreads();
  return m;
}

// cold: the wrap of a shared word, with the launch arguments read here rather than kept ready in the hot loop
__device__ __noinline__ void csc_carry(const CscHistArgs& a, int nd, int col, int b, unsigned f1, unsigned f2, int v1, int v2) {
  eng_carry(a.acc, a.descs, a.node_base + nd, a.cols, col, a.nb1, b, 0, (f1 >> 31) ? (v1 < 0 ? -1 : 1) : 0,
            (f2 >> 31) ? (v2 < 0 ? -1 : 1) : 0);
}

// (rx, ry): the row's packed record for one tree; a row in a leaf (node 0x3FF) and the all-ones record of a list entry
// past the end both fall outside [0, n)
__device__ __forceinline__ void csc_cell(const CscHistArgs& a, const CscSmem& m, unsigned hbase, int col, unsigned rx, unsigned ry,
                                         float x) {
  const unsigned nd = rec_node(rx, ry) - (unsigned)a.node_base;
  if (nd >= (unsigned)a.n_nodes) return;
  const CscNode t = m.tab[nd];
  const CscNodeB t2 = m.tab2[nd];
  const int nb = (int)(t.nbw & 0x7FFFFFFFu);
  if (nb == 0) return;
  const int v1 = ((int)rx >> 5) - t2.m1, v2 = (int)(ry >> 5) - t2.m2;  // wyq signed, wyyq unsigned
  const double mn = (t.nbw & kCscWide) ? m.mn_wide[nd] : (double)t.mn;
  const int b = bin_fast(x, mn, t.st, nb);
  const unsigned addr = hbase + (unsigned)(((int)nd * a.nbs + b) * 12);
  sh_inc(addr);
  const unsigned o1 = sh_add(addr + 4u, (unsigned)v1), o2 = sh_add(addr + 8u, (unsigned)v2);
  const unsigned f1 = wrap_bit(o1, v1), f2 = wrap_bit(o2, v2);
  if ((f1 | f2) >> 31) csc_carry(a, (int)nd, col, b, f1, f2, v1, v2);
  // enc_f32_bits as one shift and one xor; a NaN's code lies outside [enc(-inf), enc(+inf)], so it is filtered by the
  // same comparison that says "inside the range seen so far" only when that range is not empty: test it explicitly
  const unsigned xb = __float_as_uint(x);
  const unsigned ex = xb ^ ((unsigned)((int)xb >> 31) | 0x80000000u);
  if ((ex < t2.rmin) | (ex > t2.rmax)) {
    if (x == x) {
      if (ex < t2.rmin) atomicMin(&m.tab2[nd].rmin, ex);
      if (ex > t2.rmax) atomicMax(&m.tab2[nd].rmax, ex);
    }
  }
}

// copies added up, centring undone with the bin's own count
__device__ __forceinline__ void csc_flush(const CscHistArgs& a, const CscSmem& m, int col, int tid) {
  const int n = a.n_nodes, nbs = a.nbs, nbase = a.node_base;
  __syncthreads();
  for (int t = tid; t < n * nbs; t += kCscThreads) {
    const int g = t / nbs, b = t - g * nbs;
    unsigned long long nc = 0;
    long long n1 = 0, n2 = 0;
    for (int cp = 0; cp < a.copies; cp++) {
      const unsigned* w = m.h + ((cp * n + g) * nbs + b) * 3;
      nc += w[0];
      n1 += (long long)w[1] - (long long)kWyBias;
      n2 += (long long)w[2] - (long long)kWyBias;
    }
    if (nc | (unsigned long long)n1 | (unsigned long long)n2) {
      unsigned long long* gp = reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)(nbase + g) * a.cols + col) * a.nb1 + b) * 3;
      const long long s1 = n1 + (long long)nc * m.tab2[g].m1, s2 = n2 + (long long)nc * m.tab2[g].m2;
      if (nc) atomicAdd(gp, nc);
      if (s1) atomicAdd(gp + 1, (unsigned long long)s1);
      if (s2) atomicAdd(gp + 2, (unsigned long long)s2);
    }
  }
  for (int g = tid; g < n; g += kCscThreads) {
    const CscNodeB t2 = m.tab2[g];
    if (t2.rmin != kEncPosInf) atomicMin(&a.min2_enc[(int64_t)(nbase + g) * a.cols + col], t2.rmin);
    if (t2.rmax != kEncNegInf) atomicMax(&a.maxin_enc[(int64_t)(nbase + g) * a.cols + col], t2.rmax);
  }
}

// One tree.  U list entries per thread and batch, one 8-byte record gather each, and the loads of the next two batches
// in flight while a batch is added in: a batch's list entries are fetched two iterations ahead and its row records one
// iteration ahead, so neither latency sits on the chain.  Config 4, ms per level: four entries per thread without the
// look-ahead 7.7, U = 1 / 2 / 4 with it 7.6 / 6.9 / 7.0.
template <int U>
__global__ void __launch_bounds__(kCscThreads, 2) eng_hist_csc_kernel(const __grid_constant__ CscHistArgs a) {
  constexpr int THREADS = kCscThreads, STRIDE = THREADS * U;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int col = blockIdx.y;
  const int64_t len = a.clen[col];
  const int64_t i0 = len * (int64_t)blockIdx.x / gridDim.x, i1 = len * (int64_t)(blockIdx.x + 1) / gridDim.x;
  if (i0 >= i1) return;
  // (a list has fewer than 2^31 entries: 32-bit positions inside the CTA's stretch keep the loop's bookkeeping short —
  // the pass is bound by issue slots, 79 % busy in profiles/r02_hist_csc_ncu_full.csv)
  const int cnt = (int)(i1 - i0);
  const uint2* __restrict__ nz = a.cptr[col] + i0;
  const uint2* __restrict__ rec = a.rec;
  const uint2 ones = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
  uint2 ea[U], eb[U], ra[U], rb[U];  // list entries of batches t + 1 / t + 2 and records of batches t / t + 1, by turns
  float x0[U];
#pragma unroll
  for (int u = 0; u < U; u++) {
    const int i = tid + u * THREADS;
    const uint2 e = i < cnt ? __ldg(nz + i) : make_uint2(0u, 0u);
    x0[u] = __uint_as_float(e.y);
    ra[u] = i < cnt ? __ldg(rec + e.x) : ones;
    ea[u] = i + STRIDE < cnt ? __ldg(nz + i + STRIDE) : make_uint2(0u, 0u);
  }
  const CscSmem m = csc_setup(a, smem_raw, col, tid);
  const unsigned hbase = (unsigned)__cvta_generic_to_shared(m.h) + (unsigned)((lane & (a.copies - 1)) * a.n_nodes * a.nbs) * 12u;
  // one batch: gather the next batch's records, fetch the entries of the one after, add this batch in.  Written twice
  // with the buffers swapped so that nothing but the cell values moves between registers.
#define H2OK_CSC_STEP(EA, EB, RA, RB, IB)                                                            \
  {                                                                                                  \
    _Pragma("unroll") for (int u = 0; u < U; u++) {                                                  \
      const int i = (IB) + u * THREADS;                                                              \
      RB[u] = i + STRIDE < cnt ? __ldg(rec + EA[u].x) : ones;                                        \
      EB[u] = i + 2 * STRIDE < cnt ? __ldg(nz + i + 2 * STRIDE) : make_uint2(0u, 0u);                \
    }                                                                                                \
    _Pragma("unroll") for (int u = 0; u < U; u++) csc_cell(a, m, hbase, col, RA[u].x, RA[u].y, x0[u]); \
    _Pragma("unroll") for (int u = 0; u < U; u++) x0[u] = __uint_as_float(EA[u].y);                  \
  }
  // (a bit mask of the rows still in the tree, consulted before the record gather once most rows have left: slower,
  // 7.4 ms against 6.0 on config 4's levels with a tenth of the rows live — the gather is not what those levels wait for)
  // (two more variants for levels where most rows have left the tree, both measured on config 4's levels with a tenth
  // of the rows live: live cells collected in a per-warp ring and added in 32 at a time, all lanes busy — 6.0 ms against
  // 6.2; the mask and the ring say the same thing: those levels wait neither for the record gather nor for the issue
  // slots of the per-cell code; an L2 prefetch of the list 2K / 8K / 32K entries ahead made every level slower, 7.7-7.9 ms
  // per level against 7.1, so it is not the depth of the list stream either)
  for (int ib = tid; ib < cnt; ib += 2 * STRIDE) {
    H2OK_CSC_STEP(ea, eb, ra, rb, ib)
    H2OK_CSC_STEP(eb, ea, rb, ra, ib + STRIDE)  // (past the end: every record is the all-ones one, nothing is added)
  }
#undef H2OK_CSC_STEP
  csc_flush(a, m, col, tid);
}

// A forest of K class trees: a row's K records sit side by side (rec[row][K]) and every list entry is added into the
// histograms of all K trees, so the column lists are streamed once per level of the whole forest.  Measured on config 4
// (K = 10, all trees alive): 87 ms per level against 10 x 7.7 ms for the trees one after the other — no gain.  The pass
// is bound by shared-memory wavefronts and issue slots of the per-(entry, tree) work (node table, binning, three
// atomics, range), which a forest does not reduce; what it saves is the list stream (2 of ~30 wavefronts per warp and
// tree) and what it loses is gather locality (a row's 80 bytes are a line of their own, the single tree's 8-byte
// records share lines between neighbouring rows).  A variant where eight lanes fetch a row's line together and keep
// the records they fetched (trees 2p, 2p + 1 of entry q) cut the gathers tenfold and still ran at 130-185 ms: 16 steps
// of 20 live lanes per 32 entries instead of 10 steps of 32.  DESIGN.md §3.4.
template <int U>
__global__ void __launch_bounds__(kCscThreads, 2) eng_hist_csc_forest_kernel(const __grid_constant__ CscHistArgs a) {
  constexpr int THREADS = kCscThreads;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int col = blockIdx.y;
  const int64_t len = a.clen[col];
  const int64_t i0 = len * (int64_t)blockIdx.x / gridDim.x, i1 = len * (int64_t)(blockIdx.x + 1) / gridDim.x;
  if (i0 >= i1) return;
  const CscSmem m = csc_setup(a, smem_raw, col, tid);
  const unsigned hbase = (unsigned)__cvta_generic_to_shared(m.h) + (unsigned)((lane & (a.copies - 1)) * a.n_nodes * a.nbs) * 12u;
  const uint2* __restrict__ nz = a.cptr[col];
  const uint2* __restrict__ rec = a.rec;
  const int K = a.K;
  for (int64_t ib = i0 + tid; ib < i1; ib += (int64_t)THREADS * U) {
    uint2 e[U], rc[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t i = ib + (int64_t)u * THREADS;
      e[u] = i < i1 ? __ldg(nz + i) : make_uint2(0u, 0u);
    }
    for (int k = 0; k < K; k++) {
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int64_t i = ib + (int64_t)u * THREADS;
        rc[u] = i < i1 ? __ldg(rec + (int64_t)e[u].x * K + k) : make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
      }
#pragma unroll
      for (int u = 0; u < U; u++) csc_cell(a, m, hbase, col, rc[u].x, rc[u].y, __uint_as_float(e[u].y));
    }
  }
  csc_flush(a, m, col, tid);
}

// The zero cells of every tracked (node, column): totals of the node minus what the column's nonzeros added.
__global__ void __launch_bounds__(256) eng_csc_zero_kernel(int n_nodes, int cols, const h2ok_col_desc* __restrict__ descs, int nb1,
                                                           const unsigned long long* __restrict__ tot, int64_t* __restrict__ acc,
                                                           uint32_t* __restrict__ min2_enc, uint32_t* __restrict__ maxin_enc) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t pair = (int64_t)blockIdx.x * 8 + warp;
  if (pair >= (int64_t)n_nodes * cols) return;
  const h2ok_col_desc d = descs[pair];
  if (d.nbin <= 0) return;
  const int g = (int)(pair / cols);
  long long* cell = reinterpret_cast<long long*>(acc) + pair * nb1 * 3;
  long long nc = 0, n1 = 0, n2 = 0;
  for (int b = lane; b <= d.nbin; b += 32) {
    nc += cell[3 * b];
    n1 += cell[3 * b + 1];
    n2 += cell[3 * b + 2];
  }
  for (int o = 16; o; o >>= 1) {
    nc += __shfl_xor_sync(0xffffffffu, nc, o);
    n1 += __shfl_xor_sync(0xffffffffu, n1, o);
    n2 += __shfl_xor_sync(0xffffffffu, n2, o);
  }
  if (lane != 0) return;
  const long long zc = (long long)tot[3 * g] - nc;
  if (zc <= 0) return;
  const int bz = bin_of(0.0f, d.min, d.step, d.nbin);
  cell[3 * bz] += zc;
  cell[3 * bz + 1] += (long long)tot[3 * g + 1] - n1;
  cell[3 * bz + 2] += (long long)tot[3 * g + 2] - n2;
  const unsigned ez = enc_f32_bits(__float_as_uint(0.0f));
  if (ez < min2_enc[pair]) min2_enc[pair] = ez;
  if (ez > maxin_enc[pair]) maxin_enc[pair] = ez;
}

// ------------------------------------------------------------------------------------------------ a6 + regroup
// Routing walks the rows in their ORIGINAL order (row -> node map nid16), not the node-grouped records: every access
// is then a stream.  The value the split needs is not gathered from the row-major copy — a gather of one float per
// row costs a 128-byte DRAM access each (ncu: 13.2 GB per level at 100M rows, profiles/r02_route_scatter_r01code_*) —
// but decided from an 8-bit order-preserving code of the cell, kept column-major (code8, one byte per cell, written
// once per data set): code(x) < code(splat) => x < splat, code(x) > code(splat) => x > splat, and only a row whose
// code EQUALS the threshold's (~1/255 of a column's rows) reads its float and compares exactly as H2O does
// (DecidedNode: x >= splat).  A warp of 32 consecutive rows reads one sector per distinct split column among its
// rows, so a level streams (distinct split columns) x rows bytes at most.
//
// Regrouping is a stable-by-range multi-way partition in the same row order: every CTA owns one contiguous range of
// rows in both kernels; eng_route_kernel counts its rows per child, eng_scan_kernel turns the [child][CTA] counts
// into first positions, eng_scatter_kernel writes the next level's records {row, targets - child's offset} there.
// Rows of a node stay ascending up to the order inside a 512-row tile — integer sums do not depend on it.
constexpr int kRouteThreads = 512;
constexpr unsigned kNoNid = kNoNidConst;

// order-preserving code of a non-NaN float: -0 counts as +0 so that the codes agree with float comparison
__device__ __forceinline__ unsigned enc_key(float x) { return enc_f32_bits(__float_as_uint(__fadd_rn(x, 0.0f))); }
// number of thresholds <= e (thr ascending, 256 entries, unused ones 0xFFFFFFFF)
__device__ __forceinline__ unsigned code_of(const unsigned* __restrict__ thr, unsigned e) {
  int lo = 0, hi = 255;  // answer in [0, 255): at most 254 real thresholds
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int mid = (lo + hi) >> 1;
    if (thr[mid] <= e) lo = mid + 1;
    else hi = mid;
  }
  return (unsigned)lo;
}

// One CTA per column: 4096 evenly spaced sample values, sorted; thresholds = the sample's 1/255 .. 254/255 quantiles.
__global__ void __launch_bounds__(1024) eng_thresholds_kernel(const float* __restrict__ X, int64_t ld, int64_t rows,
                                                              unsigned* __restrict__ thr) {
  constexpr int S = 4096;
  __shared__ unsigned s[S];
  const float* x = X + (int64_t)blockIdx.x * ld;
  for (int i = threadIdx.x; i < S; i += 1024) {
    const float v = x[(int64_t)i * rows / S];
    s[i] = v != v ? 0xFFFFFFFFu : enc_key(v);
  }
  __syncthreads();
  for (int k = 2; k <= S; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < S; i += 1024) {
        const int l = i ^ j;
        if (l > i) {
          const unsigned a = s[i], b = s[l];
          if (((i & k) == 0) == (a > b)) {
            s[i] = b;
            s[l] = a;
          }
        }
      }
      __syncthreads();
    }
  unsigned* t = thr + (int64_t)blockIdx.x * 256;
  for (int i = threadIdx.x; i < 256; i += 1024) t[i] = i < 254 ? s[(i + 1) * S / 255] : 0xFFFFFFFFu;
}

__global__ void __launch_bounds__(256) eng_code_kernel(const float* __restrict__ X, int64_t ld, int64_t rows,
                                                       const unsigned* __restrict__ thr, uint8_t* __restrict__ code8,
                                                       int64_t ld8, int* __restrict__ na_flag) {
  __shared__ unsigned t[256];
  t[threadIdx.x] = thr[(int64_t)blockIdx.y * 256 + threadIdx.x];
  __syncthreads();
  const float* x = X + (int64_t)blockIdx.y * ld;
  uint8_t* o = code8 + (int64_t)blockIdx.y * ld8;
  const int64_t r0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;  // four rows per thread: 128-bit in, 32-bit out
  if (r0 + 3 < rows && ((ld & 3) == 0)) {
    const float4 v = *reinterpret_cast<const float4*>(x + r0);
    const float f[4] = {v.x, v.y, v.z, v.w};
    unsigned w = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) w |= (f[k] != f[k] ? 255u : code_of(t, enc_key(f[k]))) << (8 * k);
    *reinterpret_cast<unsigned*>(o + r0) = w;
    if ((v.x != v.x) | (v.y != v.y) | (v.z != v.z) | (v.w != v.w)) *na_flag = 1;
  } else {
    for (int k = 0; k < 4 && r0 + k < rows; k++) {
      const float f = x[r0 + k];
      o[r0 + k] = (uint8_t)(f != f ? 255u : code_of(t, enc_key(f)));
      if (f != f) *na_flag = 1;
    }
  }
}

constexpr int kRouteBits = 0x80;  // RouteNode.na: the split is a categorical one (bit set in the split record)
struct RouteNode {  // what routing needs of one node's split, 16 bytes
  float splat;
  short col;          // < 0: the node does not split
  unsigned char na;   // H2OK_NA_*
  unsigned char cs;   // code of splat
  short c0, c1;       // child ids (< 0: that side is a leaf)
  int leaf0;          // leaf code of "did not split"
};

// rows of route / scatter CTA b: [b * per, min(rows, (b + 1) * per)), per a multiple of the tile
__device__ __forceinline__ int64_t route_per(int64_t rows, int G) {
  const int64_t per = (rows + G - 1) / G;
  return (per + kRouteThreads - 1) / kRouteThreads * kRouteThreads;
}

__global__ void __launch_bounds__(kRouteThreads, 4) eng_route_kernel(
    const float* __restrict__ XT, int64_t ldt, const uint2* const* __restrict__ cptr, const int64_t* __restrict__ clen,
    const uint8_t* __restrict__ code8, int64_t ld8,
    const unsigned* __restrict__ thr, int64_t rows, int n_nodes, const h2ok_split* __restrict__ splits,
    const int32_t* __restrict__ child_ids, int gid_base, uint16_t* __restrict__ nid16, int32_t* __restrict__ leaf,
    int32_t* __restrict__ cta_cnt, int n_slots, int exact_codes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RouteNode* tab = reinterpret_cast<RouteNode*>(smem_raw);
  int* s_cnt = reinterpret_cast<int*>(tab + n_nodes);
  for (int g = threadIdx.x; g < n_nodes; g += kRouteThreads) {
    const h2ok_split sp = splits[g];
    RouteNode t;
    t.splat = sp.splat;
    t.col = (short)sp.col;
    t.na = (unsigned char)(sp.na_dir | (sp.col >= 0 && sp.equal >= 2 ? kRouteBits : 0));
    t.cs = 0;
    if (!exact_codes && sp.col >= 0 && sp.equal < 2 && sp.na_dir != H2OK_NA_VS_REST && sp.splat == sp.splat)
      t.cs = (unsigned char)code_of(thr + (int64_t)sp.col * 256, enc_key(sp.splat));
    t.c0 = (short)child_ids[2 * g];
    t.c1 = (short)child_ids[2 * g + 1];
    t.leaf0 = (gid_base + g) * 3;
    tab[g] = t;
  }
  for (int i = threadIdx.x; i < n_slots; i += kRouteThreads) s_cnt[i] = 0;
  __syncthreads();
  const int64_t per = route_per(rows, gridDim.x);
  const int64_t r0 = (int64_t)blockIdx.x * per;
  const int64_t r1 = r0 + per < rows ? r0 + per : rows;
  const unsigned lane = threadIdx.x & 31u;
  // RU rows per thread and iteration: the row's node id, then its code in the node's split column, are two dependent
  // loads (the second a gather whose column changes from row to row at depth); with one row per thread the pass sat on
  // them (long-scoreboard stalls 19-25 per issue).  All RU node ids are fetched first, then all RU codes.  Two rows at
  // four CTAs per SM: 0.70 ms on a tree's last level at 100M rows against 0.81; four rows at three or two CTAs per SM
  // (40 / 48 registers) 0.99 / 0.94, eight rows 0.92.
  constexpr int RU = 2;
  for (int64_t base = r0; base < r1; base += kRouteThreads * RU) {
    unsigned nn[RU], cc[RU], next[RU];
#pragma unroll
    for (int k = 0; k < RU; k++) {
      const int64_t r = base + k * kRouteThreads + threadIdx.x;
      nn[k] = r < r1 ? nid16[r] : kNoNid;
    }
#pragma unroll
    for (int k = 0; k < RU; k++) {
      const int64_t r = base + k * kRouteThreads + threadIdx.x;
      cc[k] = 0;
      if (nn[k] != kNoNid && (int)nn[k] < n_nodes) {
        const int col = tab[nn[k]].col;
        if (col >= 0) cc[k] = code8[(int64_t)col * ld8 + r];
      }
    }
#pragma unroll
    for (int k = 0; k < RU; k++) {
      const int64_t r = base + k * kRouteThreads + threadIdx.x;
      const unsigned n = nn[k];
      next[k] = kNoNid;
      if (n != kNoNid && (int)n < n_nodes) {
        const RouteNode t = tab[n];
        if (t.col < 0) {
          leaf[r] = t.leaf0;
        } else {
          const unsigned c = cc[k];
          int way;
          const int na = t.na & ~kRouteBits;
          if (c == 255u) way = (na == H2OK_NA_LEFT) ? 0 : 1;
          else if (na == H2OK_NA_VS_REST) way = 0;
          else if (t.na & kRouteBits)  // a categorical split: the row's category against the split's bit set
            way = cat_goes_right(splits[n].cat_bits,
                                 exact_codes ? (float)c : (XT ? XT[r * ldt + t.col] : csc_get(cptr, clen, r, t.col)), na);
          else if (exact_codes) way = (float)c >= t.splat ? 1 : 0;  // u8 storage: the byte is the value
          else if (c != t.cs) way = c > t.cs ? 1 : 0;
          else way = (XT ? XT[r * ldt + t.col] : csc_get(cptr, clen, r, t.col)) >= t.splat ? 1 : 0;  // code tie: the float decides
          const int ch = way ? t.c1 : t.c0;
          if (ch < 0) leaf[r] = t.leaf0 + 1 + way;
          else next[k] = (unsigned)ch;
        }
        nid16[r] = (uint16_t)next[k];
      }
    }
    // rows per child: one shared-memory add per distinct child of the warp
#pragma unroll
    for (int k = 0; k < RU; k++) {
      const unsigned grp = __match_any_sync(0xffffffffu, next[k]);
      if (next[k] != kNoNid && lane == (unsigned)(__ffs(grp) - 1)) atomicAdd(&s_cnt[next[k]], __popc(grp));
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_slots; i += kRouteThreads) cta_cnt[(int64_t)i * gridDim.x + blockIdx.x] = s_cnt[i];
}

// One CTA, one warp per child at a time: exclusive scan of the child's per-CTA counts in place; then the first record
// of every child (children are numbered in slot order, so their runs follow each other in id order), the next
// level's node table and the children's centring offsets.
__global__ void __launch_bounds__(1024) eng_scan_kernel(int32_t* __restrict__ cta_cnt, int G, int n_nodes,
                                                        const int32_t* __restrict__ child_ids, int32_t* __restrict__ seg_next,
                                                        const h2ok_split* __restrict__ splits, double scale_wy,
                                                        double scale_wyy, const h2ok_quant* __restrict__ q_dev,
                                                        const int32_t* __restrict__ mnode,
                                                        int32_t* __restrict__ mnode_next,
                                                        const unsigned char* __restrict__ node_class,
                                                        unsigned char* __restrict__ node_class_next) {
  if (q_dev && !node_class) {  // the tree's scales live on the device (no host round trip at tree start)
    scale_wy = q_dev->scale_wy;
    scale_wyy = q_dev->scale_wyy;
  }
  __shared__ int s_warp[33];
  __shared__ int s_tot[kEngMaxNodes];
  __shared__ int s_nc;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // number of children = 1 + the largest child id
  int mx = -1;
  for (int i = threadIdx.x; i < 2 * n_nodes; i += 1024) mx = max(mx, child_ids[i]);
  mx = __reduce_max_sync(0xffffffffu, mx);
  if (lane == 0) s_warp[warp] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    int m = -1;
    for (int w = 0; w < 32; w++) m = max(m, s_warp[w]);
    s_nc = m + 1;
  }
  __syncthreads();
  const int nc = s_nc;
  for (int c = warp; c < nc; c += 32) {
    int32_t* row = cta_cnt + (int64_t)c * G;
    int run = 0;
    for (int b0 = 0; b0 < G; b0 += 32) {
      const int b = b0 + lane;
      const int v = b < G ? row[b] : 0;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      if (b < G) row[b] = run + inc - v;
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) s_tot[c] = run;
  }
  __syncthreads();
  const int g = threadIdx.x;
  int live;
  const int ex = block_exscan_1024(g < nc ? s_tot[g] : 0, s_warp, &live);
  if (g < nc) seg_next[g] = ex;
  if (g == 0) {
    seg_next[nc] = live;
    seg_next[kSegNodes] = nc;
  }
  if (g < n_nodes) {
    // centring offsets of the children (see eng_root_mean_kernel): mean y and mean y^2 of the side from the split
    // record, scaled like the targets; scale 0 (weighted shards) keeps every offset at 0
    const h2ok_split sp = splits[g];
    const int pm1 = mnode[2 * g], pm2 = mnode[2 * g + 1];
    if (node_class) {  // a forest: every class tree has its own fixed-point scales; children stay in their parent's tree
      scale_wy = q_dev[node_class[g]].scale_wy;
      scale_wyy = q_dev[node_class[g]].scale_wyy;
    }
#pragma unroll
    for (int way = 0; way < 2; way++) {
      const int ch = child_ids[2 * g + way];
      if (ch < 0) continue;
      if (node_class) node_class_next[ch] = node_class[g];
      const double pred = way ? sp.pred_right : sp.pred_left, se = way ? sp.se_right : sp.se_left;
      const double n = way ? sp.n_right : sp.n_left;
      int c1 = pm1, c2 = pm2;
      if (scale_wy > 0.0 && n > 0.0 && pred == pred && se == se) {
        const double a1 = pred * scale_wy, a2 = (se / n + pred * pred) * scale_wyy;
        if (fabs(a1) < 134217728.0 && fabs(a2) < 134217728.0) {  // 2^27: the centred targets stay far inside 32 bits
          c1 = __double2int_rn(a1);
          c2 = __double2int_rn(a2);
        }
      }
      mnode_next[2 * ch] = c1;
      mnode_next[2 * ch + 1] = c2;
    }
  }
}

// Tiles of kScatTile consecutive rows are grouped by child in shared memory first (a counting sort inside the tile), so
// that what goes out to a child's run is one contiguous stretch of records per tile and child instead of one record
// per store: at depth, 32 consecutive rows belong to ~32 different children, and writing them where they fall was
// one 32-byte L2 transaction per 4-byte store (4 ms per level at 100M rows and 128 children against 0.5 ms at the root).
constexpr int kScatThreads = 1024, kScatPer = 4, kScatTile = kScatThreads * kScatPer;

template <bool WEIGHTED>
__global__ void __launch_bounds__(kScatThreads, 2) eng_scatter_kernel(
    int64_t rows, const uint16_t* __restrict__ nid16, const int32_t* __restrict__ wq_rows,
    const int32_t* __restrict__ wyq_rows, const int32_t* __restrict__ wyyq_rows, const int32_t* __restrict__ seg_next,
    const int32_t* __restrict__ mnode_next, const int32_t* __restrict__ cta_cnt, int G, int32_t* __restrict__ gidx_n,
    uint2* __restrict__ gval_n, int32_t* __restrict__ gw_n) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_warp[33];
  const int nc = seg_next[kSegNodes];
  const int ncp = (nc + 1) & ~1;
  // per child: running first free record of this CTA's share, the tile's count, its first position in the tile,
  // (global first record - tile position), centring offsets
  int* s_off = reinterpret_cast<int*>(smem_raw);
  int* s_cnt = s_off + ncp;
  int* s_lb = s_cnt + ncp;
  int* s_gb = s_lb + ncp;
  int2* s_m = reinterpret_cast<int2*>(s_gb + ncp);
  uint2* s_val = reinterpret_cast<uint2*>(s_m + ncp);
  int* s_row = reinterpret_cast<int*>(s_val + kScatTile);
  int* s_w = s_row + kScatTile;
  uint16_t* s_child = reinterpret_cast<uint16_t*>(s_w + (WEIGHTED ? kScatTile : 0));
  // every scatter CTA covers the rows of G / gridDim.x routing CTAs
  const int per_cta = G / gridDim.x;
  for (int c = threadIdx.x; c < nc; c += kScatThreads) {
    s_off[c] = seg_next[c] + cta_cnt[(int64_t)c * G + blockIdx.x * per_cta];
    s_m[c] = make_int2(mnode_next[2 * c], mnode_next[2 * c + 1]);
  }
  const int64_t per = route_per(rows, G) * per_cta;
  const int64_t r0 = (int64_t)blockIdx.x * per;
  const int64_t r1 = r0 + per < rows ? r0 + per : rows;
  const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  for (int64_t base = r0; base < r1; base += kScatTile) {
    for (int c = threadIdx.x; c < nc; c += kScatThreads) s_cnt[c] = 0;
    __syncthreads();
    unsigned ch[kScatPer];
    int rank[kScatPer];
#pragma unroll
    for (int k = 0; k < kScatPer; k++) {
      const int64_t r = base + k * kScatThreads + threadIdx.x;
      ch[k] = r < r1 ? nid16[r] : kNoNid;
    }
    // the rows' targets, on their way while the tile is counted and scanned (loaded where they are used: 1.47 ms of
    // routing + regrouping per level at 100M rows against 1.43; one CTA per SM with 64 registers and no spills: 1.51)
    int ty[kScatPer], tyy[kScatPer];
#pragma unroll
    for (int k = 0; k < kScatPer; k++) {
      const int64_t r = base + k * kScatThreads + threadIdx.x;
      ty[k] = r < r1 ? __ldg(wyq_rows + r) : 0;
      tyy[k] = r < r1 ? __ldg(wyyq_rows + r) : 0;
    }
#pragma unroll
    for (int k = 0; k < kScatPer; k++) {  // position of the row among the tile's rows of its child
      const unsigned grp = __match_any_sync(0xffffffffu, ch[k]);
      const int leader = __ffs(grp) - 1;
      int first = 0;
      if (ch[k] != kNoNid && (int)lane == leader) first = atomicAdd(&s_cnt[ch[k]], __popc(grp));
      rank[k] = __shfl_sync(0xffffffffu, first, leader) + __popc(grp & lt);
    }
    __syncthreads();
    int total = 0;
    for (int c0 = 0; c0 < nc; c0 += kScatThreads) {  // tile positions of the children (exclusive scan of the counts)
      const int c = c0 + threadIdx.x;
      const int v = c < nc ? s_cnt[c] : 0;
      int tot;
      const int ex = block_exscan_1024(v, s_warp, &tot);
      if (c < nc) {
        s_lb[c] = total + ex;
        s_gb[c] = s_off[c] - (total + ex);
        s_off[c] += v;
      }
      total += tot;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kScatPer; k++) {
      if (ch[k] != kNoNid) {
        const int64_t r = base + k * kScatThreads + threadIdx.x;
        const int pos = s_lb[ch[k]] + rank[k];
        const int2 m = s_m[ch[k]];
        s_row[pos] = (int)r;
        s_val[pos] = make_uint2((unsigned)(ty[k] - m.x), (unsigned)(tyy[k] - m.y));
        if (WEIGHTED) {
          const int wq = wq_rows[r];
          s_w[pos] = wq;
          if (wq == 0) s_val[pos] = make_uint2(0u, 0u);  // zero weight: no part in any sum (see eng_begin_kernel)
        }
        s_child[pos] = (uint16_t)ch[k];
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += kScatThreads) {
      const int d = s_gb[s_child[i]] + i;
      gidx_n[d] = s_row[i];
      gval_n[d] = s_val[i];
      if (WEIGHTED) gw_n[d] = s_w[i];
    }
    __syncthreads();
  }
}

__global__ void eng_nid_kernel(const uint16_t* __restrict__ nid16, int64_t rows, int32_t* __restrict__ nid_rows) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const unsigned n = nid16[r];
  nid_rows[r] = n == kNoNid ? -1 : (int32_t)n;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host side
cudaError_t engine_alloc(Engine& e, int64_t rows, int cols, bool weighted, bool sparse, bool u8) {
  engine_free(e);
  if ((sparse || u8) && weighted) return cudaErrorInvalidValue;
  if (sparse && u8) return cudaErrorInvalidValue;
  e.sparse = sparse;
  e.u8 = u8;
  if (rows >= (1ll << 31) - 4096) return cudaErrorInvalidValue;
  if (cols > 32767) return cudaErrorInvalidValue;  // the routing table keeps the split column in 16 bits
  e.rows = rows;
  e.cols = cols;
  e.weighted = weighted;
  e.ldt = ((int64_t)cols + 7) / 8 * 8;
  e.ld8 = (rows + 127) / 128 * 128;
  int dev = 0, sms = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (err != cudaSuccess) return err;
  e.sms = sms;
  e.route_ctas = 4 * sms;  // four 512-thread CTAs per SM, each with one contiguous range of rows
  auto A = [&](void** p, size_t bytes) {
    if (err == cudaSuccess) err = cudaMalloc(p, bytes);
  };
  if (sparse) {
    A((void**)&e.csc_ptr, sizeof(uint2*) * (size_t)cols);
    A((void**)&e.csc_len, sizeof(int64_t) * (size_t)cols);
    A((void**)&e.node_tot, sizeof(unsigned long long) * 3 * kEngMaxNodes);
    A((void**)&e.rec, sizeof(uint2) * (size_t)rows);
    for (int k = 0; k < 2; k++) A((void**)&e.node_class[k], kEngMaxNodes);
    if (err == cudaSuccess) err = cudaMemset(e.csc_ptr, 0, sizeof(uint2*) * (size_t)cols);
    if (err == cudaSuccess) err = cudaMemset(e.csc_len, 0, sizeof(int64_t) * (size_t)cols);
    e.nnz = 0;
  } else if (u8) {
    e.ldb = ((int64_t)cols + 31) / 32 * 32;  // rows start on 32-byte sectors
    A((void**)&e.XB, (size_t)rows * (size_t)e.ldb);
    A((void**)&e.load_err, sizeof(int32_t));
    if (err == cudaSuccess) err = cudaMemset(e.XB, 0, (size_t)rows * (size_t)e.ldb);
    if (err == cudaSuccess) err = cudaMemset(e.load_err, 0, sizeof(int32_t));
  } else {
    A((void**)&e.XT, sizeof(float) * (size_t)rows * (size_t)e.ldt);
  }
  A((void**)&e.code8, (size_t)cols * (size_t)e.ld8);
  A((void**)&e.thr, sizeof(uint32_t) * 256 * (size_t)cols);
  A((void**)&e.cmin, sizeof(uint32_t) * (size_t)cols);
  A((void**)&e.cmax, sizeof(uint32_t) * (size_t)cols);
  for (int k = 0; k < 2; k++) {
    if (!sparse) {  // sparse storage keeps no row records
      A((void**)&e.gidx[k], sizeof(int32_t) * (size_t)rows);
      A((void**)&e.gval[k], sizeof(uint2) * (size_t)rows);
    }
    if (weighted) A((void**)&e.gw[k], sizeof(int32_t) * (size_t)rows);
    A((void**)&e.seg[k], sizeof(int32_t) * kSegInts);
    A((void**)&e.mnode[k], sizeof(int32_t) * 2 * kEngMaxNodes);
  }
  A((void**)&e.nid16, sizeof(uint16_t) * (size_t)rows);
  A((void**)&e.cta_cnt, sizeof(int32_t) * (size_t)kEngMaxNodes * (size_t)e.route_ctas);
  A((void**)&e.leaf, sizeof(int32_t) * (size_t)rows);
  A((void**)&e.err, sizeof(int32_t));
  A((void**)&e.na_flag, sizeof(int32_t));

  if (err == cudaSuccess) err = cudaMemset(e.err, 0, sizeof(int32_t));
  if (err == cudaSuccess) err = cudaMemset(e.na_flag, 0, sizeof(int32_t));
  e.na_known = false;
  e.has_na = true;
  if (err == cudaSuccess && !sparse && !u8 && e.ldt != cols) err = cudaMemset(e.XT, 0, sizeof(float) * (size_t)rows * (size_t)e.ldt);
  if (err != cudaSuccess) engine_free(e);
  return err;
}

void engine_free(Engine& e) {
  cudaFree(e.XT); cudaFree(e.code8); cudaFree(e.thr); cudaFree(e.cmin); cudaFree(e.cmax);
  cudaFree(e.XB); cudaFree(e.load_err);
  e.XB = nullptr;
  e.load_err = nullptr;
  e.u8 = false;
  for (void* p : e.csc_blocks) cudaFree(p);
  e.csc_blocks.clear();
  cudaFree(e.csc_ptr); cudaFree(e.csc_len); cudaFree(e.node_tot); cudaFree(e.rec);
  cudaFree(e.node_class[0]); cudaFree(e.node_class[1]);
  e.node_class[0] = e.node_class[1] = nullptr;
  e.rec = nullptr;
  e.K = e.K_alloc = 1;
  e.ld_targets = 0;
  e.csc_ptr = nullptr;
  e.csc_len = nullptr;
  e.node_tot = nullptr;
  e.sparse = false;
  e.nnz = 0;
  e.cmin = e.cmax = nullptr;
  for (int k = 0; k < 2; k++) {
    cudaFree(e.gidx[k]); cudaFree(e.gval[k]); cudaFree(e.gw[k]); cudaFree(e.seg[k]); cudaFree(e.mnode[k]);
    e.gidx[k] = nullptr; e.gval[k] = nullptr; e.gw[k] = nullptr; e.seg[k] = nullptr; e.mnode[k] = nullptr;
  }
  cudaFree(e.nid16); cudaFree(e.cta_cnt); cudaFree(e.leaf); cudaFree(e.err); cudaFree(e.na_flag);
  e.na_flag = nullptr;

  e.XT = nullptr;
  e.code8 = nullptr;
  e.thr = nullptr;
  e.nid16 = nullptr;
  e.cta_cnt = e.leaf = e.err = nullptr;
  e.wq_rows = e.wyq_rows = e.wyyq_rows = nullptr;
  e.rows = e.ldt = e.ld8 = 0;
  e.n_nodes = 0;
}

cudaError_t engine_set_weighted(Engine& e, bool weighted) {
  if (weighted) {
    for (int k = 0; k < 2; k++) {
      if (!e.gw[k]) {
        cudaError_t err = cudaMalloc((void**)&e.gw[k], sizeof(int32_t) * (size_t)e.rows);
        if (err != cudaSuccess) return err;
      }
    }
  }
  e.weighted = weighted;
  return cudaSuccess;
}

cudaError_t engine_load_bins(Engine& e, int col0, int ncols, const unsigned char* B, int64_t ld, cudaStream_t s) {
  if (!e.u8) return cudaErrorInvalidValue;
  dim3 grid(nblk(e.rows, 32), nblk(ncols, 32));
  eng_load_u8_kernel<<<grid, 256, 0, s>>>(B, ld, e.rows, col0, ncols, e.XB, e.ldb, e.code8, e.ld8, e.na_flag);
  e.na_known = false;
  cudaError_t err = launch_init_enc(e.cmin + col0, e.cmax + col0, ncols, s);
  if (err != cudaSuccess) return err;
  int64_t per = (e.rows + 255) / 256;
  const unsigned bx = (unsigned)(per < 1 ? 1 : (per > 4 * e.sms ? 4 * e.sms : per));
  eng_u8_ranges_kernel<<<dim3(bx, (unsigned)ncols), 256, 0, s>>>(B, ld, e.rows, e.cmin + col0, e.cmax + col0);
  return cudaGetLastError();
}

cudaError_t engine_load_cols(Engine& e, int col0, int ncols, const float* X, int64_t ld, cudaStream_t s) {
  if (e.u8) {  // fp32 cells that already hold codes (integers 0..254, NaN = NA): converted, anything else is flagged
    unsigned char* tmp = nullptr;
    cudaError_t err = cudaMalloc((void**)&tmp, (size_t)ld * ncols);
    if (err != cudaSuccess) return err;
    const int64_t n = ld * (int64_t)ncols;
    eng_f32_codes_kernel<<<nblk(n, 256), 256, 0, s>>>(X, n, tmp, e.load_err);
    e.load_checked = false;
    err = engine_load_bins(e, col0, ncols, tmp, ld, s);
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    cudaFree(tmp);
    return err;
  }
  if (e.sparse) {
    // count the nonzeros of every (column, chunk), lay the columns out one after the other in one allocation per
    // block of columns, then compact (rows stay ascending inside a column).  One blocking round trip per block: this
    // happens once per data set.
    const int n_chunks = (int)((e.rows + kCscChunk - 1) / kCscChunk);
    const size_t n_cnt = (size_t)ncols * n_chunks;
    unsigned* d_cnt = nullptr;
    int64_t* d_start = nullptr;
    cudaError_t err = cudaMalloc((void**)&d_cnt, sizeof(unsigned) * n_cnt);
    if (err == cudaSuccess) err = cudaMalloc((void**)&d_start, sizeof(int64_t) * n_cnt);
    std::vector<unsigned> h_cnt(n_cnt);
    std::vector<int64_t> h_start(n_cnt), h_len(ncols);
    std::vector<uint2*> h_ptr(ncols);
    uint2* block = nullptr;
    if (err == cudaSuccess) {
      eng_csc_count_kernel<<<dim3((unsigned)n_chunks, (unsigned)ncols), 1024, 0, s>>>(X, ld, e.rows, n_chunks, d_cnt);
      err = cudaMemcpyAsync(h_cnt.data(), d_cnt, sizeof(unsigned) * n_cnt, cudaMemcpyDeviceToHost, s);
    }
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    int64_t total = 0;
    if (err == cudaSuccess) {
      for (int c = 0; c < ncols; c++) {
        const int64_t first = total;
        for (int k = 0; k < n_chunks; k++) {
          h_start[(size_t)c * n_chunks + k] = total;
          total += h_cnt[(size_t)c * n_chunks + k];
        }
        h_len[c] = total - first;
      }
      err = cudaMalloc((void**)&block, sizeof(uint2) * (size_t)(total > 0 ? total : 1));
    }
    if (err == cudaSuccess) {
      e.csc_blocks.push_back(block);
      int64_t off = 0;
      for (int c = 0; c < ncols; c++) {
        h_ptr[c] = block + off;
        off += h_len[c];
      }
      e.nnz += total;
      err = cudaMemcpyAsync(d_start, h_start.data(), sizeof(int64_t) * n_cnt, cudaMemcpyHostToDevice, s);
    }
    if (err == cudaSuccess) err = cudaMemcpyAsync(e.csc_ptr + col0, h_ptr.data(), sizeof(uint2*) * ncols, cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) err = cudaMemcpyAsync(e.csc_len + col0, h_len.data(), sizeof(int64_t) * ncols, cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) {
      eng_csc_fill_kernel<<<dim3((unsigned)n_chunks, (unsigned)ncols), 1024, 0, s>>>(X, ld, e.rows, n_chunks, d_start, block);
      err = cudaGetLastError();
    }
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);  // the host vectors above are read by the copies
    cudaFree(d_cnt);
    cudaFree(d_start);
    if (err != cudaSuccess) return err;
  } else {
    dim3 grid(nblk(e.rows, 32), nblk(ncols, 32));
    eng_load_kernel<<<grid, 256, 0, s>>>(X, ld, e.rows, col0, ncols, e.XT, e.ldt);
  }
  // the routing codes of the same columns: thresholds from a sorted sample, then one byte per cell
  eng_thresholds_kernel<<<ncols, 1024, 0, s>>>(X, ld, e.rows, e.thr + (int64_t)col0 * 256);
  dim3 cgrid(nblk(e.rows, 1024), (unsigned)ncols);
  eng_code_kernel<<<cgrid, 256, 0, s>>>(X, ld, e.rows, e.thr + (int64_t)col0 * 256, e.code8 + (int64_t)col0 * e.ld8, e.ld8,
                                        e.na_flag);
  e.na_known = false;
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  return launch_col_ranges(X, ld, e.rows, ncols, e.cmin + col0, e.cmax + col0, s);  // DHistogram.initialHist reads them
}

// A forest: K class trees built in lockstep on the sparse storage (one pass over the column lists per level for all of
// them).  Every row then sits in K nodes, one per tree; the node list of a level is flat, class after class.
cudaError_t engine_set_forest(Engine& e, int K) {
  if (!e.sparse || K < 1 || K > kEngMaxClasses) return cudaErrorInvalidValue;
  if (K == e.K_alloc) {
    e.K = K;
    return cudaSuccess;
  }
  cudaFree(e.nid16); cudaFree(e.leaf); cudaFree(e.rec);
  e.nid16 = nullptr;
  e.leaf = nullptr;
  e.rec = nullptr;
  cudaError_t err = cudaMalloc((void**)&e.nid16, sizeof(uint16_t) * (size_t)e.rows * K);
  if (err == cudaSuccess) err = cudaMalloc((void**)&e.leaf, sizeof(int32_t) * (size_t)e.rows * K);
  if (err == cudaSuccess) err = cudaMalloc((void**)&e.rec, sizeof(uint2) * (size_t)e.rows * K);
  if (err != cudaSuccess) return err;
  e.K = e.K_alloc = K;
  return cudaSuccess;
}

__global__ void eng_forest_classes_kernel(int K, unsigned char* __restrict__ node_class) {
  if ((int)threadIdx.x < K) node_class[threadIdx.x] = (unsigned char)threadIdx.x;
}

// wyq / wyyq: [K][ld] the class trees' targets in original row order; q_dev: [K] their scales on the device
cudaError_t engine_begin_forest(Engine& e, int K, const int32_t* wyq, const int32_t* wyyq, int64_t ld, const h2ok_quant* q_dev,
                                cudaStream_t s) {
  cudaError_t err = engine_set_forest(e, K);
  if (err != cudaSuccess) return err;
  e.cur = 0;
  e.n_nodes = K;
  e.scale_wy = e.scale_wyy = 0.0;
  e.q_dev = q_dev;
  e.wq_rows = nullptr;
  e.wyq_rows = wyq;
  e.wyyq_rows = wyyq;
  e.ld_targets = ld;
  for (int k = 0; k < K; k++) {
    eng_root_mean_kernel<<<1, 1024, 0, s>>>(e.rows, wyq + k * ld, wyyq + k * ld, false, e.mnode[0] + 2 * k);
    eng_begin_kernel<<<nblk(e.rows, 256), 256, 0, s>>>(e.rows, nullptr, wyq + k * ld, wyyq + k * ld, e.mnode[0] + 2 * k, nullptr,
                                                       nullptr, nullptr, e.leaf + (int64_t)k * e.rows,
                                                       e.nid16 + (int64_t)k * e.rows, e.seg[0], k);
  }
  eng_forest_classes_kernel<<<1, 32, 0, s>>>(K, e.node_class[0]);
  return cudaGetLastError();
}

cudaError_t engine_begin_tree(Engine& e, const int32_t* wq_rows, const int32_t* wyq_rows, const int32_t* wyyq_rows,
                              const h2ok_quant* q, cudaStream_t s, const h2ok_quant* q_dev) {
  if (e.sparse && e.K != 1) {
    cudaError_t err = engine_set_forest(e, 1);
    if (err != cudaSuccess) return err;
  }
  e.ld_targets = 0;
  e.q_dev = q_dev;
  e.cur = 0;
  e.n_nodes = 1;
  e.scale_wy = (q && !e.weighted) ? q->scale_wy : 0.0;
  e.scale_wyy = (q && !e.weighted) ? q->scale_wyy : 0.0;
  e.wq_rows = e.weighted ? wq_rows : nullptr;  // read again by every level's regrouping
  e.wyq_rows = wyq_rows;
  e.wyyq_rows = wyyq_rows;
  eng_root_mean_kernel<<<1, 1024, 0, s>>>(e.rows, wyq_rows, wyyq_rows, e.weighted, e.mnode[0]);
  eng_begin_kernel<<<nblk(e.rows, 256), 256, 0, s>>>(e.rows, e.weighted ? wq_rows : nullptr, wyq_rows, wyyq_rows,
                                                     e.mnode[0], e.gidx[0], e.gval[0], e.weighted ? e.gw[0] : nullptr,
                                                     e.leaf, e.nid16, e.seg[0]);
  return cudaGetLastError();
}

// columns per block = lanes per row x columns per lane: the widest layout whose shared histogram fits
struct EngLayout {
  int cpl, lpr;
};
static EngLayout eng_layout(int cols, int nbs) {
  if (cols > 32 && (size_t)64 * nbs * 12 <= (size_t)kHistSmemCap) return {2, 32};
  if ((size_t)32 * nbs * 12 <= (size_t)kHistSmemCap) return {1, 32};
  if ((size_t)16 * nbs * 12 <= (size_t)kHistSmemCap) return {1, 16};
  return {0, 0};
}

int engine_max_nb1(int cols) {
  (void)cols;
  return kHistSmemCap / (16 * 12);
}

template <int CPL, bool W, int LPR, bool NA, bool U8 = false>
static cudaError_t launch_eng_hist(const EngHistArgs& a, dim3 grid, size_t smem, cudaStream_t s) {
  static std::atomic<bool> attr_done[64];  // per device, per instantiation (setting it twice is harmless)
  int dev = 0;
  cudaGetDevice(&dev);
  std::atomic<bool>& attr = attr_done[dev & 63];
  if (!attr.load(std::memory_order_acquire)) {
    cudaError_t e = cudaFuncSetAttribute(eng_hist_kernel<CPL, W, LPR, NA, U8>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmemCap);
    if (e != cudaSuccess) return e;
    attr.store(true, std::memory_order_release);
  }
  eng_hist_kernel<CPL, W, LPR, NA, U8><<<grid, EngCfg<CPL, W>::threads, smem, s>>>(a);
  return cudaGetLastError();
}

cudaError_t engine_build_level(Engine& e, int n_nodes, const h2ok_col_desc* descs, int nb1, int level_nb1, int64_t* acc,
                               uint32_t* min2_enc, uint32_t* maxin_enc, cudaStream_t s) {
  if (n_nodes < 1 || n_nodes >= kEngMaxNodes) return cudaErrorInvalidValue;
  const int nbs = level_nb1 > 0 && level_nb1 < nb1 ? level_nb1 : nb1;
  if (e.sparse) {
    if (!e.wyq_rows) return cudaErrorInvalidValue;
    e.n_nodes = n_nodes;
    // nodes per pass: descriptor table + range slots + at least one histogram copy per node must fit one CTA
    const size_t per_node = kCscNodeBytes + (size_t)nbs * 12;
    // What a CTA takes when its histogram allows a choice: 56 KB.  More copies (112 KB) cut bank conflicts but leave the
    // SM 30 KB of L1 for the row-record gathers: 14 ms per level against 8.5 ms on config 4 (32 nodes, 64 bins).
    const size_t two_per_sm = 56 * 1024;
    int group = (int)((size_t)kHistSmemCap / per_node);
    if (group < 1) return cudaErrorInvalidValue;
    if (group > n_nodes) group = n_nodes;
    cudaError_t err = cudaMemsetAsync(e.node_tot, 0, sizeof(unsigned long long) * 3 * (size_t)n_nodes, s);
    if (err != cudaSuccess) return err;
    for (int k = 0; k < e.K; k++)
      eng_node_totals_kernel<<<4 * e.sms, 512, sizeof(int) * 5 * (size_t)n_nodes, s>>>(
          e.rows, e.nid16 + (int64_t)k * e.rows, e.wyq_rows + (int64_t)k * e.ld_targets, e.wyyq_rows + (int64_t)k * e.ld_targets,
          n_nodes, e.node_tot, e.rec, e.K, k);
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
      cudaError_t ea = cudaFuncSetAttribute(eng_hist_csc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmemCap);
      if (ea == cudaSuccess) ea = cudaFuncSetAttribute(eng_hist_csc_forest_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmemCap);
      if (ea != cudaSuccess) return ea;
      attr_done[dev & 63].store(true, std::memory_order_release);
    }
    CscHistArgs a{};
    a.cptr = e.csc_ptr;
    a.clen = e.csc_len;
    a.cols = e.cols;
    a.rec = e.rec;
    a.K = e.K;
    a.mnode = e.mnode[e.cur];
    a.descs = descs;
    a.nb1 = nb1;
    a.nbs = nbs;
    a.acc = acc;
    a.min2_enc = min2_enc;
    a.maxin_enc = maxin_enc;
    a.err = e.err;
    // stretches of a column's nonzero list per CTA: about eight CTAs per SM over the whole grid
    // (20 / 40 / 80 per SM: 7.2 / 7.1 / 7.4 ms per level on config 4 against 7.2)
    int chunks = (8 * e.sms + e.cols - 1) / e.cols;
    chunks = chunks < 1 ? 1 : chunks;
    a.chunks = chunks;
    for (int base = 0; base < n_nodes; base += group) {
      a.node_base = base;
      a.n_nodes = n_nodes - base < group ? n_nodes - base : group;
      // copies of the histogram (chosen by lane) spread the warp's atomics over more banks; as many as fit while two
      // CTAs still share the SM, or one CTA's worth when even a single copy needs more than that
      const size_t one = (size_t)a.n_nodes * (kCscNodeBytes + (size_t)nbs * 12);
      // (a level whose single copy is past that but under 100 KB keeps one copy, so that two CTAs still share the SM)
      const size_t cap = one <= two_per_sm ? two_per_sm : (one <= 100 * 1024 ? one : (size_t)kHistSmemCap);
      a.copies = 1;
      while (a.copies < 8 && (size_t)a.n_nodes * (kCscNodeBytes + (size_t)nbs * 12 * (a.copies * 2)) <= cap) a.copies *= 2;
      const size_t smem = (size_t)a.n_nodes * (kCscNodeBytes + (size_t)nbs * 12 * a.copies);
      const dim3 grid((unsigned)chunks, (unsigned)e.cols);
      if (e.K == 1) eng_hist_csc_kernel<2><<<grid, kCscThreads, smem, s>>>(a);
      else eng_hist_csc_forest_kernel<4><<<grid, kCscThreads, smem, s>>>(a);
    }
    const int64_t pairs = (int64_t)n_nodes * e.cols;
    eng_csc_zero_kernel<<<(unsigned)((pairs + 7) / 8), 256, 0, s>>>(n_nodes, e.cols, descs, nb1, e.node_tot, acc, min2_enc, maxin_enc);
    return cudaGetLastError();
  }
  EngLayout lay = eng_layout(e.cols, nbs);
  if (e.u8 && lay.lpr == 16) lay = {0, 0};  // u8 storage runs the 32-lane layouts only (bins <= 599)
  if (!lay.cpl) return cudaErrorInvalidValue;
  e.n_nodes = n_nodes;
  EngHistArgs a{};
  a.XT = e.XT;
  a.XB = e.XB;
  a.ldt = e.u8 ? e.ldb : e.ldt;
  a.ldt4 = e.u8 ? (unsigned)e.ldb : (unsigned)(e.ldt * 4);
  a.cols = e.cols;
  a.gidx = e.gidx[e.cur];
  a.gval = e.gval[e.cur];
  a.gw = e.weighted ? e.gw[e.cur] : nullptr;
  a.mnode = e.mnode[e.cur];
  a.seg = e.seg[e.cur];
  a.n_nodes = n_nodes;
  a.descs = descs;
  a.nb1 = nb1;
  a.nbs = nbs;
  a.acc = acc;
  a.min2_enc = min2_enc;
  a.maxin_enc = maxin_enc;
  a.err = e.err;
  const int colblk = lay.lpr * lay.cpl;
  // one CTA per SM and column block; every CTA owns the same share of the level's records
  const dim3 grid((unsigned)e.sms, (unsigned)((e.cols + colblk - 1) / colblk));
  const size_t smem = (size_t)colblk * nbs * 12;
  if (e.weighted) {
    if (lay.cpl == 2) return launch_eng_hist<2, true, 32, true>(a, grid, smem, s);
    if (lay.lpr == 32) return launch_eng_hist<1, true, 32, true>(a, grid, smem, s);
    return launch_eng_hist<1, true, 16, true>(a, grid, smem, s);
  }
  if (!e.na_known) {  // once per data set: did any cell of the shard turn out to be NaN while the codes were made?
    int flag = 1;
    cudaError_t err = cudaMemcpyAsync(&flag, e.na_flag, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    if (err != cudaSuccess) return err;
    e.has_na = flag != 0;
    e.na_known = true;
  }
  if (e.u8) {
    if (e.has_na) {
      if (lay.cpl == 2) return launch_eng_hist<2, false, 32, true, true>(a, grid, smem, s);
      return launch_eng_hist<1, false, 32, true, true>(a, grid, smem, s);
    }
    if (lay.cpl == 2) return launch_eng_hist<2, false, 32, false, true>(a, grid, smem, s);
    return launch_eng_hist<1, false, 32, false, true>(a, grid, smem, s);
  }
  if (e.has_na) {
    if (lay.cpl == 2) return launch_eng_hist<2, false, 32, true>(a, grid, smem, s);
    if (lay.lpr == 32) return launch_eng_hist<1, false, 32, true>(a, grid, smem, s);
    return launch_eng_hist<1, false, 16, true>(a, grid, smem, s);
  }
  if (lay.cpl == 2) return launch_eng_hist<2, false, 32, false>(a, grid, smem, s);
  if (lay.lpr == 32) return launch_eng_hist<1, false, 32, false>(a, grid, smem, s);
  return launch_eng_hist<1, false, 16, false>(a, grid, smem, s);
}

cudaError_t engine_apply_splits(Engine& e, const h2ok_split* splits, const int32_t* child_ids, int gid_base, int n_nodes,
                                bool regroup, cudaStream_t s) {
  if (n_nodes < 1 || n_nodes >= kEngMaxNodes) return cudaErrorInvalidValue;
  if (!e.wyq_rows) return cudaErrorInvalidValue;
  const int nx = 1 - e.cur;
  const int G = e.route_ctas;
  const int n_slots = 2 * n_nodes < kEngMaxNodes ? 2 * n_nodes : kEngMaxNodes;  // child ids are < this
  const size_t rsmem = (size_t)n_nodes * sizeof(RouteNode) + (size_t)n_slots * sizeof(int);
  for (int k = 0; k < (e.sparse ? e.K : 1); k++)  // a forest routes every class tree's row -> node map with the same flat tables
    eng_route_kernel<<<G, kRouteThreads, rsmem, s>>>(e.XT, e.ldt, e.csc_ptr, e.csc_len, e.code8, e.ld8, e.thr, e.rows, n_nodes, splits,
                                                     child_ids, gid_base, e.nid16 + (int64_t)k * e.rows,
                                                     e.leaf + (int64_t)k * e.rows, e.cta_cnt, n_slots, e.u8 ? 1 : 0);
  // every level but a tree's last: lay the children's records out for the next histogram pass
  const bool forest = e.sparse && e.ld_targets != 0;
  eng_scan_kernel<<<1, 1024, 0, s>>>(e.cta_cnt, G, n_nodes, child_ids, e.seg[nx], splits, e.scale_wy, e.scale_wyy,
                                     e.weighted ? nullptr : e.q_dev, e.mnode[e.cur], e.mnode[nx],
                                     forest ? e.node_class[e.cur] : nullptr, forest ? e.node_class[nx] : nullptr);
  if (regroup && !e.sparse) {
    const size_t ncp = (size_t)(n_slots + 1) & ~(size_t)1;
    const size_t ssmem = ncp * (4 * sizeof(int) + sizeof(int2)) +
                         (size_t)kScatTile * (sizeof(uint2) + sizeof(int) + sizeof(uint16_t) + (e.weighted ? sizeof(int) : 0));
    const int SG = G / 2;  // two 1024-thread CTAs per SM; each covers the rows of two routing CTAs
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
      cudaError_t ea = cudaFuncSetAttribute(eng_scatter_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
      if (ea == cudaSuccess)
        ea = cudaFuncSetAttribute(eng_scatter_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
      if (ea != cudaSuccess) return ea;
      attr_done[dev & 63].store(true, std::memory_order_release);
    }
    if (e.weighted)
      eng_scatter_kernel<true><<<SG, kScatThreads, ssmem, s>>>(e.rows, e.nid16, e.wq_rows, e.wyq_rows, e.wyyq_rows, e.seg[nx],
                                                               e.mnode[nx], e.cta_cnt, G, e.gidx[nx], e.gval[nx], e.gw[nx]);
    else
      eng_scatter_kernel<false><<<SG, kScatThreads, ssmem, s>>>(e.rows, e.nid16, e.wq_rows, e.wyq_rows, e.wyyq_rows, e.seg[nx],
                                                                e.mnode[nx], e.cta_cnt, G, e.gidx[nx], e.gval[nx], nullptr);
  }
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  e.cur = nx;
  e.n_nodes = 0;  // the caller says how many children there are (build_level / nid)
  return cudaSuccess;
}

cudaError_t engine_check_load(Engine& e, cudaStream_t s, bool* bad) {
  *bad = false;
  if (!e.u8) return cudaSuccess;
  if (!e.load_checked) {
    int flag = 0;
    cudaError_t err = cudaMemcpyAsync(&flag, e.load_err, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    if (err != cudaSuccess) return err;
    e.load_failed = flag != 0;
    e.load_checked = true;
  }
  *bad = e.load_failed;
  return cudaSuccess;
}

cudaError_t launch_prebin(const float* X, int64_t ld, int64_t rows, int ncols, const h2ok_col_desc* d_descs, unsigned char* B,
                          int64_t ldb, cudaStream_t s) {
  eng_prebin_kernel<<<dim3(nblk(rows, 256), (unsigned)ncols), 256, 0, s>>>(X, ld, rows, d_descs, B, ldb);
  return cudaGetLastError();
}

cudaError_t engine_nid(Engine& e, int32_t* nid_rows, cudaStream_t s) {
  eng_nid_kernel<<<nblk(e.rows, 256), 256, 0, s>>>(e.nid16, e.rows, nid_rows);
  return cudaGetLastError();
}

}  // namespace h2ok
