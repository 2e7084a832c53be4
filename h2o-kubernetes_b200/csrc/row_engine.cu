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

#include "common.cuh"
#include "engine.h"
#include "kernels.h"

namespace h2ok {

namespace {

inline unsigned nblk(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

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
                                 int32_t* __restrict__ leaf, uint16_t* __restrict__ nid16, int32_t* __restrict__ seg) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    seg[0] = 0;
    seg[1] = (int32_t)rows;
    seg[kSegNodes] = 1;
  }
  if (p >= rows) return;
  // zero-weight rows still travel down the tree (a6); the histogram pass skips them by weight
  gidx[p] = (int32_t)p;
  // a row whose quantised weight is 0 takes no part, whatever its (independently scaled) targets round to — the
  // histogram pass sends it to the NA bin with weight 0, so its targets must be 0 too (weighted shards keep M == 0)
  const bool dead = gw && wq_rows[p] == 0;
  gval[p] = dead ? make_uint2(0u, 0u) : make_uint2((unsigned)(wyq_rows[p] - mnode[0]), (unsigned)(wyyq_rows[p] - mnode[1]));
  if (gw) gw[p] = wq_rows[p];
  leaf[p] = -1;
  nid16[p] = 0;
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

template <int CPL> struct XVec;
template <> struct XVec<1> { using type = float; };
template <> struct XVec<2> { using type = float2; };

__device__ __forceinline__ float xget(const float& v, int) { return v; }
__device__ __forceinline__ float xget(const float2& v, int k) { return k ? v.y : v.x; }

// What a lane keeps in registers while it walks the rows of one node: the descriptors and running extremes of its CPL
// columns.  A (node, column) that is not tracked gets a harmless 1-bin descriptor — its rows land in words that the
// flush skips — so the hot loop has no per-column test.
template <int CPL>
struct EngLane {
  double mn[CPL], st[CPL];
  int nb[CPL];
  float fmn[CPL], fmx[CPL];
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
template <int CPL, bool WEIGHTED, int U, int LPR, bool NA>
__device__ __forceinline__ void eng_bins(EngLane<CPL>& c, const unsigned (&hb)[CPL], const typename XVec<CPL>::type (&xv)[U],
                                         const int (&w)[U], unsigned (&pa)[U * CPL]) {
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

template <int CPL, bool WEIGHTED, int LPR, bool NA>
__global__ void __launch_bounds__(EngCfg<CPL, WEIGHTED>::threads, 1) eng_hist_kernel(const __grid_constant__ EngHistArgs a) {
  static_assert(NA || !WEIGHTED, "zero-weight rows are turned into NaNs: weighted shards need the NA test");
  using V = typename XVec<CPL>::type;
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
  const char* xrow = reinterpret_cast<const char*>(a.XT + min(col_lane, (int)a.ldt - CPL));
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
        else if (d.nbin > 0) {
          tracked[k] = true;
          c.nb[k] = d.nbin;
          c.mn[k] = d.min;
          c.st[k] = d.step;
        }
      }
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
          eng_bins<CPL, WEIGHTED, 1, LPR, NA>(c, hb, x, w, pa);
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
        if (on) eng_bins<CPL, WEIGHTED, U, LPR, NA>(c, hb, x, w, pa);
        if (more) {
#pragma unroll
          for (int j = 0; j < U; j++) prefetch_l2(row_addr(xrow, (unsigned)nx[j], ldt4));
        }
        if (on) eng_accum<CPL, WEIGHTED, U, LPR>(a, hb, node, col_lane, pa, v1, v2, w);
      }
      for (; p < send; p++) one_row(p);  // the run's last, partial batch
    }

    // observed ranges: registers -> shared (integer min / max of the order-preserving encoding) -> global
#pragma unroll
    for (int k = 0; k < CPL; k++) {
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

// ------------------------------------------------------------------------------------------------ sparse storage
// For matrices that are mostly exact zeros (BASELINE config 4: 90 % zeros; H2O itself keeps such columns in sparse
// chunks, CXIChunk / CX0Chunk) the shard can be stored as one 64-byte slot per (row, block of 64 columns) holding the
// block's NONZERO cells only: byte 0 = their number, bytes 1..12 = their columns inside the block, words 4..15 = their
// values (a NaN is a nonzero).  A (row, block) with more than 12 nonzeros (1.3 % at 10 % density) says 0xFF and points
// (word 4) at a dense 64-float row in an overflow store.  256 B of dense floats become 64 B.
//
// The histogram pass then touches only the nonzeros: a warp flattens the nonzero cells of four rows into a work list
// and every lane takes one (column, value) pair, so a warp instruction updates ~26 cells instead of the ~3 nonzero
// lanes of the one-column-per-lane layout.  A lane no longer owns its columns, so descriptors come from a shared
// table, the histogram is [copy][column][bin][channel] with the copy chosen by lane, and observed ranges go through a
// shared table (compare first, atomic only on improvement).  The zeros are accounted for exactly at the flush: the
// node's totals over the CTA's rows minus what its nonzeros added is what the zero cells contribute, and it goes to
// the bin 0.0 falls into; a column with a zero row has 0.0 inside its observed range.  Integer sums make this exact.
constexpr int kSpCols = 64, kSpInline = 12, kSpRows = 4;
struct SpSlot {
  unsigned w[16];
};

__global__ void __launch_bounds__(256) eng_load_sparse_kernel(const float* __restrict__ X, int64_t ld, int64_t rows, int ncols,
                                                              SpSlot* __restrict__ slots, int nblk, int blk,
                                                              float* __restrict__ ovf, unsigned* __restrict__ ovf_count,
                                                              unsigned ovf_cap, int32_t* __restrict__ err) {
  __shared__ unsigned s[256][17];
  const int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (r >= rows) return;
  unsigned* w = s[threadIdx.x];
#pragma unroll
  for (int i = 0; i < 16; i++) w[i] = 0u;
  int cnt = 0;
  for (int c = 0; c < ncols; c++) {
    const float x = X[(int64_t)c * ld + r];
    if (!(x == 0.0f)) {  // NaN counts as a nonzero; -0.0 as a zero
      if (cnt < kSpInline) {
        w[(1 + cnt) >> 2] |= (unsigned)c << (8 * ((1 + cnt) & 3));
        w[4 + cnt] = __float_as_uint(x);
      }
      cnt++;
    }
  }
  if (cnt <= kSpInline) {
    w[0] |= (unsigned)cnt;
  } else {
    const unsigned idx = atomicAdd(ovf_count, 1u);
    if (idx < ovf_cap) {
      for (int c = 0; c < kSpCols; c++) ovf[(int64_t)idx * kSpCols + c] = c < ncols ? X[(int64_t)c * ld + r] : 0.0f;
      w[0] = 0xFFu;
      w[4] = idx;
    } else {
      *err = 1;  // denser than the overflow store was sized for: the caller must use dense storage
      w[0] = 0u;
    }
  }
  unsigned* out = slots[r * nblk + blk].w;
#pragma unroll
  for (int i = 0; i < 16; i += 4) *reinterpret_cast<uint4*>(out + i) = make_uint4(w[i], w[i + 1], w[i + 2], w[i + 3]);
}

// one cell of the sparse shard (routing's exact comparison on a code tie)
__device__ __forceinline__ float sp_get(const SpSlot* __restrict__ slots, int nblk, const float* __restrict__ ovf, int64_t r,
                                        int col) {
  const unsigned* w = slots[r * nblk + (col >> 6)].w;
  const int c = col & 63;
  const unsigned w0 = w[0], cnt = w0 & 0xFFu;
  if (cnt == 0xFFu) return ovf[(int64_t)w[4] * kSpCols + c];
  for (unsigned p = 0; p < cnt; p++) {
    const unsigned byte = (w[(1 + p) >> 2] >> (8 * ((1 + p) & 3))) & 0xFFu;
    if ((int)byte == c) return __uint_as_float(w[4 + p]);
  }
  return 0.0f;
}

struct SpHistArgs {
  const SpSlot* slots;
  int nblk;
  const float* ovf;
  int cols;
  const int32_t* gidx;
  const uint2* gval;
  const int32_t* mnode;
  const int32_t* seg;
  int n_nodes;
  const h2ok_col_desc* descs;
  int nb1, nbs, copies;
  int64_t* acc;
  uint32_t* min2_enc;
  uint32_t* maxin_enc;
  int32_t* err;
};

__global__ void __launch_bounds__(1024, 1) eng_hist_sparse_kernel(const __grid_constant__ SpHistArgs a) {
  constexpr int THREADS = 1024, NW = THREADS / 32;
  extern __shared__ __align__(16) unsigned h[];  // [copy][column][bin][channel]
  __shared__ double s_mn[kSpCols], s_st[kSpCols];
  __shared__ int s_nb[kSpCols];
  __shared__ unsigned s_rmin[kSpCols], s_rmax[kSpCols];
  __shared__ unsigned char s_tracked[kSpCols];
  __shared__ unsigned long long s_tot[3];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nb1 = a.nb1, nbs = a.nbs, copies = a.copies;
  const int col0 = (int)blockIdx.y * kSpCols;
  const int n_nodes = min(a.n_nodes, a.seg[kSegNodes]);
  if (n_nodes < 1) return;
  const int64_t live = a.seg[n_nodes];
  int p0 = (int)(live * (int64_t)blockIdx.x / gridDim.x);
  const int p1 = (int)(live * (int64_t)(blockIdx.x + 1) / gridDim.x);
  if (p0 >= p1) return;
  int node = seg_of(a.seg, n_nodes, p0);
  const int words = copies * kSpCols * nbs * 3;
  for (int i = tid; i < words; i += THREADS) h[i] = (i % 3 == 0) ? 0u : kWyBias;
  const unsigned hbase = (unsigned)__cvta_generic_to_shared(h) + (unsigned)((lane & (copies - 1)) * kSpCols * nbs) * 12u;
  const int32_t* __restrict__ gidx = a.gidx;
  const uint2* __restrict__ gval = a.gval;
  const char* sbase = reinterpret_cast<const char*>(a.slots) + (size_t)blockIdx.y * sizeof(SpSlot);
  const unsigned pitch = (unsigned)a.nblk * (unsigned)sizeof(SpSlot);

  while (p0 < p1) {
    const int nend = a.seg[node + 1];
    const int send = nend < p1 ? nend : p1;
    if (send <= p0) {
      node++;
      continue;
    }
    if (tid < kSpCols) {
      const int col = col0 + tid;
      bool tracked = false;
      double mn = 0.0, st = 0.0;
      int nb = 1;
      if (col < a.cols) {
        const h2ok_col_desc d = a.descs[(int64_t)node * a.cols + col];
        if (d.nbin >= nbs) atomicAdd(a.err, 1);
        else if (d.nbin > 0) {
          tracked = true;
          mn = d.min;
          st = d.step;
          nb = d.nbin;
        }
      }
      s_mn[tid] = mn;
      s_st[tid] = st;
      s_nb[tid] = nb;
      s_rmin[tid] = kEncPosInf;
      s_rmax[tid] = kEncNegInf;
      s_tracked[tid] = tracked ? 1 : 0;
    }
    if (tid < 3) s_tot[tid] = 0ull;
    __syncthreads();

    // one (column, value) pair of row slot j: bin from the shared descriptor table, three shared atomics, range
    auto item = [&](bool active, int col, float x, int v1, int v2) {
      if (!active) return;
      const int b = bin_fast(x, s_mn[col], s_st[col], s_nb[col]);
      const unsigned addr = hbase + (unsigned)((col * nbs + b) * 12);
      sh_inc(addr);
      const unsigned o1 = sh_add(addr + 4u, (unsigned)v1), o2 = sh_add(addr + 8u, (unsigned)v2);
      const unsigned f1 = wrap_bit(o1, v1), f2 = wrap_bit(o2, v2);
      if ((f1 | f2) >> 31)
        eng_carry(a.acc, a.descs, node, a.cols, col0 + col, nb1, b, 0, (f1 >> 31) ? (v1 < 0 ? -1 : 1) : 0,
                  (f2 >> 31) ? (v2 < 0 ? -1 : 1) : 0);
      if (x == x) {
        const unsigned e = enc_f32_bits(__float_as_uint(x));
        if (e < s_rmin[col]) atomicMin(&s_rmin[col], e);
        if (e > s_rmax[col]) atomicMax(&s_rmax[col], e);
      }
    };

    unsigned tc = 0;            // this warp's rows of the run and their (centred) target sums: what the zeros are
    long long t1 = 0, t2 = 0;   // worked out from at the flush
    for (int p = p0 + warp * kSpRows; p < send; p += NW * kSpRows) {
      const int nr = send - p < kSpRows ? send - p : kSpRows;
      int v1[kSpRows], v2[kSpRows];
      unsigned ridx[kSpRows];
#pragma unroll
      for (int j = 0; j < kSpRows; j++) {
        const int q = j < nr ? p + j : p;
        ridx[j] = (unsigned)__ldg(gidx + q);
        const uint2 r = __ldg(gval + q);
        v1[j] = (int)r.x;
        v2[j] = (int)r.y;
        if (j < nr) {
          tc += 1u;
          t1 += (int)r.x;
          t2 += (int)r.y;
        }
      }
      // rows 0/1 in the two halves of regA, rows 2/3 in regB: one 64-byte slot per half warp
      const int half = lane >> 4, word = lane & 15;
      const unsigned regA = __ldg(reinterpret_cast<const unsigned*>(row_addr(sbase, half ? ridx[1] : ridx[0], pitch)) + word);
      const unsigned regB = __ldg(reinterpret_cast<const unsigned*>(row_addr(sbase, half ? ridx[3] : ridx[2], pitch)) + word);
      unsigned cnt[kSpRows];
      cnt[0] = __shfl_sync(0xffffffffu, regA, 0) & 0xFFu;
      cnt[1] = __shfl_sync(0xffffffffu, regA, 16) & 0xFFu;
      cnt[2] = __shfl_sync(0xffffffffu, regB, 0) & 0xFFu;
      cnt[3] = __shfl_sync(0xffffffffu, regB, 16) & 0xFFu;
      unsigned dense_mask = 0;
#pragma unroll
      for (int j = 0; j < kSpRows; j++) {
        if (j >= nr) cnt[j] = 0;
        if (cnt[j] == 0xFFu) {
          dense_mask |= 1u << j;
          cnt[j] = 0;
        }
      }
      const int P1 = (int)cnt[0], P2 = P1 + (int)cnt[1], P3 = P2 + (int)cnt[2], T = P3 + (int)cnt[3];
      for (int base = 0; base < T; base += 32) {
        const int i = base + lane;
        const int j = (i >= P1) + (i >= P2) + (i >= P3);
        const int pp = i - (j == 0 ? 0 : (j == 1 ? P1 : (j == 2 ? P2 : P3)));
        const int src = (j & 1) << 4;
        const int hw = (1 + pp) >> 2;
        const unsigned cwA = __shfl_sync(0xffffffffu, regA, (src + hw) & 31), cwB = __shfl_sync(0xffffffffu, regB, (src + hw) & 31);
        const unsigned vA = __shfl_sync(0xffffffffu, regA, (src + 4 + pp) & 31), vB = __shfl_sync(0xffffffffu, regB, (src + 4 + pp) & 31);
        const unsigned cw = (j >> 1) ? cwB : cwA;
        const int col = (int)((cw >> (8 * ((1 + pp) & 3))) & 0xFFu) & 63;
        const float x = __uint_as_float((j >> 1) ? vB : vA);
        const int a1 = j == 0 ? v1[0] : (j == 1 ? v1[1] : (j == 2 ? v1[2] : v1[3]));
        const int a2 = j == 0 ? v2[0] : (j == 1 ? v2[1] : (j == 2 ? v2[2] : v2[3]));
        item(i < T, col, x, a1, a2);
      }
      if (dense_mask) {  // rows with more nonzeros than a slot holds: their dense 64-float row, two columns per lane
#pragma unroll
        for (int j = 0; j < kSpRows; j++) {
          if (!((dense_mask >> j) & 1u)) continue;
          const unsigned idx = __shfl_sync(0xffffffffu, (j >> 1) ? regB : regA, ((j & 1) << 4) + 4);
          const float* row = a.ovf + (int64_t)idx * kSpCols;
          const float x0 = __ldg(row + lane), x1 = __ldg(row + 32 + lane);
          item(!(x0 == 0.0f), lane, x0, v1[j], v2[j]);
          item(!(x1 == 0.0f), 32 + lane, x1, v1[j], v2[j]);
        }
      }
    }
    if (lane == 0 && tc) {
      atomicAdd(&s_tot[0], (unsigned long long)tc);
      atomicAdd(&s_tot[1], (unsigned long long)t1);
      atomicAdd(&s_tot[2], (unsigned long long)t2);
    }
    __syncthreads();

    // flush.  First per column (one warp each): what the nonzeros added, hence what the zero cells contribute.
    const int m1 = a.mnode[2 * node], m2 = a.mnode[2 * node + 1];
    for (int c = warp; c < kSpCols; c += NW) {
      if (!s_tracked[c]) continue;  // warp-uniform
      unsigned long long nc = 0;
      long long n1 = 0, n2 = 0;
      for (int i = lane; i < nbs * copies; i += 32) {
        const int cp = i / nbs, b = i - cp * nbs;
        const unsigned* w = h + ((cp * kSpCols + c) * nbs + b) * 3;
        nc += w[0];
        n1 += (long long)w[1] - (long long)kWyBias;
        n2 += (long long)w[2] - (long long)kWyBias;
      }
      for (int o = 16; o; o >>= 1) {
        nc += __shfl_xor_sync(0xffffffffu, nc, o);
        n1 += __shfl_xor_sync(0xffffffffu, n1, o);
        n2 += __shfl_xor_sync(0xffffffffu, n2, o);
      }
      if (lane == 0) {
        const int col = col0 + c;
        const long long zc = (long long)s_tot[0] - (long long)nc;
        const long long z1 = (long long)s_tot[1] - n1 + zc * m1, z2 = (long long)s_tot[2] - n2 + zc * m2;
        if (zc > 0) {
          const int bz = bin_of(0.0f, s_mn[c], s_st[c], s_nb[c]);
          unsigned long long* g = reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)node * a.cols + col) * nb1 + bz) * 3;
          atomicAdd(g, (unsigned long long)zc);
          if (z1) atomicAdd(g + 1, (unsigned long long)z1);
          if (z2) atomicAdd(g + 2, (unsigned long long)z2);
          const unsigned ez = enc_f32_bits(__float_as_uint(0.0f));
          if (ez < s_rmin[c]) s_rmin[c] = ez;  // only this lane touches the column's slots now
          if (ez > s_rmax[c]) s_rmax[c] = ez;
        }
        if (s_rmin[c] != kEncPosInf) atomicMin(&a.min2_enc[(int64_t)node * a.cols + col], s_rmin[c]);
        if (s_rmax[c] != kEncNegInf) atomicMax(&a.maxin_enc[(int64_t)node * a.cols + col], s_rmax[c]);
      }
    }
    __syncthreads();
    // then every (column, bin): copies added up, centring undone with the bin's own count, words re-armed
    for (int t = tid; t < kSpCols * nbs; t += THREADS) {
      const int c = t / nbs, b = t - c * nbs;
      unsigned long long nc = 0;
      long long n1 = 0, n2 = 0;
      for (int cp = 0; cp < copies; cp++) {
        unsigned* w = h + ((cp * kSpCols + c) * nbs + b) * 3;
        const unsigned c0 = w[0], u1 = w[1], u2 = w[2];
        if (c0 | (u1 ^ kWyBias) | (u2 ^ kWyBias)) {
          nc += c0;
          n1 += (long long)u1 - (long long)kWyBias;
          n2 += (long long)u2 - (long long)kWyBias;
          w[0] = 0u;
          w[1] = kWyBias;
          w[2] = kWyBias;
        }
      }
      if (s_tracked[c] && (nc | (unsigned long long)n1 | (unsigned long long)n2)) {
        unsigned long long* g = reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)node * a.cols + col0 + c) * nb1 + b) * 3;
        const long long s1 = n1 + (long long)nc * m1, s2 = n2 + (long long)nc * m2;
        if (nc) atomicAdd(g, nc);
        if (s1) atomicAdd(g + 1, (unsigned long long)s1);
        if (s2) atomicAdd(g + 2, (unsigned long long)s2);
      }
    }
    __syncthreads();
    p0 = send;
    node++;
  }
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

constexpr int kRouteThreads = 512;
constexpr unsigned kNoNid = 0xFFFFu;

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
    const float* __restrict__ XT, int64_t ldt, const SpSlot* __restrict__ slots, int nblk, const float* __restrict__ ovf,
    const uint8_t* __restrict__ code8, int64_t ld8,
    const unsigned* __restrict__ thr, int64_t rows, int n_nodes, const h2ok_split* __restrict__ splits,
    const int32_t* __restrict__ child_ids, int gid_base, uint16_t* __restrict__ nid16, int32_t* __restrict__ leaf,
    int32_t* __restrict__ cta_cnt, int n_slots) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RouteNode* tab = reinterpret_cast<RouteNode*>(smem_raw);
  int* s_cnt = reinterpret_cast<int*>(tab + n_nodes);
  for (int g = threadIdx.x; g < n_nodes; g += kRouteThreads) {
    const h2ok_split sp = splits[g];
    RouteNode t;
    t.splat = sp.splat;
    t.col = (short)sp.col;
    t.na = (unsigned char)sp.na_dir;
    t.cs = 0;
    if (sp.col >= 0 && sp.na_dir != H2OK_NA_VS_REST && sp.splat == sp.splat)
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
  for (int64_t base = r0; base < r1; base += kRouteThreads) {
    const int64_t r = base + threadIdx.x;
    unsigned next = kNoNid;
    if (r < r1) {
      const unsigned n = nid16[r];
      if (n != kNoNid && (int)n < n_nodes) {
        const RouteNode t = tab[n];
        if (t.col < 0) {
          leaf[r] = t.leaf0;
        } else {
          const unsigned c = code8[(int64_t)t.col * ld8 + r];
          int way;
          if (c == 255u) way = (t.na == H2OK_NA_LEFT) ? 0 : 1;
          else if (t.na == H2OK_NA_VS_REST) way = 0;
          else if (c != t.cs) way = c > t.cs ? 1 : 0;
          else way = (XT ? XT[r * ldt + t.col] : sp_get(slots, nblk, ovf, r, t.col)) >= t.splat ? 1 : 0;  // code tie: the float decides
          const int ch = way ? t.c1 : t.c0;
          if (ch < 0) leaf[r] = t.leaf0 + 1 + way;
          else next = (unsigned)ch;
        }
        nid16[r] = (uint16_t)next;
      }
    }
    // rows per child: one shared-memory add per distinct child of the warp
    const unsigned grp = __match_any_sync(0xffffffffu, next);
    if (next != kNoNid && lane == (unsigned)(__ffs(grp) - 1)) atomicAdd(&s_cnt[next], __popc(grp));
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
                                                        int32_t* __restrict__ mnode_next) {
  if (q_dev) {  // the tree's scales live on the device (no host round trip at tree start)
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
#pragma unroll
    for (int way = 0; way < 2; way++) {
      const int ch = child_ids[2 * g + way];
      if (ch < 0) continue;
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
        s_val[pos] = make_uint2((unsigned)(wyq_rows[r] - m.x), (unsigned)(wyyq_rows[r] - m.y));
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
cudaError_t engine_alloc(Engine& e, int64_t rows, int cols, bool weighted, bool sparse) {
  engine_free(e);
  if (sparse && weighted) return cudaErrorInvalidValue;
  e.sparse = sparse;
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
    e.nblk = (cols + kSpCols - 1) / kSpCols;
    // dense rows for the (row, block) pairs with more nonzeros than a slot holds: room for one pair in sixteen
    const size_t cap = (size_t)rows * (size_t)e.nblk / 16 + 4096;
    e.ovf_cap = cap > 0xFFFFFF00ull ? 0xFFFFFF00u : (unsigned)cap;
    A(&e.slots, sizeof(SpSlot) * (size_t)rows * (size_t)e.nblk);
    A((void**)&e.ovf, sizeof(float) * kSpCols * (size_t)e.ovf_cap);
    A((void**)&e.ovf_count, sizeof(unsigned));
    if (err == cudaSuccess) err = cudaMemset(e.ovf_count, 0, sizeof(unsigned));
  } else {
    A((void**)&e.XT, sizeof(float) * (size_t)rows * (size_t)e.ldt);
  }
  A((void**)&e.code8, (size_t)cols * (size_t)e.ld8);
  A((void**)&e.thr, sizeof(uint32_t) * 256 * (size_t)cols);
  A((void**)&e.cmin, sizeof(uint32_t) * (size_t)cols);
  A((void**)&e.cmax, sizeof(uint32_t) * (size_t)cols);
  for (int k = 0; k < 2; k++) {
    A((void**)&e.gidx[k], sizeof(int32_t) * (size_t)rows);
    A((void**)&e.gval[k], sizeof(uint2) * (size_t)rows);
    if (weighted) A((void**)&e.gw[k], sizeof(int32_t) * (size_t)rows);
    A((void**)&e.seg[k], sizeof(int32_t) * kSegInts);
    A((void**)&e.mnode[k], sizeof(int32_t) * 2 * kEngMaxNodes);
  }
  A((void**)&e.nid16, sizeof(uint16_t) * (size_t)rows);
  A((void**)&e.cta_cnt, sizeof(int32_t) * (size_t)kEngMaxNodes * (size_t)e.route_ctas);
  A((void**)&e.leaf, sizeof(int32_t) * (size_t)rows);
  A((void**)&e.err, sizeof(int32_t));
  A((void**)&e.na_flag, sizeof(int32_t));
  A((void**)&e.load_err, sizeof(int32_t));
  if (err == cudaSuccess) err = cudaMemset(e.load_err, 0, sizeof(int32_t));
  e.load_checked = false;
  if (err == cudaSuccess) err = cudaMemset(e.err, 0, sizeof(int32_t));
  if (err == cudaSuccess) err = cudaMemset(e.na_flag, 0, sizeof(int32_t));
  e.na_known = false;
  e.has_na = true;
  if (err == cudaSuccess && !sparse && e.ldt != cols) err = cudaMemset(e.XT, 0, sizeof(float) * (size_t)rows * (size_t)e.ldt);
  if (err != cudaSuccess) engine_free(e);
  return err;
}

void engine_free(Engine& e) {
  cudaFree(e.XT); cudaFree(e.code8); cudaFree(e.thr); cudaFree(e.cmin); cudaFree(e.cmax);
  cudaFree(e.slots); cudaFree(e.ovf); cudaFree(e.ovf_count);
  e.slots = nullptr;
  e.ovf = nullptr;
  e.ovf_count = nullptr;
  e.sparse = false;
  e.nblk = 0;
  e.cmin = e.cmax = nullptr;
  for (int k = 0; k < 2; k++) {
    cudaFree(e.gidx[k]); cudaFree(e.gval[k]); cudaFree(e.gw[k]); cudaFree(e.seg[k]); cudaFree(e.mnode[k]);
    e.gidx[k] = nullptr; e.gval[k] = nullptr; e.gw[k] = nullptr; e.seg[k] = nullptr; e.mnode[k] = nullptr;
  }
  cudaFree(e.nid16); cudaFree(e.cta_cnt); cudaFree(e.leaf); cudaFree(e.err); cudaFree(e.na_flag);
  e.na_flag = nullptr;
  cudaFree(e.load_err);
  e.load_err = nullptr;
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

cudaError_t engine_load_cols(Engine& e, int col0, int ncols, const float* X, int64_t ld, cudaStream_t s) {
  if (e.sparse) {
    if (col0 % kSpCols) return cudaErrorInvalidValue;  // column blocks arrive whole (the last one may be short)
    for (int c = 0; c < ncols; c += kSpCols) {
      const int nc = ncols - c < kSpCols ? ncols - c : kSpCols;
      eng_load_sparse_kernel<<<nblk(e.rows, 256), 256, 0, s>>>(X + (int64_t)c * ld, ld, e.rows, nc, static_cast<SpSlot*>(e.slots),
                                                               e.nblk, (col0 + c) / kSpCols, e.ovf, e.ovf_count, e.ovf_cap,
                                                               e.load_err);
    }
    e.load_checked = false;
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

cudaError_t engine_begin_tree(Engine& e, const int32_t* wq_rows, const int32_t* wyq_rows, const int32_t* wyyq_rows,
                              const h2ok_quant* q, cudaStream_t s, const h2ok_quant* q_dev) {
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

template <int CPL, bool W, int LPR, bool NA>
static cudaError_t launch_eng_hist(const EngHistArgs& a, dim3 grid, size_t smem, cudaStream_t s) {
  static std::atomic<bool> attr_done[64];  // per device, per instantiation (setting it twice is harmless)
  int dev = 0;
  cudaGetDevice(&dev);
  std::atomic<bool>& attr = attr_done[dev & 63];
  if (!attr.load(std::memory_order_acquire)) {
    cudaError_t e = cudaFuncSetAttribute(eng_hist_kernel<CPL, W, LPR, NA>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmemCap);
    if (e != cudaSuccess) return e;
    attr.store(true, std::memory_order_release);
  }
  eng_hist_kernel<CPL, W, LPR, NA><<<grid, EngCfg<CPL, W>::threads, smem, s>>>(a);
  return cudaGetLastError();
}

cudaError_t engine_build_level(Engine& e, int n_nodes, const h2ok_col_desc* descs, int nb1, int level_nb1, int64_t* acc,
                               uint32_t* min2_enc, uint32_t* maxin_enc, cudaStream_t s) {
  if (n_nodes < 1 || n_nodes >= kEngMaxNodes) return cudaErrorInvalidValue;
  const int nbs = level_nb1 > 0 && level_nb1 < nb1 ? level_nb1 : nb1;
  if (e.sparse) {
    if ((size_t)kSpCols * nbs * 12 > (size_t)kHistSmemCap) return cudaErrorInvalidValue;
    e.n_nodes = n_nodes;
    SpHistArgs a{};
    a.slots = static_cast<const SpSlot*>(e.slots);
    a.nblk = e.nblk;
    a.ovf = e.ovf;
    a.cols = e.cols;
    a.gidx = e.gidx[e.cur];
    a.gval = e.gval[e.cur];
    a.mnode = e.mnode[e.cur];
    a.seg = e.seg[e.cur];
    a.n_nodes = n_nodes;
    a.descs = descs;
    a.nb1 = nb1;
    a.nbs = nbs;
    a.copies = 1;
    while (a.copies < 4 && (size_t)kSpCols * nbs * 12 * (a.copies * 2) <= (size_t)kHistSmemCap) a.copies *= 2;
    a.acc = acc;
    a.min2_enc = min2_enc;
    a.maxin_enc = maxin_enc;
    a.err = e.err;
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
      cudaError_t ea = cudaFuncSetAttribute(eng_hist_sparse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmemCap);
      if (ea != cudaSuccess) return ea;
      attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const dim3 grid((unsigned)e.sms, (unsigned)e.nblk);
    eng_hist_sparse_kernel<<<grid, 1024, (size_t)a.copies * kSpCols * nbs * 12, s>>>(a);
    return cudaGetLastError();
  }
  const EngLayout lay = eng_layout(e.cols, nbs);
  if (!lay.cpl) return cudaErrorInvalidValue;
  e.n_nodes = n_nodes;
  EngHistArgs a{};
  a.XT = e.XT;
  a.ldt = e.ldt;
  a.ldt4 = (unsigned)(e.ldt * 4);
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
  eng_route_kernel<<<G, kRouteThreads, rsmem, s>>>(e.XT, e.ldt, static_cast<const SpSlot*>(e.slots), e.nblk, e.ovf, e.code8, e.ld8,
                                                   e.thr, e.rows, n_nodes, splits, child_ids,
                                                   gid_base, e.nid16, e.leaf, e.cta_cnt, n_slots);
  // every level but a tree's last: lay the children's records out for the next histogram pass
  eng_scan_kernel<<<1, 1024, 0, s>>>(e.cta_cnt, G, n_nodes, child_ids, e.seg[nx], splits, e.scale_wy, e.scale_wyy,
                                     e.weighted ? nullptr : e.q_dev,
                                     e.mnode[e.cur], e.mnode[nx]);
  if (regroup) {
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

cudaError_t engine_check_load(Engine& e, cudaStream_t s, bool* too_dense) {
  *too_dense = false;
  if (!e.sparse) return cudaSuccess;
  if (!e.load_checked) {
    int flag = 0;
    cudaError_t err = cudaMemcpyAsync(&flag, e.load_err, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    if (err != cudaSuccess) return err;
    e.load_failed = flag != 0;
    e.load_checked = true;
  }
  *too_dense = e.load_failed;
  return cudaSuccess;
}

cudaError_t engine_nid(Engine& e, int32_t* nid_rows, cudaStream_t s) {
  eng_nid_kernel<<<nblk(e.rows, 256), 256, 0, s>>>(e.nid16, e.rows, nid_rows);
  return cudaGetLastError();
}

}  // namespace h2ok
