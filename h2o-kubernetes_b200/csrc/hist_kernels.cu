// hist_kernels.cu — sm_100a kernels for SURVEY.md §8 rows a2/a3/a4 (+ generator, ranges, quantisation).
//
// build_level (the hot kernel, stateless form: rows in their original order) — a scatter-reduce, no tensor cores:
//   * one CTA = one column x one contiguous chunk of rows; 128-bit coalesced loads of x and of the 8-byte row
//     records (node + fixed-point targets, packed once per level), two register tiles in ping-pong;
//   * the CTA's histogram lives in shared memory as int32 fixed-point words [node][bin][copy][channel]; the only
//     native shared-memory atomic on sm_100a is the 32-bit integer add (fp32/u64/f64 compile to CAS loops), so the
//     sums are integers: exact, order-independent, and the cross-GPU reduce is an integer sum;
//   * `copies` privatised sub-histograms per (node,bin), selected by lane, spread a warp's updates over banks:
//     up to 32 copies per node (bank == lane: conflict-free even for 3-valued columns); copies = 64 / nodes;
//   * 32-bit words cannot hold a chunk's sum, so each add reads back the old word and the (rare) carry / borrow
//     out of bit 31 goes straight to the 64-bit global accumulator; at the end the low words are flushed once;
//   * one node: descriptor and observed range in registers; several nodes: two replicated, conflict-free
//     shared-memory tables read with one LDS.128 + one LDS.64 per row;
//   * grid = cols x chunks is a multiple of 148 SMs whenever cols % 4 == 0, column index fastest so the CTAs that
//     share a row chunk run together and the row records are served from L2 after the first column.
// Measured bound (ncu): the SM's LSU / shared-memory data pipe (92-99 % busy beyond the root), see DESIGN.md §3.1.
#include "common.cuh"
#include "kernels.h"

namespace h2ok {

// ------------------------------------------------------------------------------------------------ generator
__global__ void synth_x_kernel(int shape, uint64_t seed, int64_t row0, int64_t rows, int col0, int64_t ld,
                               int na_permille, float* __restrict__ X) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = blockIdx.y;
  if (r < ld) X[(int64_t)c * ld + r] = r < rows ? gen_x(shape, seed, row0 + r, col0 + c, na_permille) : 0.0f;
}

__global__ void synth_yw_kernel(uint64_t seed, int64_t row0, int64_t rows, int cols, int weight_mode,
                                float* __restrict__ y, float* __restrict__ w) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  if (y) y[r] = gen_y(seed, row0 + r, cols);
  if (w) w[r] = gen_w(seed, row0 + r, weight_mode);
}

// ------------------------------------------------------------------------------------------------ ranges
__global__ void init_enc_kernel(uint32_t* mn, uint32_t* mx, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    mn[i] = kEncPosInf;
    mx[i] = kEncNegInf;
  }
}

__global__ void col_ranges_kernel(const float* __restrict__ X, int64_t ld, int64_t rows, uint32_t* __restrict__ mn,
                                  uint32_t* __restrict__ mx) {
  const float* x = X + (int64_t)blockIdx.y * ld;
  uint32_t lo = kEncPosInf, hi = kEncNegInf;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    float v = x[r];
    if (v == v) {
      uint32_t e = enc_f32_bits(__float_as_uint(v));
      lo = min(lo, e);
      hi = max(hi, e);
    }
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  if ((threadIdx.x & 31) == 0) {
    atomicMin(&mn[blockIdx.y], lo);
    atomicMax(&mx[blockIdx.y], hi);
  }
}

// ------------------------------------------------------------------------------------------------ targets
__device__ inline void atomic_max_f64_nonneg(double* addr, double v) {  // v >= 0: bit pattern is monotone
  atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}

__global__ void target_absmax_kernel(const float* __restrict__ y, const float* __restrict__ w, int64_t rows,
                                     double* __restrict__ out) {
  double m0 = 0, m1 = 0, m2 = 0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    double wd = w ? (double)w[r] : 1.0, yd = (double)y[r];
    double wy = __dmul_rn(wd, yd);
    double wyy = __dmul_rn(wy, yd);
    m0 = fmax(m0, fabs(wd));
    m1 = fmax(m1, fabs(wy));
    m2 = fmax(m2, fabs(wyy));
  }
  for (int o = 16; o; o >>= 1) {
    m0 = fmax(m0, __shfl_xor_sync(0xffffffffu, m0, o));
    m1 = fmax(m1, __shfl_xor_sync(0xffffffffu, m1, o));
    m2 = fmax(m2, __shfl_xor_sync(0xffffffffu, m2, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomic_max_f64_nonneg(out + 0, m0);
    atomic_max_f64_nonneg(out + 1, m1);
    atomic_max_f64_nonneg(out + 2, m2);
  }
}

// Power-of-two fixed-point scales from max|w|, max|w*y|, max|w*y*y| on the device (the host rule of
// h2ok_hist_choose_quant, bit for bit): lets a tree start without a host round trip.
__device__ inline double pow2_scale_dev(double absmax, int qbits) {
  if (!(absmax > 0.0) || isinf(absmax)) return 1.0;
  int ex;
  frexp(absmax, &ex);
  return ldexp(1.0, qbits - ex);
}
__global__ void choose_quant_kernel(const double* __restrict__ m, int weighted, int qbits, h2ok_quant* __restrict__ q) {
  q->weighted = weighted ? 1 : 0;
  q->qbits = qbits;
  q->scale_w = weighted ? pow2_scale_dev(m[0], qbits) : 1.0;
  q->scale_wy = pow2_scale_dev(m[1], qbits);
  q->scale_wyy = pow2_scale_dev(m[2], qbits);
}

// qd != NULL: the scales come from device memory (choose_quant_kernel) instead of the by-value copy
__global__ void quantize_rows_kernel(const float* __restrict__ y, const float* __restrict__ w, int64_t rows,
                                     h2ok_quant q, const h2ok_quant* __restrict__ qd, int32_t* __restrict__ wq,
                                     int32_t* __restrict__ wyq, int32_t* __restrict__ wyyq) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  if (qd) q = *qd;
  double wd = (q.weighted && w) ? (double)w[r] : 1.0, yd = (double)y[r];
  double wy = __dmul_rn(wd, yd);
  double wyy = __dmul_rn(wy, yd);
  if (q.weighted && wq) wq[r] = __double2int_rn(__dmul_rn(wd, q.scale_w));
  wyq[r] = __double2int_rn(__dmul_rn(wy, q.scale_wy));
  wyyq[r] = __double2int_rn(__dmul_rn(wyy, q.scale_wyy));
}

// ------------------------------------------------------------------------------------------------ a2 test hook
__global__ void bin_kernel(const float* __restrict__ x, int64_t n, h2ok_col_desc d, int32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = bin_of(x[i], d.min, d.step, d.nbin);
}

// ------------------------------------------------------------------------------------------------ a3/a4 build
// Row records.  Once per level every row's node index and fixed-point targets are packed into one 8-byte word so
// that the hot kernel reads 12 B per cell from L2 instead of 16 and needs no per-cell validity / weight test:
//   lo = wyq  << 5 | node[4:0]     (wyq  signed,   |wyq|  < 2^26)
//   hi = wyyq << 5 | node[9:5]     (wyyq unsigned,  wyyq  < 2^27)
// node 0x3FF marks a row that takes no part in this level (already in a leaf, or zero weight).
__global__ void pack_rows_kernel(const int32_t* __restrict__ nid, const int32_t* __restrict__ wq,
                                 const int32_t* __restrict__ wyq, const int32_t* __restrict__ wyyq, int64_t rows,
                                 int64_t padded, int n_nodes, uint2* __restrict__ rec, int32_t* __restrict__ err) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= padded) return;
  unsigned node = kNoNode;
  int y1 = 0, y2 = 0;
  if (r < rows) {
    int g = nid[r];
    y1 = wyq[r];
    y2 = wyyq[r];
    if (g >= 0 && g < n_nodes && (!wq || wq[r] != 0)) node = (unsigned)g;
    if (y1 >= (1 << 26) || y1 < -(1 << 26) || y2 < 0 || y2 >= (1 << 27) || (wq && (wq[r] < 0 || wq[r] >= (1 << 27)))) {
      atomicOr(err, 1);  // targets outside the packed format's range (qbits > 25): reported by last_level_stats
      node = kNoNode;
    }
  }
  rec[r] = rec_pack(y1, y2, node);
}

// Shared-memory descriptor tables of the multi-node path.  A row's node is data dependent, so each lane reads its
// node's record; the tables are replicated per lane group — T1 entry (g,k) at g*8+k with k = lane&7 (16-byte records:
// the 8 lanes of an LDS.128 phase hit 8 different bank groups), T2 entry (g,k) at g*16+k with k = lane&15 (8-byte
// records, 16 lanes per LDS.64 phase) — so the reads are conflict-free whatever the nodes are.
constexpr int kRep1 = 8;
constexpr int kRep2 = 16;

// T1, fast form.  In this library a descriptor's min is always a value of the fp32 column (the observed minimum of
// the parent's rows, or the float routing threshold `splat`), so it is stored as the float it is and the record
// (step, min, nbin) fits one LDS.128.  A caller-made descriptor whose min is not exactly a float makes the CTA fall
// back to the fp64 form below (block-uniform decision, same results).
struct DescF32 {
  double step;
  float minf;
  int nbin;
};
struct DescF64 {  // fallback form: two LDS.128 per row
  double step;
  double min;
};
struct Range {  // T2: encoded observed min / max so far; a row touches them only when it beats them
  unsigned emn, emx;
};

enum { kOneNode = 0, kTableF64 = 1, kTableF32 = 2 };

template <bool WEIGHTED, int MODE>
struct RowCtx {
  const HistArgs& a;
  unsigned char* hsub;  // histogram base + this lane's copy offset
  const DescF32* t1;    // MODE kTableF32: + g*8 + (lane&7)
  const DescF64* t1d;   // MODE kTableF64
  const int* nbins;     // MODE kTableF64: per node
  Range* t2;            // + g*16 + (lane&15)
  int cstride;          // bytes between consecutive bins of one node (copies * 12)
  int nstride;          // bytes between consecutive nodes (nb1 * copies * 12)
  int col;
  // kOneNode: the single descriptor and the thread's running extremes live in registers
  double mn, step;
  int nbin;
  float fmn, fmx;
};

template <bool WEIGHTED, int MODE>
__device__ __forceinline__ void accumulate_row(RowCtx<WEIGHTED, MODE>& c, float x, unsigned lo, unsigned hi, int wq) {
  unsigned g = 0;
  int b, cellofs = 0;
  if (MODE == kOneNode) {
    if ((lo | hi) & 31u) return;  // only node 0 exists at this level
    b = bin_fast(x, c.mn, c.step, c.nbin);
    c.fmn = fminf(c.fmn, x);  // NaN-ignoring, exact: the thread's observed range stays in two registers
    c.fmx = fmaxf(c.fmx, x);
  } else {
    g = ((lo & 31u) | ((hi & 31u) << 5)) - (unsigned)c.a.node_base;
    if (g >= (unsigned)c.a.node_cnt) return;
    // No branch on "column untracked in this node" (nbin == 0): such nodes carry a harmless 1-bin record, their
    // shared words are simply never flushed, and the hot loop stays one LDS.128 + one LDS.64 per row.
    if (MODE == kTableF32) {
      const int4 q = *reinterpret_cast<const int4*>(c.t1 + g * kRep1);  // {step.lo, step.hi, minf, nbin}
      b = bin_fast(x, (double)__int_as_float(q.z), __hiloint2double(q.y, q.x), q.w);
    } else {
      const double2 d = *reinterpret_cast<const double2*>(c.t1d + g * kRep1);
      b = bin_fast(x, d.y, d.x, max(c.nbins[g], 1));
    }
    cellofs = g * c.nstride;
    Range* rg = c.t2 + g * kRep2;
    const uint2 r = *reinterpret_cast<const uint2*>(rg);
    const unsigned xb = __float_as_uint(x);
    const unsigned e = xb ^ ((unsigned)((int)xb >> 31) | 0x80000000u);  // == enc_f32_bits(xb)
    if ((e < r.x) | (e > r.y)) {  // a new record (rare), or NaN (whose encoding sorts above +inf)
      if (x == x) {
        atomicMin(&rg->emn, e);
        atomicMax(&rg->emx, e);
      }
    }
  }
  unsigned* p = reinterpret_cast<unsigned*>(c.hsub + cellofs + b * c.cstride);
  const int v1 = (int)lo >> 5;
  const unsigned v2 = hi >> 5;
  bool ev = false;
  if (WEIGHTED) {
    unsigned o0 = atomicAdd(p, (unsigned)wq);
    ev = o0 + (unsigned)wq < o0;
  } else {
    atomicAdd(p, 1u);  // < 2^32 rows per CTA: never wraps
  }
  const unsigned o1 = atomicAdd(p + 1, (unsigned)v1);
  const unsigned n1 = o1 + (unsigned)v1;
  const bool e1 = (n1 < o1) != (v1 < 0);  // carry when adding a positive value, borrow when adding a negative one
  const unsigned o2 = atomicAdd(p + 2, v2);
  const bool e2 = o2 + v2 < o2;
  if ((ev | e1 | e2) && (MODE == kOneNode || c.nbins[g] > 0)) {
    const int64_t cell = ((int64_t)(c.a.node_base + g) * c.a.cols + c.col) * c.a.nb1 + b;
    if (ev) carry_to_global(c.a.acc, cell, 0, 1);
    if (e1) carry_to_global(c.a.acc, cell, 1, v1 < 0 ? -1 : 1);
    if (e2) carry_to_global(c.a.acc, cell, 2, 1);
  }
}

// The row loop: two register tiles in ping-pong, so the next tile's loads are in flight while this tile's atomics
// issue.  128-bit loads: 4 rows of x, 2 + 2 packed row records (and 4 weights).
template <bool WEIGHTED, int MODE>
__device__ __forceinline__ void row_loop(RowCtx<WEIGHTED, MODE>& c, const float* __restrict__ x,
                                         const uint2* __restrict__ rec, const int32_t* __restrict__ wq, int n, int tid) {
  constexpr int kStep = kHistThreads * 4;
  const int nvec = n & ~3;
  const float* xp = x + tid * 4;
  const uint2* rp = rec + tid * 4;
  const int32_t* wp = WEIGHTED ? wq + tid * 4 : nullptr;
  int left = nvec - tid * 4;  // rows from this thread's position to the end of the vector part
  float4 xa = make_float4(0, 0, 0, 0), xb = xa;
  uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0, b0 = a0, b1 = a0;
  int4 wa = make_int4(0, 0, 0, 0), wb = wa;
#define H2OK_LOAD_TILE(XV, R0, R1, WV, OFF)                                        \
  {                                                                                \
    XV = __ldcs(reinterpret_cast<const float4*>(xp + (OFF))); /* streamed once */  \
    R0 = __ldg(reinterpret_cast<const uint4*>(rp + (OFF)));                        \
    R1 = __ldg(reinterpret_cast<const uint4*>(rp + (OFF) + 2));                    \
    if (WEIGHTED) WV = __ldg(reinterpret_cast<const int4*>(wp + (OFF)));           \
  }
#define H2OK_DO_TILE(XV, R0, R1, WV)           \
  {                                            \
    accumulate_row(c, XV.x, R0.x, R0.y, WV.x); \
    accumulate_row(c, XV.y, R0.z, R0.w, WV.y); \
    accumulate_row(c, XV.z, R1.x, R1.y, WV.z); \
    accumulate_row(c, XV.w, R1.z, R1.w, WV.w); \
  }
  if (left > 0) {
    H2OK_LOAD_TILE(xa, a0, a1, wa, 0);
    while (true) {
      if (left > kStep) H2OK_LOAD_TILE(xb, b0, b1, wb, kStep);
      H2OK_DO_TILE(xa, a0, a1, wa);
      left -= kStep;
      if (left <= 0) break;
      xp += 2 * kStep;
      rp += 2 * kStep;
      if (WEIGHTED) wp += 2 * kStep;
      if (left > kStep) H2OK_LOAD_TILE(xa, a0, a1, wa, 0);
      H2OK_DO_TILE(xb, b0, b1, wb);
      left -= kStep;
      if (left <= 0) break;
    }
  }
#undef H2OK_LOAD_TILE
#undef H2OK_DO_TILE
  if (tid < n - nvec) {  // at most 3 rows at the very end of the shard
    const int r = nvec + tid;
    const uint2 q = rec[r];
    accumulate_row(c, x[r], q.x, q.y, WEIGHTED ? wq[r] : 1);
  }
}

// shared-memory layout: [T1: cnt*8 x 16 B][T2: cnt*16 x 8 B][nbins: cnt x 4 B, padded to 16][histogram words]
__host__ __device__ inline size_t hist_head_bytes(int cnt) { return (size_t)cnt * 256 + (((size_t)cnt * 4 + 15) & ~(size_t)15); }

template <bool WEIGHTED, bool ONE_NODE>
__global__ void __launch_bounds__(kHistThreads, 1) hist_build_kernel(HistArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int cnt = a.node_cnt;
  DescF64* t1 = reinterpret_cast<DescF64*>(smem_raw);
  Range* t2 = reinterpret_cast<Range*>(t1 + cnt * kRep1);
  int* s_nbin = reinterpret_cast<int*>(t2 + cnt * kRep2);
  unsigned* h = reinterpret_cast<unsigned*>(smem_raw + hist_head_bytes(cnt));
  const int copies = a.copies, nb1 = a.nb1;
  const int cells = cnt * nb1;
  const int words = cells * copies * 3;
  const int col = blockIdx.x % a.cols;
  const int chunk = blockIdx.x / a.cols;
  const int tid = threadIdx.x;

  for (int i = tid; i < words; i += kHistThreads) h[i] = (i % 3 == 1) ? kWyBias : 0u;
  // every tracked min exactly a float?  (block-uniform: decides which T1 form this CTA uses)
  bool min_is_f32 = true;
  for (int g = tid; g < cnt; g += kHistThreads) {
    const h2ok_col_desc& d = a.descs[(int64_t)(a.node_base + g) * a.cols + col];
    s_nbin[g] = d.nbin;
    if (d.nbin > 0 && (double)(float)d.min != d.min) min_is_f32 = false;
  }
  const bool f32 = __syncthreads_and(min_is_f32) != 0;
  for (int gk = tid; gk < cnt * kRep1; gk += kHistThreads) {
    const h2ok_col_desc& d = a.descs[(int64_t)(a.node_base + gk / kRep1) * a.cols + col];
    const bool tracked = d.nbin > 0;
    if (f32) {
      DescF32 q;
      q.step = tracked ? d.step : 0.0;
      q.minf = tracked ? (float)d.min : 0.0f;
      q.nbin = tracked ? d.nbin : 1;
      reinterpret_cast<DescF32*>(t1)[gk] = q;
    } else {
      DescF64 q;
      q.step = tracked ? d.step : 0.0;
      q.min = tracked ? d.min : 0.0;
      t1[gk] = q;
    }
  }
  for (int gk = tid; gk < cnt * kRep2; gk += kHistThreads) {
    Range r;
    r.emn = kEncPosInf;
    r.emx = kEncNegInf;
    t2[gk] = r;
  }
  __syncthreads();

  const int64_t r0 = (int64_t)chunk * a.chunk_rows;
  const int64_t r1 = min(a.rows, r0 + a.chunk_rows);
  const int n = (int)(r1 - r0);
  const float* __restrict__ x = a.X + (int64_t)col * a.ld + r0;
  const uint2* __restrict__ rec = a.rec + r0;
  const int32_t* __restrict__ wq = WEIGHTED ? a.wq + r0 : nullptr;
  unsigned char* hsub = reinterpret_cast<unsigned char*>(h) + (tid & (copies - 1)) * 12;

  if (ONE_NODE) {
    RowCtx<WEIGHTED, kOneNode> c{a};
    c.hsub = hsub;
    c.cstride = copies * 12;
    c.col = col;
    const h2ok_col_desc& d = a.descs[(int64_t)a.node_base * a.cols + col];
    c.mn = d.min;
    c.step = d.step;
    c.nbin = d.nbin;
    c.fmn = __int_as_float(0x7f800000);
    c.fmx = __int_as_float(0xff800000);
    if (c.nbin > 0) row_loop(c, x, rec, wq, n, tid);
    unsigned emn = __reduce_min_sync(0xffffffffu, enc_f32_bits(__float_as_uint(c.fmn)));
    unsigned emx = __reduce_max_sync(0xffffffffu, enc_f32_bits(__float_as_uint(c.fmx)));
    if ((tid & 31) == 0) {
      atomicMin(&t2[0].emn, emn);
      atomicMax(&t2[0].emx, emx);
    }
  } else if (f32) {
    RowCtx<WEIGHTED, kTableF32> c{a};
    c.hsub = hsub;
    c.t1 = reinterpret_cast<const DescF32*>(t1) + (tid & (kRep1 - 1));
    c.nbins = s_nbin;
    c.t2 = t2 + (tid & (kRep2 - 1));
    c.cstride = copies * 12;
    c.nstride = nb1 * copies * 12;
    c.col = col;
    row_loop(c, x, rec, wq, n, tid);
  } else {
    RowCtx<WEIGHTED, kTableF64> c{a};
    c.hsub = hsub;
    c.t1d = t1 + (tid & (kRep1 - 1));
    c.nbins = s_nbin;
    c.t2 = t2 + (tid & (kRep2 - 1));
    c.cstride = copies * 12;
    c.nstride = nb1 * copies * 12;
    c.col = col;
    row_loop(c, x, rec, wq, n, tid);
  }
  __syncthreads();

  // flush: low words of every copy → global 64-bit accumulators (zero-extended; the wY bias is taken back out).
  // The copy order is rotated by lane so that the reads stay conflict-free.
  const int lane = tid & 31;
  for (int i = tid; i < cells; i += kHistThreads) {
    const int g = i / nb1, b = i - g * nb1;
    if (s_nbin[g] <= 0) continue;  // untracked (node, column): nothing was meant to be accumulated
    unsigned long long* gs =
        reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)(a.node_base + g) * a.cols + col) * nb1 + b) * 3;
    const unsigned* hc = h + (size_t)i * copies * 3;
    unsigned long long s0 = 0, s1 = 0, s2 = 0;
    for (int k = 0; k < copies; k++) {
      const unsigned* w = hc + ((k + lane) & (copies - 1)) * 3;
      s0 += w[0];
      s1 += w[1];
      s2 += w[2];
    }
    s1 -= (unsigned long long)copies * kWyBias;
    if (s0) atomicAdd(gs, s0);
    if (s1) atomicAdd(gs + 1, s1);
    if (s2) atomicAdd(gs + 2, s2);
  }
  for (int gk = tid; gk < cnt * kRep2; gk += kHistThreads) {
    if (s_nbin[gk / kRep2] <= 0) continue;
    const int64_t nc = (int64_t)(a.node_base + gk / kRep2) * a.cols + col;
    if (t2[gk].emn != kEncPosInf) atomicMin(&a.min2_enc[nc], t2[gk].emn);
    if (t2[gk].emx != kEncNegInf) atomicMax(&a.maxin_enc[nc], t2[gk].emx);
  }
}

__global__ void finalize_kernel(const int64_t* __restrict__ acc, int64_t n, double inv_w, double inv_wy, double inv_wyy,
                                const h2ok_quant* __restrict__ qd, double* __restrict__ hist) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (qd) {  // powers of two: the reciprocals are exact
    inv_w = 1.0 / qd->scale_w;
    inv_wy = 1.0 / qd->scale_wy;
    inv_wyy = 1.0 / qd->scale_wyy;
  }
  int ch = (int)(i % 3);
  hist[i] = __dmul_rn((double)acc[i], ch == 0 ? inv_w : (ch == 1 ? inv_wy : inv_wyy));
}

__global__ void decode_ranges_kernel(const uint32_t* __restrict__ mn, const uint32_t* __restrict__ mx, int64_t n,
                                     float* __restrict__ min2, float* __restrict__ maxin) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  min2[i] = __fadd_rn(__uint_as_float(dec_f32_bits(mn[i])), 0.0f);  // canonical +0
  maxin[i] = __fadd_rn(__uint_as_float(dec_f32_bits(mx[i])), 0.0f);
}

// ------------------------------------------------------------------------------------------------ launchers
static inline unsigned blocks_for(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

cudaError_t launch_synth(int shape, uint64_t seed, int64_t row0, int64_t rows, int cols, int64_t ld, int na_permille,
                         int weight_mode, float* X, float* y, float* w, cudaStream_t s) {
  if (X) {
    dim3 grid(blocks_for(ld, 256), (unsigned)cols);
    synth_x_kernel<<<grid, 256, 0, s>>>(shape, seed, row0, rows, 0, ld, na_permille, X);
  }
  if (y || w) synth_yw_kernel<<<blocks_for(rows, 256), 256, 0, s>>>(seed, row0, rows, cols, weight_mode, y, w);
  return cudaGetLastError();
}

cudaError_t launch_synth_cols(int shape, uint64_t seed, int64_t row0, int64_t rows, int col0, int ncols, int64_t ld,
                              int na_permille, float* X, cudaStream_t s) {
  dim3 grid(blocks_for(ld, 256), (unsigned)ncols);
  synth_x_kernel<<<grid, 256, 0, s>>>(shape, seed, row0, rows, col0, ld, na_permille, X);
  return cudaGetLastError();
}

cudaError_t launch_init_enc(uint32_t* mn, uint32_t* mx, int64_t n, cudaStream_t s) {
  init_enc_kernel<<<blocks_for(n, 256), 256, 0, s>>>(mn, mx, n);
  return cudaGetLastError();
}

cudaError_t launch_col_ranges(const float* X, int64_t ld, int64_t rows, int cols, uint32_t* mn, uint32_t* mx,
                              cudaStream_t s) {
  cudaError_t e = launch_init_enc(mn, mx, cols, s);
  if (e != cudaSuccess) return e;
  int64_t per = (rows + 255) / 256;
  unsigned bx = (unsigned)(per < 1 ? 1 : (per > 4 * kSMs ? 4 * kSMs : per));
  dim3 grid(bx, (unsigned)cols);
  col_ranges_kernel<<<grid, 256, 0, s>>>(X, ld, rows, mn, mx);
  return cudaGetLastError();
}

cudaError_t launch_target_absmax(const float* y, const float* w, int64_t rows, double* out, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(out, 0, 3 * sizeof(double), s);
  if (e != cudaSuccess) return e;
  int64_t per = (rows + 255) / 256;
  unsigned bx = (unsigned)(per < 1 ? 1 : (per > 8 * kSMs ? 8 * kSMs : per));
  target_absmax_kernel<<<bx, 256, 0, s>>>(y, w, rows, out);
  return cudaGetLastError();
}

cudaError_t launch_quantize_rows(const float* y, const float* w, int64_t rows, const h2ok_quant& q, int32_t* wq,
                                 int32_t* wyq, int32_t* wyyq, cudaStream_t s) {
  quantize_rows_kernel<<<blocks_for(rows, 256), 256, 0, s>>>(y, w, rows, q, nullptr, wq, wyq, wyyq);
  return cudaGetLastError();
}

cudaError_t launch_choose_quant_dev(const double* absmax3, int weighted, int qbits, h2ok_quant* q_dev, cudaStream_t s) {
  choose_quant_kernel<<<1, 1, 0, s>>>(absmax3, weighted, qbits, q_dev);
  return cudaGetLastError();
}

cudaError_t launch_quantize_rows_dev(const float* y, const float* w, int64_t rows, bool weighted, const h2ok_quant* q_dev,
                                     int32_t* wq, int32_t* wyq, int32_t* wyyq, cudaStream_t s) {
  h2ok_quant q{};
  q.weighted = weighted ? 1 : 0;  // overwritten from q_dev inside the kernel
  quantize_rows_kernel<<<blocks_for(rows, 256), 256, 0, s>>>(y, w, rows, q, q_dev, wq, wyq, wyyq);
  return cudaGetLastError();
}

cudaError_t launch_bin(const float* x, int64_t n, const h2ok_col_desc& d, int32_t* out, cudaStream_t s) {
  bin_kernel<<<blocks_for(n, 256), 256, 0, s>>>(x, n, d, out);
  return cudaGetLastError();
}

size_t hist_smem_bytes(int node_cnt, int nb1, int copies) {
  return hist_head_bytes(node_cnt) + (size_t)3 * node_cnt * nb1 * copies * sizeof(int);
}

cudaError_t launch_pack_rows(const int32_t* nid, const int32_t* wq, const int32_t* wyq, const int32_t* wyyq, int64_t rows,
                             int64_t padded, int n_nodes, uint2* rec, int32_t* err, cudaStream_t s) {
  pack_rows_kernel<<<blocks_for(padded, 256), 256, 0, s>>>(nid, wq, wyq, wyyq, rows, padded, n_nodes, rec, err);
  return cudaGetLastError();
}

// Chooses the node-group size and the number of privatised copies so that one CTA's histogram fits the
// 227 KB of shared memory an sm_100a CTA may opt into.
HistPlan plan_hist(int n_nodes, int cols, int64_t rows, int nb1) {
  HistPlan p;
  const size_t budget = 227 * 1024;
  const size_t slot = (size_t)nb1 * 3 * sizeof(int);
  // (node, copy) slots that fit beside the descriptor tables of 64 nodes: a first guess (64 at 256 bins) ...
  int slots = (int)((budget - hist_head_bytes(64)) / slot);
  int pow2 = 1;
  while (pow2 * 2 <= slots) pow2 *= 2;
  slots = pow2;
  p.group = n_nodes < slots ? n_nodes : slots;
  // ... which is too generous once a group has more than 64 nodes (fewer than 256 bins): the tables grow with the
  // group, so the group shrinks until tables + one histogram copy per node really fit (more passes over X instead)
  while (p.group > 1 && hist_smem_bytes(p.group, nb1, 1) > budget) p.group--;
  p.passes = (n_nodes + p.group - 1) / p.group;
  int copies = 1;
  while (copies * 2 <= 32 && hist_smem_bytes(p.group, nb1, copies * 2) <= budget) copies *= 2;
  p.copies = copies;
  // chunks: cols*chunks CTAs, a multiple of 148 when cols % 4 == 0, about 4-16 waves; never below one tile
  int64_t tile = (int64_t)kHistThreads * 4;
  int64_t max_chunks = (rows + tile - 1) / tile;
  int64_t chunks = 37;
  while (cols * chunks < 4 * kSMs) chunks += 37;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  int64_t chunk_rows = (rows + chunks - 1) / chunks;
  chunk_rows = (chunk_rows + tile - 1) / tile * tile;
  p.chunk_rows = chunk_rows;
  p.chunks = (int)((rows + chunk_rows - 1) / chunk_rows);
  if (p.chunks < 1) p.chunks = 1;
  p.smem = hist_smem_bytes(p.group, nb1, p.copies);
  return p;
}

cudaError_t launch_hist_build(const HistArgs& base, const HistPlan& plan, int n_nodes, bool weighted, cudaStream_t s) {
  static bool attr_done[64] = {};  // the opt-in to > 48 KB of dynamic shared memory is per device
  int dev = 0;
  cudaGetDevice(&dev);
  bool& attr_set = attr_done[dev & 63];
  if (!attr_set) {
    cudaError_t e = cudaSuccess;
    const int kMax = 227 * 1024;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(hist_build_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(hist_build_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(hist_build_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(hist_build_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  for (int pass = 0; pass < plan.passes; pass++) {
    HistArgs a = base;
    a.node_base = pass * plan.group;
    a.node_cnt = n_nodes - a.node_base < plan.group ? n_nodes - a.node_base : plan.group;
    a.copies = plan.copies;
    a.chunk_rows = plan.chunk_rows;
    size_t smem = hist_smem_bytes(a.node_cnt, a.nb1, a.copies);
    if (smem > (size_t)227 * 1024) return cudaErrorInvalidValue;  // a plan that does not fit must not reach the launch
    unsigned grid = (unsigned)(a.cols * plan.chunks);
    const bool one = n_nodes == 1;
    if (weighted && one) hist_build_kernel<true, true><<<grid, kHistThreads, smem, s>>>(a);
    else if (weighted) hist_build_kernel<true, false><<<grid, kHistThreads, smem, s>>>(a);
    else if (one) hist_build_kernel<false, true><<<grid, kHistThreads, smem, s>>>(a);
    else hist_build_kernel<false, false><<<grid, kHistThreads, smem, s>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_finalize(const int64_t* acc, const uint32_t* mn, const uint32_t* mx, int64_t n_node_cols, int nb1,
                            const h2ok_quant& q, double* hist, float* min2, float* maxin, cudaStream_t s) {
  int64_t n = n_node_cols * nb1 * 3;
  if (hist)
    finalize_kernel<<<blocks_for(n, 256), 256, 0, s>>>(acc, n, 1.0 / q.scale_w, 1.0 / q.scale_wy, 1.0 / q.scale_wyy, nullptr, hist);
  if (min2) decode_ranges_kernel<<<blocks_for(n_node_cols, 256), 256, 0, s>>>(mn, mx, n_node_cols, min2, maxin);
  return cudaGetLastError();
}

cudaError_t launch_finalize_dev(const int64_t* acc, int64_t n_node_cols, int nb1, const h2ok_quant* q_dev, double* hist,
                                cudaStream_t s) {
  int64_t n = n_node_cols * nb1 * 3;
  finalize_kernel<<<blocks_for(n, 256), 256, 0, s>>>(acc, n, 1.0, 1.0, 1.0, q_dev, hist);
  return cudaGetLastError();
}

}  // namespace h2ok
