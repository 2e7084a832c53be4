// hist_kernels.cu — sm_100a kernels for SURVEY.md §8 rows a2/a3/a4 (+ generator, ranges, quantisation).
//
// build_level (the hot kernel) — HBM-bound scatter-reduce, no tensor cores:
//   * one CTA = one column x one contiguous chunk of rows; 128-bit coalesced loads of x / nid / targets;
//   * the CTA's histogram lives in shared memory as int32 fixed-point words [channel][node][bin][copy]; the only
//     native shared-memory atomic on sm_100a is the 32-bit integer add (fp32/u64/f64 compile to CAS loops), so the
//     sums are integers: exact, order-independent, and the cross-GPU reduce is an integer sum;
//   * `copies` privatised sub-histograms per (node,bin), selected by lane, spread a warp's updates over banks:
//     with few nodes every lane owns a bank (conflict-free even for 3-valued columns), with many nodes the node
//     id itself spreads the addresses;
//   * 32-bit words cannot hold a chunk's sum, so each add reads back the old word and the (rare) carry / borrow
//     out of bit 31 goes straight to the 64-bit global accumulator; at the end the low words are flushed once;
//   * grid = cols x chunks is a multiple of 148 SMs whenever cols % 4 == 0, column index fastest so the CTAs that
//     share a row chunk run together and the row-indexed arrays are served from L2 after the first column.
#include "common.cuh"
#include "kernels.h"

namespace h2ok {

// ------------------------------------------------------------------------------------------------ generator
__global__ void synth_x_kernel(int shape, uint64_t seed, int64_t row0, int64_t rows, int cols, int64_t ld,
                               int na_permille, float* __restrict__ X) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = blockIdx.y;
  if (r < ld) X[(int64_t)c * ld + r] = r < rows ? gen_x(shape, seed, row0 + r, c, na_permille) : 0.0f;
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

__global__ void quantize_rows_kernel(const float* __restrict__ y, const float* __restrict__ w, int64_t rows,
                                     h2ok_quant q, int32_t* __restrict__ wq, int32_t* __restrict__ wyq,
                                     int32_t* __restrict__ wyyq) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
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
// 32-bit shared add with the carry/borrow out of bit 31 forwarded to the 64-bit global accumulator.
__device__ __forceinline__ void add_lo_carry(int* slot, int v, unsigned long long* gslot) {
  unsigned old = atomicAdd(reinterpret_cast<unsigned*>(slot), (unsigned)v);
  unsigned nw = old + (unsigned)v;
  int hi = (v >> 31) + (nw < (unsigned)v ? 1 : 0);  // sign extension of v plus carry-out: -1, 0 or +1
  if (hi != 0) atomicAdd(gslot, (unsigned long long)(long long)hi << 32);
}

template <bool WEIGHTED>
__device__ __forceinline__ void accumulate_row(const HistArgs& a, int col, float x, int nid, int wq, int wyq, int wyyq,
                                               int sub, const double* s_min, const double* s_step, const int* s_nbin,
                                               unsigned* s_mn, unsigned* s_mx, int* h0, int chstride) {
  unsigned g = (unsigned)(nid - a.node_base);
  if (g >= (unsigned)a.node_cnt) return;
  int nb = s_nbin[g];
  if (nb <= 0) return;
  if (WEIGHTED && wq == 0) return;  // updateHisto: zero-weight rows are skipped before min/max
  if (x == x) {                     // observed range (_min2/_maxIn); NaN never compares
    unsigned e = enc_f32_bits(__float_as_uint(x));
    if (e < s_mn[g]) atomicMin(&s_mn[g], e);
    if (e > s_mx[g]) atomicMax(&s_mx[g], e);
  }
  int b = bin_of(x, s_min[g], s_step[g], nb);
  int cell = (int)g * a.nb1 + b;
  int idx = cell * a.copies + sub;
  unsigned long long* gs =
      reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)(a.node_base + g) * a.cols + col) * a.nb1 + b) * 3;
  if (WEIGHTED)
    add_lo_carry(h0 + idx, wq, gs);
  else
    atomicAdd(reinterpret_cast<unsigned*>(h0 + idx), 1u);  // < 2^32 rows per CTA: never carries
  add_lo_carry(h0 + chstride + idx, wyq, gs + 1);
  add_lo_carry(h0 + 2 * chstride + idx, wyyq, gs + 2);
}

template <bool WEIGHTED>
__global__ void __launch_bounds__(kHistThreads, 1) hist_build_kernel(HistArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* s_min = reinterpret_cast<double*>(smem_raw);
  double* s_step = s_min + a.node_cnt;
  int* s_nbin = reinterpret_cast<int*>(s_step + a.node_cnt);
  unsigned* s_mn = reinterpret_cast<unsigned*>(s_nbin + a.node_cnt);
  unsigned* s_mx = s_mn + a.node_cnt;
  int* h0 = reinterpret_cast<int*>(s_mx + a.node_cnt);
  const int chstride = a.node_cnt * a.nb1 * a.copies;
  const int col = blockIdx.x % a.cols;
  const int chunk = blockIdx.x / a.cols;
  const int tid = threadIdx.x;

  for (int i = tid; i < 3 * chstride; i += kHistThreads) h0[i] = 0;
  for (int g = tid; g < a.node_cnt; g += kHistThreads) {
    const h2ok_col_desc& d = a.descs[(int64_t)(a.node_base + g) * a.cols + col];
    s_min[g] = d.min;
    s_step[g] = d.step;
    s_nbin[g] = d.nbin;
    s_mn[g] = kEncPosInf;
    s_mx[g] = kEncNegInf;
  }
  __syncthreads();

  const int64_t r0 = (int64_t)chunk * a.chunk_rows;
  const int64_t r1 = min(a.rows, r0 + a.chunk_rows);
  const float* __restrict__ x = a.X + (int64_t)col * a.ld;
  const int sub = tid & (a.copies - 1);

  for (int64_t base = r0 + (int64_t)tid * 4; base < r1; base += (int64_t)kHistThreads * 4) {
    if (base + 4 <= r1) {
      float4 xv = __ldcs(reinterpret_cast<const float4*>(x + base));  // streamed once: evict-first
      int4 nv = __ldg(reinterpret_cast<const int4*>(a.nid + base));
      int4 yv = __ldg(reinterpret_cast<const int4*>(a.wyq + base));
      int4 zv = __ldg(reinterpret_cast<const int4*>(a.wyyq + base));
      int4 wv = make_int4(1, 1, 1, 1);
      if (WEIGHTED) wv = __ldg(reinterpret_cast<const int4*>(a.wq + base));
      accumulate_row<WEIGHTED>(a, col, xv.x, nv.x, wv.x, yv.x, zv.x, sub, s_min, s_step, s_nbin, s_mn, s_mx, h0, chstride);
      accumulate_row<WEIGHTED>(a, col, xv.y, nv.y, wv.y, yv.y, zv.y, sub, s_min, s_step, s_nbin, s_mn, s_mx, h0, chstride);
      accumulate_row<WEIGHTED>(a, col, xv.z, nv.z, wv.z, yv.z, zv.z, sub, s_min, s_step, s_nbin, s_mn, s_mx, h0, chstride);
      accumulate_row<WEIGHTED>(a, col, xv.w, nv.w, wv.w, yv.w, zv.w, sub, s_min, s_step, s_nbin, s_mn, s_mx, h0, chstride);
    } else {
      for (int64_t r = base; r < r1; r++)
        accumulate_row<WEIGHTED>(a, col, x[r], a.nid[r], WEIGHTED ? a.wq[r] : 1, a.wyq[r], a.wyyq[r], sub, s_min, s_step,
                                 s_nbin, s_mn, s_mx, h0, chstride);
    }
  }
  __syncthreads();

  // flush: low words of every copy (zero-extended) → global 64-bit accumulators; lane-rotated copy order keeps the
  // shared reads conflict-free when copies == 32
  const int cells = a.node_cnt * a.nb1;
  const int lane = tid & 31;
  for (int i = tid; i < cells; i += kHistThreads) {
    int g = i / a.nb1, b = i - g * a.nb1;
    unsigned long long* gs =
        reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)(a.node_base + g) * a.cols + col) * a.nb1 + b) * 3;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
      const int* h = h0 + ch * chstride + i * a.copies;
      unsigned long long s = 0;
      for (int k = 0; k < a.copies; k++) s += (unsigned)h[(k + lane) & (a.copies - 1)];
      if (s) atomicAdd(gs + ch, s);
    }
  }
  for (int g = tid; g < a.node_cnt; g += kHistThreads) {
    int64_t nc = (int64_t)(a.node_base + g) * a.cols + col;
    if (s_mn[g] != kEncPosInf) atomicMin(&a.min2_enc[nc], s_mn[g]);
    if (s_mx[g] != kEncNegInf) atomicMax(&a.maxin_enc[nc], s_mx[g]);
  }
}

__global__ void finalize_kernel(const int64_t* __restrict__ acc, int64_t n, double inv_w, double inv_wy, double inv_wyy,
                                double* __restrict__ hist) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
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
    synth_x_kernel<<<grid, 256, 0, s>>>(shape, seed, row0, rows, cols, ld, na_permille, X);
  }
  if (y || w) synth_yw_kernel<<<blocks_for(rows, 256), 256, 0, s>>>(seed, row0, rows, cols, weight_mode, y, w);
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
  quantize_rows_kernel<<<blocks_for(rows, 256), 256, 0, s>>>(y, w, rows, q, wq, wyq, wyyq);
  return cudaGetLastError();
}

cudaError_t launch_bin(const float* x, int64_t n, const h2ok_col_desc& d, int32_t* out, cudaStream_t s) {
  bin_kernel<<<blocks_for(n, 256), 256, 0, s>>>(x, n, d, out);
  return cudaGetLastError();
}

size_t hist_smem_bytes(int node_cnt, int nb1, int copies) {
  return (size_t)node_cnt * (2 * sizeof(double) + 3 * sizeof(int)) + (size_t)3 * node_cnt * nb1 * copies * sizeof(int);
}

// Chooses the node-group size and the number of privatised copies so that one CTA's histogram fits the
// 227 KB of shared memory an sm_100a CTA may opt into.
HistPlan plan_hist(int n_nodes, int cols, int64_t rows, int nb1) {
  HistPlan p;
  const size_t budget = 227 * 1024;
  const size_t slot = (size_t)nb1 * 3 * sizeof(int);
  int slots = (int)((budget - 64 * 28) / slot);  // (node,copy) slots that fit
  int pow2 = 1;
  while (pow2 * 2 <= slots) pow2 *= 2;
  slots = pow2;  // 64 at 256 bins
  p.group = n_nodes < slots ? n_nodes : slots;
  p.passes = (n_nodes + p.group - 1) / p.group;
  int copies = 1;
  while (copies * 2 <= 32 && copies * 2 * p.group <= slots) copies *= 2;
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
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(hist_build_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(hist_build_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
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
    unsigned grid = (unsigned)(a.cols * plan.chunks);
    if (weighted)
      hist_build_kernel<true><<<grid, kHistThreads, smem, s>>>(a);
    else
      hist_build_kernel<false><<<grid, kHistThreads, smem, s>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_finalize(const int64_t* acc, const uint32_t* mn, const uint32_t* mx, int64_t n_node_cols, int nb1,
                            const h2ok_quant& q, double* hist, float* min2, float* maxin, cudaStream_t s) {
  int64_t n = n_node_cols * nb1 * 3;
  if (hist) finalize_kernel<<<blocks_for(n, 256), 256, 0, s>>>(acc, n, 1.0 / q.scale_w, 1.0 / q.scale_wy, 1.0 / q.scale_wyy, hist);
  if (min2) decode_ranges_kernel<<<blocks_for(n_node_cols, 256), 256, 0, s>>>(mn, mx, n_node_cols, min2, maxin);
  return cudaGetLastError();
}

}  // namespace h2ok
