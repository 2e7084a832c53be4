// common.cuh — shared device/host helpers for the sm_100a GBM histogram path.
// Semantics follow the normative CPU oracle (oracle/hist_oracle.c); nothing here includes or links it.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/h2ok_hist.h"

namespace h2ok {

constexpr int kSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// ---- order-preserving float <-> uint32 (observed min/max travel as integers so that blocks, ranks and NCCL can
//      combine them with integer min/max) ---------------------------------------------------------------------
__host__ __device__ inline uint32_t enc_f32_bits(uint32_t b) { return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__host__ __device__ inline uint32_t dec_f32_bits(uint32_t e) { return (e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e; }
constexpr uint32_t kEncPosInf = 0xFF800000u;  // enc(+inf): "nothing observed" for a running min
constexpr uint32_t kEncNegInf = 0x007FFFFFu;  // enc(-inf): "nothing observed" for a running max

// ---- generator: integer hash + individually rounded fp32 ops, no libm ---------------------
__host__ __device__ inline uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
__host__ __device__ inline uint64_t cell_hash(uint64_t seed, uint64_t col, uint64_t row) {
  return splitmix64(seed ^ (col * 0x9E3779B97F4A7C15ULL) ^ row);
}

#ifdef __CUDACC__
__device__ __forceinline__ float u24(uint64_t h) { return __fmul_rn((float)(uint32_t)(h >> 40), 0x1p-24f); }

__device__ inline float base28(int j, float u) {
  if (j == 8 || j == 12 || j == 16 || j == 20) return u < 0.5f ? 0.0f : (u < 0.8f ? 1.0865f : 2.173f);
  if (j >= 21) {
    float u2 = __fmul_rn(u, u);
    float u3 = __fmul_rn(u2, u);
    return __fadd_rn(0.3f, __fmul_rn(3.0f, u3));
  }
  switch (j % 3) {
    case 0:
      return __fadd_rn(0.25f, __fmul_rn(4.0f, __fmul_rn(u, u)));
    case 1:
      return __fsub_rn(__fmul_rn(5.0f, u), 2.5f);
    default:
      return __fsub_rn(__fmul_rn(6.28318548f, u), 3.14159274f);
  }
}

__device__ inline float gen_x(int shape, uint64_t seed, int64_t row, int col, int na_permille) {
  uint64_t h = cell_hash(seed, (uint64_t)col, (uint64_t)row);
  if (na_permille > 0) {
    uint64_t h2 = cell_hash(seed ^ 0xA5A5A5A5A5A5A5A5ULL, (uint64_t)col, (uint64_t)row);
    if ((int)(h2 % 1000ULL) < na_permille) return __int_as_float(0x7FC00000);
  }
  float u = u24(h);
  if (shape == H2OK_SHAPE_SPARSE) return ((h & 0xFFFFULL) < 58982ULL) ? 0.0f : u;
  float b = base28(col % 28, u);
  if (shape == H2OK_SHAPE_WIDE64) {
    float s = __fadd_rn(1.0f, __fdiv_rn((float)col, 64.0f));
    return __fadd_rn(__fmul_rn(b, s), (float)col);
  }
  return b;
}

__device__ inline float gen_y(uint64_t seed, int64_t row, int cols) {
  float uy = u24(cell_hash(seed + 1, 0, (uint64_t)row));
  float u0 = u24(cell_hash(seed, (uint64_t)(0 % cols), (uint64_t)row));
  float u1 = u24(cell_hash(seed, (uint64_t)(1 % cols), (uint64_t)row));
  float u3 = u24(cell_hash(seed, (uint64_t)(3 % cols), (uint64_t)row));
  float u5 = u24(cell_hash(seed, (uint64_t)(5 % cols), (uint64_t)row));
  float noise = __fmul_rn(0.5f, __fsub_rn(__fmul_rn(2.0f, uy), 1.0f));
  float a = __fsub_rn(u0, 0.5f);
  float b = __fsub_rn(u3, 0.5f);
  if (u5 < 0.5f) b = -b;
  float c = u1 > 0.7f ? 0.5f : 0.0f;
  float s = __fadd_rn(noise, a);
  s = __fadd_rn(s, b);
  return __fadd_rn(s, c);
}

__device__ inline float gen_w(uint64_t seed, int64_t row, int weight_mode) {
  if (weight_mode == 0) return 1.0f;
  uint64_t h = cell_hash(seed + 2, 0, (uint64_t)row);
  if ((h & 0xFFFFULL) < 1311ULL) return 0.0f;
  return __fadd_rn(0.5f, u24(h));
}

// ---- a2: DHistogram.bin, fp64, bit-exact with oracle/hist_oracle.c:h2ok_oracle_bin ---------------------------
// (x - min) * step cannot contract into an FMA; the *_rn intrinsics make that explicit.
__device__ __forceinline__ int bin_of(float x, double mn, double step, int nbin) {
  if (x != x) return nbin;  // NA bucket
  double pos = __dmul_rn(__dsub_rn((double)x, mn), step);
  // ±Inf fall out of the clamps: -inf → 0, +inf → nbin-1 (H2O's explicit rule gives the same bins)
  int idx = __double2int_rz(pos);  // saturating, so huge pos cannot wrap
  idx = pos < 0.0 ? 0 : idx;
  return idx >= nbin ? nbin - 1 : idx;
}
// a2 inside the hot loops: same value as bin_of() with fewer instructions — the saturating unsigned conversion sends
// negative and NaN positions to 0 and huge ones to UINT_MAX, one unsigned min applies the round-off clamp.
__device__ __forceinline__ int bin_fast(float x, double mn, double step, int nbin) {
  const double pos = __dmul_rn(__dsub_rn((double)x, mn), step);
  const unsigned idx = min(__double2uint_rz(pos), (unsigned)(nbin - 1));
  return x != x ? nbin : (int)idx;
}

// the same for a value known not to be NaN
__device__ __forceinline__ int bin_fast_no_na(float x, double mn, double step, int nbin) {
  const double pos = __dmul_rn(__dsub_rn((double)x, mn), step);
  return (int)min(__double2uint_rz(pos), (unsigned)(nbin - 1));
}

// Packed row record (8 B): lo = wyq << 5 | node[4:0], hi = wyyq << 5 | node[9:5]; node 0x3FF = takes no part.
constexpr unsigned kNoNode = 0x3FFu;
__device__ __forceinline__ unsigned rec_node(unsigned lo, unsigned hi) { return (lo & 31u) | ((hi & 31u) << 5); }
__device__ __forceinline__ uint2 rec_pack(int wyq, int wyyq, unsigned node) {
  return make_uint2(((unsigned)wyq << 5) | (node & 31u), ((unsigned)wyyq << 5) | (node >> 5));
}

// a6 for a categorical split: the row's category code c = (int)x goes right when bit c is set; a code outside the set's
// range follows the NAs (a category no training row had)
__device__ __forceinline__ int cat_goes_right(const uint32_t* __restrict__ bits, float x, int na) {
  const int c = (int)x;
  if (c < 0 || c >= 32 * H2OK_CAT_WORDS) return na == H2OK_NA_LEFT ? 0 : 1;
  return (int)((bits[c >> 5] >> (c & 31)) & 1u);
}

// wY words start mid-range so that sums hovering around zero never wrap
constexpr unsigned kWyBias = 0x80000000u;

// Cold path: a 32-bit shared word wrapped; the carry / borrow goes to the 64-bit global accumulator.
static __device__ __noinline__ void carry_to_global(int64_t* acc, int64_t cell_index, int ch, int delta) {
  atomicAdd(reinterpret_cast<unsigned long long*>(acc) + cell_index * 3 + ch, (unsigned long long)(long long)delta << 32);
}
#endif  // __CUDACC__

}  // namespace h2ok
