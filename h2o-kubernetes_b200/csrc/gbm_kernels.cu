// gbm_kernels.cu — SURVEY.md §8 row a9, the caller of the histogram path: GBM.buildNextKTrees' residual pass
// (ComputePredAndRes), GammaPass and AddTreeContributions for one-class trees (gaussian, bernoulli) and the K-class
// residual pass (multinomial: K one-class trees per iteration on softmax residuals).
// [H2O-3 hex/tree/gbm/GBM.java, recalled; not under /root/reference]
//
// Leaf values are ratios of sums over a leaf's rows.  The sums are taken over the same fixed-point integers the
// histograms use (numerator: wyq; denominator: a quantised w*mu*(1-mu), or the weight / row count for gaussian), so
// they are exact, order-independent and combine across ranks with an integer all-reduce.
#include "common.cuh"
#include "kernels.h"

namespace h2ok {

namespace {

__global__ void gbm_residual_kernel(int dist, const float* __restrict__ label, const float* __restrict__ f,
                                    const float* __restrict__ w, int64_t rows, double den_scale,
                                    float* __restrict__ res, int32_t* __restrict__ denq) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float fi = f[r];
  if (dist == H2OK_DIST_BERNOULLI) {
    // mu = 1/(1+exp(-f)) in fp32 (device expf: agrees with a host libm to a few ulp, not bit for bit)
    const float mu = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-fi)));
    res[r] = __fsub_rn(label[r], mu);
    const double wd = w ? (double)w[r] : 1.0;
    const double v = __dmul_rn(__dmul_rn(wd, (double)mu), (double)__fsub_rn(1.0f, mu));
    if (denq) denq[r] = __double2int_rn(__dmul_rn(v, den_scale));
  } else {
    res[r] = __fsub_rn(label[r], fi);  // gaussian: exact on host and device alike
    if (denq) denq[r] = w ? __double2int_rn(__dmul_rn((double)w[r], den_scale)) : 1;
  }
}

// K-class residuals: one thread per row walks the K scores three times (max, exp + sum, residual); K is small.
__global__ void gbm_residual_multi_kernel(int K, const float* __restrict__ label, const float* __restrict__ f, int64_t ld,
                                          const float* __restrict__ w, int64_t rows, double den_scale,
                                          float* __restrict__ res, int32_t* __restrict__ denq) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  float m = f[r];
  for (int k = 1; k < K; k++) m = fmaxf(m, f[(int64_t)k * ld + r]);
  float s = 0.0f;
  for (int k = 0; k < K; k++) s = __fadd_rn(s, expf(__fsub_rn(f[(int64_t)k * ld + r], m)));
  const int cls = (int)label[r];
  const double wd = w ? (double)w[r] : 1.0;
  for (int k = 0; k < K; k++) {
    const float p = __fdiv_rn(expf(__fsub_rn(f[(int64_t)k * ld + r], m)), s);
    const float rk = __fsub_rn(k == cls ? 1.0f : 0.0f, p);
    res[(int64_t)k * ld + r] = rk;
    const float a = fabsf(rk);
    const double v = __dmul_rn(__dmul_rn(wd, (double)a), (double)__fsub_rn(1.0f, a));
    denq[(int64_t)k * ld + r] = __double2int_rn(__dmul_rn(v, den_scale));
  }
}

// GammaPass: per leaf code, sum of numq and denq over the rows that ended in that leaf.  Rows of one warp that share
// a leaf are combined first (match + shuffle reduction), one 64-bit global add per distinct leaf and warp.
__global__ void gbm_gamma_pass_kernel(const int32_t* __restrict__ leaf, const int32_t* __restrict__ numq,
                                      const int32_t* __restrict__ denq, int64_t rows, int n_codes,
                                      int64_t* __restrict__ num_acc, int64_t* __restrict__ den_acc) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int code = -1;
  long long n = 0, d = 0;
  if (r < rows) {
    code = leaf[r];
    if (code < 0 || code >= n_codes) code = -1;
    else {
      n = numq[r];
      d = denq ? denq[r] : 1;
    }
  }
  const unsigned peers = __match_any_sync(0xffffffffu, code);
  const unsigned lane = threadIdx.x & 31u;
  const int leader = __ffs(peers) - 1;
  // reduce within the peer group: every member walks the group's lanes (at most 32 steps, usually 1-3 groups)
  long long sn = 0, sd = 0;
  for (unsigned m = peers; m; m &= m - 1) {
    const int src = __ffs(m) - 1;
    sn += __shfl_sync(peers, n, src);
    sd += __shfl_sync(peers, d, src);
  }
  if (code >= 0 && (int)lane == leader) {
    atomicAdd(reinterpret_cast<unsigned long long*>(num_acc) + code, (unsigned long long)sn);
    atomicAdd(reinterpret_cast<unsigned long long*>(den_acc) + code, (unsigned long long)sd);
  }
}

// AddTreeContributions: f += (float)(learn_rate * gamma[leaf])
__global__ void gbm_add_gamma_kernel(const int32_t* __restrict__ leaf, int64_t rows, const double* __restrict__ gamma,
                                     int n_codes, double learn_rate, float* __restrict__ f) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const int code = leaf[r];
  if (code < 0 || code >= n_codes) return;
  f[r] = __fadd_rn(f[r], (float)__dmul_rn(learn_rate, gamma[code]));
}

}  // namespace

cudaError_t launch_gbm_residual(int dist, const float* label, const float* f, const float* w, int64_t rows,
                                double den_scale, float* res, int32_t* denq, cudaStream_t s) {
  gbm_residual_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(dist, label, f, w, rows, den_scale, res, denq);
  return cudaGetLastError();
}

cudaError_t launch_gbm_residual_multi(int K, const float* label, const float* f, int64_t ld, const float* w, int64_t rows,
                                      double den_scale, float* res, int32_t* denq, cudaStream_t s) {
  gbm_residual_multi_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(K, label, f, ld, w, rows, den_scale, res, denq);
  return cudaGetLastError();
}

cudaError_t launch_gbm_gamma_pass(const int32_t* leaf, const int32_t* numq, const int32_t* denq, int64_t rows, int n_codes,
                                  int64_t* num_acc, int64_t* den_acc, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(num_acc, 0, sizeof(int64_t) * (size_t)n_codes, s);
  if (e == cudaSuccess) e = cudaMemsetAsync(den_acc, 0, sizeof(int64_t) * (size_t)n_codes, s);
  if (e != cudaSuccess) return e;
  gbm_gamma_pass_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(leaf, numq, denq, rows, n_codes, num_acc, den_acc);
  return cudaGetLastError();
}

cudaError_t launch_gbm_add_gamma(const int32_t* leaf, int64_t rows, const double* gamma, int n_codes, double learn_rate,
                                 float* f, cudaStream_t s) {
  gbm_add_gamma_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(leaf, rows, gamma, n_codes, learn_rate, f);
  return cudaGetLastError();
}

}  // namespace h2ok
