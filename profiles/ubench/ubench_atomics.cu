// ubench_atomics.cu — microbenchmarks that decide the histogram kernel's design on B200 (sm_100a):
// shared-memory atomic throughput under the address patterns the kernel produces, L2 (global) reductions,
// and the fp64 binning arithmetic.  Standalone: nvcc -gencode arch=compute_100a,code=sm_100a ubench_atomics.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

enum Pat { CONFLICT_FREE, RAND256, RAND16K, THREE, SAME, COPIES };

template <int PAT, int COPIES_N, bool RET, bool CASF>
__global__ void __launch_bounds__(1024, 1) smem_atomics(unsigned* out, int iters, uint32_t seed) {
  extern __shared__ unsigned h[];
  const int words = 50 * 1024;
  for (int i = threadIdx.x; i < words; i += blockDim.x) h[i] = 0;
  __syncthreads();
  uint32_t s = seed ^ (blockIdx.x * 1024 + threadIdx.x) * 2654435761u;
  const int lane = threadIdx.x & 31;
  unsigned accum = 0;
  for (int it = 0; it < iters; it++) {
    uint32_t r = lcg(s);
    int idx;
    if (PAT == CONFLICT_FREE) idx = lane + 32 * (r & 255);
    else if (PAT == RAND256) idx = r & 255;
    else if (PAT == RAND16K) idx = r % 16448;
    else if (PAT == THREE) idx = (r % 10 < 5) ? 0 : ((r % 10 < 8) ? 1 : 2);
    else if (PAT == SAME) idx = 7;
    else idx = (r % (16448 / COPIES_N > 256 ? 257 * (64 / COPIES_N) : 257)) * COPIES_N + (lane & (COPIES_N - 1));
    if (CASF) {
      atomicAdd(reinterpret_cast<float*>(h) + idx, 1.0f);
    } else if (RET) {
      unsigned old = atomicAdd(&h[idx], r | 1u);
      accum += (old + (r | 1u) < old);
    } else {
      atomicAdd(&h[idx], 1u);
    }
  }
  __syncthreads();
  unsigned t = accum;
  for (int i = threadIdx.x; i < words; i += blockDim.x) t += h[i];
  if (t == 0xdeadbeef) out[0] = t;
}

// three channels per "cell" like the real kernel: idx, idx+stride, idx+2*stride
template <int COPIES_N, int NODES>
__global__ void __launch_bounds__(1024, 1) smem_cell3(unsigned* out, int iters, uint32_t seed) {
  extern __shared__ unsigned h[];
  const int words = 50 * 1024;
  for (int i = threadIdx.x; i < words; i += blockDim.x) h[i] = 0;
  __syncthreads();
  uint32_t s = seed ^ (blockIdx.x * 1024 + threadIdx.x) * 2654435761u;
  const int sub = threadIdx.x & (COPIES_N - 1);
  const int stride = NODES * 257 * COPIES_N;
  unsigned accum = 0;
  for (int it = 0; it < iters; it++) {
    uint32_t r = lcg(s);
    int cell = (r >> 8) % NODES * 257 + (r & 255);
    int idx = cell * COPIES_N + sub;
    atomicAdd(&h[idx], 1u);
    unsigned v = r | 1u;
    unsigned o1 = atomicAdd(&h[stride + idx], v);
    unsigned o2 = atomicAdd(&h[2 * stride + idx], v >> 3);
    accum += (o1 + v < v) + (o2 + (v >> 3) < (v >> 3));
  }
  __syncthreads();
  unsigned t = accum;
  for (int i = threadIdx.x; i < words; i += blockDim.x) t += h[i];
  if (t == 0xdeadbeef) out[0] = t;
}

template <int MODE>  // 0: u64 red, 1: f32x4 red, 2: f32 red
__global__ void __launch_bounds__(1024, 1) gmem_red(void* buf, uint32_t entries, int iters, uint32_t seed) {
  uint32_t s = seed ^ (blockIdx.x * 1024 + threadIdx.x) * 2654435761u;
  for (int it = 0; it < iters; it++) {
    uint32_t r = (lcg(s) ^ (lcg(s) << 12)) % entries;
    if (MODE == 0) atomicAdd(reinterpret_cast<unsigned long long*>(buf) + r, 1ull);
    else if (MODE == 2) atomicAdd(reinterpret_cast<float*>(buf) + r, 1.0f);
    else {
      float* p = reinterpret_cast<float*>(buf) + 4 * (size_t)r;
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(0.f) : "memory");
    }
  }
}

__global__ void __launch_bounds__(1024, 1) bin_math(int* out, int iters, uint32_t seed, double mn, double step) {
  uint32_t s = seed ^ (blockIdx.x * 1024 + threadIdx.x) * 2654435761u;
  int acc = 0;
  for (int it = 0; it < iters; it++) {
    float x = __uint_as_float(0x3f800000u | (lcg(s) & 0x7fffff));
    double pos = __dmul_rn(__dsub_rn((double)x, mn), step);
    int idx = __double2int_rz(pos);
    idx = pos < 0.0 ? 0 : idx;
    acc += idx >= 256 ? 255 : idx;
  }
  if (acc == 0x7fffffff) out[0] = acc;
}

template <typename F>
static float time_ms(F f, int reps = 3) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  f();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int i = 0; i < reps; i++) {
    cudaEventRecord(a);
    f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  return best;
}

#define RUN_SMEM(name, ...)                                                                              \
  {                                                                                                      \
    cudaFuncSetAttribute(__VA_ARGS__, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);                \
    float ms = time_ms([&] { __VA_ARGS__<<<sms, 1024, smem>>>(d_out, iters, 12345u); });                 \
    double ops = (double)sms * 1024 * iters;                                                             \
    printf("%-34s %8.3f ms  %8.1f Gops/s  %6.2f lane-ops/clk/SM @%d MHz\n", name, ms, ops / ms * 1e-6,   \
           ops / ms * 1e-3 / sms / (clk_khz * 1e-3), clk_khz / 1000);                                    \
  }

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  int sms = prop.multiProcessorCount, clk_khz = prop.clockRate;
  printf("device %s, %d SMs, clockRate %d MHz (nominal max; per-clk figures assume it)\n", prop.name, sms, clk_khz / 1000);
  unsigned* d_out;
  CK(cudaMalloc(&d_out, 1024));
  const int smem = 50 * 1024 * 4, iters = 4096;
  RUN_SMEM("u32 add conflict-free", smem_atomics<CONFLICT_FREE, 1, false, false>);
  RUN_SMEM("u32 add random 256 bins", smem_atomics<RAND256, 1, false, false>);
  RUN_SMEM("u32 add random 16448 entries", smem_atomics<RAND16K, 1, false, false>);
  RUN_SMEM("u32 add 3-valued", smem_atomics<THREE, 1, false, false>);
  RUN_SMEM("u32 add same address", smem_atomics<SAME, 1, false, false>);
  RUN_SMEM("u32 add 32 copies (1 node)", smem_atomics<COPIES, 32, false, false>);
  RUN_SMEM("u32 add 16 copies (4 nodes)", smem_atomics<COPIES, 16, false, false>);
  RUN_SMEM("u32 add 8 copies (8 nodes)", smem_atomics<COPIES, 8, false, false>);
  RUN_SMEM("u32 add 4 copies (16 nodes)", smem_atomics<COPIES, 4, false, false>);
  RUN_SMEM("u32 add 2 copies (32 nodes)", smem_atomics<COPIES, 2, false, false>);
  RUN_SMEM("u32 add+return conflict-free", smem_atomics<CONFLICT_FREE, 1, true, false>);
  RUN_SMEM("u32 add+return random 16448", smem_atomics<RAND16K, 1, true, false>);
  RUN_SMEM("f32 CAS add conflict-free", smem_atomics<CONFLICT_FREE, 1, false, true>);
  RUN_SMEM("f32 CAS add random 256", smem_atomics<RAND256, 1, false, true>);
  RUN_SMEM("f32 CAS add random 16448", smem_atomics<RAND16K, 1, false, true>);
  printf("-- 3 atomics per cell (count, wy+ret, wyy+ret): ops below are CELLS\n");
  RUN_SMEM("cell3 32 copies x 1 node", smem_cell3<32, 1>);
  RUN_SMEM("cell3 32 copies x 2 nodes", smem_cell3<32, 2>);
  RUN_SMEM("cell3 16 copies x 4 nodes", smem_cell3<16, 4>);
  RUN_SMEM("cell3 8 copies x 8 nodes", smem_cell3<8, 8>);
  RUN_SMEM("cell3 4 copies x 16 nodes", smem_cell3<4, 16>);
  RUN_SMEM("cell3 2 copies x 32 nodes", smem_cell3<2, 32>);
  RUN_SMEM("cell3 1 copy x 64 nodes", smem_cell3<1, 64>);
  {
    void* buf;
    size_t bytes = (size_t)64 << 20;
    CK(cudaMalloc(&buf, bytes));
    CK(cudaMemset(buf, 0, bytes));
    const int it2 = 512;
    struct { const char* n; int mode; uint32_t entries; } cs[] = {
        {"REDG u64 random over 50 MB", 0, 6316032}, {"REDG u64 random over 1 MB", 0, 131072},
        {"REDG f32x4 random over 64 MB", 1, 4194304}, {"REDG f32x4 random over 1 MB", 1, 65536},
        {"REDG f32 random over 25 MB", 2, 6316032}};
    for (auto& c : cs) {
      float ms = time_ms([&] {
        if (c.mode == 0) gmem_red<0><<<sms * 2, 1024>>>(buf, c.entries, it2, 99u);
        else if (c.mode == 1) gmem_red<1><<<sms * 2, 1024>>>(buf, c.entries, it2, 99u);
        else gmem_red<2><<<sms * 2, 1024>>>(buf, c.entries, it2, 99u);
      });
      double ops = (double)sms * 2 * 1024 * it2;
      printf("%-34s %8.3f ms  %8.1f Gops/s  %6.2f lane-ops/clk/SM\n", c.n, ms, ops / ms * 1e-6,
             ops / ms * 1e-3 / sms / (clk_khz * 1e-3));
    }
    cudaFree(buf);
  }
  {
    float ms = time_ms([&] { bin_math<<<sms * 2, 1024>>>((int*)d_out, 8192, 5u, 1.0, 256.0); });
    double ops = (double)sms * 2 * 1024 * 8192;
    printf("%-34s %8.3f ms  %8.1f Gcells/s  %6.2f cells/clk/SM\n", "fp64 bin math (cvt,sub,mul,cvt)", ms, ops / ms * 1e-6,
           ops / ms * 1e-3 / sms / (clk_khz * 1e-3));
  }
  CK(cudaDeviceSynchronize());
  printf("done\n");
  return 0;
}
