// Microbenchmark of the softmax exponent phase: each thread holds 128 scores in registers and runs
// x = fma(s, sc, -m); p = ex2(x); sum += p (8 chains), the way attn_fa.cu does, with W warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o expphase expphase.cu && ./expphase
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <bool SUM, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k(float* out, float sc, float m, int iters, long long* cyc) {
  float s[128];
#pragma unroll
  for (int i = 0; i < 128; ++i) s[i] = (float)((threadIdx.x * 131 + i * 17) & 255) * -0.01f;
  float sum8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 128; ++i) {
      const float x = fmaf(s[i], sc, m);
      s[i] = ex2(x);
      if (SUM) sum8[i & 7] += s[i];
    }
    // feed the results back so the next iteration depends on this one (like a new score tile would not, but it
    // keeps the compiler from hoisting); scale to stay in range
#pragma unroll
    for (int i = 0; i < 128; ++i) s[i] = s[i] * -4.0f;
  }
  const long long t1 = clock64();
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 128; ++i) acc += s[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc += sum8[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  float* out; cudaMalloc(&out, p.multiProcessorCount * 512 * 4);
  long long* cyc; cudaMallocManaged(&cyc, 8);
  const int iters = 200;
#define RUN(W)                                                                                                   \
  for (int sum = 0; sum < 2; ++sum) {                                                                            \
    if (sum) k<true, 128 * W><<<p.multiProcessorCount, 128 * W>>>(out, 0.9f, -0.3f, iters, cyc);                  \
    else k<false, 128 * W><<<p.multiProcessorCount, 128 * W>>>(out, 0.9f, -0.3f, iters, cyc);                     \
    cudaDeviceSynchronize();                                                                                     \
    printf("%d warp(s) per SM sub-partition, row sum %s: %.0f cycles per 128-element row phase (incl. 128 FMUL feedback)\n", \
           W, sum ? "on " : "off", (double)*cyc / iters);                                                        \
  }
  RUN(1) RUN(2) RUN(3)
  printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
