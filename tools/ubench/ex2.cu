// Microbenchmark: per-SM throughput of the exponent paths a softmax can use on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ex2 ex2.cu && ./ex2
// Reports results (elements) per clock per SM for: ex2.approx.ftz.f32, ex2.approx.ftz.bf16x2, ex2.approx.ftz.f16x2,
// cvt.rn.bf16x2.f32 (pack), and a degree-3 polynomial exp2 on the FMA pipe.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int ILP = 8;

template <int MODE>
__global__ void k(uint32_t* out, float seed) {
  float x[ILP];
  uint32_t h[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) { x[i] = seed * (threadIdx.x + i) * 1e-3f; h[i] = __float_as_uint(x[i]) | 0x3c003c00u; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
      if (MODE == 1) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 2) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 3) {   // F2FP pack with a data dependence (h feeds back into x) so it cannot be hoisted
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(x[i]), "f"(x[(i + 1) % ILP]));
        x[i] = __uint_as_float(h[i]);
      }
      if (MODE == 6) {   // pack by truncation: PRMT of the two high halves
        asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(h[i]) : "r"(__float_as_uint(x[i])), "r"(__float_as_uint(x[(i + 1) % ILP])));
        x[i] = __uint_as_float(h[i]);
      }
      if (MODE == 7) {   // 3-input max
        x[i] = fmaxf(fmaxf(x[i], x[(i + 1) % ILP]), x[(i + 2) % ILP]);
      }
      if (MODE == 8) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x[i]));
      if (MODE == 4) {  // Cody-Waite + cubic on the FMA/ALU pipes (one element)
        float t = x[i];
        float fl = floorf(t);
        float f = t - fl;
        float p = fmaf(f, 0.0555041f, 0.2402265f);
        p = fmaf(p, f, 0.6931472f);
        p = fmaf(p, f, 1.0f);
        x[i] = __uint_as_float(__float_as_uint(p) + ((int)fl << 23));
      }
      if (MODE == 5) {  // fma + pack + bf16x2 ex2: the proposed softmax inner step for TWO elements
        float a = fmaf(x[i], 1.0001f, -0.5f), b = fmaf(x[(i + 1) % ILP], 1.0001f, -0.5f);
        uint32_t pk;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk) : "f"(a), "f"(b));
        asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(pk));
        h[i] ^= pk;
      }
    }
  }
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) acc = (acc * 31u + h[i]) * 31u + __float_as_uint(x[i]);   // (an XOR would cancel h against x in the pack modes)
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
void run(const char* name, int elems_per_op, uint32_t* out, int sms, int khz) {
  const int blocks = sms * 4, threads = 512;
  k<MODE><<<blocks, threads>>>(out, 1.0f);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int r = 0; r < 5; ++r) k<MODE><<<blocks, threads>>>(out, 1.0f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double ops = 5.0 * blocks * threads * (double)ITERS * ILP * elems_per_op;
  printf("%-34s %8.3f ms  %8.2f Gelem/s  %6.2f elem/clk/SM at %d MHz (nominal max clock)\n", name, ms / 5,
         ops / (ms * 1e-3) / 1e9, ops / (ms * 1e-3) / sms / (khz * 1e3), khz / 1000);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  uint32_t* out; cudaMalloc(&out, p.multiProcessorCount * 4 * 512 * 4);
  run<0>("ex2.approx.ftz.f32", 1, out, p.multiProcessorCount, khz);
  run<1>("ex2.approx.ftz.bf16x2", 2, out, p.multiProcessorCount, khz);
  run<2>("ex2.approx.f16x2", 2, out, p.multiProcessorCount, khz);
  run<3>("cvt.rn.bf16x2.f32 (pack 2)", 2, out, p.multiProcessorCount, khz);
  run<4>("poly3 exp2 (FMA pipe)", 1, out, p.multiProcessorCount, khz);
  run<5>("2x fma + pack + ex2.bf16x2", 2, out, p.multiProcessorCount, khz);
  run<6>("prmt pack by truncation (pack 2)", 2, out, p.multiProcessorCount, khz);
  run<7>("fmax3", 2, out, p.multiProcessorCount, khz);
  run<8>("tanh.approx.f32", 1, out, p.multiProcessorCount, khz);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
