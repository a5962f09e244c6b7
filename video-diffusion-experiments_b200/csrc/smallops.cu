// Small kernels at the edges of the UNet: timestep sinusoid, small-M Linear (time MLP, temb projections),
// conv_in (4 -> C0 straight from the [B,C,F,h,w] latents), conv_out (C0 -> 4 back to [B,C,F,h,w]) and the
// fused classifier-free-guidance + scheduler update.
#include "../../include/b200ad.h"
#include "ops_common.cuh"

namespace b200ad {

// Timesteps(flip_sin_to_cos=True, freq_shift=0): out[b] = [cos(t*f_i) | sin(t*f_i)], f_i = exp(-ln(1e4)*i/half)
__global__ void timestep_sinusoid_kernel(const float* __restrict__ t, int B, int dim, float* __restrict__ out) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int half = dim >> 1;
  if (i >= B * half) return;
  const int b = i / half, k = i % half;
  const float freq = expf(-9.210340371976184f * (float)k / (float)half);
  const float e = t[b] * freq;
  out[b * dim + k] = cosf(e);
  out[b * dim + half + k] = sinf(e);
}

// y[M,N] = act_out(W[N,K] x[M,K]^T + b); one warp per output column, M <= 8 rows kept in registers.
__global__ void small_linear_kernel(const float* __restrict__ x, int M, int K, const float* __restrict__ W,
                                    const float* __restrict__ bias, int N, int silu_in, int silu_out,
                                    float* __restrict__ y) {
  pdl_enter();
  const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float acc[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) acc[m] = 0.f;
  const float* wr = W + (long long)n * K;
  for (int k = lane * 4; k < K; k += 128) {
    const float4 w4 = *reinterpret_cast<const float4*>(wr + k);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      if (m < M) {
        float4 x4 = *reinterpret_cast<const float4*>(x + (long long)m * K + k);
        if (silu_in) { x4.x = silu_f(x4.x); x4.y = silu_f(x4.y); x4.z = silu_f(x4.z); x4.w = silu_f(x4.w); }
        acc[m] += w4.x * x4.x + w4.y * x4.y + w4.z * x4.z + w4.w * x4.w;
      }
    }
  }
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    if (m < M) {
      float v = warp_sum(acc[m]);
      if (lane == 0) {
        v += bias ? bias[n] : 0.f;
        if (silu_out) v = silu_f(v);
        y[(long long)m * N + n] = v;
      }
    }
  }
}

// conv_in: 3x3, pad 1, tiny cin.  Block = PIX consecutive output pixels of one frame-image x all cout
// (thread = cout, weights for that cout in registers, input patch shared through smem).
static constexpr int CIN_PIX = 16;
static constexpr int CIN_MAXK = 72;     // cin*9 <= 72 (cin <= 8)
// ones_ch = 1: one extra input channel (index cin) that is 1 inside the image and 0 in the padding (b200ad_conv_in_affine)
__global__ void conv_in_kernel(const float* __restrict__ sample, int B, int cin, int F, int h, int w,
                               const float* __restrict__ W, const float* __restrict__ bias, int cout,
                               float* __restrict__ out, int ones_ch) {
  pdl_enter();
  __shared__ float patch[CIN_PIX][CIN_MAXK];
  const int kk = (cin + ones_ch) * 9;
  const long long pix0 = (long long)blockIdx.x * CIN_PIX;
  const long long npix = (long long)B * F * h * w;
  for (int i = threadIdx.x; i < CIN_PIX * kk; i += blockDim.x) {
    const int p = i / kk, k = i % kk;
    const long long pix = pix0 + p;
    float v = 0.f;
    if (pix < npix) {
      const int x = (int)(pix % w), y = (int)((pix / w) % h);
      const int n = (int)(pix / ((long long)w * h));
      const int b = n / F, f = n % F;
      const int c = k / 9, ky = (k % 9) / 3, kx = k % 3;
      const int yy = y + ky - 1, xx = x + kx - 1;
      if (yy >= 0 && yy < h && xx >= 0 && xx < w)
        v = c < cin ? sample[((((long long)b * cin + c) * F + f) * h + yy) * w + xx] : 1.0f;
    }
    patch[p][k] = v;
  }
  __syncthreads();
  for (int co = threadIdx.x; co < cout; co += blockDim.x) {
    float wreg[CIN_MAXK];
#pragma unroll
    for (int k = 0; k < CIN_MAXK; ++k) wreg[k] = (k < kk) ? W[(long long)co * kk + k] : 0.f;
    const float bv = bias ? bias[co] : 0.f;
    for (int p = 0; p < CIN_PIX; ++p) {
      if (pix0 + p >= npix) break;
      float acc = bv;
#pragma unroll
      for (int k = 0; k < CIN_MAXK; ++k)
        if (k < kk) acc += wreg[k] * patch[p][k];
      out[(pix0 + p) * cout + co] = acc;
    }
  }
}

// conv_in, compile-time cin (the UNet's cin = 4): KK = cin*9 weights per output channel live in registers, the
// input patch of PIX pixels is staged once in smem as KK/4 float4 per pixel and read with broadcast LDS.128, so the
// inner loop is 1 LDS.128 per 4 FFMA (FMA-bound, ~1.5 GFMA at 16f 512^2) and every store is a coalesced 4-byte
// lane-contiguous write of the NHWC row.
template <int CIN, int PIX>
__global__ void __launch_bounds__(320) conv_in_fast_kernel(const float* __restrict__ sample, int B, int F, int h,
                                                           int w, const float* __restrict__ W,
                                                           const float* __restrict__ bias, int cout,
                                                           float* __restrict__ out) {
  pdl_enter();
  constexpr int KK = CIN * 9;
  static_assert(KK % 4 == 0, "patch rows are read as float4");
  __shared__ __align__(16) float patch[PIX][KK];
  const long long pix0 = (long long)blockIdx.x * PIX;
  const long long npix = (long long)B * F * h * w;
  for (int i = threadIdx.x; i < PIX * KK; i += blockDim.x) {
    const int p = i / KK, k = i % KK;
    const long long pix = pix0 + p;
    float v = 0.f;
    if (pix < npix) {
      const int x = (int)(pix % w), y = (int)((pix / w) % h);
      const int n = (int)(pix / ((long long)w * h));
      const int b = n / F, f = n % F;
      const int c = k / 9, ky = (k % 9) / 3, kx = k % 3;
      const int yy = y + ky - 1, xx = x + kx - 1;
      if (yy >= 0 && yy < h && xx >= 0 && xx < w)
        v = sample[((((long long)b * CIN + c) * F + f) * h + yy) * w + xx];
    }
    patch[p][k] = v;
  }
  __syncthreads();
  const int npx = (int)((npix - pix0) < PIX ? (npix - pix0) : PIX);
  for (int co = threadIdx.x; co < cout; co += blockDim.x) {
    float wreg[KK];
#pragma unroll
    for (int k = 0; k < KK; ++k) wreg[k] = W[(long long)co * KK + k];
    const float bv = bias ? bias[co] : 0.f;
    float* o = out + pix0 * cout + co;
#pragma unroll 2
    for (int p = 0; p < npx; ++p) {
      const float4* pr = reinterpret_cast<const float4*>(patch[p]);
      float a0 = bv, a1 = 0.f;
#pragma unroll
      for (int q = 0; q < KK / 4; ++q) {
        const float4 v = pr[q];
        a0 = fmaf(wreg[4 * q], v.x, a0);
        a1 = fmaf(wreg[4 * q + 1], v.y, a1);
        a0 = fmaf(wreg[4 * q + 2], v.z, a0);
        a1 = fmaf(wreg[4 * q + 3], v.w, a1);
      }
      o[(long long)p * cout] = a0 + a1;
    }
  }
}

// conv_out: 3x3, pad 1, tiny cout (<= 4).  One warp per output pixel; lanes split the channel axis.
template <typename T>
__global__ void conv_out_kernel(const T* __restrict__ act, int B, int F, int h, int w, int cin,
                                const float* __restrict__ W, const float* __restrict__ bias, int cout,
                                float* __restrict__ out) {
  pdl_enter();
  extern __shared__ float wsm[];          // [cout][9][cin]  (re-ordered from torch [cout][cin][3][3])
  for (int i = threadIdx.x; i < cout * 9 * cin; i += blockDim.x) {
    const int c = i % cin, t = (i / cin) % 9, co = i / (9 * cin);
    wsm[i] = W[((long long)co * cin + c) * 9 + t];
  }
  __syncthreads();
  const long long npix = (long long)B * F * h * w;
  const long long pix = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (pix >= npix) return;
  const int x = (int)(pix % w), y = (int)((pix / w) % h);
  const int n = (int)(pix / ((long long)w * h));
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int t = 0; t < 9; ++t) {
    const int yy = y + t / 3 - 1, xx = x + t % 3 - 1;
    if (yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
    const T* ar = act + (((long long)n * h + yy) * w + xx) * cin;
    for (int c = lane * 2; c < cin; c += 64) {
      const float2 a = load2<T>(ar + c);
#pragma unroll
      for (int co = 0; co < 4; ++co) {
        if (co < cout) {
          const float* wr = wsm + ((long long)co * 9 + t) * cin + c;
          acc[co] += a.x * wr[0] + a.y * wr[1];
        }
      }
    }
  }
  const int b = n / F, f = n % F;
#pragma unroll
  for (int co = 0; co < 4; ++co) {
    if (co < cout) {
      const float v = warp_sum(acc[co]);
      if (lane == 0) out[((((long long)b * cout + co) * F + f) * h + y) * w + x] = v + (bias ? bias[co] : 0.f);
    }
  }
}

// conv_out on warp-level tensor-core MMAs (bf16 operand mode): the 3x3 conv with cout <= 4 is the implicit GEMM
// [pixels, 9*cin] x [9*cin, 8] (couts 4..7 are zero columns).  One warp owns 16 consecutive output pixels; per filter
// tap and 32-channel block each lane fetches 16 bytes of two pixel rows (rows g and g+8 of the m16 tile) straight
// from the NHWC activations and feeds them to two m16n8k16 MMAs -- the K order inside a block is permuted so that
// one 16-byte load is the A fragment of both (lane q holds channels q*8..q*8+7: MMA s takes {4s,4s+1 | 4s+2,4s+3});
// the weights are staged once per (persistent) block in smem in exactly that order, rows padded against bank
// conflicts.  HBM-bound by construction: reads the activations once (taps hit L1/L2), writes [B,cout,F,h,w] fp32.
// weight-row length in bf16 elements: rows 16 words apart (mod 32 banks), so the 8 lanes of an LDS.128 phase
// (two rows x four 16-byte columns) cover all 32 banks once
__host__ __device__ __forceinline__ int conv_out_wrow(int cin) { return cin + (((cin / 2) % 32 == 16) ? 0 : 32); }
__device__ __forceinline__ void mma_bf16_m16n8k16(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                  uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__global__ void __launch_bounds__(256) conv_out_mma_kernel(const bf16* __restrict__ act, int B, int F, int h, int w,
                                                           int cin, const float* __restrict__ W,
                                                           const float* __restrict__ bias, int cout,
                                                           float* __restrict__ out) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t cosm[];
  bf16* wsm = reinterpret_cast<bf16*>(cosm);              // [9][8][wrow]
  const int wrow = conv_out_wrow(cin);
  for (int i = threadIdx.x; i < 9 * 8 * cin; i += blockDim.x) {
    const int c = i % cin, n = (i / cin) % 8, t = i / (8 * cin);
    const float v = n < cout ? W[((long long)n * cin + c) * 9 + t] : 0.f;
    wsm[(t * 8 + n) * wrow + c] = __float2bfloat16(v);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, q = lane & 3;
  const long long npix = (long long)B * F * h * w;
  const long long ngroups = (npix + 15) / 16;
  const long long gstride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long grp = (long long)blockIdx.x * (blockDim.x >> 5) + warp; grp < ngroups; grp += gstride) {
    long long pix[2];
    int px[2], py[2];
    long long pn[2];
    bool pv[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      pix[r] = grp * 16 + g + 8 * r;
      pv[r] = pix[r] < npix;
      const long long pp = pv[r] ? pix[r] : 0;
      px[r] = (int)(pp % w);
      py[r] = (int)((pp / w) % h);
      pn[r] = pp / ((long long)w * h);
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
    for (int t = 0; t < 9; ++t) {
      const int dy = t / 3 - 1, dx = t % 3 - 1;
      const bf16* ar[2];
      bool ok[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int yy = py[r] + dy, xx = px[r] + dx;
        ok[r] = pv[r] && yy >= 0 && yy < h && xx >= 0 && xx < w;
        ar[r] = act + ((pn[r] * h + (ok[r] ? yy : 0)) * w + (ok[r] ? xx : 0)) * cin + q * 8;
      }
      const bf16* wr = wsm + (t * 8 + g) * wrow + q * 8;
#pragma unroll 2
      for (int c0 = 0; c0 < cin; c0 += 32) {
        uint4 a_lo = make_uint4(0u, 0u, 0u, 0u), a_hi = make_uint4(0u, 0u, 0u, 0u);
        if (ok[0]) a_lo = *reinterpret_cast<const uint4*>(ar[0] + c0);
        if (ok[1]) a_hi = *reinterpret_cast<const uint4*>(ar[1] + c0);
        const uint4 bw = *reinterpret_cast<const uint4*>(wr + c0);
        mma_bf16_m16n8k16(acc, a_lo.x, a_hi.x, a_lo.y, a_hi.y, bw.x, bw.y);
        mma_bf16_m16n8k16(acc, a_lo.z, a_hi.z, a_lo.w, a_hi.w, bw.z, bw.w);
      }
    }
    // acc[0..1] = row g, couts 2q, 2q+1; acc[2..3] = row g+8
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (!pv[r]) continue;
      const int b = (int)(pn[r] / F), f = (int)(pn[r] % F);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int co = 2 * q + j;
        if (co < cout)
          out[((((long long)b * cout + co) * F + f) * h + py[r]) * w + px[r]] = acc[2 * r + j] + (bias ? bias[co] : 0.f);
      }
    }
  }
}

// guided = u + g*(c-u); x' = cx*x + ce*guided; model_in[k] = x'*in_scale for k < copies
// dev_coef (optional): [g, cx, ce, in_scale] in device memory, so that one captured CUDA graph serves every step.
__global__ void cfg_step_kernel(const float4* __restrict__ eu, const float4* __restrict__ ec,
                                const float4* __restrict__ lat, float g, float cx, float ce, float in_scale,
                                const float* __restrict__ dev_coef, long long n4, float4* __restrict__ lat_out,
                                float4* __restrict__ model_in, int copies) {
  pdl_enter();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  if (dev_coef) {
    const float4 k = *reinterpret_cast<const float4*>(dev_coef);
    g = k.x; cx = k.y; ce = k.z; in_scale = k.w;
  }
  const float4 u = eu[i], c = ec[i], x = lat[i];
  float4 r;
  r.x = cx * x.x + ce * (u.x + g * (c.x - u.x));
  r.y = cx * x.y + ce * (u.y + g * (c.y - u.y));
  r.z = cx * x.z + ce * (u.z + g * (c.z - u.z));
  r.w = cx * x.w + ce * (u.w + g * (c.w - u.w));
  lat_out[i] = r;
  if (model_in) {
    const float4 s = make_float4(r.x * in_scale, r.y * in_scale, r.z * in_scale, r.w * in_scale);
    for (int k = 0; k < copies; ++k) model_in[(long long)k * n4 + i] = s;
  }
}

}  // namespace b200ad

using namespace b200ad;

extern "C" int b200ad_timestep_sinusoid(const float* t, int B, int dim, float* out, void* stream) {
  B200AD_REQUIRE(dim % 2 == 0, "timestep_sinusoid: dim=%d", dim);
  const int total = B * dim / 2;
  pdl_launch(timestep_sinusoid_kernel, dim3((total + 127) / 128), dim3(128), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), t, B, dim, out);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_small_linear(const float* x, int M, int K, const float* W, const float* b, int N, int silu_in,
                                   int silu_out, float* y, void* stream) {
  B200AD_REQUIRE(M >= 1 && M <= 8, "small_linear: M=%d (1..8)", M);
  B200AD_REQUIRE(K % 4 == 0, "small_linear: K=%d must be a multiple of 4", K);
  const int threads = 256;
  const int blocks = (int)(((long long)N * 32 + threads - 1) / threads);
  pdl_launch(small_linear_kernel, dim3(blocks), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), x, M, K, W, b, N, silu_in, silu_out, y);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_conv_in(const float* sample, int B, int cin, int F, int h, int w, const float* W,
                              const float* bias, int cout, float* out_nhwc, void* stream) {
  B200AD_REQUIRE(cin * 9 <= CIN_MAXK, "conv_in: cin=%d too large for the direct kernel", cin);
  const long long npix = (long long)B * F * h * w;
  const int threads = cout >= 320 ? 320 : ((cout + 31) / 32 * 32);
  if (cin == 4) {
    constexpr int PIX = 32;
    pdl_launch(conv_in_fast_kernel<4, PIX>, dim3((unsigned)((npix + PIX - 1) / PIX)), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
        sample, B, F, h, w, W, bias, cout, out_nhwc);
  } else {
    pdl_launch(conv_in_kernel, dim3((unsigned)((npix + CIN_PIX - 1) / CIN_PIX)), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
        sample, B, cin, F, h, w, W, bias, cout, out_nhwc, 0);
  }
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_conv_in_affine(const float* sample, int B, int cin, int F, int h, int w, const float* W,
                                     const float* bias, int cout, float* out_nhwc, void* stream) {
  B200AD_REQUIRE((cin + 1) * 9 <= CIN_MAXK, "conv_in_affine: cin=%d too large for the direct kernel", cin);
  const long long npix = (long long)B * F * h * w;
  const int threads = cout >= 256 ? 256 : ((cout + 31) / 32 * 32);
  pdl_launch(conv_in_kernel, dim3((unsigned)((npix + CIN_PIX - 1) / CIN_PIX)), dim3(threads), (size_t)(0), (cudaStream_t)stream,
             sample, B, cin, F, h, w, W, bias, cout, out_nhwc, 1);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_conv_out(int op_dtype, const void* act_nhwc, int B, int F, int h, int w, int cin,
                               const float* W, const float* bias, int cout, float* out_bcfhw, void* stream) {
  B200AD_REQUIRE(cout >= 1 && cout <= 4 && cin % 2 == 0, "conv_out: cin=%d cout=%d unsupported", cin, cout);
  const size_t smem = (size_t)cout * 9 * cin * sizeof(float);
  B200AD_REQUIRE(smem <= 96 * 1024, "conv_out: weights need %zu bytes of shared memory", smem);
  static bool attr_set = false;
  if (!attr_set) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(conv_out_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(conv_out_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set = true;
  }
  const long long npix = (long long)B * F * h * w;
  const int threads = 512;     // 16 pixels per block amortise the weight staging
  const long long blocks = (npix * 32 + threads - 1) / threads;
  if (op_dtype == 0 && cin % 32 == 0) {
    static int sms = 0;
    if (!sms) {
      int dev = 0;
      B200AD_CHECK_CUDA(cudaGetDevice(&dev));
      B200AD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      B200AD_CHECK_CUDA(cudaFuncSetAttribute(conv_out_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    }
    const size_t wbytes = (size_t)9 * 8 * conv_out_wrow(cin) * sizeof(bf16);
    B200AD_REQUIRE(wbytes <= 96 * 1024, "conv_out: weights need %zu bytes of shared memory", wbytes);
    const long long ngroups = (npix + 15) / 16;
    long long nb = (ngroups + 7) / 8;
    if (nb > 2LL * sms) nb = 2LL * sms;
    pdl_launch(conv_out_mma_kernel, dim3((unsigned)nb), dim3(256), (size_t)(wbytes), (cudaStream_t)((cudaStream_t)stream), (const bf16*)act_nhwc, B, F, h, w, cin, W,
                                                                            bias, cout, out_bcfhw);
  } else if (op_dtype == 0)
    pdl_launch(conv_out_kernel<bf16>, dim3((unsigned)blocks), dim3(threads), (size_t)(smem), (cudaStream_t)((cudaStream_t)stream), (const bf16*)act_nhwc, B, F, h, w, cin, W, bias, cout, out_bcfhw);
  else
    pdl_launch(conv_out_kernel<float>, dim3((unsigned)blocks), dim3(threads), (size_t)(smem), (cudaStream_t)((cudaStream_t)stream), (const float*)act_nhwc, B, F, h, w, cin, W, bias, cout, out_bcfhw);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_cfg_step(const float* eps_u, const float* eps_c, const float* latents, float g, float cx,
                               float ce, float in_scale, long long n, float* latents_out, float* model_in,
                               int model_in_copies, void* stream) {
  B200AD_REQUIRE(n % 4 == 0, "cfg_step: n=%lld must be a multiple of 4", n);
  const long long n4 = n / 4;
  const int threads = 256;
  pdl_launch(cfg_step_kernel, dim3((unsigned)((n4 + threads - 1) / threads)), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
      (const float4*)eps_u, (const float4*)eps_c, (const float4*)latents, g, cx, ce, in_scale, nullptr, n4,
      (float4*)latents_out, (float4*)model_in, model_in_copies);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_cfg_step_dev(const float* eps_u, const float* eps_c, const float* latents,
                                   const float* dev_coef4, long long n, float* latents_out, float* model_in,
                                   int model_in_copies, void* stream) {
  B200AD_REQUIRE(n % 4 == 0, "cfg_step_dev: n=%lld must be a multiple of 4", n);
  B200AD_REQUIRE(dev_coef4 && ((uintptr_t)dev_coef4 & 15) == 0, "cfg_step_dev: coefficient pointer must be 16-byte aligned");
  const long long n4 = n / 4;
  const int threads = 256;
  pdl_launch(cfg_step_kernel, dim3((unsigned)((n4 + threads - 1) / threads)), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
      (const float4*)eps_u, (const float4*)eps_c, (const float4*)latents, 0.f, 0.f, 0.f, 0.f, dev_coef4, n4,
      (float4*)latents_out, (float4*)model_in, model_in_copies);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}
