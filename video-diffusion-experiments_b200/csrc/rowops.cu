// LayerNorm (+positional embedding), re-layout casts and the motion module's temporal attention.
#include "../../include/b200ad.h"
#include "ops_common.cuh"

namespace b200ad {

// ----------------------------------------------------------------------------- LayerNorm (+PE)
// Persistent warps: each warp walks rows warp, warp + nwarps, ... with the row in registers (float2 per lane,
// lane-strided => coalesced 256 B requests) and the NEXT row's loads already in flight while the current row goes
// through its two shuffle reductions and its stores.  (One short-lived warp per row reached 3.9-4.5 TB/s: a warp's
// 1.3 KB of loads, two serial 5-step reductions and the block turn-over left the memory pipe idle in between.)
template <typename T, int MAXP, bool PE>   // MAXP = max float2 per lane (C <= 64*MAXP); PE: positional embedding added
                                           // (compile-time: a run-time `if (pe)` inside the store loop is a branch per pair)
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, int rows, int C,
                                                        const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, float eps,
                                                        const float* __restrict__ pe, int hw, int frames,
                                                        T* __restrict__ out) {
  pdl_enter();
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int npairs = C >> 1;
  const float inv_c = 1.0f / (float)C;
  int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  constexpr bool CACHE_GB = MAXP <= 5;       // narrow rows keep gamma / beta in registers; wide rows re-read them (L1)
  float2 g[CACHE_GB ? MAXP : 1], b[CACHE_GB ? MAXP : 1];
  if (CACHE_GB) {
#pragma unroll
    for (int j = 0; j < MAXP; ++j) {
      const int pr = lane + 32 * j;
      if (pr < npairs) {
        g[j] = *reinterpret_cast<const float2*>(gamma + 2 * pr);
        b[j] = *reinterpret_cast<const float2*>(beta + 2 * pr);
      }
    }
  }
  float2 nxt[MAXP];
#pragma unroll
  for (int j = 0; j < MAXP; ++j) {
    const int pr = lane + 32 * j;
    if (pr < npairs) nxt[j] = *reinterpret_cast<const float2*>(x + (long long)row * C + 2 * pr);
  }
  for (; row < rows; row += nwarps) {
    float2 v[MAXP];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < MAXP; ++j) {
      v[j] = nxt[j];
      if (lane + 32 * j < npairs) s += v[j].x + v[j].y;
    }
    const int row_n = row + nwarps;
    if (row_n < rows) {
#pragma unroll
      for (int j = 0; j < MAXP; ++j) {
        const int pr = lane + 32 * j;
        if (pr < npairs) nxt[j] = *reinterpret_cast<const float2*>(x + (long long)row_n * C + 2 * pr);
      }
    }
    const float mean = warp_sum(s) * inv_c;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < MAXP; ++j)
      if (lane + 32 * j < npairs) {
        const float a0 = v[j].x - mean, a1 = v[j].y - mean;
        q += a0 * a0 + a1 * a1;
      }
    const float rstd = rsqrtf(warp_sum(q) * inv_c + eps);
    const float* per = PE ? pe + (long long)((row / hw) % frames) * C : nullptr;
    T* orow = out + (long long)row * C;
#pragma unroll
    for (int j = 0; j < MAXP; ++j) {
      const int pr = lane + 32 * j;
      if (pr < npairs) {
        const float2 gj = CACHE_GB ? g[CACHE_GB ? j : 0] : *reinterpret_cast<const float2*>(gamma + 2 * pr);
        const float2 bj = CACHE_GB ? b[CACHE_GB ? j : 0] : *reinterpret_cast<const float2*>(beta + 2 * pr);
        float a0 = (v[j].x - mean) * rstd * gj.x + bj.x;
        float a1 = (v[j].y - mean) * rstd * gj.y + bj.y;
        if (PE) {
          const float2 p2 = *reinterpret_cast<const float2*>(per + 2 * pr);
          a0 += p2.x; a1 += p2.y;
        }
        store2<T>(orow + 2 * pr, a0, a1);
      }
    }
  }
}

// ----------------------------------------------------------------------------- casts with re-layout
template <typename T>
__global__ void cast_kernel(const float* __restrict__ x, int nimg, int h, int w, int C, int mode, T* __restrict__ out,
                            long long total4) {
  pdl_enter();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int c4 = C >> 2;
  const int c = (int)(i % c4) * 4;
  long long pix = i / c4;
  long long src;
  if (mode == 0) {
    src = pix;
  } else if (mode == 1) {             // out [4*nimg, h/2, w/2, C]; phase p = py*2+px stacked on the image axis
    const int wo = w >> 1, ho = h >> 1;
    const int xo = (int)(pix % wo); pix /= wo;
    const int yo = (int)(pix % ho); pix /= ho;
    const int n = (int)(pix % nimg);
    const int p = (int)(pix / nimg);
    src = ((long long)n * h + (2 * yo + (p >> 1))) * w + (2 * xo + (p & 1));
  } else {                            // nearest x2: out [nimg, 2h, 2w, C]
    const int wo = w << 1, ho = h << 1;
    const int xo = (int)(pix % wo); pix /= wo;
    const int yo = (int)(pix % ho);
    const int n = (int)(pix / ho);
    src = ((long long)n * h + (yo >> 1)) * w + (xo >> 1);
  }
  const float4 v = *reinterpret_cast<const float4*>(x + src * C + c);
  store4<T>(out + (i / c4) * C + c, v);
}

// ----------------------------------------------------------------------------- temporal attention
// One block per (spatial position (b, hw), group of `hpb` heads); warp = head.  The F rows of q|k|v for this
// position are F strided segments (row stride hw*3C): no permute copy is ever made.
template <typename T>
__global__ void temporal_attn_kernel(const T* __restrict__ qkv, int F, int hw, int C, int heads, int hpb,
                                     T* __restrict__ out) {
  pdl_enter();
  extern __shared__ float sm[];
  const int d = C / heads;
  const int W = hpb * d;                    // columns of this head group
  const int pos = blockIdx.x;               // b*hw + s
  const int b = pos / hw, s = pos % hw;
  const int col0 = blockIdx.y * W;
  const int C3 = 3 * C;
  const int ROW = 3 * W + 1;                // smem row: [q(W) | k(W) | v(W)] + 1 pad word (bank spread)
  float* sq = sm;                           // [F][ROW]
  float* sp = sm + (size_t)F * ROW;         // [hpb][F][F]
  const long long row0 = ((long long)b * F) * hw + s;
  const int pairs_per_part = W >> 1;
  for (int i = threadIdx.x; i < F * 3 * pairs_per_part; i += blockDim.x) {
    const int p = i % pairs_per_part;
    const int part = (i / pairs_per_part) % 3;
    const int f = i / (3 * pairs_per_part);
    const float2 v = load2<T>(qkv + (row0 + (long long)f * hw) * C3 + part * C + col0 + 2 * p);
    sq[f * ROW + part * W + 2 * p] = v.x;
    sq[f * ROW + part * W + 2 * p + 1] = v.y;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float scale = rsqrtf((float)d);
  for (int hl = warp; hl < hpb; hl += (blockDim.x >> 5)) {
    float* P = sp + (size_t)hl * F * F;
    const float* Q = sq + hl * d;
    const float* K = sq + W + hl * d;
    const float* V = sq + 2 * W + hl * d;
    for (int e = lane; e < F * F; e += 32) {
      const int i = e / F, j = e % F;
      float acc = 0.f;
      for (int c = 0; c < d; ++c) acc += Q[i * ROW + c] * K[j * ROW + c];
      P[e] = acc * scale;
    }
    __syncwarp();
    for (int i = lane; i < F; i += 32) {      // one lane per query row (F <= 32)
      float m = -INFINITY;
      for (int j = 0; j < F; ++j) m = fmaxf(m, P[i * F + j]);
      float l = 0.f;
      for (int j = 0; j < F; ++j) { const float p = __expf(P[i * F + j] - m); P[i * F + j] = p; l += p; }
      const float inv = 1.f / l;
      for (int j = 0; j < F; ++j) P[i * F + j] *= inv;
    }
    __syncwarp();
    for (int e = lane; e < F * (d >> 1); e += 32) {
      const int i = e / (d >> 1), c = (e % (d >> 1)) * 2;
      float a0 = 0.f, a1 = 0.f;
      for (int j = 0; j < F; ++j) {
        const float p = P[i * F + j];
        a0 += p * V[j * ROW + c];
        a1 += p * V[j * ROW + c + 1];
      }
      store2<T>(out + (row0 + (long long)i * hw) * C + col0 + hl * d + c, a0, a1);
    }
  }
}

int temporal_attention_bf16_mma(const bf16* qkv, int B, int F, int hw, int C, int heads, bf16* out,
                                cudaStream_t st);   // temporal_mma.cu

}  // namespace b200ad

using namespace b200ad;

extern "C" int b200ad_layernorm(int op_dtype, const float* x, int rows, int C, const float* gamma, const float* beta,
                                float eps, const float* pe, int hw, int frames, void* out_op, void* stream) {
  B200AD_REQUIRE(C % 2 == 0 && C <= 64 * 20, "layernorm: C=%d unsupported (even, <= 1280)", C);
  if (!pe) { hw = 1; frames = 1; }
  const int threads = 256;
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    B200AD_CHECK_CUDA(cudaGetDevice(&dev));
    B200AD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  // persistent: at most one wave of blocks (8 per SM at <= 64 registers, fewer for the wide rows)
  const long long want = ((long long)rows * 32 + threads - 1) / threads;
  const int per_sm = C <= 320 ? 8 : (C <= 640 ? 3 : 2);      // resident blocks at 64 / 80 / 128 registers
  const int blocks = (int)(want < (long long)sms * per_sm ? want : (long long)sms * per_sm);
  cudaStream_t st = (cudaStream_t)stream;
#define LN_LAUNCH(TT, MP)                                                                                            \
  do {                                                                                                               \
    if (pe)                                                                                                          \
      pdl_launch(layernorm_kernel<TT, MP, true>, dim3(blocks), dim3(threads), 0, st, x, rows, C, gamma, beta, eps, pe, hw, \
                 frames, (TT*)out_op);                                                                               \
    else                                                                                                             \
      pdl_launch(layernorm_kernel<TT, MP, false>, dim3(blocks), dim3(threads), 0, st, x, rows, C, gamma, beta, eps, pe, hw, \
                 frames, (TT*)out_op);                                                                               \
  } while (0)
  if (op_dtype == 0) {
    if (C <= 320) LN_LAUNCH(bf16, 5); else if (C <= 640) LN_LAUNCH(bf16, 10); else LN_LAUNCH(bf16, 20);
  } else {
    if (C <= 320) LN_LAUNCH(float, 5); else if (C <= 640) LN_LAUNCH(float, 10); else LN_LAUNCH(float, 20);
  }
#undef LN_LAUNCH
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_cast(int op_dtype, const float* x, int nimg, int h, int w, int C, int mode, void* out_op,
                           void* stream) {
  B200AD_REQUIRE(C % 4 == 0, "cast: C=%d must be a multiple of 4", C);
  B200AD_REQUIRE(mode >= 0 && mode <= 2, "cast: mode %d", mode);
  if (mode == 1) B200AD_REQUIRE(h % 2 == 0 && w % 2 == 0, "cast: space-to-depth needs even h,w (%d,%d)", h, w);
  long long out_pix = (long long)nimg * h * w;
  if (mode == 2) out_pix *= 4;
  const long long total4 = out_pix * (C / 4);
  const int threads = 256;
  const long long blocks = (total4 + threads - 1) / threads;
  if (op_dtype == 0)
    pdl_launch(cast_kernel<bf16>, dim3((unsigned)blocks), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), x, nimg, h, w, C, mode, (bf16*)out_op, total4);
  else
    pdl_launch(cast_kernel<float>, dim3((unsigned)blocks), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), x, nimg, h, w, C, mode, (float*)out_op, total4);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_temporal_attention(int op_dtype, const void* qkv, int B, int F, int hw, int C, int heads,
                                         void* out_op, void* stream) {
  B200AD_REQUIRE(F >= 1 && F <= 32, "temporal_attention: F=%d (motion_max_seq_length is 32)", F);
  B200AD_REQUIRE(C % heads == 0 && (C / heads) % 2 == 0, "temporal_attention: C=%d heads=%d", C, heads);
  const int d = C / heads;
  if (op_dtype == 0)
    return temporal_attention_bf16_mma((const bf16*)qkv, B, F, hw, C, heads, (bf16*)out_op, (cudaStream_t)stream);
  int hpb = heads;
  auto smem_for = [&](int h) { return ((size_t)F * (3 * h * d + 1) + (size_t)h * F * F) * sizeof(float); };
  while (hpb > 1 && (smem_for(hpb) > 64 * 1024 || heads % hpb != 0)) --hpb;
  const size_t smem = smem_for(hpb);
  B200AD_REQUIRE(smem <= 200 * 1024, "temporal_attention: needs %zu bytes of shared memory", smem);
  static bool attr_set = false;
  if (!attr_set) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(temporal_attn_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(temporal_attn_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  const int threads = 32 * (hpb < 4 ? 4 : hpb);
  dim3 grid(B * hw, heads / hpb);
  cudaStream_t st = (cudaStream_t)stream;
  if (op_dtype == 0)
    pdl_launch(temporal_attn_kernel<bf16>, dim3(grid), dim3(threads), (size_t)(smem), (cudaStream_t)(st), (const bf16*)qkv, F, hw, C, heads, hpb, (bf16*)out_op);
  else
    pdl_launch(temporal_attn_kernel<float>, dim3(grid), dim3(threads), (size_t)(smem), (cudaStream_t)(st), (const float*)qkv, F, hw, C, heads, hpb, (float*)out_op);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}
