// Pointwise / row kernels outside the fused denoise step:
//  * eltwise: SiLU, broadcast add, GEGLU gate -- the glue of the hook-visible module mode, where every diffusers
//    sub-module (nn.SiLU, the residual adds, GEGLU) is called on its own so that forward hooks fire
//    (/root/reference/utils/forward_tracer.py:177-194);
//  * softmax_rows: row softmax of an fp32 score matrix into the operand type (the VAE decoder's single-head
//    attention, d = 512, is run as S = Q K^T (GEMM) -> softmax_rows -> P V (GEMM));
//  * video_to_u8: diffusers' postprocess_video, (x/2 + 0.5).clamp(0,1) * 255 -> round -> uint8, [C,F,h,w] -> [F,h,w,C].
// All memory-bound: coalesced, 16-byte accesses where the layout allows, fp32 math.
#include "../../include/b200ad.h"
#include "ops_common.cuh"

namespace b200ad {

// mode 0: out[i] = silu(a[i]);  mode 1: out[i] = a[i] + b[(i / div) % mod];  mode 3: quick_gelu, a * sigmoid(1.702 a)
__global__ void eltwise_kernel(int mode, const float* __restrict__ a, const float* __restrict__ b, long long n,
                               long long div, long long mod, float* __restrict__ out) {
  pdl_enter();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = a[i];
    out[i] = mode == 0 ? silu_f(v) : (mode == 3 ? v / (1.0f + expf(-1.702f * v)) : v + b[(i / div) % mod]);
  }
}

// CLIP text embeddings: out[t, :] = token_table[ids[t], :] + position_table[t % seq, :]
__global__ void embed_tokens_kernel(const long long* __restrict__ ids, long long ntok, int seq, int vocab,
                                    const float* __restrict__ tok, const float* __restrict__ pos, int dim,
                                    float* __restrict__ out) {
  pdl_enter();
  const long long n = ntok * dim;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const long long t = i / dim;
    const int c = (int)(i % dim);
    long long id = ids[t];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    out[i] = tok[id * dim + c] + pos[(t % seq) * (long long)dim + c];
  }
}

// Short-sequence attention with an optional causal mask (CLIP text encoder: 77 tokens, 12 heads of 64): one warp per
// (batch item, head, query row); lanes split the keys for the scores and the head width for P V.  fp32 math.
template <typename T>
__global__ void __launch_bounds__(128) attn_small_kernel(const T* __restrict__ q, long long ldq, const T* __restrict__ k,
                                                         long long ldk, const T* __restrict__ v, long long ldv, int heads,
                                                         int L, int d, int causal, T* __restrict__ out, long long ldo,
                                                         long long nrows) {
  pdl_enter();
  extern __shared__ float sm_attn[];              // [warps][d] query row + [warps][L] probabilities
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const long long w = (long long)blockIdx.x * nw + warp;     // (n * heads + h) * L + i
  if (w >= nrows) return;
  const int i = (int)(w % L);
  const int h = (int)((w / L) % heads);
  const long long n = w / ((long long)L * heads);
  float* sq = sm_attn + (size_t)warp * d;
  float* sp = sm_attn + (size_t)nw * d + (size_t)warp * L;
  const float scale = rsqrtf((float)d);
  const T* qr = q + (n * L + i) * ldq + (long long)h * d;
  for (int c = lane; c < d; c += 32) sq[c] = to_f<T>(qr[c]) * scale;
  __syncwarp();
  const int nk = causal ? i + 1 : L;
  float mx = -INFINITY;
  for (int j = lane; j < nk; j += 32) {
    const T* kr = k + (n * L + j) * ldk + (long long)h * d;
    float s = 0.f;
    for (int c = 0; c < d; ++c) s = fmaf(sq[c], to_f<T>(kr[c]), s);
    sp[j] = s;
    mx = fmaxf(mx, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
  for (int j = lane; j < nk; j += 32) {
    const float p = __expf(sp[j] - mx);
    sp[j] = p;
    sum += p;
  }
  sum = warp_sum(sum);
  __syncwarp();
  const float inv = 1.0f / sum;
  T* orow = out + (n * L + i) * ldo + (long long)h * d;
  for (int c = lane; c < d; c += 32) {
    float acc = 0.f;
    for (int j = 0; j < nk; ++j) acc = fmaf(sp[j], to_f<T>(v[(n * L + j) * ldv + (long long)h * d + c]), acc);
    if (sizeof(T) == 2) orow[c] = (T)__float2bfloat16(acc * inv);
    else *reinterpret_cast<float*>(orow + c) = acc * inv;
  }
}

// hg [rows, 2C] = (h | g)  ->  out [rows, C] = h * gelu_erf(g)   (diffusers GEGLU after its Linear)
__global__ void geglu_gate_kernel(const float* __restrict__ hg, long long rows, int C, float* __restrict__ out) {
  pdl_enter();
  const long long n = rows * C;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const long long r = i / C;
    const int c = (int)(i % C);
    const float* row = hg + r * 2 * C;
    out[i] = row[c] * gelu_erf_f(row[C + c]);
  }
}

// One block per row: out[r, :] = softmax(scale * s[r, :]).  cols <= 4 * 256 * SM_VPT; the row lives in registers.
static constexpr int SM_THREADS = 256;
static constexpr int SM_VPT = 8;           // float4 vectors per thread -> cols <= 8192
template <typename T>
__global__ void __launch_bounds__(SM_THREADS) softmax_rows_kernel(const float* __restrict__ s, int cols, long long ld,
                                                                  float scale, T* __restrict__ out, long long ldo) {
  pdl_enter();
  __shared__ float red[SM_THREADS / 32];
  const long long r = blockIdx.x;
  const float4* src = reinterpret_cast<const float4*>(s + r * ld);
  const int nvec = cols >> 2;
  float4 v[SM_VPT];
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < SM_VPT; ++j) {
    const int i = threadIdx.x + j * SM_THREADS;
    if (i < nvec) {
      v[j] = __ldg(src + i);
      m = fmaxf(m, fmaxf(fmaxf(v[j].x, v[j].y), fmaxf(v[j].z, v[j].w)));
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < SM_THREADS / 32; ++w) m = fmaxf(m, red[w]);
  __syncthreads();
  const float k = scale * 1.4426950408889634f;      // exp(scale*(x-m)) = exp2(k*(x-m))
  const float mk = m * k;
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < SM_VPT; ++j) {
    const int i = threadIdx.x + j * SM_THREADS;
    if (i < nvec) {
      v[j].x = exp2f(fmaf(v[j].x, k, -mk)); v[j].y = exp2f(fmaf(v[j].y, k, -mk));
      v[j].z = exp2f(fmaf(v[j].z, k, -mk)); v[j].w = exp2f(fmaf(v[j].w, k, -mk));
      sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
    }
  }
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int w = 0; w < SM_THREADS / 32; ++w) sum += red[w];
  const float inv = 1.0f / sum;
  T* dst = out + r * ldo;
#pragma unroll
  for (int j = 0; j < SM_VPT; ++j) {
    const int i = threadIdx.x + j * SM_THREADS;
    if (i < nvec) store4<T>(dst + 4 * i, make_float4(v[j].x * inv, v[j].y * inv, v[j].z * inv, v[j].w * inv));
  }
}

// video [C, F, h, w] fp32 (one batch item of [B, C, F, h, w]) -> frames [F, h, w, C] uint8
__global__ void video_to_u8_kernel(const float* __restrict__ video, int C, int F, int h, int w,
                                   uint8_t* __restrict__ out) {
  pdl_enter();
  const long long npix = (long long)F * h * w;
  const long long plane = (long long)h * w;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += stride) {
    const long long f = p / plane, q = p % plane;
    for (int c = 0; c < C; ++c) {
      float v = video[((long long)c * F + f) * plane + q] * 0.5f + 0.5f;
      v = fminf(fmaxf(v, 0.f), 1.f);
      out[p * C + c] = (uint8_t)rintf(v * 255.0f);       // torch.round: half to even
    }
  }
}

}  // namespace b200ad

using namespace b200ad;

static int grid_for(long long n, int threads) {
  long long b = (n + threads - 1) / threads;
  const long long cap = 148LL * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

extern "C" int b200ad_eltwise(int mode, const float* a, const float* b, long long n, long long div, long long mod,
                              float* out, void* stream) {
  B200AD_REQUIRE(a && out && n >= 0, "eltwise: null operand");
  if (n == 0) return 0;
  if (mode == 2) {
    // geglu gate: a = hg [rows, 2C], n = rows*C, mod = C
    B200AD_REQUIRE(mod > 0 && n % mod == 0, "eltwise(geglu): n=%lld C=%lld", n, mod);
    pdl_launch(geglu_gate_kernel, dim3(grid_for(n, 256)), dim3(256), (size_t)0, (cudaStream_t)stream, a, n / mod, (int)mod,
               out);
  } else {
    B200AD_REQUIRE(mode == 0 || mode == 1 || mode == 3, "eltwise: mode %d", mode);
    B200AD_REQUIRE(mode != 1 || (b && div > 0 && mod > 0), "eltwise(add): needs b, div > 0, mod > 0");
    pdl_launch(eltwise_kernel, dim3(grid_for(n, 256)), dim3(256), (size_t)0, (cudaStream_t)stream, mode, a, b, n,
               div > 0 ? div : 1, mod > 0 ? mod : 1, out);
  }
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_softmax_rows(int op_dtype, const float* s, long long rows, int cols, long long ld, float scale,
                                   void* out_op, long long ldo, void* stream) {
  B200AD_REQUIRE(s && out_op, "softmax_rows: null operand");
  B200AD_REQUIRE(cols % 4 == 0 && cols <= 4 * SM_THREADS * SM_VPT && ld % 4 == 0 && ldo % 4 == 0,
                 "softmax_rows: cols=%d ld=%lld ldo=%lld unsupported", cols, ld, ldo);
  B200AD_REQUIRE(rows >= 1 && rows <= 0x7fffffffLL, "softmax_rows: rows=%lld", rows);
  if (op_dtype == 0)
    pdl_launch(softmax_rows_kernel<bf16>, dim3((unsigned)rows), dim3(SM_THREADS), (size_t)0, (cudaStream_t)stream, s, cols,
               ld, scale, (bf16*)out_op, ldo);
  else
    pdl_launch(softmax_rows_kernel<float>, dim3((unsigned)rows), dim3(SM_THREADS), (size_t)0, (cudaStream_t)stream, s, cols,
               ld, scale, (float*)out_op, ldo);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_video_to_u8(const float* video_cfhw, int C, int F, int h, int w, unsigned char* out_fhwc,
                                  void* stream) {
  B200AD_REQUIRE(video_cfhw && out_fhwc && C >= 1 && C <= 4, "video_to_u8: C=%d", C);
  const long long npix = (long long)F * h * w;
  pdl_launch(video_to_u8_kernel, dim3(grid_for(npix, 256)), dim3(256), (size_t)0, (cudaStream_t)stream, video_cfhw, C, F, h,
             w, (uint8_t*)out_fhwc);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_embed_tokens(const long long* ids, long long ntok, int seq, int vocab, const float* tok_table,
                                   const float* pos_table, int dim, float* out, void* stream) {
  B200AD_REQUIRE(ids && tok_table && pos_table && out && seq > 0 && dim > 0 && vocab > 0, "embed_tokens: bad arguments");
  pdl_launch(embed_tokens_kernel, dim3(grid_for(ntok * dim, 256)), dim3(256), (size_t)0, (cudaStream_t)stream, ids, ntok, seq,
             vocab, tok_table, pos_table, dim, out);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_attention_small(int op_dtype, const void* q, long long ldq, const void* k, long long ldk,
                                      const void* v, long long ldv, int nbatch, int heads, int L, int d, int causal,
                                      void* out_op, long long ldo, void* stream) {
  B200AD_REQUIRE(q && k && v && out_op, "attention_small: null operand");
  B200AD_REQUIRE(L >= 1 && L <= 1024 && d >= 1 && d <= 256, "attention_small: L=%d d=%d unsupported", L, d);
  const int nw = 4;
  const long long nrows = (long long)nbatch * heads * L;
  const size_t smem = (size_t)nw * (d + L) * sizeof(float);
  const unsigned blocks = (unsigned)((nrows + nw - 1) / nw);
  if (op_dtype == 0)
    pdl_launch(attn_small_kernel<bf16>, dim3(blocks), dim3(nw * 32), smem, (cudaStream_t)stream, (const bf16*)q, ldq,
               (const bf16*)k, ldk, (const bf16*)v, ldv, heads, L, d, causal, (bf16*)out_op, ldo, nrows);
  else
    pdl_launch(attn_small_kernel<float>, dim3(blocks), dim3(nw * 32), smem, (cudaStream_t)stream, (const float*)q, ldq,
               (const float*)k, ldk, (const float*)v, ldv, heads, L, d, causal, (float*)out_op, ldo, nrows);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}
