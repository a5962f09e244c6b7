// fp32 kernel mode ("<= 1e-4 against the fp32 oracle"): SIMT FFMA GEMM / implicit-GEMM conv and a plain
// online-softmax attention, sharing the argument structures and epilogue of the tcgen05 path.  These are
// correctness-mode kernels (true fp32 products and accumulation), not the performance path.
#include "../../include/b200ad.h"
#include "gemm.cuh"
#include "ops_common.cuh"

namespace b200ad {

static constexpr int SB = 64;     // tile M = N = 64
static constexpr int SK = 16;

__device__ __forceinline__ float simt_load_a(const GemmArgs& a, int row, int k) {
  if (row >= a.M || k >= a.K) return 0.f;
  const float* A = reinterpret_cast<const float*>(a.A);
  if (!a.conv.enabled) return A[(long long)row * a.lda + k];
  const ConvGeom& g = a.conv;
  const int x = row % g.w, y = (row / g.w) % g.h, n = row / (g.w * g.h);
  const int tap = k / g.cin, c = k % g.cin;
  const int yy = y + g.tap_dy[tap], xx = x + g.tap_dx[tap], nn = n + g.tap_dn[tap];
  if (yy < 0 || yy >= g.src_h || xx < 0 || xx >= g.src_w) return 0.f;
  return A[(((long long)nn * g.src_h + yy) * g.src_w + xx) * g.cin + c];
}

__global__ void __launch_bounds__(256) gemm_simt_kernel(const __grid_constant__ GemmArgs a) {
  pdl_enter();
  __shared__ float sA[SK][SB + 1];
  __shared__ float sW[SK][SB + 1];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * SB, n0 = blockIdx.x * SB;
  const float* W = reinterpret_cast<const float*>(a.W);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < a.K; k0 += SK) {
    for (int i = threadIdx.x; i < SB * SK; i += 256) {
      const int r = i / SK, k = i % SK;
      sA[k][r] = simt_load_a(a, m0 + r, k0 + k);
      const int n = n0 + r;
      sW[k][r] = (n < a.N && k0 + k < a.K) ? W[(long long)n * a.K + k0 + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SK; ++k) {
      float av[4], wv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = sA[k][ty * 4 + i]; wv[i] = sW[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
  const Epilogue& e = a.epi;
  const int col = n0 + tx * 4;
  if (col >= a.N) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long row = m0 + ty * 4 + i;
    if (row >= a.M) continue;
    float add[4];
    epilogue_load_addends(e, row, col, add);
    if (e.geglu) {
      float* out = reinterpret_cast<float*>(e.out_op) + row * e.ld_oop + (col >> 1);
      const float h0 = acc[i][0] + add[0], g0 = acc[i][1] + add[1];
      const float h1 = acc[i][2] + add[2], g1 = acc[i][3] + add[3];
      out[0] = h0 * gelu_erf_f(g0);
      out[1] = h1 * gelu_erf_f(g1);
      continue;
    }
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = acc[i][j] + add[j];
    if (e.residual) {
      const float4 r = *reinterpret_cast<const float4*>(e.residual + row * e.ld_res + col);
      v[0] += r.x; v[1] += r.y; v[2] += r.z; v[3] += r.w;
    }
    if (e.out_f32) *reinterpret_cast<float4*>(e.out_f32 + row * e.ld_o32 + col) = make_float4(v[0], v[1], v[2], v[3]);
    if (e.out_op)
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out_op) + row * e.ld_oop + col) =
          make_float4(v[0], v[1], v[2], v[3]);
  }
}

int gemm_f32_simt(const GemmArgs& a, cudaStream_t stream) {
  B200AD_REQUIRE(a.N % 4 == 0, "gemm(fp32): N=%d must be a multiple of 4", a.N);
  dim3 grid((a.N + SB - 1) / SB, (a.M + SB - 1) / SB);
  pdl_launch(gemm_simt_kernel, dim3(grid), dim3(256), (size_t)(0), (cudaStream_t)(stream), a);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ----------------------------------------------------------------------------- fp32 attention
// One warp per query row; keys streamed through shared memory in chunks of 32 with an online softmax.
template <int DMAX>   // per-lane output columns = DMAX/32
__global__ void attn_simt_kernel(const float* __restrict__ q, long long ldq, const float* __restrict__ k,
                                 long long ldk, const float* __restrict__ v, long long ldv, int heads, int Lq, int Lk,
                                 int d, int dpad, int kv_div, float* __restrict__ out, long long ldo) {
  pdl_enter();
  extern __shared__ float sm[];
  const int KC = 32;
  float* sK = sm;                    // [KC][d+1]
  float* sV = sm + KC * (d + 1);     // [KC][d+1]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int n = blockIdx.z, hd = blockIdx.y;
  const int i = blockIdx.x * nw + warp;
  const bool active = i < Lq;
  const float scale = rsqrtf((float)d);
  float* sQ = sV + KC * (d + 1) + warp * d;   // [nw][d]
  if (active)
    for (int c = lane; c < d; c += 32) sQ[c] = q[((long long)n * Lq + i) * ldq + hd * dpad + c] * scale;
  float m = -INFINITY, l = 0.f;
  float o[DMAX / 32];
#pragma unroll
  for (int c = 0; c < DMAX / 32; ++c) o[c] = 0.f;
  const long long kvb = (long long)(n / kv_div) * Lk;
  for (int j0 = 0; j0 < Lk; j0 += KC) {
    __syncthreads();
    for (int e = threadIdx.x; e < KC * d; e += blockDim.x) {
      const int j = e / d, c = e % d;
      const bool ok = j0 + j < Lk;
      sK[j * (d + 1) + c] = ok ? k[(kvb + j0 + j) * ldk + hd * dpad + c] : 0.f;
      sV[j * (d + 1) + c] = ok ? v[(kvb + j0 + j) * ldv + hd * dpad + c] : 0.f;
    }
    __syncthreads();
    if (!active) continue;
    float s = -INFINITY;
    if (j0 + lane < Lk) {
      s = 0.f;
      for (int c = 0; c < d; ++c) s = fmaf(sQ[c], sK[lane * (d + 1) + c], s);
    }
    const float mnew = fmaxf(m, warp_max(s));
    const float alpha = __expf(m - mnew);
    const float p = (j0 + lane < Lk) ? __expf(s - mnew) : 0.f;
    l = l * alpha + warp_sum(p);
    m = mnew;
#pragma unroll
    for (int c = 0; c < DMAX / 32; ++c) o[c] *= alpha;
    for (int j = 0; j < KC; ++j) {
      const float pj = __shfl_sync(0xffffffffu, p, j);
#pragma unroll
      for (int c = 0; c < DMAX / 32; ++c) {
        const int col = lane + 32 * c;
        if (col < d) o[c] = fmaf(pj, sV[j * (d + 1) + col], o[c]);
      }
    }
  }
  if (!active) return;
  const float inv = 1.f / l;
#pragma unroll
  for (int c = 0; c < DMAX / 32; ++c) {
    const int col = lane + 32 * c;
    if (col < d) out[((long long)n * Lq + i) * ldo + hd * d + col] = o[c] * inv;
  }
}

int attention_f32_simt(const float* q, long long ldq, const float* k, long long ldk, const float* v, long long ldv,
                       int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, float* out, long long ldo,
                       cudaStream_t stream) {
  B200AD_REQUIRE(d <= 192, "attention(fp32): d=%d > 192", d);
  const int nw = 8;
  const size_t smem = ((size_t)2 * 32 * (d + 1) + (size_t)nw * d) * sizeof(float);
  dim3 grid((Lq + nw - 1) / nw, heads, nbatch);
  static bool attr = false;
  if (!attr) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(attn_simt_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    attr = true;
  }
  pdl_launch(attn_simt_kernel<192>, dim3(grid), dim3(nw * 32), (size_t)(smem), (cudaStream_t)(stream), q, ldq, k, ldk, v, ldv, heads, Lq, Lk, d, dpad, kv_div, out, ldo);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200ad
