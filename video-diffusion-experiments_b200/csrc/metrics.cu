// Temporal-consistency metrics of a generated video on the device (the reference measures
// them on the CPU frame pair by frame pair in experiments/06_measure_grid_search.py:209-284,290-352):
//   * pair sums: per consecutive frame pair the sum of squared differences (-> MSE, PSNR) and per frame triplet the sum
//     of |I_t - 2 I_{t+1} + I_{t+2}| (-> flicker index), in EXACT integer arithmetic on the uint8 frames;
//   * warp error: backward warp of frame t by the optical flow (bilinear, border padding, align_corners=True, the
//     reference's F.grid_sample call) fused with the squared error against frame t+1 -- no warped image is stored --
//     plus the flow-magnitude sums (mean / std of |flow|);
//   * gray_u8: the uint8 luminance the reference feeds to OpenCV's Farneback flow ((mean over RGB of x/255) * 255,
//     truncated), so the only host work left is the flow itself.
// All are one-read reductions: per-block partial sums (fixed order => deterministic), reduced by a second tiny kernel.
#include "../../include/b200ad.h"
#include "ops_common.cuh"

namespace b200ad {

static constexpr int MT = 256;

__device__ __forceinline__ double block_sum_d(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < MT / 32; ++w) t += sh[w];
  return t;                                        // valid in thread 0
}

// grid (nb, F-1).  partial[(pair*2 + k)*nb + block]: k = 0 sum (a-b)^2, k = 1 sum |a - 2b + c| (0 for the last pair)
__global__ void __launch_bounds__(MT) pair_sums_kernel(const uint8_t* __restrict__ fr, int F, long long n,
                                                      double* __restrict__ partial) {
  pdl_enter();
  __shared__ double sh[MT / 32];
  const int pair = blockIdx.y, nb = gridDim.x;
  const uint8_t* a = fr + (long long)pair * n;
  const uint8_t* b = a + n;
  const bool tri = pair + 2 < F;
  const uint8_t* c = tri ? b + n : b;
  unsigned long long sq = 0, fl = 0;
  const long long n16 = n >> 4;
  const long long stride = (long long)nb * MT;
  for (long long i = (long long)blockIdx.x * MT + threadIdx.x; i < n16; i += stride) {
    const uint4 va = __ldg(reinterpret_cast<const uint4*>(a) + i);
    const uint4 vb = __ldg(reinterpret_cast<const uint4*>(b) + i);
    const uint4 vc = tri ? __ldg(reinterpret_cast<const uint4*>(c) + i) : vb;
    const uint32_t wa[4] = {va.x, va.y, va.z, va.w}, wb[4] = {vb.x, vb.y, vb.z, vb.w}, wc[4] = {vc.x, vc.y, vc.z, vc.w};
#pragma unroll
    for (int w = 0; w < 4; ++w)
#pragma unroll
      for (int s = 0; s < 32; s += 8) {
        const int x = (wa[w] >> s) & 255, y = (wb[w] >> s) & 255, z = (wc[w] >> s) & 255;
        const int d = x - y;
        sq += (unsigned)(d * d);
        const int e = x - 2 * y + z;
        fl += (unsigned)(e < 0 ? -e : e);
      }
  }
  if (blockIdx.x == 0)                             // tail (n % 16 elements)
    for (long long i = (n16 << 4) + threadIdx.x; i < n; i += MT) {
      const int x = a[i], y = b[i], z = c[i];
      sq += (unsigned)((x - y) * (x - y));
      const int e = x - 2 * y + z;
      fl += (unsigned)(e < 0 ? -e : e);
    }
  const double s0 = block_sum_d((double)sq, sh);
  const double s1 = block_sum_d(tri ? (double)fl : 0.0, sh);
  if (threadIdx.x == 0) {
    partial[((long long)pair * 2 + 0) * nb + blockIdx.x] = s0;
    partial[((long long)pair * 2 + 1) * nb + blockIdx.x] = s1;
  }
}

// out[r] = sum_b partial[r*nb + b]   (one warp per row, fixed order)
__global__ void reduce_rows_kernel(const double* __restrict__ partial, int rows, int nb, double* __restrict__ out) {
  pdl_enter();
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= rows) return;
  double s = 0.0;
  for (int b = lane; b < nb; b += 32) s += partial[(long long)r * nb + b];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[r] = s;
}

// grid (nb, F-1).  partial[(pair*3 + k)*nb + block]: k = 0 sum (warp(frame_t) - frame_{t+1})^2 over C*H*W in [0,1]
// units, k = 1 sum |flow|, k = 2 sum |flow|^2 over H*W.  flow: [F-1, H, W, 2] (dx, dy), as cv2 returns it.
__global__ void __launch_bounds__(MT) warp_error_kernel(const uint8_t* __restrict__ fr, int H, int W, int C,
                                                       const float* __restrict__ flow, double* __restrict__ partial) {
  pdl_enter();
  __shared__ double sh[MT / 32];
  const int pair = blockIdx.y, nb = gridDim.x;
  const long long npix = (long long)H * W;
  const uint8_t* f1 = fr + (long long)pair * npix * C;
  const uint8_t* f2 = f1 + npix * C;
  const float* fl = flow + (long long)pair * npix * 2;
  double e = 0.0, m1 = 0.0, m2 = 0.0;
  const float inv255 = 1.0f / 255.0f;
  for (long long p = (long long)blockIdx.x * MT + threadIdx.x; p < npix; p += (long long)nb * MT) {
    const int x = (int)(p % W), y = (int)(p / W);
    const float2 d = *reinterpret_cast<const float2*>(fl + 2 * p);
    const float mag = sqrtf(d.x * d.x + d.y * d.y);
    m1 += mag;
    m2 += (double)mag * mag;
    // the reference's coordinate round trip (06:262-271 + grid_sample's align_corners=True un-normalisation), in fp32
    float sx = (float)x + d.x, sy = (float)y + d.y;
    sx = 2.0f * sx / (float)(W - 1) - 1.0f;
    sy = 2.0f * sy / (float)(H - 1) - 1.0f;
    sx = (sx + 1.0f) * 0.5f * (float)(W - 1);
    sy = (sy + 1.0f) * 0.5f * (float)(H - 1);
    sx = fminf(fmaxf(sx, 0.0f), (float)(W - 1));         // padding_mode='border'
    sy = fminf(fmaxf(sy, 0.0f), (float)(H - 1));
    const int x0 = (int)floorf(sx), y0 = (int)floorf(sy);
    const int x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1);
    const float ax = sx - (float)x0, ay = sy - (float)y0;
    const float w00 = (1.f - ax) * (1.f - ay), w01 = ax * (1.f - ay), w10 = (1.f - ax) * ay, w11 = ax * ay;
    const uint8_t* p00 = f1 + ((long long)y0 * W + x0) * C;
    const uint8_t* p01 = f1 + ((long long)y0 * W + x1) * C;
    const uint8_t* p10 = f1 + ((long long)y1 * W + x0) * C;
    const uint8_t* p11 = f1 + ((long long)y1 * W + x1) * C;
    for (int c = 0; c < C; ++c) {
      const float wv = (p00[c] * inv255) * w00 + (p01[c] * inv255) * w01 + (p10[c] * inv255) * w10 + (p11[c] * inv255) * w11;
      const float df = wv - f2[p * C + c] * inv255;
      e += (double)df * df;
    }
  }
  const double s0 = block_sum_d(e, sh), s1 = block_sum_d(m1, sh), s2 = block_sum_d(m2, sh);
  if (threadIdx.x == 0) {
    partial[((long long)pair * 3 + 0) * nb + blockIdx.x] = s0;
    partial[((long long)pair * 3 + 1) * nb + blockIdx.x] = s1;
    partial[((long long)pair * 3 + 2) * nb + blockIdx.x] = s2;
  }
}

// gray[f, y, x] = uint8(((r/255 + g/255 + b/255) / 3) * 255), truncated -- the reference's
// (frame.mean(dim=0) * 255).numpy().astype(np.uint8) on [C,H,W] floats in [0,1] (06:163-164), same fp32 op order
__global__ void gray_u8_kernel(const uint8_t* __restrict__ fr, long long npix, int C, uint8_t* __restrict__ gray) {
  pdl_enter();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += stride) {
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += (float)fr[p * C + c] / 255.0f;
    const float m = (s / (float)C) * 255.0f;
    gray[p] = (uint8_t)m;
  }
}


// ----------------------------------------------------------------------------- LPIPS (AlexNet) pieces
// uint8 frames [N,H,W,3] -> fp32 NHWC, LPIPS' ScalingLayer applied: ((v/255*2 - 1) - shift_c) / scale_c
__global__ void lpips_input_kernel(const uint8_t* __restrict__ fr, long long n, float* __restrict__ out) {
  pdl_enter();
  const float shift[3] = {-0.030f, -0.088f, -0.188f}, scale[3] = {0.458f, 0.448f, 0.450f};
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int c = (int)(i % 3);
    const float x = (float)fr[i] / 255.0f * 2.0f - 1.0f;
    out[i] = (x - shift[c]) / scale[c];
  }
}

// Direct convolution as an fp32 SIMT implicit GEMM over NHWC (any kernel size / stride / padding; bias + optional ReLU):
// out[n, oy, ox, co] = act(b[co] + sum_{ky,kx,c} x[n, oy*s - p + ky, ox*s - p + kx, c] * w[co, ky, kx, c]).
// Tile 64 output pixels x 64 output channels per block of 256 threads (4 x 4 outputs per thread), K staged 16 at a time.
static constexpr int DC_B = 64, DC_K = 16;
__global__ void __launch_bounds__(256) conv_direct_kernel(const float* __restrict__ x, int N, int H, int W, int Cin,
                                                         const float* __restrict__ w, const float* __restrict__ bias,
                                                         int Cout, int KH, int KW, int stride, int pad, int OH, int OW,
                                                         int relu, float* __restrict__ out) {
  pdl_enter();
  __shared__ float sA[DC_K][DC_B + 1];
  __shared__ float sW[DC_K][DC_B + 1];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const long long m0 = (long long)blockIdx.x * DC_B;
  const int n0 = blockIdx.y * DC_B;
  const long long M = (long long)N * OH * OW;
  const int K = KH * KW * Cin;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += DC_K) {
    for (int i = threadIdx.x; i < DC_B * DC_K; i += 256) {
      const int r = i / DC_K, kk = i % DC_K;
      const int k = k0 + kk;
      float a = 0.f;
      const long long m = m0 + r;
      if (m < M && k < K) {
        const int ox = (int)(m % OW), oy = (int)((m / OW) % OH);
        const long long n = m / ((long long)OW * OH);
        const int c = k % Cin, kx = (k / Cin) % KW, ky = k / (Cin * KW);
        const int iy = oy * stride - pad + ky, ix = ox * stride - pad + kx;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) a = x[((n * H + iy) * W + ix) * Cin + c];
      }
      sA[kk][r] = a;
      const int co = n0 + r;
      sW[kk][r] = (co < Cout && k < K) ? w[(long long)co * K + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < DC_K; ++kk) {
      float av[4], wv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = sA[kk][ty * 4 + i]; wv[i] = sW[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= Cout) continue;
      float v = acc[i][j] + (bias ? bias[co] : 0.f);
      if (relu) v = fmaxf(v, 0.f);
      out[m * Cout + co] = v;
    }
  }
}

// MaxPool2d(kernel, stride) without padding, NHWC
__global__ void maxpool_kernel(const float* __restrict__ x, int N, int H, int W, int C, int ks, int stride, int OH, int OW,
                               float* __restrict__ out) {
  pdl_enter();
  const long long total = (long long)N * OH * OW * C;
  const long long gs = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gs) {
    const int c = (int)(i % C);
    long long p = i / C;
    const int ox = (int)(p % OW); p /= OW;
    const int oy = (int)(p % OH);
    const long long n = p / OH;
    float m = -INFINITY;
    for (int ky = 0; ky < ks; ++ky)
      for (int kx = 0; kx < ks; ++kx) {
        const int iy = oy * stride + ky, ix = ox * stride + kx;
        if (iy < H && ix < W) m = fmaxf(m, x[((n * H + iy) * W + ix) * C + c]);
      }
    out[i] = m;
  }
}

// LPIPS layer distance between the features of consecutive images: per pixel both feature vectors are normalised to
// unit length (x / (|x| + 1e-10)), the squared differences are weighted by lin[c] and summed; the spatial SUM per pair
// goes to partial[pair*nb + block] (the caller divides by h*w).  grid (nb, N-1); one warp per pixel.
__global__ void __launch_bounds__(MT) lpips_layer_kernel(const float* __restrict__ f, int hw, int C,
                                                        const float* __restrict__ lin, double* __restrict__ partial) {
  pdl_enter();
  __shared__ double sh[MT / 32];
  const int pair = blockIdx.y, nb = gridDim.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = MT / 32;
  const float* f0 = f + (long long)pair * hw * C;
  const float* f1 = f0 + (long long)hw * C;
  double acc = 0.0;
  for (int p = blockIdx.x * nw + warp; p < hw; p += nb * nw) {
    const float* a = f0 + (long long)p * C;
    const float* b = f1 + (long long)p * C;
    float sa = 0.f, sb = 0.f;
    for (int c = lane; c < C; c += 32) { sa = fmaf(a[c], a[c], sa); sb = fmaf(b[c], b[c], sb); }
    sa = warp_sum(sa);
    sb = warp_sum(sb);
    const float ia = 1.0f / (sqrtf(sa) + 1e-10f), ib = 1.0f / (sqrtf(sb) + 1e-10f);
    float d = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float t = a[c] * ia - b[c] * ib;
      d = fmaf(lin[c] * t, t, d);
    }
    d = warp_sum(d);
    if (lane == 0) acc += d;
  }
  const double s = block_sum_d(acc, sh);
  if (threadIdx.x == 0) partial[(long long)pair * nb + blockIdx.x] = s;
}

}  // namespace b200ad

using namespace b200ad;

extern "C" int b200ad_metric_blocks(void) { return 148; }

extern "C" int b200ad_video_pair_sums(const unsigned char* frames_fhwc, int F, long long n_per_frame, double* workspace,
                                      double* out, void* stream) {
  B200AD_REQUIRE(frames_fhwc && workspace && out && F >= 2 && n_per_frame > 0, "video_pair_sums: bad arguments");
  B200AD_REQUIRE(((uintptr_t)frames_fhwc & 15) == 0 && n_per_frame % 16 == 0,
                 "video_pair_sums: frames must be 16-byte aligned with a multiple of 16 bytes per frame");
  const int nb = b200ad_metric_blocks();
  pdl_launch(pair_sums_kernel, dim3(nb, F - 1), dim3(MT), (size_t)0, (cudaStream_t)stream, (const uint8_t*)frames_fhwc, F,
             n_per_frame, workspace);
  const int rows = 2 * (F - 1);
  pdl_launch(reduce_rows_kernel, dim3((rows * 32 + 127) / 128), dim3(128), (size_t)0, (cudaStream_t)stream,
             (const double*)workspace, rows, nb, out);
  count_launch(2);
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_video_warp_error(const unsigned char* frames_fhwc, int F, int H, int W, int C, const float* flow,
                                       double* workspace, double* out, void* stream) {
  B200AD_REQUIRE(frames_fhwc && flow && workspace && out && F >= 2 && H >= 2 && W >= 2 && C >= 1 && C <= 4,
                 "video_warp_error: bad arguments");
  const int nb = b200ad_metric_blocks();
  pdl_launch(warp_error_kernel, dim3(nb, F - 1), dim3(MT), (size_t)0, (cudaStream_t)stream, (const uint8_t*)frames_fhwc, H,
             W, C, flow, workspace);
  const int rows = 3 * (F - 1);
  pdl_launch(reduce_rows_kernel, dim3((rows * 32 + 127) / 128), dim3(128), (size_t)0, (cudaStream_t)stream,
             (const double*)workspace, rows, nb, out);
  count_launch(2);
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_video_gray_u8(const unsigned char* frames_fhwc, long long npix, int C, unsigned char* gray,
                                    void* stream) {
  B200AD_REQUIRE(frames_fhwc && gray && npix > 0 && C >= 1 && C <= 4, "video_gray_u8: bad arguments");
  long long blocks = (npix + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pdl_launch(gray_u8_kernel, dim3((unsigned)blocks), dim3(256), (size_t)0, (cudaStream_t)stream,
             (const uint8_t*)frames_fhwc, npix, C, (uint8_t*)gray);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_lpips_input(const unsigned char* frames_nhwc, long long n_elems, float* out, void* stream) {
  B200AD_REQUIRE(frames_nhwc && out && n_elems % 3 == 0, "lpips_input: expects RGB uint8 frames");
  long long blocks = (n_elems + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pdl_launch(lpips_input_kernel, dim3((unsigned)blocks), dim3(256), (size_t)0, (cudaStream_t)stream,
             (const uint8_t*)frames_nhwc, n_elems, out);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_conv2d_direct(const float* x_nhwc, int N, int H, int W, int Cin, const float* w_ohwi,
                                    const float* bias, int Cout, int KH, int KW, int stride, int pad, int relu,
                                    float* out_nhwc, void* stream) {
  B200AD_REQUIRE(x_nhwc && w_ohwi && out_nhwc && stride >= 1 && KH >= 1 && KW >= 1, "conv2d_direct: bad arguments");
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  B200AD_REQUIRE(OH >= 1 && OW >= 1, "conv2d_direct: empty output");
  const long long M = (long long)N * OH * OW;
  dim3 grid((unsigned)((M + DC_B - 1) / DC_B), (Cout + DC_B - 1) / DC_B);
  pdl_launch(conv_direct_kernel, grid, dim3(256), (size_t)0, (cudaStream_t)stream, x_nhwc, N, H, W, Cin, w_ohwi, bias, Cout,
             KH, KW, stride, pad, OH, OW, relu, out_nhwc);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_maxpool2d(const float* x_nhwc, int N, int H, int W, int C, int ks, int stride, float* out_nhwc,
                                void* stream) {
  B200AD_REQUIRE(x_nhwc && out_nhwc && ks >= 1 && stride >= 1 && H >= ks && W >= ks, "maxpool2d: bad arguments");
  const int OH = (H - ks) / stride + 1, OW = (W - ks) / stride + 1;
  const long long total = (long long)N * OH * OW * C;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pdl_launch(maxpool_kernel, dim3((unsigned)blocks), dim3(256), (size_t)0, (cudaStream_t)stream, x_nhwc, N, H, W, C, ks,
             stride, OH, OW, out_nhwc);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200ad_lpips_layer(const float* feat_nhwc, int N, int hw, int C, const float* lin, double* workspace,
                                  double* out, void* stream) {
  B200AD_REQUIRE(feat_nhwc && lin && workspace && out && N >= 2, "lpips_layer: bad arguments");
  const int nb = b200ad_metric_blocks();
  pdl_launch(lpips_layer_kernel, dim3(nb, N - 1), dim3(MT), (size_t)0, (cudaStream_t)stream, feat_nhwc, hw, C, lin,
             workspace);
  pdl_launch(reduce_rows_kernel, dim3(((N - 1) * 32 + 127) / 128), dim3(128), (size_t)0, (cudaStream_t)stream,
             (const double*)workspace, N - 1, nb, out);
  count_launch(2);
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ----------------------------------------------------------------------------- no-allocation variant
// b200ad_video_pair_sums_static: b200ad_video_pair_sums for callers that cannot allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  The per-block partial sums live in a file-scope device array
// instead of a caller-provided workspace; nothing is allocated, by this library or by any library it calls.
//   * limit: F <= B200AD_STATIC_MAX_FRAMES (64) frames -> 148 blocks * 2 sums * 63 pairs doubles = 149 184 bytes,
//     held by every GPU that loads the library until the process exits;
//   * in host code `g_pair_scratch` is only a host-side stand-in for the array, NOT a device address: the address is
//     taken once per device, with that device current, by cudaGetSymbolAddress, and that pointer goes to the kernels;
//   * all calls on a device share the array.  pair_sums_kernel writes every partial that reduce_rows_kernel reads
//     (each block stores its two sums unconditionally), so nothing has to start at zero and nothing is cleared; the
//     array keeps the last call's partials.  Two calls that overlap on the GPU (different streams without an ordering
//     event) race on it -- serialise them.
static constexpr int kStaticMaxFrames = 64;
static constexpr int kStaticBlocks = 148;          // == b200ad_metric_blocks()
static __device__ __align__(256) double g_pair_scratch[kStaticBlocks * 2 * (kStaticMaxFrames - 1)];

static constexpr int kMaxDevices = 64;
static double* g_pair_scratch_ptr[kMaxDevices];    // per device, filled on first use (0 = not yet)

static int pair_scratch_for_current_device(double** ptr) {
  int dev = -1;
  B200AD_CHECK_CUDA(cudaGetDevice(&dev));
  B200AD_REQUIRE(dev >= 0 && dev < kMaxDevices, "video_pair_sums_static: device ordinal %d is beyond %d", dev, kMaxDevices);
  double* p = __atomic_load_n(&g_pair_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    // the array's device address ON THE CURRENT DEVICE (each device that loads the module has its own copy)
    B200AD_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_pair_scratch));
    B200AD_REQUIRE(p != nullptr, "video_pair_sums_static: cudaGetSymbolAddress returned a null pointer");
    __atomic_store_n(&g_pair_scratch_ptr[dev], p, __ATOMIC_RELEASE);   // same value from every thread: benign if raced
  }
  *ptr = p;
  return 0;
}

extern "C" int b200ad_static_max_frames(void) { return kStaticMaxFrames; }

extern "C" int b200ad_video_pair_sums_static(const unsigned char* frames_fhwc, int F, long long n_per_frame, double* out,
                                             void* stream) {
  static_assert(kStaticBlocks == 148, "keep in step with b200ad_metric_blocks()");
  B200AD_REQUIRE(frames_fhwc && out && F >= 2 && n_per_frame > 0, "video_pair_sums_static: bad arguments");
  B200AD_REQUIRE(F <= kStaticMaxFrames,
                 "video_pair_sums_static: F = %d frames is beyond the static scratch limit of %d frames; use "
                 "b200ad_video_pair_sums with a caller-provided workspace", F, kStaticMaxFrames);
  B200AD_REQUIRE(((uintptr_t)frames_fhwc & 15) == 0 && n_per_frame % 16 == 0,
                 "video_pair_sums_static: frames must be 16-byte aligned with a multiple of 16 bytes per frame");
  B200AD_REQUIRE(b200ad_metric_blocks() == kStaticBlocks, "video_pair_sums_static: block count changed");
  double* scratch = nullptr;
  const int rc = pair_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  const int nb = kStaticBlocks;
  B200AD_CHECK_CUDA(pdl_launch(pair_sums_kernel, dim3(nb, F - 1), dim3(MT), (size_t)0, (cudaStream_t)stream,
                               (const uint8_t*)frames_fhwc, F, n_per_frame, scratch));
  const int rows = 2 * (F - 1);
  B200AD_CHECK_CUDA(pdl_launch(reduce_rows_kernel, dim3((rows * 32 + 127) / 128), dim3(128), (size_t)0,
                               (cudaStream_t)stream, (const double*)scratch, rows, nb, out));
  count_launch(2);
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}
