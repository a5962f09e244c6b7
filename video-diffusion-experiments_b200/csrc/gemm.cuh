// Argument structures shared by the tcgen05 (bf16 operand) and SIMT (fp32 operand) GEMM / conv kernels.
//   C[M,N] = A[M,K] * W[N,K]^T  (+ epilogue), both operands K-major ("row-major A, nn.Linear weight").
// Convolution is the same contraction with A gathered from an NHWC image: K = ntaps * cin, k = tap*cin + c.
#pragma once
#include "common.cuh"

namespace b200ad {

struct ConvGeom {
  int enabled;                 // 0: dense GEMM, A is [M, K] with leading dimension lda
  int nimg, h, w;              // output image geometry; output row = (n*h + y)*w + x
  int cin;                     // channels per tap
  int ntaps;                   // 9 for 3x3
  int tap_dy[9], tap_dx[9];    // source pixel = (y + dy, x + dx) in the source image ...
  int tap_dn[9];               // ... of source image index n + dn (phase-split stride-2 input stacks 4 phases on n)
  int src_nimg, src_h, src_w;  // source tensor geometry [src_nimg, src_h, src_w, cin]
  int n_box, h_box, w_box;     // tcgen05 path: M tile (128 rows) = n_box*h_box*w_box output pixels
  // Sub-pixel output (nearest-x2 upsample + conv3x3 as four 2x2 convolutions, one per output phase): pixel (n, y, x) of
  // this launch is pixel (n, out_scale*y + out_py, out_scale*x + out_px) of an [nimg, out_scale*h, out_scale*w] image.
  // out_scale = 0 or 1: plain output.  Applies to out_f32 / out_op / residual alike (tcgen05 path only).
  int out_scale, out_py, out_px;
  int nphase;                  // 4: all four sub-pixel phases in ONE launch (W = [4][N][K] stacked, taps of phase (0,0)
                               // shifted by (py, px) per phase, out_py/out_px ignored); 0 / 1: the single phase above
  // Second K segment (a 1x1 convolution accumulated into the same tile: ResnetBlock2D.conv2 + conv_shortcut): after the
  // ntaps*cin columns of the taps, K continues with sc_cin channels of `sc_src` [nimg, h, w, sc_cin] (operand-typed,
  // output geometry, no shift).  W is [N, ntaps*cin + sc_cin].  tcgen05 path only.
  const void* sc_src;
  int sc_cin;
  // column statistics: slots per image of the whole output (0: the tiles of this launch) and this launch's first slot
  int stats_slots, stats_slot0;
};

struct Epilogue {
  const float* bias;           // [N] or null
  const float* row_bias;       // row r adds row_bias[(r / rb_div) * ld_rb + col]   (timestep-embedding add)
  int rb_div;
  int ld_rb;
  const float* residual;       // fp32 [M, ld_res] or null (may alias out_f32)
  long long ld_res;
  float* out_f32;              // fp32 output or null
  long long ld_o32;
  void* out_op;                // operand-typed output (bf16 on the tcgen05 path, fp32 on the SIMT path) or null
  long long ld_oop;
  int geglu;                   // 1: columns are interleaved (h0,g0,h1,g1,...); writes N/2 columns h*gelu(g)
  // GroupNorm statistics of the fp32 OUTPUT, emitted by the epilogue (tcgen05 path only): per 32-row slot and column
  // (sum, sum of squares) as float2 col_stats[(image * stats_spi + k) * N + col]; the GroupNorm that consumes out_f32
  // then needs no statistics pass over the tensor (b200ad_gn_apply_stats).  Dense rows: image = row / stats_hw.
  float* col_stats;
  int stats_hw;                // rows per image (dense GEMM; conv uses its own geometry)
};

struct GemmArgs {
  const void* A;               // dense: [M, lda]; conv: NHWC source image
  long long lda;
  const void* W;               // [N, K] K-major
  int M, N, K;
  ConvGeom conv;
  Epilogue epi;
};

int gemm_bf16_tc(const GemmArgs& a, cudaStream_t stream);    // gemm_tc.cu  (tcgen05 + TMA)
int gemm_f32_simt(const GemmArgs& a, cudaStream_t stream);   // simt.cu     (fp32 reference mode)

// Epilogue math on 4 consecutive columns of one row (shared by both kernels).
__device__ __forceinline__ void epilogue_load_addends(const Epilogue& e, long long row, int col, float add[4]) {
  add[0] = add[1] = add[2] = add[3] = 0.f;
  if (e.bias) {
    float4 b = *reinterpret_cast<const float4*>(e.bias + col);
    add[0] += b.x; add[1] += b.y; add[2] += b.z; add[3] += b.w;
  }
  if (e.row_bias) {
    float4 b = *reinterpret_cast<const float4*>(e.row_bias + (row / e.rb_div) * (long long)e.ld_rb + col);
    add[0] += b.x; add[1] += b.y; add[2] += b.z; add[3] += b.w;
  }
}

}  // namespace b200ad
