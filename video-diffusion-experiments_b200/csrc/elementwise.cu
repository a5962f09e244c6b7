// Memory-bound kernels of the denoise step: GroupNorm statistics / apply (+SiLU, + skip concat), LayerNorm
// (+positional embedding), re-layout casts, timestep embedding, conv_in / conv_out, CFG + scheduler update.
// All are coalesced, vectorised where channel grouping allows, with fp32 math; outputs that feed a
// contraction are written in the operand type T (bf16 fast mode, float in the fp32 kernel mode).
#include "../../include/b200ad.h"
#include "ops_common.cuh"

namespace b200ad {

// ----------------------------------------------------------------------------- GroupNorm
// Thread <-> channel-pair mapping is fixed for a whole block (pairs tid + j*T), so every thread keeps its
// group's statistics / affine parameters in registers while the block walks over its rows.
static constexpr int GN_ROWS_PER_BLOCK = 32;
static constexpr int GN_MAX_PPT = 2;       // channel pairs per thread (C <= 4096)

__host__ __device__ inline int gn_nblk(int hw) { return (hw + GN_ROWS_PER_BLOCK - 1) / GN_ROWS_PER_BLOCK; }

__device__ __forceinline__ float2 gn_load_pair(const float* x0, int c0, const float* x1, int c1, long long row,
                                               int c) {
  return (c < c0) ? *reinterpret_cast<const float2*>(x0 + row * c0 + c)
                  : *reinterpret_cast<const float2*>(x1 + row * c1 + (c - c0));
}

__global__ void gn_stats_kernel(const float* __restrict__ x0, int c0, const float* __restrict__ x1, int c1, int hw,
                                int groups, float2* __restrict__ partial) {
  pdl_enter();
  extern __shared__ float2 sh[];          // [npairs]
  const int C = c0 + c1;
  const int npairs = C >> 1;
  const int cpg = C / groups;
  const int n = blockIdx.y;
  const int blk = blockIdx.x;
  const int r0 = blk * GN_ROWS_PER_BLOCK;
  const int r1 = min(hw, r0 + GN_ROWS_PER_BLOCK);
  const int T = blockDim.x;
#pragma unroll
  for (int j = 0; j < GN_MAX_PPT; ++j) {
    const int pr = threadIdx.x + j * T;
    if (pr < npairs) {
      float s = 0.f, q = 0.f;
      const int c = pr * 2;
#pragma unroll 4
      for (int r = r0; r < r1; ++r) {
        float2 v = gn_load_pair(x0, c0, x1, c1, (long long)n * hw + r, c);
        s += v.x + v.y;
        q += v.x * v.x + v.y * v.y;
      }
      sh[pr] = make_float2(s, q);
    }
  }
  __syncthreads();
  if (threadIdx.x < groups) {
    const int g = threadIdx.x;
    const int p0 = g * cpg / 2, p1 = (g + 1) * cpg / 2;
    float s = 0.f, q = 0.f;
    for (int pr = p0; pr < p1; ++pr) { s += sh[pr].x; q += sh[pr].y; }
    partial[((long long)n * gridDim.x + blk) * groups + g] = make_float2(s, q);
  }
}

// One warp per (statistic index, group): deterministic fp64 reduction of the per-block partial sums.
__global__ void gn_finalize_kernel(const float2* __restrict__ partial, int nblk, int frames_per_stat, int groups,
                                   int hw, int cpg, float eps, float2* __restrict__ meanrstd, int nstat) {
  pdl_enter();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nstat * groups) return;
  const int st = warp / groups, g = warp % groups;
  const int total = frames_per_stat * nblk;
  const float2* base = partial + (long long)st * total * groups + g;
  double s = 0.0, q = 0.0;
  for (int i = lane; i < total; i += 32) {
    float2 v = base[(long long)i * groups];
    s += v.x; q += v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) {
    const double cnt = (double)frames_per_stat * hw * cpg;
    const double mean = s / cnt;
    double var = q / cnt - mean * mean;
    if (var < 0.0) var = 0.0;
    meanrstd[warp] = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)eps)));
  }
}

// Group statistics from the per-(slot, column) sums the producing GEMM / conv epilogue emitted (gemm_tc.cu STATS):
// float2 st[(image*spi + k)*c + col].  The input is the channel concatenation of up to two tensors with their own
// statistics buffers.  One block per (image, group) sums that image's slots x the group's channels (fixed order: fp32
// per thread, fp64 across threads => deterministic).  frames_per_stat == 1: writes (mean, rstd) at once; otherwise (the
// motion module: statistics over all frames of a batch item) writes the image's (sum, sum of squares) as doubles and
// gn_combine_frames_kernel adds the frames_per_stat images of a statistic -- one block per (statistic, group) over all
// frames took 57 us at level 0 (64 blocks, 20480 elements each), per image it is a few microseconds.
__global__ void __launch_bounds__(128) gn_finalize_colstats_kernel(const float2* __restrict__ st0, int c0, int spi0,
                                                                   const float2* __restrict__ st1, int c1, int spi1,
                                                                   int groups, int frames_per_stat, int hw, float eps,
                                                                   float2* __restrict__ meanrstd,
                                                                   double2* __restrict__ img_sums) {
  pdl_enter();
  __shared__ double red_s[4], red_q[4];
  const int img = blockIdx.y, g = blockIdx.x;
  const int C = c0 + c1, cpg = C / groups;
  const int ch0 = g * cpg, ch1 = ch0 + cpg;
  double s = 0.0, q = 0.0;
  for (int src = 0; src < 2; ++src) {
    const float2* st = src ? st1 : st0;
    const int cs = src ? c1 : c0, off = src ? c0 : 0, spi = src ? spi1 : spi0;
    if (!st) continue;
    const int a = max(ch0, off) - off, b = min(ch1, off + cs) - off;      // the group's channels inside this source
    if (b <= a) continue;
    const int wd = b - a;
    const float2* base = st + ((long long)img * spi) * cs + a;
    const int total = spi * wd;
    float fs = 0.f, fq = 0.f;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
      const int sl = i / wd, ch = i - sl * wd;
      const float2 v = __ldg(base + (long long)sl * cs + ch);
      fs += v.x; fq += v.y;
    }
    s += fs; q += fq;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red_s[warp] = s; red_q[warp] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    s = (red_s[0] + red_s[1]) + (red_s[2] + red_s[3]);
    q = (red_q[0] + red_q[1]) + (red_q[2] + red_q[3]);
    if (frames_per_stat == 1) {
      const double cnt = (double)hw * cpg;
      const double mean = s / cnt;
      double var = q / cnt - mean * mean;
      if (var < 0.0) var = 0.0;
      meanrstd[img * groups + g] = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)eps)));
    } else {
      img_sums[img * groups + g] = make_double2(s, q);
    }
  }
}

__global__ void gn_combine_frames_kernel(const double2* __restrict__ img_sums, int nstat, int groups, int frames_per_stat,
                                         int hw, int cpg, float eps, float2* __restrict__ meanrstd) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nstat * groups) return;
  const int stat = i / groups, g = i % groups;
  double s = 0.0, q = 0.0;
  for (int f = 0; f < frames_per_stat; ++f) {
    const double2 v = img_sums[(stat * frames_per_stat + f) * groups + g];
    s += v.x; q += v.y;
  }
  const double cnt = (double)frames_per_stat * hw * cpg;
  const double mean = s / cnt;
  double var = q / cnt - mean * mean;
  if (var < 0.0) var = 0.0;
  meanrstd[i] = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)eps)));
}

template <typename T>
__global__ void gn_apply_kernel(const float* __restrict__ x0, int c0, const float* __restrict__ x1, int c1, int hw,
                                int groups, const float2* __restrict__ meanrstd, int frames_per_stat,
                                const float* __restrict__ gamma, const float* __restrict__ beta, int silu,
                                T* __restrict__ out, T* __restrict__ raw) {
  pdl_enter();
  __shared__ float s_mean[64], s_rstd[64];
  const int C = c0 + c1;
  const int npairs = C >> 1;
  const int cpg = C / groups;
  const int n = blockIdx.y;
  const int blk = blockIdx.x;
  if (threadIdx.x < groups) {
    float2 mr = meanrstd[(n / frames_per_stat) * groups + threadIdx.x];
    s_mean[threadIdx.x] = mr.x;
    s_rstd[threadIdx.x] = mr.y;
  }
  __syncthreads();
  const int r0 = blk * GN_ROWS_PER_BLOCK;
  const int r1 = min(hw, r0 + GN_ROWS_PER_BLOCK);
  const int T_ = blockDim.x;
#pragma unroll
  for (int j = 0; j < GN_MAX_PPT; ++j) {
    const int pr = threadIdx.x + j * T_;
    if (pr < npairs) {
      const int c = pr * 2;
      const int g = c / cpg;
      const float mean = s_mean[g], rstd = s_rstd[g];
      const float ga = gamma[c] * rstd, gb = gamma[c + 1] * rstd;
      const float ba = beta[c] - mean * ga, bb = beta[c + 1] - mean * gb;
#pragma unroll 4
      for (int r = r0; r < r1; ++r) {
        const long long row = (long long)n * hw + r;
        float2 v = gn_load_pair(x0, c0, x1, c1, row, c);
        float a = v.x * ga + ba, b = v.y * gb + bb;
        if (silu) { a = silu_f(a); b = silu_f(b); }
        store2<T>(out + row * C + c, a, b);
        if (raw) store2<T>(raw + row * C + c, v.x, v.y);
      }
    }
  }
}

// ---- vectorised variants (C % 4 == 0, c0 % 4 == 0): one float4 (4 channels) per thread and row, R rows of the
// block in flight side by side (thread = (row lane r, vector v)), 8 rows unrolled per thread => 128 bytes of loads
// in flight per thread.  The float2 kernels above were latency-bound at 2.4-3.2 TB/s (32 B in flight per thread,
// profiles/ncu_full_r01f.md: 29 % DRAM utilisation at 52 % occupancy).
__device__ __forceinline__ float4 gn_load_vec(const float* x0, int c0, const float* x1, int c1, long long row, int c) {
  return (c < c0) ? __ldg(reinterpret_cast<const float4*>(x0 + row * c0 + c))
                  : __ldg(reinterpret_cast<const float4*>(x1 + row * c1 + (c - c0)));
}
static constexpr int GN_UNROLL = 8;
static constexpr int GN_SB = 1;            // 32-row sub-blocks per thread block of the vectorised kernels.  4 (128 rows,
                                           // several load batches per thread) did not help at 64^2 (106.9 vs 106.1 us)
                                           // and starved the low-resolution levels of blocks (hw = 256: 44 -> 105 us)

// FIN (one statistic per image, i.e. frames_per_stat == 1): the LAST block of an image to finish -- ticket from a
// self-resetting per-image counter -- reduces the image's partial sums itself, in the order of gn_finalize_kernel
// (same bits), so the separate finalize launch (82 per step, ~6.7 us each for a few KB of work) disappears.
template <bool FIN>
__global__ void gn_stats_vec_kernel(const float* __restrict__ x0, int c0, const float* __restrict__ x1, int c1, int hw,
                                    int groups, int R, float2* __restrict__ partial, unsigned int* __restrict__ counters,
                                    float2* __restrict__ meanrstd, float eps) {
  pdl_enter();
  extern __shared__ float2 sh[];          // [R][C]
  const int C = c0 + c1;
  const int nvec = C >> 2;
  const int cpg = C / groups;
  const int n = blockIdx.y, blk = blockIdx.x * GN_SB;          // first 32-row sub-block of this block
  const int r0 = blk * GN_ROWS_PER_BLOCK;
  const int r1 = min(hw, r0 + GN_SB * GN_ROWS_PER_BLOCK);
  const int v = threadIdx.x % nvec, rl = threadIdx.x / nvec;
  const int c = v * 4;
  const float* base = (c < c0) ? x0 + c : x1 + (c - c0);
  const long long ld = (c < c0) ? c0 : c1;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = r0 + rl; r < r1; r += R * GN_UNROLL) {
    float4 t[GN_UNROLL];
#pragma unroll
    for (int u = 0; u < GN_UNROLL; ++u) {
      const int rr = r + u * R;
      t[u] = rr < r1 ? __ldg(reinterpret_cast<const float4*>(base + ((long long)n * hw + rr) * ld))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < GN_UNROLL; ++u) {
      s.x += t[u].x; s.y += t[u].y; s.z += t[u].z; s.w += t[u].w;
      q.x += t[u].x * t[u].x; q.y += t[u].y * t[u].y; q.z += t[u].z * t[u].z; q.w += t[u].w * t[u].w;
    }
  }
  float2* row = sh + (size_t)rl * C + c;
  row[0] = make_float2(s.x, q.x); row[1] = make_float2(s.y, q.y);
  row[2] = make_float2(s.z, q.z); row[3] = make_float2(s.w, q.w);
  __syncthreads();
  if (threadIdx.x < groups) {
    const int g = threadIdx.x;
    float ss = 0.f, qq = 0.f;
    for (int rr = 0; rr < R; ++rr)
      for (int cc = g * cpg; cc < (g + 1) * cpg; ++cc) { const float2 e = sh[(size_t)rr * C + cc]; ss += e.x; qq += e.y; }
    // the block's sums go to the slot of its first sub-block, the other slots it covers are zero (the finalize
    // kernel adds all nblk slots of an image)
    const int nblk = gn_nblk(hw);
    partial[((long long)n * nblk + blk) * groups + g] = make_float2(ss, qq);
    for (int sb = 1; sb < GN_SB && blk + sb < nblk; ++sb)
      partial[((long long)n * nblk + blk + sb) * groups + g] = make_float2(0.f, 0.f);
  }
  if (FIN) {
    __shared__ int s_last;
    __threadfence();                       // this block's partial sums are visible device-wide before its ticket
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&counters[n], 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last) {
      __threadfence();
      const int nblk = gn_nblk(hw);
      // FULL warps only: the block has R*C/4 threads, which need not be a multiple of 32 (C = 960: 240 threads).  The
      // trailing partial warp used to take part with warp index nw: it re-did groups warp 0 also owns with half of its
      // lanes missing (and a full-mask shuffle over them) and raced warp 0 for the same (mean, rstd) slot -- one
      // image's statistics of four groups came out wrong every few dozen launches (tools/stress_determinism.py).
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
      for (int g = warp; warp < nw && g < groups; g += nw) {
        const float2* base = partial + (long long)n * nblk * groups + g;
        double sd = 0.0, qd = 0.0;
        for (int i = lane; i < nblk; i += 32) {
          const float2 e = __ldcg(&base[(long long)i * groups]);      // written by other SMs: bypass L1
          sd += e.x; qd += e.y;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          sd += __shfl_xor_sync(0xffffffffu, sd, o);
          qd += __shfl_xor_sync(0xffffffffu, qd, o);
        }
        if (lane == 0) {
          const double cnt = (double)hw * cpg;
          const double mean = sd / cnt;
          double var = qd / cnt - mean * mean;
          if (var < 0.0) var = 0.0;
          meanrstd[n * groups + g] = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)eps)));
        }
      }
      if (threadIdx.x == 0) counters[n] = 0;       // ready for the next call (the library runs on one stream)
    }
  }
}

template <typename T>
__global__ void gn_apply_vec_kernel(const float* __restrict__ x0, int c0, const float* __restrict__ x1, int c1, int hw,
                                    int groups, int R, const float2* __restrict__ meanrstd, int frames_per_stat,
                                    const float* __restrict__ gamma, const float* __restrict__ beta, int silu,
                                    T* __restrict__ out, T* __restrict__ raw) {
  pdl_enter();
  const int C = c0 + c1;
  const int nvec = C >> 2;
  const int cpg = C / groups;
  const int n = blockIdx.y, blk = blockIdx.x;
  const int r0 = blk * GN_SB * GN_ROWS_PER_BLOCK;
  const int r1 = min(hw, r0 + GN_SB * GN_ROWS_PER_BLOCK);
  const int v = threadIdx.x % nvec, rl = threadIdx.x / nvec;
  const int c = v * 4;
  const float2* mr = meanrstd + (n / frames_per_stat) * groups;
  float sc[4], sf[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 m = mr[(c + j) / cpg];
    sc[j] = gamma[c + j] * m.y;
    sf[j] = beta[c + j] - m.x * sc[j];
  }
  const float* base = (c < c0) ? x0 + c : x1 + (c - c0);
  const long long ld = (c < c0) ? c0 : c1;
  for (int r = r0 + rl; r < r1; r += R * GN_UNROLL) {
    float4 t[GN_UNROLL];
#pragma unroll
    for (int u = 0; u < GN_UNROLL; ++u) {
      const int rr = r + u * R;
      if (rr < r1) t[u] = __ldg(reinterpret_cast<const float4*>(base + ((long long)n * hw + rr) * ld));
    }
#pragma unroll
    for (int u = 0; u < GN_UNROLL; ++u) {
      const int rr = r + u * R;
      if (rr >= r1) break;
      const long long o = ((long long)n * hw + rr) * C + c;
      float4 y = make_float4(fmaf(t[u].x, sc[0], sf[0]), fmaf(t[u].y, sc[1], sf[1]), fmaf(t[u].z, sc[2], sf[2]),
                             fmaf(t[u].w, sc[3], sf[3]));
      if (silu) {
        y.x = __fdividef(y.x, 1.0f + __expf(-y.x)); y.y = __fdividef(y.y, 1.0f + __expf(-y.y));
        y.z = __fdividef(y.z, 1.0f + __expf(-y.z)); y.w = __fdividef(y.w, 1.0f + __expf(-y.w));
      }
      store4<T>(out + o, y);
      if (raw) store4<T>(raw + o, t[u]);
    }
  }
}

// Persistent apply: a block keeps its thread <-> (row lane, 4-channel vector) mapping for the whole launch and walks
// row blocks of R*GN_PU rows (grid-stride), with the NEXT row block's loads in flight while the current one is
// normalised and stored -- the short-lived blocks of gn_apply_vec_kernel (32 rows each, ~60 KB of traffic, a couple of
// microseconds of life) left the memory pipe idle at every block turn-over (49 % DRAM utilisation at 28 % occupancy).
static constexpr int GN_PU = 4;            // rows per thread and row block
template <typename T>
__global__ void __launch_bounds__(640) gn_apply_persist_kernel(const float* __restrict__ x0, int c0,
                                                               const float* __restrict__ x1, int c1, int hw, int groups,
                                                               int R, const float2* __restrict__ meanrstd,
                                                               int frames_per_stat, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, int silu,
                                                               T* __restrict__ out, T* __restrict__ raw, int nrb,
                                                               long long total) {
  pdl_enter();
  const int C = c0 + c1;
  const int nvec = C >> 2;
  const int cpg = C / groups;
  const int v = threadIdx.x % nvec, rl = threadIdx.x / nvec;
  const int c = v * 4;
  const float4 g4 = *reinterpret_cast<const float4*>(gamma + c);
  const float4 b4 = *reinterpret_cast<const float4*>(beta + c);
  const int gi0 = c / cpg, gi1 = (c + 1) / cpg, gi2 = (c + 2) / cpg, gi3 = (c + 3) / cpg;
  const float* base = (c < c0) ? x0 + c : x1 + (c - c0);
  const long long ld = (c < c0) ? c0 : c1;
  const int rows_per_rb = R * GN_PU;
  float4 cur[GN_PU], nxt[GN_PU];
  auto load = [&](long long wb, float4* t) {
    const int n = (int)(wb / nrb), r0 = (int)(wb % nrb) * rows_per_rb;
#pragma unroll
    for (int u = 0; u < GN_PU; ++u) {
      const int rr = r0 + rl + u * R;
      t[u] = rr < hw ? __ldg(reinterpret_cast<const float4*>(base + ((long long)n * hw + rr) * ld))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  long long wb = blockIdx.x;
  if (wb < total) load(wb, cur);
  for (; wb < total; wb += gridDim.x) {
    const long long wn = wb + gridDim.x;
    if (wn < total) load(wn, nxt);
    const int n = (int)(wb / nrb), r0 = (int)(wb % nrb) * rows_per_rb;
    const float2* mr = meanrstd + (n / frames_per_stat) * groups;
    const float2 m0 = mr[gi0], m1 = mr[gi1], m2 = mr[gi2], m3 = mr[gi3];
    const float s0 = g4.x * m0.y, s1 = g4.y * m1.y, s2 = g4.z * m2.y, s3 = g4.w * m3.y;
    const float f0 = b4.x - m0.x * s0, f1 = b4.y - m1.x * s1, f2 = b4.z - m2.x * s2, f3 = b4.w - m3.x * s3;
#pragma unroll
    for (int u = 0; u < GN_PU; ++u) {
      const int rr = r0 + rl + u * R;
      if (rr < hw) {
        const long long o = ((long long)n * hw + rr) * C + c;
        float4 y = make_float4(fmaf(cur[u].x, s0, f0), fmaf(cur[u].y, s1, f1), fmaf(cur[u].z, s2, f2),
                               fmaf(cur[u].w, s3, f3));
        if (silu) {
          y.x = __fdividef(y.x, 1.0f + __expf(-y.x)); y.y = __fdividef(y.y, 1.0f + __expf(-y.y));
          y.z = __fdividef(y.z, 1.0f + __expf(-y.z)); y.w = __fdividef(y.w, 1.0f + __expf(-y.w));
        }
        store4<T>(out + o, y);
        if (raw) store4<T>(raw + o, cur[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < GN_PU; ++u) cur[u] = nxt[u];
  }
}

// rows of a block processed side by side by the vectorised kernels (block = R * C/4 threads, ~256..640)
static int gn_vec_R(int C) {
  const int nvec = C / 4;
  int R = (288 + nvec / 2) / nvec;
  if (R < 1) R = 1;
  if (R > 4) R = 4;
  return R;
}
static bool gn_vec_ok(int C, int c0, int groups) { return C % 4 == 0 && c0 % 4 == 0 && C / 4 <= 1024 && C >= 128; }

static int gn_threads(int C) {
  const int npairs = C / 2;
  int t = (npairs + GN_MAX_PPT - 1) / GN_MAX_PPT;
  if (npairs <= 1024) t = npairs;
  t = (t + 31) / 32 * 32;
  return t < 64 ? 64 : t;
}

}  // namespace b200ad

using namespace b200ad;

extern "C" int b200ad_gn_nblk(int hw) { return gn_nblk(hw); }

static constexpr int GN_MAX_COUNTERS = 8192;
// Per-device ticket counters of the fused finalize (zeroed once, self-resetting).  Not created while a stream capture
// is in progress (cudaMalloc is illegal there): the caller then keeps the separate finalize kernel.
static unsigned int* gn_counters(cudaStream_t st) {
  static unsigned int* buf[16] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  if (!buf[dev]) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return nullptr;
    unsigned int* p = nullptr;
    if (cudaMalloc(&p, GN_MAX_COUNTERS * sizeof(unsigned int)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cudaMemset(p, 0, GN_MAX_COUNTERS * sizeof(unsigned int)) != cudaSuccess) { cudaGetLastError(); cudaFree(p); return nullptr; }
    buf[dev] = p;
  }
  return buf[dev];
}

// diagnostics: number of non-zero ticket counters on the current device (0 whenever no GroupNorm is in flight);
// synchronises the device.  -1: counters not created yet.
extern "C" int b200ad_gn_counters_nonzero(void) {
  unsigned int* c = gn_counters(nullptr);
  if (!c) return -1;
  static unsigned int host[GN_MAX_COUNTERS];
  if (cudaDeviceSynchronize() != cudaSuccess) return -2;
  if (cudaMemcpy(host, c, sizeof(host), cudaMemcpyDeviceToHost) != cudaSuccess) return -2;
  int nz = 0;
  for (int i = 0; i < GN_MAX_COUNTERS; ++i) nz += host[i] != 0;
  return nz;
}

static int gn_stats_launch(const float* x0, int c0, const float* x1, int c1, int nimg, int hw, int groups,
                           float* partial, void* stream, unsigned int* counters, float2* meanrstd, float eps);

extern "C" int b200ad_gn_stats(const float* x0, int c0, const float* x1, int c1, int nimg, int hw, int groups,
                               float* partial, void* stream) {
  return gn_stats_launch(x0, c0, x1, c1, nimg, hw, groups, partial, stream, nullptr, nullptr, 0.f);
}

static int gn_stats_launch(const float* x0, int c0, const float* x1, int c1, int nimg, int hw, int groups,
                           float* partial, void* stream, unsigned int* counters, float2* meanrstd, float eps) {
  if (!x1) c1 = 0;
  const int C = c0 + c1;
  B200AD_REQUIRE(groups <= 64 && C % groups == 0 && (C / groups) % 2 == 0 && c0 % 2 == 0,
                 "gn_stats: unsupported channels C=%d (c0=%d) groups=%d", C, c0, groups);
  B200AD_REQUIRE(C / 2 <= 1024 * GN_MAX_PPT, "gn_stats: C=%d too large", C);
  dim3 grid(gn_nblk(hw), nimg);
  if (gn_vec_ok(C, c0, groups)) {
    const int R = gn_vec_R(C);
    const dim3 vgrid((gn_nblk(hw) + GN_SB - 1) / GN_SB, nimg);
    // the ticket counters must be zero when the kernel starts: clear them in stream order (a residue -- e.g. from a
    // launch on another stream that overlapped this one -- would make an image finalize early on stale partial sums)
    if (counters && GN_SB == 1)
      B200AD_CHECK_CUDA(cudaMemsetAsync(counters, 0, (size_t)nimg * sizeof(unsigned int), (cudaStream_t)stream));
    if (counters && GN_SB == 1)
      pdl_launch(gn_stats_vec_kernel<true>, dim3(vgrid), dim3(R * (C / 4)), (size_t)R * C * sizeof(float2), (cudaStream_t)stream,
                 x0, c0, x1, c1, hw, groups, R, reinterpret_cast<float2*>(partial), counters, meanrstd, eps);
    else
      pdl_launch(gn_stats_vec_kernel<false>, dim3(vgrid), dim3(R * (C / 4)), (size_t)R * C * sizeof(float2), (cudaStream_t)stream,
                 x0, c0, x1, c1, hw, groups, R, reinterpret_cast<float2*>(partial), (unsigned int*)nullptr, (float2*)nullptr, 0.f);
  } else
    pdl_launch(gn_stats_kernel, dim3(grid), dim3(gn_threads(C)), (size_t)((C / 2) * sizeof(float2)), (cudaStream_t)((cudaStream_t)stream), 
        x0, c0, x1, c1, hw, groups, reinterpret_cast<float2*>(partial));
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int gn_apply_launch(int op_dtype, const float* x0, int c0, const float* x1, int c1, int nimg, int hw,
                           int groups, const float* partial, int frames_per_stat, float eps, const float* gamma,
                           const float* beta, int silu, void* out_op, void* raw_op, void* stream, bool finalize,
                           const float2* meanrstd_in = nullptr);

extern "C" int b200ad_gn_apply(int op_dtype, const float* x0, int c0, const float* x1, int c1, int nimg, int hw,
                               int groups, const float* partial, int frames_per_stat, float eps, const float* gamma,
                               const float* beta, int silu, void* out_op, void* raw_op, void* stream) {
  return gn_apply_launch(op_dtype, x0, c0, x1, c1, nimg, hw, groups, partial, frames_per_stat, eps, gamma, beta, silu,
                         out_op, raw_op, stream, true);
}

static int gn_apply_launch(int op_dtype, const float* x0, int c0, const float* x1, int c1, int nimg, int hw,
                           int groups, const float* partial, int frames_per_stat, float eps, const float* gamma,
                           const float* beta, int silu, void* out_op, void* raw_op, void* stream, bool finalize,
                           const float2* meanrstd_in) {
  if (!x1) c1 = 0;
  const int C = c0 + c1;
  B200AD_REQUIRE(groups <= 64 && C % groups == 0 && (C / groups) % 2 == 0 && c0 % 2 == 0,
                 "gn_apply: unsupported channels C=%d (c0=%d) groups=%d", C, c0, groups);
  B200AD_REQUIRE(frames_per_stat >= 1 && nimg % frames_per_stat == 0, "gn_apply: nimg=%d frames_per_stat=%d", nimg,
                 frames_per_stat);
  dim3 grid(gn_nblk(hw), nimg);
  const int nblk = gn_nblk(hw);
  const int nstat = nimg / frames_per_stat;
  // mean / rstd live in the tail of the caller's `partial` buffer (see b200ad.h)
  float2* part = reinterpret_cast<float2*>(const_cast<float*>(partial));
  float2* meanrstd = meanrstd_in ? const_cast<float2*>(meanrstd_in) : part + (long long)nimg * nblk * groups;
  if (finalize) {
    const int warps = nstat * groups;
    const int threads = 256;
    pdl_launch(gn_finalize_kernel, dim3((warps * 32 + threads - 1) / threads), dim3(threads), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
        part, nblk, frames_per_stat, groups, hw, C / groups, eps, meanrstd, nstat);
    count_launch();
  }
  // measured (tools/kbench.py norm): the persistent kernel wins on the large level-0 tensors (68.7 vs 73.0 us at
  // 32 x 64^2 x 320) and loses on the small ones (45.1 vs 43.8 us at 32^2 x 640, 31.7 vs 28.3 us at 16^2 x 1280)
  const bool persist = (double)nimg * hw * C * 4.0 >= 128e6;
  if (gn_vec_ok(C, c0, groups) && persist) {
    const int R = gn_vec_R(C);
    const int threads = R * (C / 4);
    const int nrb = (hw + R * GN_PU - 1) / (R * GN_PU);
    const long long total = (long long)nimg * nrb;
    static int sms = 0;
    if (!sms) {
      int dev = 0;
      B200AD_CHECK_CUDA(cudaGetDevice(&dev));
      B200AD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    int per_sm = 2048 / threads;                    // resident blocks (<= 64 registers per thread)
    if (per_sm > 6) per_sm = 6;
    if (per_sm < 1) per_sm = 1;
    const long long want = (long long)sms * per_sm;
    const int blocks = (int)(total < want ? total : want);
    if (op_dtype == 0)
      pdl_launch(gn_apply_persist_kernel<bf16>, dim3(blocks), dim3(threads), (size_t)0, (cudaStream_t)stream, x0, c0, x1, c1,
                 hw, groups, R, (const float2*)meanrstd, frames_per_stat, gamma, beta, silu, (bf16*)out_op, (bf16*)raw_op,
                 nrb, total);
    else
      pdl_launch(gn_apply_persist_kernel<float>, dim3(blocks), dim3(threads), (size_t)0, (cudaStream_t)stream, x0, c0, x1,
                 c1, hw, groups, R, (const float2*)meanrstd, frames_per_stat, gamma, beta, silu, (float*)out_op,
                 (float*)raw_op, nrb, total);
  } else if (gn_vec_ok(C, c0, groups)) {
    const int R = gn_vec_R(C);
    const dim3 vgrid((gn_nblk(hw) + GN_SB - 1) / GN_SB, nimg);
    if (op_dtype == 0)
      pdl_launch(gn_apply_vec_kernel<bf16>, dim3(vgrid), dim3(R * (C / 4)), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
          x0, c0, x1, c1, hw, groups, R, meanrstd, frames_per_stat, gamma, beta, silu, (bf16*)out_op, (bf16*)raw_op);
    else
      pdl_launch(gn_apply_vec_kernel<float>, dim3(vgrid), dim3(R * (C / 4)), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
          x0, c0, x1, c1, hw, groups, R, meanrstd, frames_per_stat, gamma, beta, silu, (float*)out_op, (float*)raw_op);
  } else if (op_dtype == 0)
    pdl_launch(gn_apply_kernel<bf16>, dim3(grid), dim3(gn_threads(C)), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
        x0, c0, x1, c1, hw, groups, meanrstd, frames_per_stat, gamma, beta, silu, (bf16*)out_op, (bf16*)raw_op);
  else
    pdl_launch(gn_apply_kernel<float>, dim3(grid), dim3(gn_threads(C)), (size_t)(0), (cudaStream_t)((cudaStream_t)stream), 
        x0, c0, x1, c1, hw, groups, meanrstd, frames_per_stat, gamma, beta, silu, (float*)out_op, (float*)raw_op);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" long long b200ad_gn_stats_ws_bytes(int nimg, int groups, int frames_per_stat) {
  const long long nstat = nimg / (frames_per_stat > 0 ? frames_per_stat : 1);
  const long long mr = ((nstat * groups + 1) & ~1LL) * (long long)sizeof(float2);
  return mr + (frames_per_stat > 1 ? (long long)nimg * groups * (long long)sizeof(double2) : 0);
}

// GroupNorm whose statistics come from the producer's epilogue: finalize (tiny) + apply; the fp32 input is read ONCE.
extern "C" int b200ad_gn_apply_stats(int op_dtype, const float* x0, int c0, const float* st0, int spi0,
                                     const float* x1, int c1, const float* st1, int spi1, int nimg, int hw, int groups,
                                     int frames_per_stat, float eps, const float* gamma, const float* beta, int silu,
                                     float* meanrstd_ws, void* out_op, void* raw_op, void* stream) {
  if (!x1) { c1 = 0; st1 = nullptr; }
  const int C = c0 + c1;
  B200AD_REQUIRE(st0 && (c1 == 0 || st1) && meanrstd_ws, "gn_apply_stats: null statistics / workspace");
  B200AD_REQUIRE(groups <= 64 && C % groups == 0, "gn_apply_stats: C=%d groups=%d", C, groups);
  B200AD_REQUIRE(frames_per_stat >= 1 && nimg % frames_per_stat == 0, "gn_apply_stats: nimg=%d frames_per_stat=%d", nimg,
                 frames_per_stat);
  const int nstat = nimg / frames_per_stat;
  // workspace: [nstat*groups] float2 (mean, rstd), then (frames_per_stat > 1) [nimg*groups] double2 image sums
  float2* mr = reinterpret_cast<float2*>(meanrstd_ws);
  double2* sums = reinterpret_cast<double2*>(mr + (((size_t)nstat * groups + 1) & ~(size_t)1));
  pdl_launch(gn_finalize_colstats_kernel, dim3(groups, nimg), dim3(128), (size_t)0, (cudaStream_t)stream,
             reinterpret_cast<const float2*>(st0), c0, spi0, reinterpret_cast<const float2*>(st1), c1, spi1, groups,
             frames_per_stat, hw, eps, mr, sums);
  count_launch();
  if (frames_per_stat > 1) {
    const int n = nstat * groups;
    pdl_launch(gn_combine_frames_kernel, dim3((n + 127) / 128), dim3(128), (size_t)0, (cudaStream_t)stream,
               (const double2*)sums, nstat, groups, frames_per_stat, hw, C / groups, eps, mr);
    count_launch();
  }
  B200AD_CHECK_CUDA(cudaGetLastError());
  return gn_apply_launch(op_dtype, x0, c0, x1, c1, nimg, hw, groups, nullptr, frames_per_stat, eps, gamma, beta, silu,
                         out_op, raw_op, stream, false, reinterpret_cast<const float2*>(meanrstd_ws));
}

// Fused GroupNorm: statistics + finalize + apply in one call (same workspace layout as gn_stats/gn_apply).  (A walk
// over L2-sized sub-batches measured slower -- 145 vs 106 us at 32 x 64^2 x 320, profiles/kbench_r01j.txt -- and is gone;
// the two-pass path itself now only serves inputs whose producer emits no column statistics: conv_in's output.)
extern "C" int b200ad_group_norm(int op_dtype, const float* x0, int c0, const float* x1, int c1, int nimg, int hw,
                                 int groups, float* partial, int frames_per_stat, float eps, const float* gamma,
                                 const float* beta, int silu, void* out_op, void* raw_op, void* stream) {
  if (!x1) c1 = 0;
  const int C = c0 + c1;
  B200AD_REQUIRE(frames_per_stat >= 1 && nimg % frames_per_stat == 0, "group_norm: nimg=%d frames_per_stat=%d", nimg,
                 frames_per_stat);
  const int chunk = nimg;
  const int nblk = gn_nblk(hw);
  const size_t esz = op_dtype == 0 ? 2 : 4;
  // per-call layout of `partial`: [nimg*nblk*groups] float2 partial sums, then [nimg*groups] float2 (mean, rstd);
  // a sub-batch uses a prefix of each region
  for (int n0 = 0; n0 < nimg; n0 += chunk) {
    const int nn = (nimg - n0 < chunk) ? nimg - n0 : chunk;
    const long long roff = (long long)n0 * hw;
    const float* a0 = x0 + roff * c0;
    const float* a1 = x1 ? x1 + roff * c1 : nullptr;
    // one statistic per image: the last statistics block of each image finalizes it (no finalize launch)
    unsigned int* counters = nullptr;
    if (frames_per_stat == 1 && nn <= GN_MAX_COUNTERS && gn_vec_ok(C, c0, groups) && groups <= 64)
      counters = gn_counters((cudaStream_t)stream);
    float2* meanrstd = reinterpret_cast<float2*>(partial) + (long long)nn * nblk * groups;
    if (int rc = gn_stats_launch(a0, c0, a1, c1, nn, hw, groups, partial, stream, counters, meanrstd, eps)) return rc;
    char* o = reinterpret_cast<char*>(out_op) + (size_t)roff * C * esz;
    char* r = raw_op ? reinterpret_cast<char*>(raw_op) + (size_t)roff * C * esz : nullptr;
    if (int rc = gn_apply_launch(op_dtype, a0, c0, a1, c1, nn, hw, groups, partial, frames_per_stat, eps, gamma, beta,
                                 silu, o, r, stream, counters == nullptr))
      return rc;
  }
  return 0;
}
