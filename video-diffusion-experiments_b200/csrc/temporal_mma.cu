// Motion-module temporal self-attention (bf16 operand mode): batched small-sequence attention over the frame axis.
//   For every spatial position (b, s) and head: O[F,d] = softmax(Q K^T / sqrt(d)) V with F <= 32 frames.
// The kernel is HBM-bound (read q|k|v once, write o once: 8*C bytes per row); the F x F contraction is 0.14 % of
// the step's FLOPs, so it runs on warp-level mma (m16n8k16, fp32 accumulate) out of shared memory rather than on
// a 128-row tcgen05 tile that would be 7/8 padding.  The frame-axis "permute" of diffusers
// ([B*F, hw, C] -> [B*hw, F, C]) is never materialised: the F rows of one position are F strided row segments of the
// (b, f, s)-ordered activation matrix, gathered with 16-byte cp.async and scattered back with 16-byte stores.
// Block = PB positions x `hpb` heads, one warp per (position, head).
#include "../../include/b200ad.h"
#include "ops_common.cuh"

namespace b200ad {

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2_t(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// FT = number of 16-row frame tiles (F <= 16*FT).  D = head width as a compile-time constant (40 / 80 / 160: the
// k-step, n-tile and half-step logic unrolls and folds away) or 0 for the run-time value `d_rt`.  The kernel is
// issue-bound (ncu: 86 % issue-active at 48 % DRAM, ~1040 warp instructions per (position, head)), so every loop
// counter, select and slow-path math call (exp2f, IEEE division) removed is time.
template <int FT, int D>
__global__ void __launch_bounds__(256)
temporal_attn_mma_kernel(const bf16* __restrict__ qkv, int F, int hw, int C, int d_rt, int hpb, int PB, int npos,
                         float scale_log2, bf16* __restrict__ out) {
  pdl_enter();
  const int d = D ? D : d_rt;
  extern __shared__ __align__(16) uint8_t tsm[];
  constexpr int FP = 16 * FT;
  const int W = hpb * d;                 // columns of this head group, per part
  const int RS = 3 * W + 8;              // smem row stride in elements (+16 B: ldmatrix rows land in distinct banks)
  bf16* sm = reinterpret_cast<bf16*>(tsm);
  const int col0 = blockIdx.y * W;
  const int pos0 = blockIdx.x * PB;
  const int C3 = 3 * C;
  const int cpp = W >> 3;                // 16-byte chunks per part per row

  // ---- gather: PB positions x FP rows x (q|k|v) segments; rows >= F are zero (finite padding for the MMAs).
  // Index math is incremental (one division per thread, then adds and compares): with `i / cpp`, `% 3`, `/ hw`
  // per 16-byte chunk the integer divisions were 60 % of the kernel's instructions and made it issue-bound
  // (profiles/ncu_full_r01f.md: 118 M warp instructions, SM 75 % busy at 26 % DRAM utilisation).
  __shared__ long long pos_base[8], out_base[8];
  if (threadIdx.x < PB) {
    const int pos = pos0 + threadIdx.x;
    const int b = pos < npos ? pos / hw : 0, sp = pos < npos ? pos - b * hw : 0;
    pos_base[threadIdx.x] = pos < npos ? ((long long)b * F * hw + sp) * C3 + col0 : -1;
    out_base[threadIdx.x] = ((long long)b * F * hw + sp) * C + col0;
  }
  __syncthreads();
  const int rowlen = 3 * cpp;                          // 16-byte chunks per smem row
  if ((int)blockDim.x >= rowlen) {
    // every thread keeps ONE 16-byte column of the (q|k|v) row and walks down the rows, blockDim / rowlen rows per
    // pass (the last blockDim % rowlen threads idle): part / chunk / addresses are computed once, the loop body is
    // an add, a compare and the copy
    const int nrows = PB * FP;
    const int drow = (int)blockDim.x / rowlen;
    int row = (int)threadIdx.x / rowlen;
    const int col = (int)threadIdx.x - row * rowlen;
    if (row < drow) {
      const int part = (col >= cpp) + (col >= 2 * cpp);
      const int ch = col - part * cpp;
      const int soff = part * W + ch * 8;              // element offset inside an smem row
      const int goff = part * C + ch * 8;              // element offset inside a global (q|k|v) row
      const long long fstride = (long long)hw * C3;
      for (; row < nrows; row += drow) {
        const int f = row & (FP - 1), pp = row / FP;
        bf16* dst = sm + (size_t)row * RS + soff;
        const long long pb = pos_base[pp];
        if (f < F && pb >= 0) cp_async16(dst, qkv + pb + f * fstride + goff);
        else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  } else {
    const int nrows = PB * FP;
    const int dcol = (int)blockDim.x % rowlen, drow = (int)blockDim.x / rowlen;
    int row = (int)threadIdx.x / rowlen, col = (int)threadIdx.x - row * rowlen;
    const long long fstride = (long long)hw * C3;
    for (; row < nrows; ) {
      const int part = (col >= cpp) + (col >= 2 * cpp);
      const int ch = col - part * cpp;
      const int f = row & (FP - 1), pp = row / FP;     // FP is a compile-time power of two
      bf16* dst = sm + (size_t)row * RS + part * W + ch * 8;
      const long long pb = pos_base[pp];
      if (f < F && pb >= 0) cp_async16(dst, qkv + pb + f * fstride + part * C + ch * 8);
      else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
      col += dcol; row += drow;
      if (col >= rowlen) { col -= rowlen; ++row; }
    }
  }
  cp_async_wait_all();
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = warp / hpb, hl = warp - p * hpb;
  if (p < PB && pos0 + p < npos) {
    bf16* base = sm + (size_t)p * FP * RS;
    const uint32_t q_s = smem_u32(base + hl * d);
    const uint32_t k_s = smem_u32(base + W + hl * d);
    const uint32_t v_s = smem_u32(base + 2 * W + hl * d);
    const int g = lane >> 2, t = lane & 3;
    // ---- S = Q K^T: FT m-tiles x 2FT n-tiles, k over d in steps of 16 (a trailing half step is zeroed)
    float S[FT][2 * FT][4];
#pragma unroll
    for (int mt = 0; mt < FT; ++mt)
#pragma unroll
      for (int nt = 0; nt < 2 * FT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) S[mt][nt][i] = 0.f;
    const int ksteps = (d + 15) >> 4;                 // compile-time when D != 0
    const uint32_t a_off = (uint32_t)(((lane & 15) * RS + (lane >> 4) * 8) * 2);
    const uint32_t b_off = (uint32_t)((((lane & 7) + ((lane >> 4) << 3)) * RS + ((lane >> 3) & 1) * 8) * 2);
#pragma unroll
    for (int ks = 0; ks < (D ? (D + 15) / 16 : ksteps); ++ks) {
      const bool half = (ks * 16 + 8 >= d);           // only columns [16ks, 16ks+8) belong to this head
      uint32_t a[FT][4];
#pragma unroll
      for (int mt = 0; mt < FT; ++mt) {
        ldsm_x4(q_s + a_off + (uint32_t)((mt * 16 * RS + ks * 16) * 2), a[mt][0], a[mt][1], a[mt][2], a[mt][3]);
        if (half) { a[mt][2] = 0u; a[mt][3] = 0u; }
      }
#pragma unroll
      for (int np = 0; np < FT; ++np) {               // pairs of key n-tiles
        uint32_t b0, b1, b2, b3;
        ldsm_x4(k_s + b_off + (uint32_t)((np * 16 * RS + ks * 16) * 2), b0, b1, b2, b3);
        if (half) { b1 = 0u; b3 = 0u; }
#pragma unroll
        for (int mt = 0; mt < FT; ++mt) {
          mma_bf16_16816(S[mt][2 * np], a[mt], b0, b1);
          mma_bf16_16816(S[mt][2 * np + 1], a[mt], b2, b3);
        }
      }
    }
    // ---- softmax over the keys (row = frame g / g+8 of each m-tile; a row lives in one quad of lanes)
    uint32_t P[FT][FT][4];                            // bf16 A fragments of exp(S - max), [m-tile][key k-step]
    float inv_l[FT][2];
#pragma unroll
    for (int mt = 0; mt < FT; ++mt) {
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 2 * FT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int col = nt * 8 + 2 * t + (i & 1);
          if (F < FP && col >= F) S[mt][nt][i] = -INFINITY;
          if (i < 2) mx0 = fmaxf(mx0, S[mt][nt][i]); else mx1 = fmaxf(mx1, S[mt][nt][i]);
        }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float o0 = mx0 * scale_log2, o1 = mx1 * scale_log2;
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 2 * FT; ++nt) {
        // arguments are <= 0 (the row maximum is subtracted); -inf (masked column) -> 0
        const float p0 = ex2_fast(S[mt][nt][0] * scale_log2 - o0), p1 = ex2_fast(S[mt][nt][1] * scale_log2 - o0);
        const float p2 = ex2_fast(S[mt][nt][2] * scale_log2 - o1), p3 = ex2_fast(S[mt][nt][3] * scale_log2 - o1);
        l0 += p0 + p1;
        l1 += p2 + p3;
        P[mt][nt >> 1][(nt & 1) * 2 + 0] = pack_bf16x2(p0, p1);
        P[mt][nt >> 1][(nt & 1) * 2 + 1] = pack_bf16x2(p2, p3);
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
      l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      inv_l[mt][0] = rcp_fast(l0);                  // l >= 1 (the maximum's own term)
      inv_l[mt][1] = rcp_fast(l1);
    }
    __syncwarp();                                     // every lane's Q fragments are loaded: the Q slot may be reused
    // ---- O = P V, two 8-column n-tiles at a time; results go to this warp's own (dead) Q slot in smem
    const uint32_t v_off = (uint32_t)((((lane & 7) + ((lane >> 3) & 1) * 8) * RS + (lane >> 4) * 8) * 2);
    bf16* o_s = base + hl * d;
    const int ntiles = d >> 3;
#pragma unroll
    for (int n0 = 0; n0 < (D ? D / 8 : ntiles); n0 += 2) {
      const bool pair = (n0 + 1 < ntiles);
      float O[FT][2][4];
#pragma unroll
      for (int mt = 0; mt < FT; ++mt)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int i = 0; i < 4; ++i) O[mt][j][i] = 0.f;
#pragma unroll
      for (int ks = 0; ks < FT; ++ks) {               // key k-steps of 16
        uint32_t b0, b1, b2 = 0u, b3 = 0u;
        const uint32_t addr = v_s + (uint32_t)((ks * 16 * RS + n0 * 8) * 2);
        if (pair) ldsm_x4_t(addr + v_off, b0, b1, b2, b3);
        else ldsm_x2_t(addr + (uint32_t)(((lane & 15) * RS) * 2), b0, b1);
#pragma unroll
        for (int mt = 0; mt < FT; ++mt) {
          mma_bf16_16816(O[mt][0], P[mt][ks], b0, b1);
          if (pair) mma_bf16_16816(O[mt][1], P[mt][ks], b2, b3);
        }
      }
#pragma unroll
      for (int mt = 0; mt < FT; ++mt)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (j == 1 && !pair) continue;
          bf16* dst = o_s + (size_t)(mt * 16 + g) * RS + (n0 + j) * 8 + 2 * t;
          *reinterpret_cast<uint32_t*>(dst) = pack_bf16x2(O[mt][j][0] * inv_l[mt][0], O[mt][j][1] * inv_l[mt][0]);
          *reinterpret_cast<uint32_t*>(dst + 8 * RS) =
              pack_bf16x2(O[mt][j][2] * inv_l[mt][1], O[mt][j][3] * inv_l[mt][1]);
        }
    }
  }
  __syncthreads();
  // ---- scatter: the q part of smem now holds O for all heads of the group (same incremental indexing)
  {
    const int dcol = (int)blockDim.x % cpp, drow = (int)blockDim.x / cpp;
    const long long fstride = (long long)hw * C;
    for (int pp = 0; pp < PB; ++pp) {
      const long long pb = pos_base[pp];
      if (pb < 0) break;
      const long long ob = out_base[pp];
      int f = (int)threadIdx.x / cpp, ch = (int)threadIdx.x - f * cpp;
      for (; f < F; ) {
        const uint4 v = *reinterpret_cast<const uint4*>(sm + (size_t)(pp * FP + f) * RS + ch * 8);
        *reinterpret_cast<uint4*>(out + ob + f * fstride + ch * 8) = v;
        ch += dcol; f += drow;
        if (ch >= cpp) { ch -= cpp; ++f; }
      }
    }
  }
}

int temporal_attention_bf16_mma(const bf16* qkv, int B, int F, int hw, int C, int heads, bf16* out,
                                cudaStream_t st) {
  const int d = C / heads;
  B200AD_REQUIRE(d % 8 == 0, "temporal_attention: head width %d must be a multiple of 8 on the bf16 path", d);
  int hpb = heads > 8 ? 8 : heads;
  while (hpb > 1 && (hpb * d > 320 || heads % hpb != 0)) --hpb;
  const int FT = F <= 16 ? 1 : 2;
  const size_t per_pos = (size_t)(16 * FT) * (3 * hpb * d + 8) * sizeof(bf16);
  int PB = 8 / hpb;
  if (PB < 1) PB = 1;
  while (PB > 1 && PB * per_pos > 64 * 1024) PB >>= 1;
  const size_t smem = PB * per_pos;
  B200AD_REQUIRE(smem <= 200 * 1024, "temporal_attention: needs %zu bytes of shared memory", smem);
  static bool attr_set = false;
  if (!attr_set) {
#define TA_ATTR(FTV, DV) B200AD_CHECK_CUDA((cudaFuncSetAttribute(temporal_attn_mma_kernel<FTV, DV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)))
    TA_ATTR(1, 0); TA_ATTR(2, 0); TA_ATTR(1, 40); TA_ATTR(1, 80); TA_ATTR(1, 160); TA_ATTR(2, 40); TA_ATTR(2, 80); TA_ATTR(2, 160);
#undef TA_ATTR
    attr_set = true;
  }
  const int npos = B * hw;
  const int threads = 32 * PB * hpb;
  dim3 grid((npos + PB - 1) / PB, heads / hpb);
  const float sl2 = 1.4426950408889634f / sqrtf((float)d);
#define TA_LAUNCH(FTV, DV) pdl_launch(temporal_attn_mma_kernel<FTV, DV>, dim3(grid), dim3(threads), (size_t)(smem), (cudaStream_t)(st), qkv, F, hw, C, d, hpb, PB, npos, sl2, out)
  // the UNet's head widths (SD1.5 motion modules: 320 / 640 / 1280 channels over 8 heads) get their own instantiation
  if (FT == 1) {
    if (d == 40) TA_LAUNCH(1, 40); else if (d == 80) TA_LAUNCH(1, 80); else if (d == 160) TA_LAUNCH(1, 160); else TA_LAUNCH(1, 0);
  } else {
    if (d == 40) TA_LAUNCH(2, 40); else if (d == 80) TA_LAUNCH(2, 80); else if (d == 160) TA_LAUNCH(2, 160); else TA_LAUNCH(2, 0);
  }
#undef TA_LAUNCH
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200ad
