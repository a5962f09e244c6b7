// tcgen05 flash attention for sm_100a: spatial self-attention and text cross-attention of the UNet.
//   O = softmax(Q K^T / sqrt(d)) V per (frame-image n, head), bf16 operands, fp32 S / O accumulators in TMEM.
// One CTA = 128 query rows of one (n, head).  warp 0: TMA producer (Q once, K/V tiles of 128 keys through a
// ring), warp 1: UMMA issuer (S = Q K^T into TMEM; O += P V with V as an MN-major operand), warps 2..5: online
// softmax, one thread per query row (tcgen05.ld of the S row, exp2, bf16 P written to swizzled smem, running
// rescale of the TMEM O accumulator).  Head width is padded to DP in {64,128,192} (pad columns are zeros that
// the QKV projection produces from zero weight rows), keys beyond Lk are zero-filled by TMA and masked.
#include "../../include/b200ad.h"
#include "tmap.cuh"

namespace b200ad {

int attention_f32_simt(const float* q, long long ldq, const float* k, long long ldk, const float* v, long long ldv,
                       int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, float* out, long long ldo,
                       cudaStream_t stream);   // simt.cu

int attention_bf16_fa(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                      int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, int v_ones, void* out,
                      long long ldo, cudaStream_t stream);   // attn_fa.cu: ping-pong kernel for the long self-attention layers
int attention_bf16_x(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                     int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, void* out, long long ldo,
                     cudaStream_t stream);   // attn_x.cu: persistent kernel for Lk <= 128 (text cross-attention)

static constexpr int ATT_THREADS = 192;
static constexpr int SLAB = 128 * 128;   // bytes of one [128 rows x 64 bf16] swizzled slab

struct AttnParams {
  int Lq, Lk, d, kv_div, nkv;
  int dk16;                // k-steps of 16 for Q K^T
  int hs;                  // head stride in the q/k/v rows (see FaParams::hs in attn_fa.cu)
  float scale_log2;        // d^-1/2 * log2(e)
  bf16* out;
  long long ldo;
};

// NST: K/V ring depth.  Text cross-attention (Lk = 77: one key tile) runs with NST = 1, which at DP = 64 brings the
// CTA to 81 KB of shared memory so that TWO CTAs share an SM (TMEM: 2 x 256 columns): the kernel is a serial
// load -> S -> softmax -> PV -> store chain per CTA, so residency is its only source of overlap (one CTA per SM
// measured 9 % occupancy and 6 % of HBM bandwidth, profiles/ncu_full_r01f.md).
template <int DP, int NST>
__global__ void __launch_bounds__(ATT_THREADS, (DP == 64) ? 2 : 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const __grid_constant__ AttnParams p) {
  pdl_trigger();
  constexpr int NS = DP / 64;                  // 64-column slabs per operand tile
  constexpr uint32_t TMEM_COLS = (128 + DP <= 256) ? 256u : 512u;
  constexpr uint32_t O_COL = 128;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + NS * SLAB;
  uint8_t* sV = sK + NST * NS * SLAB;
  uint8_t* sP = sV + NST * NS * SLAB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * SLAB);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;            // [NST]
  uint64_t* kv_empty = kv_full + NST;      // [NST]
  uint64_t* s_full = kv_empty + NST;
  uint64_t* p_full = s_full + 1;
  uint64_t* o_full = p_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int qt = blockIdx.x, hd = blockIdx.y, n = blockIdx.z;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int s = 0; s < NST; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init(s_full, 1);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                            // everything above overlapped the previous kernel's tail

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, NS * SLAB);
#pragma unroll
      for (int s = 0; s < NS; ++s) tma_load_3d(sQ + s * SLAB, &tmQ, q_full, hd * p.hs + s * 64, qt * 128, n);
      const int nkvb = n / p.kv_div;
      for (int j = 0; j < p.nkv; ++j) {
        const int st = j % NST;
        mbar_wait(&kv_empty[st], ((j / NST) & 1) ^ 1);
        mbar_arrive_expect_tx(&kv_full[st], 2 * NS * SLAB);
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          tma_load_3d(sK + (st * NS + s) * SLAB, &tmK, &kv_full[st], hd * p.hs + s * 64, j * 128, nkvb);
          tma_load_3d(sV + (st * NS + s) * SLAB, &tmV, &kv_full[st], hd * p.hs + s * 64, j * 128, nkvb);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc(128u, false, 1u);
      const uint32_t idesc_o = umma_idesc((uint32_t)DP, true, 1u);
      const uint32_t s_tmem = tmem_base;
      const uint32_t o_tmem = tmem_base + O_COL;
      mbar_wait(q_full, 0);
      for (int j = 0; j < p.nkv; ++j) {
        const int st = j % NST;
        mbar_wait(&kv_full[st], (j / NST) & 1);
        tc_fence_after();
        // S = Q K^T  (K-major A and B; K dimension = head width DP, 16 per instruction)
#pragma unroll
        for (int kk = 0; kk < p.dk16; ++kk) {
          const uint64_t adesc = umma_desc_k_sw128(smem_u32(sQ + (kk / 4) * SLAB)) + (uint64_t)(2 * (kk % 4));
          const uint64_t bdesc =
              umma_desc_k_sw128(smem_u32(sK + (st * NS + kk / 4) * SLAB)) + (uint64_t)(2 * (kk % 4));
          umma_f16_ss(s_tmem, adesc, bdesc, idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit(s_full);
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        // O += P V  (A = P [128 x 128 keys] K-major in 2 slabs; B = V [128 keys x DP] MN-major, 16 keys per step)
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t adesc = umma_desc_k_sw128(smem_u32(sP + (kk / 4) * SLAB)) + (uint64_t)(2 * (kk % 4));
          const uint64_t bdesc = umma_desc_mn_sw128(smem_u32(sV + st * NS * SLAB + kk * 2048), (uint32_t)SLAB);
          umma_f16_ss(o_tmem, adesc, bdesc, idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&kv_empty[st]);
        umma_commit(o_full);
      }
    }
  } else {
    // ===================== softmax: one thread per query row =====================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const uint32_t s_tmem = tmem_base + lane_addr;
    const uint32_t o_tmem = tmem_base + lane_addr + O_COL;
    float m = -INFINITY, l = 0.f;
    uint8_t* prow = sP + row * 128;
    const int sw = row & 7;
    for (int j = 0; j < p.nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const int kbase = j * 128;
      const bool tail = kbase + 128 > p.Lk;
      // pass 1: row maximum
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(s_tmem + c, r);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = __uint_as_float(r[i]);
          if (tail && kbase + c + i >= p.Lk) s = -INFINITY;
          mx = fmaxf(mx, s);
        }
      }
      const float m_new = fmaxf(m, mx * p.scale_log2);
      const float alpha = exp2f(m - m_new);
      // PV(j-1) must have finished reading P and writing O before either is touched again
      if (j > 0) {
        mbar_wait(o_full, (j - 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < DP; c += 32) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(o_tmem + c, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32b_x32(o_tmem + c, r);
        }
        tmem_wait_st();
      }
      // pass 2: probabilities -> bf16 P tile in swizzled smem
      float sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(s_tmem + c, r);
        tmem_wait_ld();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = exp2f(__uint_as_float(r[i]) * p.scale_log2 - m_new);
          float p1 = exp2f(__uint_as_float(r[i + 1]) * p.scale_log2 - m_new);
          if (tail) {
            if (kbase + c + i >= p.Lk) p0 = 0.f;
            if (kbase + c + i + 1 >= p.Lk) p1 = 0.f;
          }
          sum += p0 + p1;
          pk[i >> 1] = pack_bf16x2(p0, p1);
        }
        // 32 keys = 4 x 16-byte chunks of slab (c/64), chunk index ((c%64)/8 + q) swizzled with row&7
        uint8_t* slab = prow + (c >> 6) * SLAB;
        const int chunk0 = (c & 63) >> 3;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const int chunk = (chunk0 + q4) ^ sw;
          *reinterpret_cast<uint4*>(slab + chunk * 16) =
              make_uint4(pk[4 * q4], pk[4 * q4 + 1], pk[4 * q4 + 2], pk[4 * q4 + 3]);
        }
      }
      l = l * alpha + sum;
      m = m_new;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    // ---- epilogue: O / l -> bf16 ----
    mbar_wait(o_full, (p.nkv - 1) & 1);
    tc_fence_after();
    const float inv = 1.f / l;
    const int qrow = qt * 128 + row;
    bf16* orow = p.out + ((long long)n * p.Lq + qrow) * p.ldo + hd * p.d;
#pragma unroll 1
    for (int c = 0; c < DP; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(o_tmem + c, r);
      tmem_wait_ld();
      if (qrow < p.Lq) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          if (c + i < p.d)
            *reinterpret_cast<uint32_t*>(orow + c + i) =
                pack_bf16x2(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int DP, int NST>
static int launch_attn_tc(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                          int nbatch, int heads, int Lq, int Lk, int d, int hs, int kv_div, void* out, long long ldo,
                          cudaStream_t stream) {
  constexpr int NS = DP / 64;
  CUtensorMap tmQ, tmK, tmV;
  const int nkvb = (nbatch + kv_div - 1) / kv_div;
  {
    uint64_t dims[3] = {(uint64_t)heads * hs, (uint64_t)Lq, (uint64_t)nbatch};
    uint64_t str[2] = {(uint64_t)ldq * 2, (uint64_t)Lq * ldq * 2};
    uint32_t box[3] = {64, 128, 1};
    if (int rc = make_tmap_bf16(&tmQ, q, 3, dims, str, box)) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)heads * hs, (uint64_t)Lk, (uint64_t)nkvb};
    uint64_t str[2] = {(uint64_t)ldk * 2, (uint64_t)Lk * ldk * 2};
    uint32_t box[3] = {64, 128, 1};
    if (int rc = make_tmap_bf16(&tmK, k, 3, dims, str, box)) return rc;
    uint64_t strv[2] = {(uint64_t)ldv * 2, (uint64_t)Lk * ldv * 2};
    if (int rc = make_tmap_bf16(&tmV, v, 3, dims, strv, box)) return rc;
  }
  AttnParams p;
  p.Lq = Lq; p.Lk = Lk; p.d = d; p.kv_div = kv_div;
  p.nkv = (Lk + 127) / 128;
  p.dk16 = (d + 15) / 16;
  p.hs = hs;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)d);
  p.out = reinterpret_cast<bf16*>(out);
  p.ldo = ldo;
  const size_t smem = (size_t)(NS + 2 * NST * NS + 2) * SLAB + 1024 + 128;
  static bool attr = false;
  if (!attr) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel<DP, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = true;
  }
  dim3 grid((Lq + 127) / 128, heads, nbatch);
  pdl_launch(attn_tc_kernel<DP, NST>, dim3(grid), dim3(ATT_THREADS), (size_t)(smem), (cudaStream_t)(stream), tmQ, tmK, tmV, p);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200ad

using namespace b200ad;

extern "C" int b200ad_attention2(int op_dtype, const void* q, long long ldq, const void* k, long long ldk,
                                 const void* v, long long ldv, int nbatch, int heads, int Lq, int Lk, int d, int dpad,
                                 int kv_div, int flags, void* out_op, long long ldo, void* stream);
extern "C" int b200ad_attention(int op_dtype, const void* q, long long ldq, const void* k, long long ldk,
                                const void* v, long long ldv, int nbatch, int heads, int Lq, int Lk, int d, int dpad,
                                int kv_div, void* out_op, long long ldo, void* stream) {
  return b200ad_attention2(op_dtype, q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, 0, out_op, ldo,
                           stream);
}

extern "C" int b200ad_attention2(int op_dtype, const void* q, long long ldq, const void* k, long long ldk,
                                 const void* v, long long ldv, int nbatch, int heads, int Lq, int Lk, int d, int dpad,
                                 int kv_div, int flags, void* out_op, long long ldo, void* stream) {
  B200AD_REQUIRE(q && k && v && out_op, "attention: null pointer");
  B200AD_REQUIRE(Lq > 0 && Lk > 0 && d > 0 && d % 2 == 0 && dpad >= d && kv_div >= 1, "attention: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  if (op_dtype == 1)
    return attention_f32_simt((const float*)q, ldq, (const float*)k, ldk, (const float*)v, ldv, nbatch, heads, Lq, Lk,
                              d, dpad, kv_div, (float*)out_op, ldo, st);
  B200AD_REQUIRE(op_dtype == 0, "attention: unknown op_dtype %d", op_dtype);
  B200AD_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0, "attention: leading dims must be multiples of 8");
  B200AD_REQUIRE((d * 2) % 4 == 0 && (ldo % 2) == 0, "attention: output alignment");
  {
    const int rc = attention_bf16_fa(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, flags & 1, out_op, ldo, st);
    if (rc <= 0) return rc;          // served (0) or failed (<0); 1 = not a shape of the ping-pong kernel
  }
  {
    const int rc = attention_bf16_x(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out_op, ldo, st);
    if (rc <= 0) return rc;          // single key tile: the persistent kernel of attn_x.cu
  }
  const bool one_tile = Lk <= 128;
  B200AD_REQUIRE(dpad % 8 == 0 && dpad >= 16 * ((d + 15) / 16),
                 "attention: head stride %d must be a multiple of 8 and >= %d", dpad, 16 * ((d + 15) / 16));
  if (dpad <= 64)
    return one_tile ? launch_attn_tc<64, 1>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out_op, ldo, st)
                    : launch_attn_tc<64, 2>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out_op, ldo, st);
  if (dpad <= 128)
    return one_tile ? launch_attn_tc<128, 1>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out_op, ldo, st)
                    : launch_attn_tc<128, 2>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out_op, ldo, st);
  if (dpad <= 192)
    return launch_attn_tc<192, 1>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out_op, ldo, st);
  set_error("attention: head stride %d unsupported on the tcgen05 path (<= 192)", dpad);
  return -1;
}
