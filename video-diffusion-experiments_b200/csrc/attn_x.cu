// Text cross-attention (and any attention with Lk <= 128 keys: one key tile) on tcgen05, persistent over query tiles.
//   O = softmax(Q K^T / sqrt(d)) V per (frame-image n, head); K/V = the 77 text tokens, shared by all frames (kv_div).
// The layer is HBM-bound on paper -- Q in, O out, 168 MB at 64^2 = 26 us; 13 GFLOP -- but the one-CTA-per-query-tile
// kernel (attn_tc.cu) ran it as 8192 short-lived CTAs, each a serial chain TMA -> S -> softmax -> PV -> store behind a
// barrier / TMEM-allocation prologue: 224 us.  Here a CTA keeps K and V of its (n, head) in shared memory and walks the
// query tiles of that image:
//   warp 0      TMA producer: K, V once (box of KB = ceil16(Lk) key rows; rows >= Lk are zero-filled), then the Q tiles
//               through a two-stage ring
//   warp 1      UMMA issuer: S(i) = Q_i K^T (N = KB), O(i) = P(i) V with P read from TENSOR MEMORY; S(i+1) is issued as
//               soon as the softmax threads hold S(i) in registers, so it overlaps their exponentials and the
//               epilogue of tile i
//   warps 2..5  one thread per query row: tcgen05.ld of the S row, max, exp2, row sum, bf16 P -> tcgen05.st; the
//               epilogue of the PREVIOUS tile (O(i-1) / l -> bf16 -> global, 16-byte stores) sits between the
//               exponentials and the P store, so no MMA round trip is exposed
// No online softmax (one key tile), no O rescale.  TMEM: O DP + S KB + P KB/2 columns (184 at d=40: two CTAs per SM).
#include "../../include/b200ad.h"
#include "tmap.cuh"

namespace b200ad {

static constexpr int AX_THREADS = 192;
// Q ring stages: four at DP = 64 (a query tile lasts ~2 us, about one TMA round trip under load: two stages left the
// issuer waiting for Q), two above (shared memory for two CTAs per SM at DP = 128)
__host__ __device__ constexpr int ax_qst(int dp) { return dp == 64 ? 4 : 2; }
static constexpr int QSLAB_X = 128 * 128;      // [128 rows x 64 bf16] swizzled slab

struct AxParams {
  int Lq, Lk, d, kv_div, KB, ntiles, dk16;
  int hs;                  // head stride in the q/k/v rows (see FaParams::hs in attn_fa.cu)
  float scale_log2;        // d^-1/2 * log2(e)
  bf16* out;
  long long ldo;
};

// KBT: compile-time bound of the key-tile height (80 for the 77 text tokens, 128 in general): sizes the per-thread
// score row so that the softmax threads stay inside the 168 registers two CTAs per SM allow.
template <int DP, int KBT, bool LAST_ONLY>
__global__ void __launch_bounds__(AX_THREADS, (DP <= 128) ? 2 : 1)
attn_x_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
              const __grid_constant__ CUtensorMap tmV, const __grid_constant__ AxParams p) {
  pdl_trigger();
  constexpr int NS = DP / 64;                  // 64-column slabs per operand row
  constexpr int AX_QST = ax_qst(DP);
  // Key rows per tile = KBT, a compile-time constant (TMA zero-fills the rows beyond Lk): with a run-time height every
  // 16-column chunk of the softmax below sat behind its own branch, and branches are scheduling barriers.
  // LAST_ONLY: Lk > KBT - 16, i.e. only the last chunk holds keys to mask.
  constexpr int KB = KBT;
  // TMEM columns: O [0, DP), S [DP, DP + KB), P [DP + KB, DP + 3 KB / 2) -- 184 at d=40, 248 at d=80 (KB = 80): two
  // CTAs per SM with 256 columns each; wider cases take all 512 (a second resident CTA then waits in tcgen05.alloc)
  constexpr uint32_t O_COL = 0, S_COL = DP;
  constexpr uint32_t P_COL = (uint32_t)(DP + KB);
  constexpr uint32_t TMEM_COLS = (DP + KB + KB / 2 <= 256) ? 256u : 512u;
  constexpr int KSLAB = KB * 128;                  // bytes of one [KB rows x 64 bf16] slab

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                  // [AX_QST][NS] slabs
  uint8_t* sK = sQ + AX_QST * NS * QSLAB_X;            // [NS] slabs of KB rows (1024-byte multiples: KB % 8 == 0)
  uint8_t* sV = sK + NS * KSLAB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + NS * KSLAB);
  uint64_t* kv_full = bars;                            // [1]
  uint64_t* q_full = bars + 1;                         // [AX_QST]
  uint64_t* q_empty = q_full + AX_QST;                 // [AX_QST]
  uint64_t* s_full = q_empty + AX_QST;                 // [1]
  uint64_t* s_read = s_full + 1;                       // [1] 128 arrivals: the S tile is in registers
  uint64_t* p_full = s_read + 1;                       // [1] 128 arrivals (also: the previous O tile has been read)
  uint64_t* o_full = p_full + 1;                       // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int hd = blockIdx.y, n = blockIdx.z;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(kv_full, 1);
    for (int s = 0; s < AX_QST; ++s) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
    mbar_init(s_full, 1);
    mbar_init(s_read, 128);
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
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      const int nkvb = n / p.kv_div;
      mbar_arrive_expect_tx(kv_full, 2 * NS * KSLAB);
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        tma_load_3d(sK + s * KSLAB, &tmK, kv_full, hd * p.hs + s * 64, 0, nkvb);
        tma_load_3d(sV + s * KSLAB, &tmV, kv_full, hd * p.hs + s * 64, 0, nkvb);
      }
      int st = 0;
      uint32_t ph = 0;
      for (int qt = blockIdx.x; qt < p.ntiles; qt += gridDim.x) {
        mbar_wait(&q_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&q_full[st], NS * QSLAB_X);
#pragma unroll
        for (int s = 0; s < NS; ++s)
          tma_load_3d(sQ + (st * NS + s) * QSLAB_X, &tmQ, &q_full[st], hd * p.hs + s * 64, qt * 128, n);
        if (++st == AX_QST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    const uint32_t idesc_s = umma_idesc((uint32_t)KB, false, 1u);
    const uint32_t idesc_o = umma_idesc((uint32_t)DP, true, 1u);
    const uint32_t sQ_a = smem_u32(sQ), sK_a = smem_u32(sK), sV_a = smem_u32(sV);
    const int my_tiles = (p.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    auto issue_s = [&](int st) {
      if (elect_one()) {
        for (int kk = 0; kk < p.dk16; ++kk) {
          const uint64_t adesc = umma_desc_k_sw128(sQ_a + (uint32_t)((st * NS + (kk >> 2)) * QSLAB_X)) + (uint64_t)(2 * (kk & 3));
          const uint64_t bdesc = umma_desc_k_sw128(sK_a + (uint32_t)((kk >> 2) * KSLAB)) + (uint64_t)(2 * (kk & 3));
          umma_f16_ss(tmem_base + S_COL, adesc, bdesc, idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit(s_full);
        umma_commit(&q_empty[st]);       // the Q stage is free once S has been computed
      }
      __syncwarp();
    };
    mbar_wait(kv_full, 0);
    if (my_tiles > 0) {
      mbar_wait(&q_full[0], 0);
      tc_fence_after();
      issue_s(0);
    }
    int st = 0;
    uint32_t ph = 0;
    for (int i = 0; i < my_tiles; ++i) {
      int st_n = st + 1;
      uint32_t ph_n = ph;
      if (st_n == AX_QST) { st_n = 0; ph_n ^= 1; }
      if (i + 1 < my_tiles) {
        mbar_wait(&q_full[st_n], ph_n);
        mbar_wait(s_read, i & 1);
        tc_fence_after();
        issue_s(st_n);
      }
      mbar_wait(p_full, i & 1);
      tc_fence_after();
      if (elect_one()) {
        for (int kk = 0; kk < KB / 16; ++kk) {
          const uint64_t bdesc = umma_desc_mn_sw128(sV_a + (uint32_t)(kk * 2048), (uint32_t)KSLAB);
          umma_f16_ts(tmem_base + O_COL, tmem_base + P_COL + (uint32_t)(kk * 8), bdesc, idesc_o, kk > 0 ? 1u : 0u);
        }
        umma_commit(o_full);
      }
      __syncwarp();
      st = st_n;
      ph = ph_n;
    }
  } else {
    // ===================== softmax + epilogue: one thread per query row =====================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const uint32_t s_tmem = tmem_base + lane_addr + S_COL;
    const uint32_t o_tmem = tmem_base + lane_addr + O_COL;
    const uint32_t p_tmem = tmem_base + lane_addr + P_COL;
    const float sc = p.scale_log2;
    // O(t) / l -> bf16 -> global, for the tile whose PV has completed
    auto epilogue = [&](int qt, float inv) {
      const int qrow = qt * 128 + row;
      bf16* orow = p.out + ((long long)n * p.Lq + qrow) * p.ldo + hd * p.d;
#pragma unroll
      for (int c = 0; c < DP; c += 32) {
        if (c < p.d) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(o_tmem + c, r);
          tmem_wait_ld();
          if (qrow < p.Lq) {
#pragma unroll
            for (int k = 0; k < 32; k += 8) {
              if (c + k + 8 <= p.d) {
                uint4 v4;
                v4.x = pack_bf16x2(__uint_as_float(r[k]) * inv, __uint_as_float(r[k + 1]) * inv);
                v4.y = pack_bf16x2(__uint_as_float(r[k + 2]) * inv, __uint_as_float(r[k + 3]) * inv);
                v4.z = pack_bf16x2(__uint_as_float(r[k + 4]) * inv, __uint_as_float(r[k + 5]) * inv);
                v4.w = pack_bf16x2(__uint_as_float(r[k + 6]) * inv, __uint_as_float(r[k + 7]) * inv);
                *reinterpret_cast<uint4*>(orow + c + k) = v4;
              }
            }
          }
        }
      }
    };
    // Software pipeline: the epilogue of tile i-1 runs AFTER the exponentials of tile i, so the P -> PV -> O round trip
    // of a tile (hand-off to the issuer warp, MMA, commit) is hidden behind the next tile's softmax arithmetic.
    int i = 0;
    float inv_prev = 0.f;
    int qt_prev = 0;
    for (int qt = blockIdx.x; qt < p.ntiles; qt += gridDim.x, ++i) {
      mbar_wait(s_full, i & 1);
      tc_fence_after();
      float s[KBT];
#pragma unroll
      for (int c = 0; c < KBT; c += 16) tmem_ld_32x32b_x16(s_tmem + c, reinterpret_cast<uint32_t*>(s + c));
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(s_read);
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < KBT; c += 16) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          if ((!LAST_ONLY || c + 16 == KBT) && c + k >= p.Lk) s[c + k] = -INFINITY;     // a select, not a branch
          mx = fmaxf(mx, s[c + k]);
        }
      }
      const float neg_m = -mx * sc;
      float sum = 0.f;
      uint32_t pk[KBT / 2];
#pragma unroll
      for (int c = 0; c < KBT; c += 16) {
#pragma unroll
        for (int k = 0; k < 16; k += 2) {
          float e0, e1;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(fmaf(s[c + k], sc, neg_m)));
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(fmaf(s[c + k + 1], sc, neg_m)));
          sum += e0 + e1;
          pk[(c + k) >> 1] = pack_bf16x2(e0, e1);
        }
      }
      if (i > 0) {
        // PV(i-1) complete: its O can be read and its P overwritten
        mbar_wait(o_full, (i - 1) & 1);
        tc_fence_after();
        epilogue(qt_prev, inv_prev);
      }
#pragma unroll
      for (int c = 0; c < KBT; c += 16) tmem_st_32x32b_x8(p_tmem + (c >> 1), pk + (c >> 1));
      tmem_wait_st();
      tc_fence_before();       // orders the O loads above and the P stores before the arrive, i.e. before PV(i)
      mbar_arrive(p_full);
      inv_prev = 1.f / sum;
      qt_prev = qt;
    }
    if (i > 0) {
      mbar_wait(o_full, (i - 1) & 1);
      tc_fence_after();
      epilogue(qt_prev, inv_prev);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int DP, int KBT, bool LAST_ONLY>
static int launch_attn_x(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                         int nbatch, int heads, int Lq, int Lk, int d, int hs, int kv_div, void* out, long long ldo,
                         cudaStream_t stream) {
  constexpr int NS = DP / 64;
  constexpr int KB = KBT;
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
    uint32_t box[3] = {64, (uint32_t)KB, 1};
    if (int rc = make_tmap_bf16(&tmK, k, 3, dims, str, box)) return rc;
    uint64_t strv[2] = {(uint64_t)ldv * 2, (uint64_t)Lk * ldv * 2};
    if (int rc = make_tmap_bf16(&tmV, v, 3, dims, strv, box)) return rc;
  }
  AxParams p;
  p.Lq = Lq; p.Lk = Lk; p.d = d; p.kv_div = kv_div; p.KB = KB;
  p.ntiles = (Lq + 127) / 128;
  p.dk16 = (d + 15) / 16;
  p.hs = hs;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)d);
  p.out = reinterpret_cast<bf16*>(out);
  p.ldo = ldo;
  const size_t smem = (size_t)ax_qst(DP) * NS * QSLAB_X + (size_t)2 * NS * KB * 128 + 1024 + 128;
  static bool attr = false;
  if (!attr) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(attn_x_kernel<DP, KBT, LAST_ONLY>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 1024));
    attr = true;
  }
  // one wave of CTAs: as many query-tile walkers per (n, head) as fill the SMs' CTA slots
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int slots = sms * ((DP <= 128) ? 2 : 1);
  int gx = slots / (heads * nbatch);
  if (gx < 1) gx = 1;
  if (gx > p.ntiles) gx = p.ntiles;
  dim3 grid(gx, heads, nbatch);
  pdl_launch(attn_x_kernel<DP, KBT, LAST_ONLY>, grid, dim3(AX_THREADS), smem, stream, tmQ, tmK, tmV, p);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Served here: bf16, Lk <= 128, d % 8 == 0 (16-byte output stores), 16-byte aligned output rows.  Returns 1 otherwise.
int attention_bf16_x(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                     int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, void* out, long long ldo,
                     cudaStream_t stream) {
  if (Lk > 128 || d % 8 != 0 || ldo % 8 != 0 || (reinterpret_cast<uintptr_t>(out) & 15)) return 1;
#define AX_LAUNCH(DPV)                                                                                              \
  do {                                                                                                              \
    if (Lk <= 80)                                                                                                   \
      return Lk > 64 ? launch_attn_x<DPV, 80, true>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out, ldo, stream)   \
                     : launch_attn_x<DPV, 80, false>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out, ldo, stream); \
    return Lk > 112 ? launch_attn_x<DPV, 128, true>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out, ldo, stream)   \
                    : launch_attn_x<DPV, 128, false>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, Lk, d, dpad, kv_div, out, ldo, stream); \
  } while (0)
  if (dpad % 8 != 0 || dpad < 16 * ((d + 15) / 16)) return 1;
  if (dpad <= 64) AX_LAUNCH(64);
  if (dpad <= 128) AX_LAUNCH(128);
  if (dpad <= 192) AX_LAUNCH(192);
#undef AX_LAUNCH
  return 1;
}

}  // namespace b200ad
