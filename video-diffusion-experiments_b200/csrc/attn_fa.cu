// Ping-pong tcgen05 flash attention for the long spatial self-attention layers (L = 4096 / 1024 at 512^2,
// 9216 / 2304 at 768^2; head width 40 or 80).
//   O = softmax(Q K^T / sqrt(d)) V per (frame-image n, head); bf16 operands, fp32 S / O accumulators in TMEM.
// One CTA owns TWO query tiles of 128 rows (A and B) of one (n, head) and walks the keys in tiles of BN:
//   warps 0..3  softmax of tile A, warps 4..7 softmax of tile B: one thread per query row; the whole score row is
//               held in registers (tcgen05.ld), exp2 with the running maximum, bf16 P written to TENSOR MEMORY
//               (tcgen05.st; the PV MMA takes its A operand from there) -- or to swizzled smem (PT = false)
//   warp 8      TMA producer: Q_A, Q_B once, then K_j / V_j through a ring of NST stages
//   warps 9,10  UMMA issuers, one per query tile t: S_t(j)=Q_t K_j^T, O_t+=P_t(j)V_j; S_t(j+1) is issued as soon as
//               the softmax threads of tile t hold S_t(j) in registers
//   warp 11     idle (completes the third warpgroup so that setmaxnreg can move registers: the softmax
//               warpgroups run with 232 registers per thread, the third one with 40)
// The layers this kernel serves are bound by the exponential (128 x 128 exp2 per 2.6 MFLOP of useful MMA at d=40),
// so the structure aims at keeping both softmax groups busy all the time:
//   * the O accumulator is rescaled lazily -- only when a row maximum grew by more than 2^8 since the last
//     rescale (FlashAttention-4's trick); P then stays <= 256, well inside bf16/fp32 range;
//   * QK^T only contracts over ceil(d/16) k-steps (the head-width padding columns are zero).
#include "../../include/b200ad.h"
#include "tmap.cuh"

namespace b200ad {

// threads per CTA: 8*TPR softmax warps + one warpgroup holding the TMA producer, the UMMA issuer and two idle warps
__host__ __device__ constexpr int fa_threads(int tpr, int nt = 2) { return (4 * nt * tpr + 4) * 32; }
__host__ __device__ constexpr int fa_nst(int tpr, bool pt, int nt = 2) { return nt == 3 ? 6 : pt ? 5 : (tpr == 1 ? 4 : 3); }

// Cycle counters of the softmax loop for tools/fa_waits.py (compiled in with -DB200AD_GEMM_PROFILE only).
#ifdef B200AD_GEMM_PROFILE
__device__ __forceinline__ long long fa_clock() { return clock64(); }
#else
__device__ __forceinline__ long long fa_clock() { return 0; }
#endif
static long long* g_fa_dbg = nullptr;
void set_fa_debug_buffer(long long* p) { g_fa_dbg = p; }

struct FaParams {
  long long* dbg;          // optional [8] counters of CTA (0,0,0), tile A, warp 0 (null in production)
  int Lq, Lk, d, nkv;
  int stagger;             // tile B starts its first exponent phase after tile A's (initial anti-phase offset)
  int dk16;                // k-steps of 16 for Q K^T
  int hs;                  // head stride in the q/k/v rows (elements): 16*dk16 <= hs <= DP; a 64-column slab may run
                           // into the next head, whose columns the MMAs never use (QK^T stops at 16*dk16, O beyond d is not read)
  float scale_log2;        // d^-1/2 * log2(e)
  bf16* out;
  long long ldo;
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x on the FMA/ALU pipes (Cody-Waite split + degree-3 polynomial, FlashAttention-4's coefficients; relative error
// ~1e-4, far below the bf16 rounding of P).  Valid for x in [-126, 126]; callers clamp below.
__device__ __forceinline__ float poly_exp2(float x) {
  x = fmaxf(x, -126.0f);
  float j;
  asm("add.rm.ftz.f32 %0, %1, %2;" : "=f"(j) : "f"(x), "f"(12582912.0f));   // 1.5 * 2^23: integer part in the mantissa
  const float f = x - (j - 12582912.0f);                                     // fractional part in [0, 1)
  float r = fmaf(0.077119089f, f, 0.227564f);
  r = fmaf(r, f, 0.695146f);
  r = fmaf(r, f, 1.0f);
  return __int_as_float(__float_as_int(r) + (__float_as_int(j) << 23));
}
// Probabilities -> bf16 pair.  Default: F2FP (cvt.rn.bf16x2.f32).  Opt-in (-DB200AD_FA_TRUNC_P=1): pack on the ALU pipe --
//   VS  (denominator from the V ones column): TRUNCATE with one PRMT; numerator and denominator are sums of the SAME
//       bf16 values, so the common relative bias of truncation cancels in the normalisation;
//   !VS: round half up (add 0x8000 to the bit patterns of the positive finite P), then the same PRMT.
// Tested on the suspicion that F2FP shares the 16-lane XU pipe with MUFU.EX2: it does not bound the loop -- 1.560 ms
// (PRMT) vs 1.548 ms (F2FP) at L=4096 d=40 with three query tiles -- so the exact rounding stays.
#ifndef B200AD_FA_TRUNC_P
#define B200AD_FA_TRUNC_P 0
#endif
template <bool VS>
__device__ __forceinline__ uint32_t pack_p(float lo, float hi) {
  if (B200AD_FA_TRUNC_P) {
    uint32_t a = __float_as_uint(lo), b = __float_as_uint(hi), r;
    if (!VS) { a += 0x8000u; b += 0x8000u; }
    asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  return pack_bf16x2(lo, hi);
}
#ifndef B200AD_FA_SPEC
#define B200AD_FA_SPEC 0           // speculative exponentials against the previous reference maximum (see the softmax loop):
                                   // measured SLOWER (1.77 vs 1.62 ms at L=4096 d=40: the fused loop schedules worse) -> opt-in
#endif
#ifndef B200AD_FA_EMU_EVERY
#define B200AD_FA_EMU_EVERY 8      // every 8th exponential runs on the FMA pipes instead of the MUFU unit
#endif

// DP: padded head width (64 or 128).  BN: keys per tile (128 for DP=64, 64 for DP=128).
// VS: the caller guarantees that pad column d of every V row is 1.0 (the QKV GEMM writes it through its bias), so the
// softmax denominator comes out of the PV MMA as O[:, d] -- the sum of the bf16-ROUNDED probabilities, i.e. exactly
// the weights the numerator used -- and the 128 FADDs per row and key tile of the explicit row sum disappear
// (17 % of the softmax loop's instructions, which is issue-bound at d = 40); the lazy rescale covers it for free.
// TPR: softmax threads per score row.  1: one thread holds the whole BN-wide score row (8 softmax warps, two per
// scheduler: the loop is bound by MUFU / dependency latency, profiles/ncu_full_r01f.md).  2: two threads split the
// row's columns (16 softmax warps, four per scheduler, 104 registers each); they exchange the row maximum through
// shared memory once per key tile (double-buffered by tile parity, one named barrier per query tile) and split the O
// columns of the lazy rescale and of the final normalisation.
// PT: the bf16 probabilities go to TENSOR MEMORY (tcgen05.st, two per 32-bit column) and the PV MMA reads its A operand
// from there.  Per key tile and query tile that removes 32 KB of st.shared and 32 KB of operand reads from a shared
// memory pipe that otherwise moves 256 KB per key tile (K/V fill 32, Q+K reads 64, P writes 64, P+V reads 96) --
// 2048 cycles at 128 B/clk, as much as the MUFU floor of the exponentials.  The freed 64 KB deepen the K/V ring.
// NT: query tiles per CTA.  With two, each scheduler hosts two softmax warps and a warp's chain per key tile (S wait,
// tcgen05.ld, max, exp, pack, tcgen05.st: ~2400 cycles uncontended at BN = 128) is longer than the 2 x 896 cycles of
// MUFU work the scheduler has per key tile, so the MUFU pipe idles a third of the time however the tiles are phased.
// Three tiles (BN = 64 so that 3 x (S 64 + O 64 + P 32) = 480 TMEM columns fit; 16 warps, 128 registers each, no
// setmaxnreg) put three softmax warps on every scheduler: the chain of one is covered by the exponentials of the
// other two.
template <int DP, int BN, bool VS, int TPR, bool PT, int NT>
__global__ void __launch_bounds__(fa_threads(TPR, NT), 1)
attn_fa_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const __grid_constant__ FaParams p) {
  pdl_trigger();
  constexpr int NS = DP / 64;                    // 64-column slabs per operand row
  // K/V ring depth.  Four stages (everything the 227 KB allow) when one thread owns a row: with three, a K/V tile is
  // requested only ~1.4 key tiles before its S MMA, which pulls the two query tiles back into lock-step whenever one
  // runs half a tile behind the other (the anti-phase schedule needs that slack).
  constexpr int NST = fa_nst(TPR, PT, NT);
  constexpr int QSLAB = 128 * 128;               // [128 rows x 64 bf16]
  constexpr int KSLAB = BN * 128;                // [BN rows x 64 bf16]
  constexpr int PS = BN / 64;                    // P slabs (each [128 q rows x 64 keys])
  // TMEM columns: S_t at t*BN, O_t behind the S tiles, P_t (PT: [128 lanes x BN/2 columns] bf16x2) behind the O tiles
  auto S_COL = [](int t) { return (uint32_t)(t * BN); };
  auto O_COL = [](int t) { return (uint32_t)(NT * BN + t * DP); };
  auto P_COL = [](int t) { return (uint32_t)(NT * (BN + DP) + t * (BN / 2)); };
  static_assert(NT == 2 || (NT == 3 && PT && TPR == 1), "three query tiles need P in tensor memory");
  constexpr uint32_t TMEM_COLS = 512;
  static_assert(NT * (BN + DP) + (PT ? NT * BN / 2 : 0) <= 512, "TMEM budget");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                  // [NT][NS] slabs
  uint8_t* sK = sQ + NT * NS * QSLAB;                   // [NST][NS]
  uint8_t* sV = sK + NST * NS * KSLAB;                 // [NST][NS]
  uint8_t* sP = sV + NST * NS * KSLAB;                 // [2][PS]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + (PT ? 0 : 2 * PS * QSLAB));
  uint64_t* q_full = bars;                             // [1]
  uint64_t* kv_full = bars + 1;                        // [NST]
  uint64_t* kv_empty = kv_full + NST;                  // [NST]
  uint64_t* s_full = kv_empty + NST;                   // [NT]
  uint64_t* p_full = s_full + NT;                      // [NT]  128 arrivals
  uint64_t* o_full = p_full + NT;                      // [NT]
  uint64_t* s_read = o_full + NT;                      // [NT]  128 arrivals: the S tile is in registers
  uint64_t* stag = s_read + NT;                        // [NT] tile t's first exponent phase is over (stagger)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stag + NT);
  static_assert((1 + 2 * NST + 5 * NT + 1 + 4 * NT) * 8 <= 512, "barrier block");
  float* smx = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);   // [2 parities][NT tiles][TPR][128]
  constexpr int PW = 4 * NT * TPR, MW = PW + 1;        // producer warp / first of the NT UMMA-issuer warps

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int qt = blockIdx.x, hd = blockIdx.y, n = blockIdx.z;     // qt: pair of 128-row query tiles

  if (warp == PW && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int s = 0; s < NST; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], NT); }
    for (int s = 0; s < NT; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&p_full[s], 128 * TPR);
      mbar_init(&o_full[s], 1);
      mbar_init(&s_read[s], 128 * TPR);
      mbar_init(&stag[s], 128 * TPR);
    }
    fence_barrier_init();
  }
  if (warp == MW) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                            // everything above overlapped the previous kernel's tail

  if (warp >= PW) {
  if (NT == 2 && !(TPR == 2 && PT)) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
  if (warp == PW) {
    // ===================== TMA producer (whole warp loops; one elected lane issues) =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, NT * NS * QSLAB);
#pragma unroll
      for (int t = 0; t < NT; ++t)
#pragma unroll
        for (int s = 0; s < NS; ++s)
          tma_load_3d(sQ + (t * NS + s) * QSLAB, &tmQ, q_full, hd * p.hs + s * 64, (NT * qt + t) * 128, n);
    }
    __syncwarp();
    int st = 0;
    uint32_t ph = 0;
    for (int j = 0; j < p.nkv; ++j) {
      mbar_wait(&kv_empty[st], ph ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&kv_full[st], 2 * NS * KSLAB);
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          tma_load_3d(sK + (st * NS + s) * KSLAB, &tmK, &kv_full[st], hd * p.hs + s * 64, j * BN, n);
          tma_load_3d(sV + (st * NS + s) * KSLAB, &tmV, &kv_full[st], hd * p.hs + s * 64, j * BN, n);
        }
      }
      __syncwarp();
      if (++st == NST) { st = 0; ph ^= 1; }
    }
  } else if (warp >= MW && warp < MW + NT) {
    // ===================== UMMA issuers =====================
    const uint32_t idesc_s = umma_idesc((uint32_t)BN, false, 1u);
    const uint32_t idesc_o = umma_idesc((uint32_t)DP, true, 1u);
    const uint32_t sQ_a = smem_u32(sQ), sK_a = smem_u32(sK), sV_a = smem_u32(sV), sP_a = smem_u32(sP);
    // S_t = Q_t K^T for the K tile in ring stage `st`
    auto issue_s = [&](int t, int st) {
      if (elect_one()) {
        for (int kk = 0; kk < p.dk16; ++kk) {
          const uint64_t adesc = umma_desc_k_sw128(sQ_a + (uint32_t)((t * NS + (kk >> 2)) * QSLAB)) + (uint64_t)(2 * (kk & 3));
          const uint64_t bdesc = umma_desc_k_sw128(sK_a + (uint32_t)((st * NS + (kk >> 2)) * KSLAB)) + (uint64_t)(2 * (kk & 3));
          umma_f16_ss(tmem_base + S_COL(t), adesc, bdesc, idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit(&s_full[t]);
      }
      __syncwarp();
    };
    // O_t (+)= P_t V for the V tile in ring stage `st`; 16 keys per instruction
    auto issue_pv = [&](int t, int st, bool first, bool release_kv) {
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          const uint64_t bdesc = umma_desc_mn_sw128(sV_a + (uint32_t)(st * NS * KSLAB + kk * 2048), (uint32_t)KSLAB);
          if (PT) {
            umma_f16_ts(tmem_base + O_COL(t), tmem_base + P_COL(t) + (uint32_t)(kk * 8), bdesc,
                        idesc_o, (first && kk == 0) ? 0u : 1u);
          } else {
            const uint64_t adesc = umma_desc_k_sw128(sP_a + (uint32_t)((t * PS + (kk >> 2)) * QSLAB)) + (uint64_t)(2 * (kk & 3));
            umma_f16_ss(tmem_base + O_COL(t), adesc, bdesc, idesc_o, (first && kk == 0) ? 0u : 1u);
          }
        }
        umma_commit(&o_full[t]);
        if (release_kv) umma_commit(&kv_empty[st]);
      }
      __syncwarp();
    };
    // One issuer warp PER query tile (warp MW -> tile A, warp MW+1 -> tile B).  With a single in-order issuer the two
    // tiles were coupled: S_A(j+1) could not be issued while the warp sat in its wait for P_B(j-1), so the tiles ran
    // in lock-step, both in their MUFU-bound exponent phase at the same time (tools/fa_waits.py: exp 1890 cycles per
    // key tile against 896 uncontended, 300-1000 cycles waiting for S).  Each issuer now only follows its own tile:
    //   S_t(j+1) as soon as the softmax threads of tile t hold S_t(j) in registers, PV_t(j) when P_t(j) is in smem;
    // a K/V ring stage is released by both (kv_empty counts two arrivals).
    const int mt = warp - MW;
    mbar_wait(q_full, 0);
    mbar_wait(&kv_full[0], 0);
    tc_fence_after();
    issue_s(mt, 0);
    int st = 0;
    uint32_t ph = 0;
    for (int j = 0; j < p.nkv; ++j) {
      int st_n = st + 1;
      uint32_t ph_n = ph;
      if (st_n == NST) { st_n = 0; ph_n ^= 1; }
      if (j + 1 < p.nkv) {
        mbar_wait(&kv_full[st_n], ph_n);
        mbar_wait(&s_read[mt], j & 1);
        tc_fence_after();
        issue_s(mt, st_n);
      }
      mbar_wait(&p_full[mt], j & 1);
      tc_fence_after();
      issue_pv(mt, st, j == 0, true);
      st = st_n;
      ph = ph_n;
    }
  }
  } else {
    // ===================== softmax: 4*TPR warps per query tile; TPR threads share a query row ====
    if (NT == 3 || (TPR == 2 && PT)) {}          // 16 warps x 128 / 20 or 28 warps x 72-96 registers: nothing to move
    else if (TPR == 1) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");   // 512*104 + 128*40 <= 640*96 (the CTA's own allocation)
    constexpr int HC = BN / TPR;                 // score columns per thread
    constexpr int OC = DP / TPR;                 // O columns per thread (rescale, normalisation)
    constexpr bool SPEC = PT && TPR == 1 && B200AD_FA_SPEC;
    const int t = warp / (4 * TPR);
    const int hf = (warp >> 2) % TPR;            // which part of the row's columns
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const uint32_t s_tmem = tmem_base + lane_addr + S_COL(t) + (uint32_t)(hf * HC);
    const uint32_t o_tmem = tmem_base + lane_addr + O_COL(t);
    const uint32_t p_tmem = tmem_base + lane_addr + P_COL(t) + (uint32_t)(hf * (HC / 2));
    const uint32_t p_row = smem_u32(sP + t * PS * QSLAB) + (uint32_t)(row * 128);
    const int sw = row & 7;
    const float sc = p.scale_log2;
    float m_used = -INFINITY;       // maximum (scaled, log2 units) the stored O / l are relative to
    float l = 0.f;
    long long c_sfull = 0, c_ld = 0, c_max = 0, c_exp = 0, c_ofull = 0, c_store = 0;
    // anti-phase start: query tile t enters its loop when tile t-1 has finished its first exponent phase
    if (p.stagger && t > 0) mbar_wait(&stag[t - 1], 0);
    const long long c_begin = fa_clock();
    for (int j = 0; j < p.nkv; ++j) {
      const long long k0 = fa_clock();
      mbar_wait(&s_full[t], j & 1);
      tc_fence_after();
      const long long k1 = fa_clock();
      c_sfull += k1 - k0;
      float s[HC];
#pragma unroll
      for (int c = 0; c < HC; c += 32) tmem_ld_32x32b_x32(s_tmem + c, reinterpret_cast<uint32_t*>(s + c));
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&s_read[t]);
      const long long k2 = fa_clock();
      c_ld += k2 - k1;
      // no key masking: the shapes served here have Lk % 256 == 0 (attention_bf16_fa), i.e. only whole key tiles.
      // (Every branch inside this loop is a scheduling barrier for the compiler: two extra ones around the exponent
      // loop cost 7 % of the kernel, profiles/kbench_attention_variants_r01.txt.)
      long long k3 = k2;
      uint32_t pk[PT ? HC / 2 : 1];                // PT: the packed bf16 probabilities of this row
      bool need;
      float m_new;
      float sum8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      auto exp_one = [&](int i, float x) {
        return (B200AD_FA_EMU_EVERY > 0 && (i % B200AD_FA_EMU_EVERY) == B200AD_FA_EMU_EVERY - 1) ? poly_exp2(x) : fast_exp2(x);
      };
      if (SPEC) {
        // Speculate that the reference maximum stays (true for all but the first few key tiles with the 2^8 slack of
        // the lazy rescale): exponentials against the OLD maximum, packed at once, with the row maximum computed in
        // the shadow of the MUFU pipe instead of in a phase of its own.  The scores stay in registers; a row whose
        // maximum did grow past the slack recomputes from them below -- same values as the non-speculative order.
        const float neg_old = -m_used;
        float mx8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) mx8[i] = s[i];
#pragma unroll
        for (int i = 0; i < HC; i += 2) {
          const float e0 = exp_one(i, fmaf(s[i], sc, neg_old)), e1 = exp_one(i + 1, fmaf(s[i + 1], sc, neg_old));
          pk[i / 2] = pack_p<VS>(e0, e1);
          if (!VS) { sum8[i & 7] += e0; sum8[(i + 1) & 7] += e1; }
          if (i >= 8) { mx8[i & 7] = fmaxf(mx8[i & 7], s[i]); mx8[(i + 1) & 7] = fmaxf(mx8[(i + 1) & 7], s[i + 1]); }
        }
        const float mx = sc * fmaxf(fmaxf(fmaxf(mx8[0], mx8[1]), fmaxf(mx8[2], mx8[3])),
                                    fmaxf(fmaxf(mx8[4], mx8[5]), fmaxf(mx8[6], mx8[7])));
        need = mx > m_used + 8.0f;
        m_new = need ? mx : m_used;
        if (need) {
          const float neg_m = -m_new;
#pragma unroll
          for (int i = 0; i < 8; ++i) sum8[i] = 0.f;
#pragma unroll
          for (int i = 0; i < HC; i += 2) {
            const float e0 = exp_one(i, fmaf(s[i], sc, neg_m)), e1 = exp_one(i + 1, fmaf(s[i + 1], sc, neg_m));
            pk[i / 2] = pack_p<VS>(e0, e1);
            if (!VS) { sum8[i & 7] += e0; sum8[(i + 1) & 7] += e1; }
          }
        }
      } else {
      // 8 independent chains: a single max / sum chain is HC dependent 4-cycle ops per row
      float mx8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx8[i] = s[i];
#pragma unroll
      for (int i = 8; i < HC; ++i) mx8[i & 7] = fmaxf(mx8[i & 7], s[i]);
      float mx = fmaxf(fmaxf(fmaxf(mx8[0], mx8[1]), fmaxf(mx8[2], mx8[3])),
                       fmaxf(fmaxf(mx8[4], mx8[5]), fmaxf(mx8[6], mx8[7])));
      if (TPR == 2) {
        // row maximum across the two column halves: both threads of a row end up with the same value, hence the
        // same rescale decision and the same reference maximum
        float* slot = smx + (((j & 1) * NT + t) * 2) * 128;
        slot[hf * 128 + row] = mx;
        named_bar_sync(1 + t, 128 * TPR);
        mx = fmaxf(mx, slot[(hf ^ 1) * 128 + row]);
      }
      mx *= sc;
      k3 = fa_clock();
      c_max += k3 - k2;
      // lazy rescale: keep the old reference maximum unless this row's maximum grew by more than 2^8
      need = mx > m_used + 8.0f;
      m_new = need ? mx : m_used;
      const float neg_m = -m_new;
#pragma unroll
      for (int i = 0; i < HC; ++i) {
        s[i] = exp_one(i, fmaf(s[i], sc, neg_m));
        if (!VS) sum8[i & 7] += s[i];
      }
      if (PT) {
#pragma unroll
        for (int i = 0; i < HC / 2; ++i) pk[i] = pack_p<VS>(s[2 * i], s[2 * i + 1]);
      }
      }
      mbar_arrive(&stag[t]);       // (every key tile, branch-free; the next query tile only waits for the first phase)
      const float sum = ((sum8[0] + sum8[1]) + (sum8[2] + sum8[3])) + ((sum8[4] + sum8[5]) + (sum8[6] + sum8[7]));
      const long long k4 = fa_clock();
      c_exp += k4 - k3;
#ifdef B200AD_GEMM_PROFILE
      // timeline of the exponent phases of key tiles 8..15 (start, end per query tile) for tools/fa_waits.py
      if (p.dbg && j >= 8 && j < 16 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (warp & 3) == 0 && lane == 0) {
        p.dbg[64 + t * 16 + (j - 8) * 2] = k3;
        p.dbg[64 + t * 16 + (j - 8) * 2 + 1] = k4;
      }
#endif
      // PV(j-1) must be complete before P is overwritten or O rescaled.  Branch-free: at j = 0 the wait is for the
      // parity opposite to the barrier's initial phase and returns at once.
      mbar_wait(&o_full[t], (j - 1) & 1);
      tc_fence_after();
      const long long k5 = fa_clock();
      c_ofull += k5 - k4;
      if (__any_sync(0xffffffffu, need) && j > 0) {
        const float alpha = need ? fast_exp2(m_used - m_new) : 1.0f;
#pragma unroll 1
        for (int c = hf * OC; c < (hf + 1) * OC; c += 32) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(o_tmem + c, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32b_x32(o_tmem + c, r);
        }
        tmem_wait_st();
        l *= alpha;
      }
      m_used = m_new;
      l += sum;
      if (PT) {
        // bf16 P row -> this thread's TMEM lane, two keys per column
#pragma unroll
        for (int c = 0; c < HC / 2; c += 32) tmem_st_32x32b_x32(p_tmem + c, pk + c);
        tmem_wait_st();
      } else {
      // bf16 P row -> swizzled smem: 8 keys per 16-byte chunk, chunk index XOR (row & 7) inside each 64-key slab
#pragma unroll
      for (int c8 = 0; c8 < HC / 8; ++c8) {
        const int g8 = hf * (HC / 8) + c8;       // 8-key group within the BN-wide tile
        const int slab = g8 >> 3, chunk = (g8 & 7) ^ sw;
        const uint32_t a = p_row + (uint32_t)(slab * QSLAB + chunk * 16);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pack_p<VS>(s[8 * c8], s[8 * c8 + 1])),
                     "r"(pack_p<VS>(s[8 * c8 + 2], s[8 * c8 + 3])), "r"(pack_p<VS>(s[8 * c8 + 4], s[8 * c8 + 5])),
                     "r"(pack_p<VS>(s[8 * c8 + 6], s[8 * c8 + 7]))
                     : "memory");
      }
      fence_proxy_async_smem();
      }
      tc_fence_before();
      mbar_arrive(&p_full[t]);
      c_store += fa_clock() - k5;
    }
    if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (warp == 0 || warp == 4 * TPR) && lane == 0) {
      long long* d = p.dbg + (warp ? 8 : 0);
      d[0] = fa_clock() - c_begin; d[1] = c_sfull; d[2] = c_ld; d[3] = c_max; d[4] = c_exp; d[5] = c_ofull;
      d[6] = c_store; d[7] = p.nkv;
    }
    // ---- epilogue: O / l -> bf16 ----
    mbar_wait(&o_full[t], (p.nkv - 1) & 1);
    tc_fence_after();
    if (TPR == 2 && !VS) {
      // the two halves of a row each summed their own probabilities
      float* slot = smx + ((p.nkv & 1) * NT + t) * 2 * 128;
      slot[hf * 128 + row] = l;
      named_bar_sync(1 + t, 128 * TPR);
      l += slot[(hf ^ 1) * 128 + row];
    }
    float inv = 1.f / l;
    if (VS) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(o_tmem + (uint32_t)(p.d & ~31), r);
      tmem_wait_ld();
      float lv = 1.f;
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i == (p.d & 31)) lv = __uint_as_float(r[i]);
      inv = 1.f / lv;
    }
    const int qrow = (NT * qt + t) * 128 + row;
    bf16* orow = p.out + ((long long)n * p.Lq + qrow) * p.ldo + hd * p.d;
#pragma unroll 1
    for (int c = hf * OC; c < (hf + 1) * OC; c += 32) {
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
  if (warp == MW) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int DP, int BN, bool VS, int TPR, bool PT, int NT = 2>
static int launch_attn_fa(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                          int nbatch, int heads, int L, int d, int hs, void* out, long long ldo, cudaStream_t stream) {
  constexpr int NS = DP / 64;
  CUtensorMap tmQ, tmK, tmV;
  uint64_t dims[3] = {(uint64_t)heads * hs, (uint64_t)L, (uint64_t)nbatch};
  {
    uint64_t str[2] = {(uint64_t)ldq * 2, (uint64_t)L * ldq * 2};
    uint32_t box[3] = {64, 128, 1};
    if (int rc = make_tmap_bf16(&tmQ, q, 3, dims, str, box)) return rc;
  }
  {
    uint32_t box[3] = {64, (uint32_t)BN, 1};
    uint64_t strk[2] = {(uint64_t)ldk * 2, (uint64_t)L * ldk * 2};
    if (int rc = make_tmap_bf16(&tmK, k, 3, dims, strk, box)) return rc;
    uint64_t strv[2] = {(uint64_t)ldv * 2, (uint64_t)L * ldv * 2};
    if (int rc = make_tmap_bf16(&tmV, v, 3, dims, strv, box)) return rc;
  }
  FaParams p;
  p.Lq = L; p.Lk = L; p.d = d;
  p.nkv = (L + BN - 1) / BN;
  // measured (profiles/kbench_attention_variants_r01.txt, L=4096 d=40, V ones column): with one issuer warp per query
  // tile, P in tensor memory and a 5-stage K/V ring the anti-phase start gives 1.62 ms against 1.71 ms without
  // (it was a loss with the single in-order issuer) -> on by default
  p.stagger = 1;
  p.dbg = g_fa_dbg;
  p.dk16 = (d + 15) / 16;
  p.hs = hs;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)d);
  p.out = reinterpret_cast<bf16*>(out);
  p.ldo = ldo;
  // slabs + alignment + barriers + row-maximum exchange of the two-threads-per-row variant
  const size_t smem = (size_t)(NT * NS * 128 * 128 + 2 * fa_nst(TPR, PT, NT) * NS * BN * 128 + (PT ? 0 : 2 * (BN / 64) * 128 * 128)) +
                      1024 + 512 + (TPR == 2 ? 2 * NT * 2 * 128 * 4 : 0);
  B200AD_REQUIRE(smem <= 232448, "attention: %zu bytes of shared memory", smem);
  static bool attr = false;
  if (!attr) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(attn_fa_kernel<DP, BN, VS, TPR, PT, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = true;
  }
  dim3 grid((L + 128 * NT - 1) / (128 * NT), heads, nbatch);
  pdl_launch(attn_fa_kernel<DP, BN, VS, TPR, PT, NT>, dim3(grid), dim3(fa_threads(TPR, NT)), (size_t)(smem), (cudaStream_t)(stream), tmQ, tmK, tmV, p);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Self-attention fast path; returns 1 when the shape is not served here (the caller falls back to attn_tc.cu).
int attention_bf16_fa(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                      int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, int v_ones, void* out,
                      long long ldo, cudaStream_t stream) {
  if (Lq != Lk || kv_div != 1 || Lq % 256 != 0 || Lq < 512) return 1;
  const bool vs = v_ones && d < dpad;
  // Instantiated: the variants that measured fastest (profiles/kbench_attention_variants_r01.txt keeps the numbers of
  // the ones that did not -- two softmax threads per row, P through shared memory, speculative exponentials, other tile
  // counts; their launch paths and environment switches are gone).  One softmax thread per row, P in tensor memory;
  // d <= 64 with the V ones column: three query tiles per CTA (1.56 vs 1.62-1.70 ms at L=4096 d=40), else two.
  if (dpad % 8 != 0 || dpad < 16 * ((d + 15) / 16)) return 1;       // head stride: see FaParams::hs
  if (dpad <= 64)
    return vs ? launch_attn_fa<64, 64, true, 1, true, 3>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, d, dpad, out, ldo, stream)
              : launch_attn_fa<64, 128, false, 1, true, 2>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, d, dpad, out, ldo, stream);
  if (dpad <= 128)
    return vs ? launch_attn_fa<128, 64, true, 1, true, 2>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, d, dpad, out, ldo, stream)
              : launch_attn_fa<128, 64, false, 1, true, 2>(q, ldq, k, ldk, v, ldv, nbatch, heads, Lq, d, dpad, out, ldo, stream);
  return 1;
}

}  // namespace b200ad
