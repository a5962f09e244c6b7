// tcgen05 GEMM / implicit-GEMM conv3x3 for sm_100a.
//   C[M,N] = A[M,K] * W[N,K]^T, bf16 operands, fp32 accumulation in TMEM, fused fp32 epilogue.
// Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = UMMA issuer (+TMEM alloc),
// warps 2..5 = epilogue (TMEM -> registers -> global).  Two TMEM accumulator stages so the epilogue
// of tile i overlaps the mainloop of tile i+1.  Operand tiles are [rows x 64] bf16, 128-byte swizzle.
// Conv mode loads the A tile with a 4-D tensor map over the NHWC image: one shifted box per filter
// tap, out-of-bounds (halo) pixels zero-filled by TMA.
#include "gemm.cuh"
#include "tmap.cuh"

namespace b200ad {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KiB
static constexpr int GEMM_THREADS = 192;
static constexpr int SMEM_TILE_BUDGET = 192 * 1024;

struct TcParams {
  int M, N, K;
  int BN;
  int num_stages;
  int stage_bytes;
  int num_m_tiles, num_n_tiles, num_kb;
  int conv;
  int nimg, h, w;
  int n_box, h_box, w_box;
  int tiles_x, tiles_y;
  int cin_blocks;
  int tap_dy[9], tap_dx[9], tap_dn[9];
  Epilogue epi;
};

__device__ __forceinline__ void epilogue_chunk(const TcParams& p, long long out_row, bool row_ok, int col0,
                                               const uint32_t* acc) {
  if (!row_ok || col0 >= p.N) return;
  const Epilogue& e = p.epi;
  if (e.geglu) {
    // interleaved (h, g) pairs -> 16 outputs
    bf16* out = reinterpret_cast<bf16*>(e.out_op) + out_row * e.ld_oop + (col0 >> 1);
    uint32_t packed[8];
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      float add[4];
      epilogue_load_addends(e, out_row, col0 + j, add);
      float h0 = __uint_as_float(acc[j + 0]) + add[0];
      float g0 = __uint_as_float(acc[j + 1]) + add[1];
      float h1 = __uint_as_float(acc[j + 2]) + add[2];
      float g1 = __uint_as_float(acc[j + 3]) + add[3];
      packed[j >> 2] = pack_bf16x2(h0 * gelu_erf_f(g0), h1 * gelu_erf_f(g1));
    }
    uint4* o4 = reinterpret_cast<uint4*>(out);
    o4[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    o4[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    return;
  }
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    float add[4];
    epilogue_load_addends(e, out_row, col0 + j, add);
    if (e.residual) {
      float4 r = *reinterpret_cast<const float4*>(e.residual + out_row * e.ld_res + col0 + j);
      add[0] += r.x; add[1] += r.y; add[2] += r.z; add[3] += r.w;
    }
    v[j + 0] = __uint_as_float(acc[j + 0]) + add[0];
    v[j + 1] = __uint_as_float(acc[j + 1]) + add[1];
    v[j + 2] = __uint_as_float(acc[j + 2]) + add[2];
    v[j + 3] = __uint_as_float(acc[j + 3]) + add[3];
  }
  if (e.out_f32) {
    float4* o = reinterpret_cast<float4*>(e.out_f32 + out_row * e.ld_o32 + col0);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  }
  if (e.out_op) {
    uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out_op) + out_row * e.ld_oop + col0);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      o[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                        pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
  }
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)p.num_stages * p.stage_bytes);
  uint64_t* full_bar = bars;                       // [num_stages]
  uint64_t* empty_bar = bars + p.num_stages;       // [num_stages]
  uint64_t* tfull_bar = bars + 2 * p.num_stages;   // [2]
  uint64_t* tempty_bar = tfull_bar + 2;            // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const uint32_t tmem_cols = (2 * p.BN <= 256) ? 256u : 512u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 4);     // one elected lane per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = A_STAGE_BYTES + p.BN * BK * 2;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_tile = tile / p.num_n_tiles;
        const int n_tile = tile % p.num_n_tiles;
        int x0 = 0, y0 = 0, n0 = 0;
        if (p.conv) {
          x0 = (m_tile % p.tiles_x) * p.w_box;
          y0 = ((m_tile / p.tiles_x) % p.tiles_y) * p.h_box;
          n0 = (m_tile / (p.tiles_x * p.tiles_y)) * p.n_box;
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
          uint8_t* sb = sa + A_STAGE_BYTES;
          mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
          if (p.conv) {
            const int tap = kb / p.cin_blocks;
            const int cb = kb - tap * p.cin_blocks;
            tma_load_4d(sa, &tmA, &full_bar[stage], cb * BK, x0 + p.tap_dx[tap], y0 + p.tap_dy[tap],
                        n0 + p.tap_dn[tap]);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], kb * BK, m_tile * BM);
          }
          tma_load_2d(sb, &tmB, &full_bar[stage], kb * BK, n_tile * p.BN);
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc((uint32_t)p.BN, false, 1u);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * p.BN);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * p.stage_bytes);
          const uint64_t adesc = umma_desc_k_sw128(sa);
          const uint64_t bdesc = umma_desc_k_sw128(sa + A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle atom: +2 in 16-byte units
            umma_f16_ss(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                        (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[as]);
      }
    }
  } else {
    // ===================== epilogue warps (2..5) =====================
    const int quarter = warp & 3;                // TMEM lane quarter this warp may access
    const int r_in_tile = quarter * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      const int m_tile = tile / p.num_n_tiles;
      const int n_tile = tile % p.num_n_tiles;
      long long out_row;
      bool row_ok;
      if (p.conv) {
        const int x = (m_tile % p.tiles_x) * p.w_box + (r_in_tile % p.w_box);
        const int y = ((m_tile / p.tiles_x) % p.tiles_y) * p.h_box + ((r_in_tile / p.w_box) % p.h_box);
        const int n = (m_tile / (p.tiles_x * p.tiles_y)) * p.n_box + r_in_tile / (p.w_box * p.h_box);
        row_ok = (x < p.w) && (y < p.h) && (n < p.nimg);
        out_row = ((long long)n * p.h + y) * p.w + x;
      } else {
        out_row = (long long)m_tile * BM + r_in_tile;
        row_ok = out_row < p.M;
      }
      mbar_wait(&tfull_bar[as], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * p.BN);
      for (int c = 0; c < p.BN; c += 32) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(taddr + (uint32_t)c, acc);
        tmem_wait_ld();
        epilogue_chunk(p, out_row, row_ok, n_tile * p.BN + c, acc);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ----------------------------------------------------------------------------- host side
static int pick_bn(int N, long long m_tiles, int num_sms) {
  // Prefer wide tiles (smem bandwidth per MMA) unless that leaves SMs idle on small problems.
  const int cands[4] = {256, 192, 160, 128};
  int best = 128;
  double best_cost = 1e30;
  for (int i = 0; i < 4; ++i) {
    const int bn = cands[i];
    const long long n_tiles = (N + bn - 1) / bn;
    const long long tiles = m_tiles * n_tiles;
    const long long waves = (tiles + num_sms - 1) / num_sms;
    // cost ~ waves * per-tile time (proportional to bn, with a fixed per-tile overhead), padded N included
    const double cost = (double)waves * (bn + 24.0);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = bn; }
  }
  return best;
}

static int g_num_sms = 0;
static bool g_attr_set = false;

int gemm_bf16_tc(const GemmArgs& a, cudaStream_t stream) {
  B200AD_REQUIRE(a.M > 0 && a.N > 0 && a.K > 0, "gemm: empty problem M=%d N=%d K=%d", a.M, a.N, a.K);
  B200AD_REQUIRE(a.N % 32 == 0, "gemm: N=%d must be a multiple of 32", a.N);
  B200AD_REQUIRE(a.K % 8 == 0, "gemm: K=%d must be a multiple of 8", a.K);
  if (g_num_sms == 0) {
    int dev = 0;
    B200AD_CHECK_CUDA(cudaGetDevice(&dev));
    B200AD_CHECK_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.epi = a.epi;
  CUtensorMap tmA, tmB;
  const ConvGeom& g = a.conv;
  if (g.enabled) {
    B200AD_REQUIRE(g.cin % BK == 0, "conv: cin=%d must be a multiple of %d on the tcgen05 path", g.cin, BK);
    B200AD_REQUIRE(g.n_box * g.h_box * g.w_box == BM, "conv: tile box %dx%dx%d != %d rows", g.n_box, g.h_box,
                   g.w_box, BM);
    B200AD_REQUIRE(a.K == g.ntaps * g.cin, "conv: K=%d != ntaps*cin=%d", a.K, g.ntaps * g.cin);
    B200AD_REQUIRE(a.M == g.nimg * g.h * g.w, "conv: M=%d != nimg*h*w", a.M);
    p.conv = 1;
    p.nimg = g.nimg; p.h = g.h; p.w = g.w;
    p.n_box = g.n_box; p.h_box = g.h_box; p.w_box = g.w_box;
    p.tiles_x = (g.w + g.w_box - 1) / g.w_box;
    p.tiles_y = (g.h + g.h_box - 1) / g.h_box;
    p.num_m_tiles = p.tiles_x * p.tiles_y * ((g.nimg + g.n_box - 1) / g.n_box);
    p.cin_blocks = g.cin / BK;
    for (int t = 0; t < 9; ++t) { p.tap_dy[t] = g.tap_dy[t]; p.tap_dx[t] = g.tap_dx[t]; p.tap_dn[t] = g.tap_dn[t]; }
    uint64_t dims[4] = {(uint64_t)g.cin, (uint64_t)g.src_w, (uint64_t)g.src_h, (uint64_t)g.src_nimg};
    uint64_t strides[3] = {(uint64_t)g.cin * 2, (uint64_t)g.src_w * g.cin * 2, (uint64_t)g.src_h * g.src_w * g.cin * 2};
    uint32_t box[4] = {(uint32_t)BK, (uint32_t)g.w_box, (uint32_t)g.h_box, (uint32_t)g.n_box};
    if (int rc = make_tmap_bf16(&tmA, a.A, 4, dims, strides, box)) return rc;
  } else {
    p.num_m_tiles = (a.M + BM - 1) / BM;
    uint64_t dims[2] = {(uint64_t)a.K, (uint64_t)a.M};
    uint64_t strides[1] = {(uint64_t)a.lda * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)BM};
    if (int rc = make_tmap_bf16(&tmA, a.A, 2, dims, strides, box)) return rc;
  }
  p.BN = pick_bn(a.N, p.num_m_tiles, g_num_sms);
  if (a.epi.geglu) B200AD_REQUIRE(a.epi.out_op && !a.epi.out_f32 && !a.epi.residual, "geglu epilogue writes out_op only");
  p.num_n_tiles = (a.N + p.BN - 1) / p.BN;
  p.num_kb = (a.K + BK - 1) / BK;
  p.stage_bytes = A_STAGE_BYTES + p.BN * BK * 2;
  p.num_stages = SMEM_TILE_BUDGET / p.stage_bytes;
  if (p.num_stages > 8) p.num_stages = 8;
  {
    uint64_t dims[2] = {(uint64_t)a.K, (uint64_t)a.N};
    uint64_t strides[1] = {(uint64_t)a.K * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)p.BN};
    if (int rc = make_tmap_bf16(&tmB, a.W, 2, dims, strides, box)) return rc;
  }
  const size_t smem_bytes = (size_t)p.num_stages * p.stage_bytes + 1024 /*align*/ + 256 /*barriers*/;
  if (!g_attr_set) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    g_attr_set = true;
  }
  const long long tiles = (long long)p.num_m_tiles * p.num_n_tiles;
  const int grid = (int)(tiles < g_num_sms ? tiles : g_num_sms);
  gemm_tc_kernel<<<grid, GEMM_THREADS, smem_bytes, stream>>>(tmA, tmB, p);
  count_launch();
  B200AD_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200ad
