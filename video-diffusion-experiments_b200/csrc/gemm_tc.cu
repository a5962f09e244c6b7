// tcgen05 GEMM / implicit-GEMM conv3x3 for sm_100a.
//   C[M,N] = A[M,K] * W[N,K]^T, bf16 operands, fp32 accumulation in TMEM, fused fp32 epilogue.
//
// CTA pair (cluster of 2, tcgen05 cta_group::2): one UMMA covers a 256 x BN output tile; each CTA stages its own
// 128 rows of A and its own BN/2 rows of W, so every operand byte leaves L2 once per pair (the single-CTA version
// of this kernel was bound by L2->SM operand traffic at ~49 % tensor-pipe activity, profiles/ncu_conv_r01b.md).
// Persistent, warp-specialised:  warp 0 = TMA producer of A, warp 1 = TMA producer of W (issuing the two loads of a
// stage from one thread made the producer the bottleneck: ~575 cycles per conv k-block against 320 cycles of UMMA
// work, profiles/gemm_waits_r01.txt), warp 2 = UMMA issuer (leader CTA) + TMEM allocation,
// warps 3..10 = epilogue math in two groups of four warps that take alternate 32-column chunks of the tile,
// warp 11 = epilogue I/O: TMA stores of finished chunks, buffer recycling and residual prefetch, so the math warps
// never wait for a store to drain and never block-synchronise (chunk hand-off is mbarrier arrive/wait only).
// Two TMEM accumulator stages: the epilogue of tile i overlaps the mainloop of tile i+1.
// Epilogue data path: the fp32 residual chunk is fetched by TMA into swizzled smem ahead of time, each thread adds
// its accumulator row (tcgen05.ld) + bias, the result is written back to smem and leaves through a TMA store --
// no uncoalesced per-thread global traffic, and the bytes in flight per SM are set by the number of chunk buffers.
// Conv mode loads the A tile with a 4-D tensor map over the NHWC image: one shifted box per filter tap,
// out-of-bounds (halo) pixels zero-filled by TMA; outputs use the same 4-D box so the pixel tile maps back.
#include <cooperative_groups.h>

#include "gemm.cuh"
#include "tmap.cuh"

namespace b200ad {

static constexpr int BM = 128;                      // rows per CTA (256 per pair)
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KiB
static constexpr int GEMM_THREADS = 384;
static constexpr int EPI_THREADS = 256;
static constexpr int F32_BUF_BYTES = 128 * 32 * 4;  // one [128 x 32] fp32 chunk, 128-byte rows, SWIZZLE_128B
static constexpr int OP_BUF_BYTES = 128 * 32 * 2;   // one [128 x 32] bf16 chunk, 64-byte rows, SWIZZLE_64B
static constexpr int MAX_STAGES = 8;
static constexpr int MAX_NBUF = 4;
static constexpr int SMEM_LIMIT = 232448;           // 227 KiB opt-in limit per CTA

// Cycle counters for tools/gemm_waits.py: compiled in only with -DB200AD_GEMM_PROFILE (reading the clock a dozen
// times per chunk is itself a measurable cost in the epilogue loops).
#ifdef B200AD_GEMM_PROFILE
__device__ __forceinline__ long long prof_clock() { return clock64(); }
#else
__device__ __forceinline__ long long prof_clock() { return 0; }
#endif

struct TcParams {
  int M, N, K;
  int BN;                      // tile width of the pair (each CTA stages BN/2 rows of W)
  int num_stages, stage_bytes;
  int num_m_pairs, num_n_tiles, num_kb;
  int conv;
  int nimg, h, w;
  int n_box, h_box, w_box;
  int tiles_x, tiles_y;
  int cin_blocks;
  int kb_main;                 // conv: k-blocks that come from the taps; the rest (if any) from EpiMaps::a2
  int nphase;                  // conv: 1, or 4 sub-pixel output phases in one launch (tile = (phase, m pair, n tile))
  int slots_per_phase;         // conv statistics: slots one phase fills per image (tiles per image)
  int tap_dy[9], tap_dx[9], tap_dn[9];
  // epilogue
  int has_res, has_o32, has_oop, geglu;
  int nbuf;                    // chunk buffers per epilogue group
  int off_f32, off_op, off_bias, off_bars;   // smem offsets (from the 1024-aligned base)
  const float* bias;
  const float* row_bias;
  int rb_div, ld_rb;
  long long* dbg;              // optional [grid][16] cycle counters (b200ad_debug_buffer), null in production
  // GroupNorm column statistics of the fp32 output (STATS instantiations): float2 [(image*spi + k)*N + col]
  float2* stats;
  int stats_spi;               // statistic slots per image
  int stats_mq;                // 32-row quarters of a tile merged into one slot (4: the whole tile lies in one image)
  int stats_slot0;             // conv: first slot of this launch inside an image's stats_spi slots (sub-pixel phases)
  int off_stat;                // smem: [2 groups][MAX_NBUF][4 quarters][32 columns] float2 hand-over to the I/O warp
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_smem_addr` in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default semantics (.release at CTA scope): the accumulator hand-off is ordered by tcgen05.wait::ld +
  // tcgen05.fence::before_thread_sync; a cluster-scope release compiles to MEMBAR + ERRBAR (~1000 cycles per tile
  // and warp in profiles/ncu_full_r01f)
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA loads of a CTA pair: data lands in this CTA's smem, the transaction bytes are reported to the mbarrier at
// cluster address `bar` (the leader CTA's full barrier).
__device__ __forceinline__ void tma2_load_2d(void* dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_4d(void* dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma2_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all previously issued MMAs of this thread complete -> one arrive on the barrier at this smem offset in BOTH CTAs
__device__ __forceinline__ void umma2_commit_mc(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__host__ __device__ __forceinline__ uint32_t umma_idesc_m(uint32_t m, uint32_t n, uint32_t ab_fmt) {
  uint32_t d = 0;
  d |= 1u << 4;               // c_format = F32
  d |= ab_fmt << 7;           // a_format (1 = bf16)
  d |= ab_fmt << 10;          // b_format
  d |= (n >> 3) << 17;        // n_dim
  d |= (m >> 4) << 24;        // m_dim
  return d;
}

// Coordinates of one tile of this CTA.
struct TileCoord {
  int m_tile, n_tile;
  int x0, y0, n0;
  int ph;                      // sub-pixel phase py*2 + px (0 unless TcParams::nphase == 4)
};
__device__ __forceinline__ TileCoord tile_coord(const TcParams& p, int tile, int rank) {
  TileCoord t;
  t.ph = 0;
  if (p.nphase > 1) {
    const int per_phase = p.num_m_pairs * p.num_n_tiles;
    t.ph = tile / per_phase;
    tile -= t.ph * per_phase;
  }
  t.n_tile = tile % p.num_n_tiles;
  t.m_tile = (tile / p.num_n_tiles) * 2 + rank;
  t.x0 = t.y0 = t.n0 = 0;
  if (p.conv) {
    t.x0 = (t.m_tile % p.tiles_x) * p.w_box;
    t.y0 = ((t.m_tile / p.tiles_x) % p.tiles_y) * p.h_box;
    t.n0 = (t.m_tile / (p.tiles_x * p.tiles_y)) * p.n_box;
  }
  return t;
}
// accumulator columns per output chunk (a chunk is always 32 OUTPUT columns)
__device__ __forceinline__ int acc_width(const TcParams& p) { return p.geglu ? 64 : 32; }
__device__ __forceinline__ int tile_chunks(const TcParams& p, int n_tile) {
  int cols = p.N - n_tile * p.BN;
  if (cols > p.BN) cols = p.BN;
  return cols >> (p.geglu ? 6 : 5);
}

// Iterator over the chunks one epilogue group processes: tiles blockIdx-strided, chunks c = g, g+2, ...
// The tile's coordinates are cached so the integer divisions happen once per tile, not once per chunk.
// Chunk c of the cluster's it-th tile belongs to epilogue group (c + it) & 1: with an odd number of chunks per tile
// (BN = 160: five) the group that gets three of them alternates from tile to tile.
struct ChunkCursor {
  int tile, c, nchunks, it;
  TileCoord t;
};
__device__ __forceinline__ void cursor_load(const TcParams& p, ChunkCursor& cur, int num_tiles, int rank) {
  if (cur.tile < num_tiles) {
    cur.t = tile_coord(p, cur.tile, rank);
    cur.nchunks = tile_chunks(p, cur.t.n_tile);
  }
}
__device__ __forceinline__ void cursor_settle(const TcParams& p, ChunkCursor& cur, int g, int num_tiles, int stride,
                                              int rank) {
  while (cur.tile < num_tiles && cur.c >= cur.nchunks) {
    cur.tile += stride;
    cur.it += 1;
    cur.c = (g ^ cur.it) & 1;
    cursor_load(p, cur, num_tiles, rank);
  }
}

// explicit shared-space accesses (the aligned dynamic-smem base is an integer-derived generic pointer: plain
// dereferences compile to generic LD.E / ST.E)
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f4(uint32_t a, float x, float y, float z, float w) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ void sts_u4(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

struct EpiMaps {
  CUtensorMap res, o32, oop;
  CUtensorMap a2;              // conv: source of the second K segment (ConvGeom::sc_src)
  CUtensorMap o32p[3];         // fp32 output maps of sub-pixel phases 1..3 (phase 0 uses o32)
};

// Elected epilogue thread: make chunk buffer `buf` of group `g` ready for chunk `cur` (fetch its residual, or
// just publish that the buffer is free).
__device__ __forceinline__ void chunk_prefetch(const TcParams& p, const EpiMaps& em, uint8_t* f32buf, uint64_t* bar,
                                               const ChunkCursor& cur, int rank) {
  if (!p.has_res) {
    mbar_arrive(bar);
    return;
  }
  const TileCoord& t = cur.t;
  const int col0 = t.n_tile * p.BN + cur.c * 32;          // residual is never combined with geglu
  mbar_arrive_expect_tx(bar, F32_BUF_BYTES);
  if (p.conv) tma_load_4d(f32buf, &em.res, bar, col0, t.x0, t.y0, t.n0);
  else tma_load_2d(f32buf, &em.res, bar, col0, t.m_tile * BM);
}

// Per-group state of the epilogue I/O warp.
struct IoGroup {
  uint8_t* f32;            // this group's fp32 chunk buffers
  uint8_t* op;             // this group's bf16 staging buffers
  uint64_t* ready;         // res_full barriers of the group
  uint64_t* cdone;         // chunk_done barriers of the group
  int buf, prev_buf;       // buffer of the chunk being stored / of the previous chunk (the next one to recycle)
  uint32_t phase;
  ChunkCursor pre;         // next chunk whose buffer has not been handed out yet
};
__device__ __forceinline__ void io_group_init(const TcParams& p, const EpiMaps& em, IoGroup& G, int g, uint8_t* smem,
                                              uint64_t* res_full, uint64_t* chunk_done, int cluster_id,
                                              int num_tiles, int num_clusters, int rank, bool io_lane) {
  G.f32 = smem + p.off_f32 + (size_t)g * p.nbuf * F32_BUF_BYTES;
  G.op = smem + p.off_op + (size_t)g * p.nbuf * OP_BUF_BYTES;
  G.ready = res_full + g * MAX_NBUF;
  G.cdone = chunk_done + g * MAX_NBUF;
  G.buf = 0;
  G.prev_buf = p.nbuf - 1;           // the one buffer not handed out below
  G.phase = 0;
  G.pre.tile = cluster_id;
  G.pre.it = 0;
  G.pre.c = g;
  cursor_load(p, G.pre, num_tiles, rank);
  cursor_settle(p, G.pre, g, num_tiles, num_clusters, rank);
  for (int b = 0; b < p.nbuf - 1 && G.pre.tile < num_tiles; ++b) {
    if (io_lane) chunk_prefetch(p, em, G.f32 + (size_t)b * F32_BUF_BYTES, &G.ready[b], G.pre, rank);
    __syncwarp();
    G.pre.c += 2;
    cursor_settle(p, G.pre, g, num_tiles, num_clusters, rank);
  }
}

// EPI: the epilogue's output combination as a compile-time constant -- a bit mask (8 GEGLU, 4 residual, 2 fp32 output,
// 1 operand output) for the combinations the UNet uses, or -1 for run-time flags.  With run-time flags every chunk of
// the epilogue loops went through five uniform branches, and a branch is a scheduling barrier (as in attn_fa.cu).
__device__ __forceinline__ float lds_f1(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}

// STATS: the math warps also emit per-(32-row slot, column) sum / sum of squares of the fp32 output they just wrote
// to the chunk buffer (each warp owns rows quarter*32 .. +31 of the tile: lane l re-reads column l of those rows from
// the swizzled buffer, conflict-free), so the GroupNorm that follows needs no statistics pass over the tensor.
template <int EPI, bool STATS = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ EpiMaps em, const __grid_constant__ TcParams p) {
  const bool f_geglu = EPI < 0 ? (bool)p.geglu : (EPI & 8) != 0;
  const bool f_res = EPI < 0 ? (bool)p.has_res : (EPI & 4) != 0;
  const bool f_o32 = EPI < 0 ? (bool)p.has_o32 : (EPI & 2) != 0;
  const bool f_oop = EPI < 0 ? (bool)p.has_oop : (EPI & 1) != 0;
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bars);
  uint64_t* full_bar = bars;                            // [MAX_STAGES]  (used in the leader CTA)
  uint64_t* empty_bar = full_bar + MAX_STAGES;          // [MAX_STAGES]  (both CTAs, signalled by multicast commit)
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;         // [2]           (both CTAs)
  uint64_t* tempty_bar = tfull_bar + 2;                 // [2]           (leader; 16 remote/local warp arrivals)
  uint64_t* res_full = tempty_bar + 2;                  // [2][MAX_NBUF] chunk buffer ready (per epilogue group)
  uint64_t* chunk_done = res_full + 2 * MAX_NBUF;       // [2][MAX_NBUF] chunk written to smem by its 128 threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(chunk_done + 2 * MAX_NBUF);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = p.num_m_pairs * p.num_n_tiles * p.nphase;
  const uint32_t tmem_cols = (2 * p.BN <= 256) ? 256u : 512u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (f_res) tma_prefetch_desc(&em.res);
    if (f_o32) tma_prefetch_desc(&em.o32);
    if (f_oop) tma_prefetch_desc(&em.oop);
    if (p.conv && p.kb_main < p.num_kb) tma_prefetch_desc(&em.a2);
    if (p.nphase > 1)
      for (int i = 0; i < 3; ++i) tma_prefetch_desc(&em.o32p[i]);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(&full_bar[s], 2);        // the leader's A producer and W producer each arrive with their tx bytes
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 16);
    }
    for (int i = 0; i < 2 * MAX_NBUF; ++i) {
      mbar_init(&res_full[i], 1);
      mbar_init(&chunk_done[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc2(tmem_slot, tmem_cols);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();                    // barrier inits + TMEM allocation visible in both CTAs
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                            // everything above overlapped the previous kernel's tail

  if (warp == 0) {
    // ===================== TMA producer of A (both CTAs; transaction bytes go to the leader's barrier) =====
    // The whole warp runs the loop so that every operand of the TMA / UMMA instructions is warp-uniform (uniform
    // registers); only the issue itself is predicated on one elected lane.  Issuing from inside an `if (lane == 0)`
    // branch made ptxas wrap every UTMALDG / UTCHMMA in an ELECT + R2UR.BROADCAST loop (~100 cycles each).
    {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t fb0 = mapa_u32(smem_u32(&full_bar[0]), 0);
      long long t_wait = 0;
      const long long t_begin = prof_clock();
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const TileCoord t = tile_coord(p, tile, rank);
        int tap = 0, cb = 0;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          const long long w0 = prof_clock();
          mbar_wait(&empty_bar[stage], phase ^ 1);
          t_wait += prof_clock() - w0;
          uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
          if (elect_one()) {
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2u * A_STAGE_BYTES);
            if (p.conv) {
              if (kb < p.kb_main)
                tma2_load_4d(sa, &tmA, fb0 + 8u * stage, cb * BK, t.x0 + p.tap_dx[tap] + (t.ph & 1),
                             t.y0 + p.tap_dy[tap] + (t.ph >> 1), t.n0 + p.tap_dn[tap]);
              else
                tma2_load_4d(sa, &em.a2, fb0 + 8u * stage, (kb - p.kb_main) * BK, t.x0, t.y0, t.n0);
            } else
              tma2_load_2d(sa, &tmA, fb0 + 8u * stage, kb * BK, t.m_tile * BM);
          }
          __syncwarp();
          if (p.conv && kb < p.kb_main && ++cb == p.cin_blocks) { cb = 0; ++tap; }
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
      if (p.dbg && lane == 0) {
        p.dbg[blockIdx.x * 16 + 0] = prof_clock() - t_begin;
        p.dbg[blockIdx.x * 16 + 1] = t_wait;
      }
    }
  } else if (warp == 1) {
    // ===================== TMA producer of W: this CTA's BN/2 rows of the tile =====================
    {
      int stage = 0;
      uint32_t phase = 0;
      const int half_bn = p.BN >> 1;
      const uint32_t b_bytes = (uint32_t)(p.stage_bytes - A_STAGE_BYTES);
      const uint32_t fb0 = mapa_u32(smem_u32(&full_bar[0]), 0);
      long long t_wait = 0;
      const long long t_begin = prof_clock();
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const TileCoord tw = tile_coord(p, tile, rank);
        const int n_row = tw.ph * p.N + tw.n_tile * p.BN + rank * half_bn;     // phase weights are stacked on the rows
        for (int kb = 0; kb < p.num_kb; ++kb) {
          const long long w0 = prof_clock();
          mbar_wait(&empty_bar[stage], phase ^ 1);
          t_wait += prof_clock() - w0;
          uint8_t* sb = smem + (size_t)stage * p.stage_bytes + A_STAGE_BYTES;
          if (elect_one()) {
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2u * b_bytes);
            tma2_load_2d(sb, &tmB, fb0 + 8u * stage, kb * BK, n_row);
          }
          __syncwarp();
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
      if (p.dbg && lane == 0) {
        p.dbg[blockIdx.x * 16 + 5] = prof_clock() - t_begin;
        p.dbg[blockIdx.x * 16 + 6] = t_wait;
      }
    }
  } else if (warp == 2) {
    // ===================== UMMA issuer (leader CTA only) =====================
    if (rank == 0) {
      const uint32_t idesc = umma_idesc_m(256u, (uint32_t)p.BN, 1u);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      long long t_full = 0, t_tempty = 0;
      const long long t_begin = prof_clock();
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
        const int as = it & 1;
        const long long w0 = prof_clock();
        mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
        t_tempty += prof_clock() - w0;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * p.BN);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          const long long w1 = prof_clock();
          mbar_wait(&full_bar[stage], phase);
          t_full += prof_clock() - w1;
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * p.stage_bytes);
          const uint64_t adesc = umma_desc_k_sw128(sa);
          const uint64_t bdesc = umma_desc_k_sw128(sa + A_STAGE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle atom: +2 in 16-byte units
              umma2_f16_ss(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                           (kb | k) != 0 ? 1u : 0u);
            }
            umma2_commit_mc(&empty_bar[stage]);
            if (kb == p.num_kb - 1) umma2_commit_mc(&tfull_bar[as]);
          }
          __syncwarp();
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
      if (p.dbg && lane == 0) {
        p.dbg[blockIdx.x * 16 + 2] = prof_clock() - t_begin;
        p.dbg[blockIdx.x * 16 + 3] = t_full;
        p.dbg[blockIdx.x * 16 + 4] = t_tempty;
      }
    }
  } else if (warp == 11) {
    // ===================== epilogue I/O warp =====================
    // Walks the chunks of this CTA in tile order (chunk c belongs to math group c & 1).  For each: wait until the
    // group's 128 threads have written it, issue its TMA store(s), and once the store of the group's previous chunk
    // has drained hand that buffer to the chunk nbuf-1 ahead (fetch its residual / publish that it is free).
    // All per-group state lives in scalars (no indexed arrays, no divisions: this loop is the epilogue's pacemaker).
    const bool io_lane = elect_one();
    const int AW = (f_geglu ? 64 : 32);
    const int nbuf = p.nbuf;
    IoGroup G0, G1;
    io_group_init(p, em, G0, 0, smem, res_full, chunk_done, cluster_id, num_tiles, num_clusters, rank, io_lane);
    io_group_init(p, em, G1, 1, smem, res_full, chunk_done, cluster_id, num_tiles, num_clusters, rank, io_lane);
    int last_g = -1;
    long long io_wait_done = 0, io_wait_read = 0, io_issue = 0;
    const long long io_begin = prof_clock();
    int io_it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++io_it) {
      const TileCoord t = tile_coord(p, tile, rank);
      const int nchunks = tile_chunks(p, t.n_tile);
      const int ocol0 = (t.n_tile * p.BN) >> (f_geglu ? 1 : 0);
      // column statistics: element offset (slot * N) of the up to four slots this tile feeds, -1 = no such image / rows
      const int st_mq = STATS ? p.stats_mq : 4;
      long long st_off[4] = {-1, -1, -1, -1};
      if (STATS) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j * st_mq >= 4) continue;
          if (p.conv) {
            const int img = t.n0 + j;
            if (img < p.nimg) st_off[j] = ((long long)img * p.stats_spi + p.stats_slot0 + t.ph * p.slots_per_phase + (t.m_tile % (p.tiles_x * p.tiles_y))) * p.N;
          } else {
            const long long row0 = (long long)t.m_tile * BM + j * 32 * st_mq;
            if (row0 < p.M) st_off[j] = (row0 / (32 * st_mq)) * p.N;
          }
        }
      }
      for (int c = 0; c < nchunks; ++c) {
        const int g = (c + io_it) & 1;
        IoGroup& G = g ? G1 : G0;
        const long long i0 = prof_clock();
        mbar_wait(&G.cdone[G.buf], G.phase);
        const long long i1 = prof_clock();
        io_wait_done += i1 - i0;
        if (io_lane) {
          uint8_t* fb = G.f32 + (size_t)G.buf * F32_BUF_BYTES;
          uint8_t* ob = G.op + (size_t)G.buf * OP_BUF_BYTES;
          const int ocol = ocol0 + c * 32;
          if (p.conv) {
            if (f_o32) tma_store_4d(t.ph == 0 ? &em.o32 : &em.o32p[t.ph - 1], fb, ocol, t.x0, t.y0, t.n0);
            if (f_oop) tma_store_4d(&em.oop, ob, ocol, t.x0, t.y0, t.n0);
          } else {
            if (f_o32) tma_store_2d(&em.o32, fb, ocol, t.m_tile * BM);
            if (f_oop) tma_store_2d(&em.oop, ob, ocol, t.m_tile * BM);
          }
          tma_store_commit();
        }
        if (STATS) {
          // (after the chunk's TMA store is on its way) lane = column of the chunk: merge the quarters that belong to
          // one image and store one float2 per slot and column
          const uint32_t sa = smem_u32(smem + p.off_stat) + (uint32_t)((((g * MAX_NBUF + G.buf) * 4) * 32 + lane) * 8);
          float2 qs[4];
#pragma unroll
          for (int qq = 0; qq < 4; ++qq)
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(qs[qq].x), "=f"(qs[qq].y) : "r"(sa + (uint32_t)(qq * 256)));
          float2* col = p.stats + (t.n_tile * p.BN + c * 32 + lane);
          if (st_mq == 4) {
            if (st_off[0] >= 0) col[st_off[0]] = make_float2((qs[0].x + qs[1].x) + (qs[2].x + qs[3].x), (qs[0].y + qs[1].y) + (qs[2].y + qs[3].y));
          } else if (st_mq == 2) {
            if (st_off[0] >= 0) col[st_off[0]] = make_float2(qs[0].x + qs[1].x, qs[0].y + qs[1].y);
            if (st_off[1] >= 0) col[st_off[1]] = make_float2(qs[2].x + qs[3].x, qs[2].y + qs[3].y);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (st_off[j] >= 0) col[st_off[j]] = qs[j];
          }
        }
        if (io_lane) {
          const long long i2 = prof_clock();
          io_issue += i2 - i1;
          // the store of this group's previous chunk is the 2nd (or, at an odd tile boundary, the 1st) most recent
          if (last_g == g) tma_store_wait_read1(); else tma_store_wait_read2();
          io_wait_read += prof_clock() - i2;
          // no residual: the hand-off is a bare arrive, no look-ahead cursor to maintain (a surplus arrive after
          // the group's last chunk completes a phase nobody waits for)
          if (!f_res) mbar_arrive(&G.ready[G.prev_buf]);
          else if (G.pre.tile < num_tiles)
            chunk_prefetch(p, em, G.f32 + (size_t)G.prev_buf * F32_BUF_BYTES, &G.ready[G.prev_buf], G.pre, rank);
        }
        __syncwarp();
        if (f_res && G.pre.tile < num_tiles) {
          G.pre.c += 2;
          cursor_settle(p, G.pre, g, num_tiles, num_clusters, rank);
        }
        G.prev_buf = G.buf;
        if (++G.buf == nbuf) { G.buf = 0; G.phase ^= 1u; }
        last_g = g;
      }
    }
    (void)AW;
    if (io_lane) tma_store_wait_all();
    if (p.dbg && io_lane) {
      long long* d = p.dbg + blockIdx.x * 16;
      d[7] = prof_clock() - io_begin; d[11] = io_wait_done; d[12] = io_wait_read; d[15] = io_issue;
    }
  } else {
    // ===================== epilogue: 2 groups x 4 warps =====================
    const int ew = warp - 3;                     // 0..7
    const int g = ew >> 2;                       // group: chunks c = g, g+2, ...
    const int quarter = warp & 3;                // TMEM lane quarter this warp may access
    const int r = quarter * 32 + lane;           // row of the CTA tile owned by this thread
    const int et = threadIdx.x - 96;             // 0..255
    const int AW = (f_geglu ? 64 : 32);
    uint8_t* f32bufs = smem + p.off_f32 + (size_t)g * p.nbuf * F32_BUF_BYTES;
    uint8_t* opbufs = smem + p.off_op + (size_t)g * p.nbuf * OP_BUF_BYTES;
    uint64_t* rbar = res_full + g * MAX_NBUF;
    uint64_t* cdone = chunk_done + g * MAX_NBUF;
    const uint32_t tempty_remote = mapa_u32(smem_u32(&tempty_bar[0]), 0);
    int k = 0;                                   // chunks processed by this group so far
    int buf = 0;                                 // chunk buffer of chunk k, and its barrier phase
    uint32_t bphase = 0;
    int it = 0;
    long long e_tfull = 0, e_rbar = 0, e_bar = 0, e_store = 0, e_bias = 0;
    const long long e_begin = prof_clock();
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int as = it & 1;
      const TileCoord t = tile_coord(p, tile, rank);
      const int nchunks = tile_chunks(p, t.n_tile);
      // ---- bias / timestep-embedding row: read straight from global memory in the chunk loop (every thread of a
      // warp reads the same 128 bytes of `bias`: one L1-resident broadcast line per chunk).  Staging them in smem
      // cost a block-wide named barrier per tile, which showed up as ~5 % of the epilogue at K = 320.
      long long out_row = 0;
      if (p.row_bias) {
        if (p.conv) {
          const int x = t.x0 + (r % p.w_box), y = t.y0 + ((r / p.w_box) % p.h_box);
          int n = t.n0 + r / (p.w_box * p.h_box);
          if (n > p.nimg - 1) n = p.nimg - 1;
          out_row = ((long long)n * p.h + (y < p.h ? y : p.h - 1)) * p.w + (x < p.w ? x : p.w - 1);
        } else {
          out_row = (long long)t.m_tile * BM + r;
          if (out_row > p.M - 1) out_row = p.M - 1;
        }
      }
      // column statistics: which rows of this warp's quarter exist, and the slot they are summed into
      uint32_t st_mask = 0;
      if (STATS) {
        bool valid;
        if (p.conv) {
          const int x = t.x0 + (r % p.w_box), y = t.y0 + ((r / p.w_box) % p.h_box);
          const int n = t.n0 + r / (p.w_box * p.h_box);
          valid = x < p.w && y < p.h && n < p.nimg;
        } else {
          valid = (long long)t.m_tile * BM + r < p.M;
        }
        st_mask = __ballot_sync(0xffffffffu, valid);
      }
      const long long q0 = prof_clock();
      const float* bias_t = p.bias ? p.bias + t.n_tile * p.BN : nullptr;
      const float* rb_row =
          p.row_bias ? p.row_bias + (out_row / p.rb_div) * (long long)p.ld_rb + t.n_tile * p.BN : nullptr;

      const long long q1 = prof_clock();
      e_bias += q1 - q0;
      mbar_wait(&tfull_bar[as], (it >> 1) & 1);
      e_tfull += prof_clock() - q1;
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * p.BN);
      for (int c = (g ^ it) & 1; c < nchunks; c += 2, ++k) {
        uint8_t* fb = f32bufs + (size_t)buf * F32_BUF_BYTES;
        uint8_t* ob = opbufs + (size_t)buf * OP_BUF_BYTES;
        float v[32];
        {
          uint32_t acc[32];
          tmem_ld_32x32b_x32(taddr + (uint32_t)(c * AW), acc);
          tmem_wait_ld();
          const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
          const float4* bp = reinterpret_cast<const float4*>(bias_t + c * AW);
          if (f_geglu) {
            uint32_t acc2[32];
            tmem_ld_32x32b_x32(taddr + (uint32_t)(c * AW + 32), acc2);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b0 = bias_t ? __ldg(bp + j) : z4, b1 = bias_t ? __ldg(bp + 8 + j) : z4;
              v[2 * j] = (__uint_as_float(acc[4 * j]) + b0.x) * gelu_tanhfit_f(__uint_as_float(acc[4 * j + 1]) + b0.y);
              v[2 * j + 1] = (__uint_as_float(acc[4 * j + 2]) + b0.z) * gelu_tanhfit_f(__uint_as_float(acc[4 * j + 3]) + b0.w);
              v[16 + 2 * j] = (__uint_as_float(acc2[4 * j]) + b1.x) * gelu_tanhfit_f(__uint_as_float(acc2[4 * j + 1]) + b1.y);
              v[17 + 2 * j] =
                  (__uint_as_float(acc2[4 * j + 2]) + b1.z) * gelu_tanhfit_f(__uint_as_float(acc2[4 * j + 3]) + b1.w);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b = bias_t ? __ldg(bp + j) : z4;
              v[4 * j] = __uint_as_float(acc[4 * j]) + b.x;
              v[4 * j + 1] = __uint_as_float(acc[4 * j + 1]) + b.y;
              v[4 * j + 2] = __uint_as_float(acc[4 * j + 2]) + b.z;
              v[4 * j + 3] = __uint_as_float(acc[4 * j + 3]) + b.w;
            }
            if (rb_row) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 q = __ldg(reinterpret_cast<const float4*>(rb_row + c * 32 + j));
                v[j] += q.x; v[j + 1] += q.y; v[j + 2] += q.z; v[j + 3] += q.w;
              }
            }
          }
        }
        // the chunk buffer is free (its previous TMA store has read it) and holds the residual chunk if any
        const long long q2 = prof_clock();
        mbar_wait(&rbar[buf], bphase);
        e_rbar += prof_clock() - q2;
        if (f_res || f_o32) {
          const uint32_t rowa = smem_u32(fb) + (uint32_t)(r * 128);
          const int sw = r & 7;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t qa = rowa + (uint32_t)((j ^ sw) << 4);
            if (f_res) {
              const float4 x = lds_f4(qa);
              v[4 * j] += x.x; v[4 * j + 1] += x.y; v[4 * j + 2] += x.z; v[4 * j + 3] += x.w;
            }
            if (f_o32) sts_f4(qa, v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
        }
        if (f_oop) {
          const uint32_t rowa = smem_u32(ob) + (uint32_t)(r * 64);
          const int sw = (r >> 1) & 3;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            sts_u4(rowa + (uint32_t)((j ^ sw) << 4), pack_bf16x2(v[8 * j], v[8 * j + 1]),
                   pack_bf16x2(v[8 * j + 2], v[8 * j + 3]), pack_bf16x2(v[8 * j + 4], v[8 * j + 5]),
                   pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        }
        if (STATS) {
          __syncwarp();                       // this warp's 32 rows of the chunk are in the buffer
          const uint32_t colw = (uint32_t)((lane & 3) << 2);
          const uint32_t base = smem_u32(fb) + (uint32_t)(quarter * 32 * 128);
          float s0 = 0.f, s1 = 0.f, q0_ = 0.f, q1_ = 0.f;
#pragma unroll
          for (int rr = 0; rr < 32; rr += 2) {
            float x0 = lds_f1(base + (uint32_t)(rr * 128) + (uint32_t)((((lane >> 2) ^ (rr & 7)) << 4)) + colw);
            float x1 = lds_f1(base + (uint32_t)((rr + 1) * 128) + (uint32_t)((((lane >> 2) ^ ((rr + 1) & 7)) << 4)) + colw);
            if (!((st_mask >> rr) & 1u)) x0 = 0.f;
            if (!((st_mask >> (rr + 1)) & 1u)) x1 = 0.f;
            s0 += x0; q0_ = fmaf(x0, x0, q0_);
            s1 += x1; q1_ = fmaf(x1, x1, q1_);
          }
          // hand the quarter's column sums to the I/O warp, which merges the quarters of one image and stores them
          const uint32_t sa = smem_u32(smem + p.off_stat) + (uint32_t)((((g * MAX_NBUF + buf) * 4 + quarter) * 32 + lane) * 8);
          asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(sa), "f"(s0 + s1), "f"(q0_ + q1_) : "memory");
        }
        fence_proxy_async_smem();             // generic-proxy smem writes -> visible to the TMA store
        mbar_arrive(&cdone[buf]);
        if (++buf == p.nbuf) { buf = 0; bphase ^= 1u; }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tempty_remote + (uint32_t)(as * 8));
    }
    if (p.dbg && ew == 0 && lane == 0) {
      long long* d = p.dbg + blockIdx.x * 16;
      d[8] = prof_clock() - e_begin; d[9] = e_tfull; d[10] = e_rbar; d[13] = e_bias;
      d[14] = k;
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, tmem_cols);
  }
}

// ----------------------------------------------------------------------------- host side
static int g_num_sms = 0;
static bool g_attr_set = false;
static long long* g_dbg = nullptr;
void set_gemm_debug_buffer(long long* p) { g_dbg = p; }

static int pick_bn(int N, long long m_pairs, int num_clusters, bool geglu) {
  // Prefer wide tiles (operand bytes per MMA) unless that leaves SM pairs idle on small problems.
  const int cands[4] = {256, 192, 160, 128};
  int best = 128;
  double best_cost = 1e30;
  for (int i = 0; i < 4; ++i) {
    const int bn = cands[i];
    if (geglu && bn % 64 != 0) continue;
    const long long n_tiles = (N + bn - 1) / bn;
    const long long tiles = m_pairs * n_tiles;
    const long long waves = (tiles + num_clusters - 1) / num_clusters;
    // cost ~ waves * per-tile time (proportional to bn, with a fixed per-tile overhead), padded N included
    const double cost = (double)waves * (bn + 24.0);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = bn; }
  }
  return best;
}

// Output / residual tensor map: dense [M, ld] rows, or the NHWC pixel box of a conv tile.  `elem` = 4 (fp32) or 2.
static int make_out_map(CUtensorMap* tm, const void* ptr, long long ld, int ncols, int elem, const GemmArgs& a,
                        int phase = -1) {
  ConvGeom g = a.conv;
  if (phase >= 0) { g.out_py = phase >> 1; g.out_px = phase & 1; }
  const CUtensorMapDataType dt = elem == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const CUtensorMapSwizzle sw = elem == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  if (g.enabled) {
    // sub-pixel phase (out_scale = 2): the launch's pixels are every 2nd pixel of every 2nd row of the full image
    const uint64_t sc = g.out_scale > 1 ? (uint64_t)g.out_scale : 1u;
    const uint64_t wf = sc * g.w, hf = sc * g.h;
    const char* base = reinterpret_cast<const char*>(ptr) + ((uint64_t)g.out_py * wf + (uint64_t)g.out_px) * ld * elem;
    uint64_t dims[4] = {(uint64_t)ncols, (uint64_t)g.w, (uint64_t)g.h, (uint64_t)g.nimg};
    uint64_t str[3] = {sc * ld * elem, sc * wf * ld * elem, hf * wf * ld * elem};
    uint32_t box[4] = {32u, (uint32_t)g.w_box, (uint32_t)g.h_box, (uint32_t)g.n_box};
    return make_tmap(tm, dt, base, 4, dims, str, box, sw);
  }
  uint64_t dims[2] = {(uint64_t)ncols, (uint64_t)a.M};
  uint64_t str[1] = {(uint64_t)ld * elem};
  uint32_t box[2] = {32u, (uint32_t)BM};
  return make_tmap(tm, dt, ptr, 2, dims, str, box, sw);
}

int gemm_bf16_tc(const GemmArgs& a, cudaStream_t stream) {
  B200AD_REQUIRE(a.M > 0 && a.N > 0 && a.K > 0, "gemm: empty problem M=%d N=%d K=%d", a.M, a.N, a.K);
  B200AD_REQUIRE(a.N % 32 == 0, "gemm: N=%d must be a multiple of 32", a.N);
  B200AD_REQUIRE(a.K % 8 == 0, "gemm: K=%d must be a multiple of 8", a.K);
  if (g_num_sms == 0) {
    int dev = 0;
    B200AD_CHECK_CUDA(cudaGetDevice(&dev));
    B200AD_CHECK_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const Epilogue& e = a.epi;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.has_res = e.residual != nullptr;
  p.has_o32 = e.out_f32 != nullptr;
  p.has_oop = e.out_op != nullptr;
  p.geglu = e.geglu;
  p.dbg = g_dbg;
  p.stats = reinterpret_cast<float2*>(e.col_stats);
  p.bias = e.bias; p.row_bias = e.row_bias; p.rb_div = e.rb_div > 0 ? e.rb_div : 1; p.ld_rb = e.ld_rb;
  if (e.geglu) {
    B200AD_REQUIRE(e.out_op && !e.out_f32 && !e.residual && !e.row_bias, "geglu epilogue writes out_op only");
    B200AD_REQUIRE(a.N % 64 == 0, "geglu: N=%d must be a multiple of 64", a.N);
  }
  if (e.residual) B200AD_REQUIRE(e.ld_res % 4 == 0, "gemm: residual leading dimension %lld not a multiple of 4", e.ld_res);
  if (e.out_f32) B200AD_REQUIRE(e.ld_o32 % 4 == 0, "gemm: out_f32 leading dimension %lld not a multiple of 4", e.ld_o32);
  if (e.out_op) B200AD_REQUIRE(e.ld_oop % 8 == 0, "gemm: out_op leading dimension %lld not a multiple of 8", e.ld_oop);
  CUtensorMap tmA, tmB;
  EpiMaps em;
  memset(&em, 0, sizeof(em));
  const ConvGeom& g = a.conv;
  int num_m_tiles;
  if (g.enabled) {
    B200AD_REQUIRE(g.cin % BK == 0, "conv: cin=%d must be a multiple of %d on the tcgen05 path", g.cin, BK);
    B200AD_REQUIRE(g.n_box * g.h_box * g.w_box == BM, "conv: tile box %dx%dx%d != %d rows", g.n_box, g.h_box,
                   g.w_box, BM);
    B200AD_REQUIRE(a.K == g.ntaps * g.cin + g.sc_cin, "conv: K=%d != ntaps*cin + sc_cin = %d", a.K,
                   g.ntaps * g.cin + g.sc_cin);
    B200AD_REQUIRE(g.sc_cin == 0 || (g.sc_src && g.sc_cin % BK == 0),
                   "conv: the second K segment needs a source and a multiple of %d channels (got %d)", BK, g.sc_cin);
    B200AD_REQUIRE(a.M == g.nimg * g.h * g.w, "conv: M=%d != nimg*h*w", a.M);
    B200AD_REQUIRE(g.out_scale <= 1 || !e.row_bias, "conv: a row bias cannot be combined with a sub-pixel output");
    p.conv = 1;
    p.nimg = g.nimg; p.h = g.h; p.w = g.w;
    p.n_box = g.n_box; p.h_box = g.h_box; p.w_box = g.w_box;
    p.tiles_x = (g.w + g.w_box - 1) / g.w_box;
    p.tiles_y = (g.h + g.h_box - 1) / g.h_box;
    num_m_tiles = p.tiles_x * p.tiles_y * ((g.nimg + g.n_box - 1) / g.n_box);
    p.cin_blocks = g.cin / BK;
    p.kb_main = g.ntaps * p.cin_blocks;
    if (g.nphase == 4) {
      B200AD_REQUIRE(g.out_scale == 2 && !e.residual && !e.out_op && e.out_f32 && g.sc_cin == 0,
                     "conv: the four-phase launch writes an fp32 output only (no residual / operand copy / second K segment)");
      p.nphase = 4;
    }
    if (g.sc_cin > 0) {
      uint64_t d2[4] = {(uint64_t)g.sc_cin, (uint64_t)g.w, (uint64_t)g.h, (uint64_t)g.nimg};
      uint64_t s2[3] = {(uint64_t)g.sc_cin * 2, (uint64_t)g.w * g.sc_cin * 2, (uint64_t)g.h * g.w * g.sc_cin * 2};
      uint32_t b2[4] = {(uint32_t)BK, (uint32_t)g.w_box, (uint32_t)g.h_box, (uint32_t)g.n_box};
      if (int rc = make_tmap_bf16(&em.a2, g.sc_src, 4, d2, s2, b2)) return rc;
    }
    for (int t = 0; t < 9; ++t) { p.tap_dy[t] = g.tap_dy[t]; p.tap_dx[t] = g.tap_dx[t]; p.tap_dn[t] = g.tap_dn[t]; }
    uint64_t dims[4] = {(uint64_t)g.cin, (uint64_t)g.src_w, (uint64_t)g.src_h, (uint64_t)g.src_nimg};
    uint64_t strides[3] = {(uint64_t)g.cin * 2, (uint64_t)g.src_w * g.cin * 2, (uint64_t)g.src_h * g.src_w * g.cin * 2};
    uint32_t box[4] = {(uint32_t)BK, (uint32_t)g.w_box, (uint32_t)g.h_box, (uint32_t)g.n_box};
    if (int rc = make_tmap_bf16(&tmA, a.A, 4, dims, strides, box)) return rc;
  } else {
    num_m_tiles = (a.M + BM - 1) / BM;
    uint64_t dims[2] = {(uint64_t)a.K, (uint64_t)a.M};
    uint64_t strides[1] = {(uint64_t)a.lda * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)BM};
    if (int rc = make_tmap_bf16(&tmA, a.A, 2, dims, strides, box)) return rc;
  }
  if (p.stats) {
    B200AD_REQUIRE(e.out_f32 && !e.geglu, "gemm: column statistics need an fp32 output");
    if (g.enabled) {
      p.stats_mq = g.h_box * g.w_box / 32;                 // = 4 / n_box: quarters of one image inside a tile
      B200AD_REQUIRE(p.stats_mq >= 1, "conv: column statistics need >= 32 pixels of one image per tile (box %dx%dx%d)",
                     g.n_box, g.h_box, g.w_box);
      p.stats_spi = g.stats_slots > 0 ? g.stats_slots : p.tiles_x * p.tiles_y;
      p.stats_slot0 = g.stats_slot0;
      p.slots_per_phase = p.tiles_x * p.tiles_y;
      B200AD_REQUIRE(p.stats_slot0 >= 0 && p.stats_slot0 + (g.nphase == 4 ? 4 : 1) * p.tiles_x * p.tiles_y <= p.stats_spi,
                     "conv: statistic slots %d + %d beyond %d", p.stats_slot0, p.tiles_x * p.tiles_y, p.stats_spi);
    } else {
      B200AD_REQUIRE(e.stats_hw > 0 && e.stats_hw % 32 == 0 && a.M % e.stats_hw == 0,
                     "gemm: column statistics need rows per image (%d) to be a multiple of 32 dividing M=%d", e.stats_hw, a.M);
      p.stats_mq = e.stats_hw % 128 == 0 ? 4 : (e.stats_hw % 64 == 0 ? 2 : 1);
      p.stats_spi = e.stats_hw / (32 * p.stats_mq);
    }
  }
  p.num_m_pairs = (num_m_tiles + 1) / 2;
  if (p.nphase == 0) p.nphase = 1;
  const int max_clusters = g_num_sms / 2;
  p.BN = pick_bn(a.N, (long long)p.num_m_pairs * p.nphase, max_clusters, e.geglu != 0);
  p.num_n_tiles = (a.N + p.BN - 1) / p.BN;
  p.num_kb = (a.K + BK - 1) / BK;
  p.stage_bytes = A_STAGE_BYTES + (p.BN / 2) * BK * 2;
  {
    uint64_t dims[2] = {(uint64_t)a.K, (uint64_t)a.N * p.nphase};      // phase weights stacked on the rows
    uint64_t strides[1] = {(uint64_t)a.K * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)(p.BN / 2)};
    if (int rc = make_tmap_bf16(&tmB, a.W, 2, dims, strides, box)) return rc;
  }
  const int n_out = e.geglu ? a.N / 2 : a.N;
  if (p.has_res) if (int rc = make_out_map(&em.res, e.residual, e.ld_res, n_out, 4, a)) return rc;
  if (p.has_o32) if (int rc = make_out_map(&em.o32, e.out_f32, e.ld_o32, n_out, 4, a, p.nphase == 4 ? 0 : -1)) return rc;
  if (p.nphase == 4)
    for (int ph = 1; ph < 4; ++ph)
      if (int rc = make_out_map(&em.o32p[ph - 1], e.out_f32, e.ld_o32, n_out, 4, a, ph)) return rc;
  if (p.has_oop) if (int rc = make_out_map(&em.oop, e.out_op, e.ld_oop, n_out, 2, a)) return rc;
  // ---- shared memory plan: mainloop stages first, the rest goes to epilogue chunk buffers (bytes in flight)
  const int per_buf = ((p.has_res || p.has_o32) ? F32_BUF_BYTES : 0) + (p.has_oop ? OP_BUF_BYTES : 0);
  const int stat_bytes = p.stats ? 2 * MAX_NBUF * 4 * 32 * 8 : 0;       // quarter sums handed to the I/O warp
  const int fixed = 2 * 256 * 4 + 512 + stat_bytes;          // bias staging + barriers (+ statistics hand-over)
  const int budget = SMEM_LIMIT - 1024 - fixed;
  int want_stages = p.num_kb < 6 ? (p.num_kb < 3 ? 3 : p.num_kb) : 6;
  {
    // HBM-bound shapes (the small-K GEMMs over the fp32 residual stream) are paced by the epilogue's bytes in flight,
    // not by the mainloop: give the shared memory to chunk buffers (residual prefetch depth nbuf-1) before stages.
    // With 6 stages the K=320 res32 GEMMs were left with nbuf=2, i.e. ONE residual chunk prefetched per group
    // (epilogue waiting on the residual 38 % of its time, tools/gemm_waits.py).
    const double t_mma = 2.0 * a.M * (double)a.N * a.K / 1.4e15;
    const double bytes = (double)a.M * a.K * 2 + (double)a.M * n_out * ((p.has_res ? 4 : 0) + (p.has_o32 ? 4 : 0) + (p.has_oop ? 2 : 0));
    if (bytes / 6.5e12 > t_mma) {
      // stages cover the A operand's share of the bytes in flight (~180 KB per SM at 2 us of loaded latency):
      // 3 stages when A is 20 % of the traffic (K = 320 res32), 6 when it is half of it (K = 1280 res32: with 3
      // stages the MMA warp waited for operands 64 % of the time).
      const double a_frac = (double)a.M * a.K * 2 / bytes;
      int st = (int)(11.0 * a_frac + 0.999);
      if (st < 3) st = 3;
      if (st < want_stages) want_stages = st;
    }
  }
  p.nbuf = MAX_NBUF;
  while (p.nbuf > 2 && want_stages * p.stage_bytes + 2 * p.nbuf * per_buf > budget) --p.nbuf;
  p.num_stages = (budget - 2 * p.nbuf * per_buf) / p.stage_bytes;
  if (p.num_stages > MAX_STAGES) p.num_stages = MAX_STAGES;
  if (getenv("B200AD_GEMM_VERBOSE"))
    fprintf(stderr, "gemm M=%d N=%d K=%d BN=%d stages=%d nbuf=%d res=%d o32=%d oop=%d geglu=%d\n", a.M, a.N, a.K, p.BN,
            p.num_stages, p.nbuf, p.has_res, p.has_o32, p.has_oop, p.geglu);
  B200AD_REQUIRE(p.num_stages >= 2, "gemm: shared memory plan failed (BN=%d)", p.BN);
  p.off_f32 = p.num_stages * p.stage_bytes;
  p.off_op = p.off_f32 + ((p.has_res || p.has_o32) ? 2 * p.nbuf * F32_BUF_BYTES : 0);
  p.off_bias = p.off_op + (p.has_oop ? 2 * p.nbuf * OP_BUF_BYTES : 0);
  p.off_stat = p.off_bias + 2 * 256 * 4;
  p.off_bars = p.off_stat + stat_bytes;
  const size_t smem_bytes = (size_t)p.off_bars + 512 + 1024 /*align*/;
  B200AD_REQUIRE(smem_bytes <= (size_t)SMEM_LIMIT, "gemm: %zu bytes of shared memory", smem_bytes);
  const int epi_mask = (p.geglu ? 8 : 0) | (p.has_res ? 4 : 0) | (p.has_o32 ? 2 : 0) | (p.has_oop ? 1 : 0);
  // instantiated: operand only (1), fp32 only (2), residual + operand (5), residual + fp32 (6), all three (7), GEGLU (9)
  const bool epi_known = (epi_mask == 1 || epi_mask == 2 || epi_mask == 5 || epi_mask == 6 ||
                                          epi_mask == 7 || epi_mask == 9);
  if (!g_attr_set) {
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200AD_CHECK_CUDA((cudaFuncSetAttribute(gemm_tc_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)));
    B200AD_CHECK_CUDA((cudaFuncSetAttribute(gemm_tc_kernel<6, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)));
    g_attr_set = true;
  }
  const long long tiles = (long long)p.num_m_pairs * p.num_n_tiles * p.nphase;
  const int clusters = (int)(tiles < max_clusters ? tiles : max_clusters);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // see pdl_trigger / pdl_wait in common.cuh
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  if (p.stats) {
    B200AD_REQUIRE(epi_mask == 2 || epi_mask == 6, "gemm: column statistics are implemented for fp32-only outputs (mask %d)", epi_mask);
    if (epi_mask == 2) B200AD_CHECK_CUDA((cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2, true>, tmA, tmB, em, p)));
    else B200AD_CHECK_CUDA((cudaLaunchKernelEx(&cfg, gemm_tc_kernel<6, true>, tmA, tmB, em, p)));
    count_launch();
    return 0;
  }
  switch (epi_known ? epi_mask : -1) {
    case 1: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<1>, tmA, tmB, em, p)); break;
    case 2: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2>, tmA, tmB, em, p)); break;
    case 5: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<5>, tmA, tmB, em, p)); break;
    case 6: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<6>, tmA, tmB, em, p)); break;
    case 7: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<7>, tmA, tmB, em, p)); break;
    case 9: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<9>, tmA, tmB, em, p)); break;
    default: B200AD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<-1>, tmA, tmB, em, p)); break;
  }
  count_launch();
  return 0;
}

}  // namespace b200ad
