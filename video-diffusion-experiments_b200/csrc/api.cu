// C-ABI glue: error reporting, launch accounting, driver entry point lookup, and the GEMM / conv entry
// points that build GemmArgs and dispatch on the operand dtype (bf16 -> tcgen05, fp32 -> SIMT).
#include <atomic>
#include <stdarg.h>

#include "../../include/b200ad.h"
#include "gemm.cuh"
#include "tmap.cuh"

namespace b200ad {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static void fill_epilogue(Epilogue& e, const float* bias, const float* row_bias, int rb_div, int ld_rb,
                          const float* residual, long long ld_res, float* out_f32, long long ld_o32, void* out_op,
                          long long ld_oop, int geglu) {
  e.bias = bias;
  e.row_bias = row_bias;
  e.rb_div = rb_div > 0 ? rb_div : 1;
  e.ld_rb = ld_rb;
  e.residual = residual;
  e.ld_res = ld_res;
  e.out_f32 = out_f32;
  e.ld_o32 = ld_o32;
  e.out_op = out_op;
  e.ld_oop = ld_oop;
  e.geglu = geglu;
  e.col_stats = nullptr;
  e.stats_hw = 0;
}

}  // namespace b200ad

namespace b200ad { void set_gemm_debug_buffer(long long* p); void set_fa_debug_buffer(long long* p); }
using namespace b200ad;

extern "C" {

int b200ad_debug_buffer(void* dev_i64) {
  set_gemm_debug_buffer((long long*)dev_i64);
  set_fa_debug_buffer((long long*)dev_i64);
  return 0;
}

const char* b200ad_last_error(void) { return g_err; }
int b200ad_version(void) { return 100; }
long long b200ad_launch_count(void) { return g_launches.load(); }

static int gemm_impl(int op_dtype, const void* A, long long lda, const void* W, int M, int N, int K, const float* bias,
                     const float* row_bias, int rb_div, int ld_rb, const float* residual, long long ld_res,
                     float* out_f32, long long ld_o32, void* out_op, long long ld_oop, int geglu, float* col_stats,
                     int stats_hw, void* stream) {
  B200AD_REQUIRE(A && W, "gemm: null operand");
  B200AD_REQUIRE(out_f32 || out_op, "gemm: no output buffer");
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.A = A; a.lda = lda; a.W = W; a.M = M; a.N = N; a.K = K;
  fill_epilogue(a.epi, bias, row_bias, rb_div, ld_rb, residual, ld_res, out_f32, ld_o32, out_op, ld_oop, geglu);
  a.epi.col_stats = col_stats;
  a.epi.stats_hw = stats_hw;
  B200AD_REQUIRE(!col_stats || op_dtype == 0, "gemm: column statistics exist on the tcgen05 path only");
  if (op_dtype == 0) return gemm_bf16_tc(a, (cudaStream_t)stream);
  if (op_dtype == 1) return gemm_f32_simt(a, (cudaStream_t)stream);
  set_error("gemm: unknown op_dtype %d", op_dtype);
  return -1;
}

int b200ad_gemm(int op_dtype, const void* A, long long lda, const void* W, int M, int N, int K, const float* bias,
                const float* row_bias, int rb_div, int ld_rb, const float* residual, long long ld_res,
                float* out_f32, long long ld_o32, void* out_op, long long ld_oop, int geglu, void* stream) {
  return gemm_impl(op_dtype, A, lda, W, M, N, K, bias, row_bias, rb_div, ld_rb, residual, ld_res, out_f32, ld_o32,
                   out_op, ld_oop, geglu, nullptr, 0, stream);
}

int b200ad_gemm_stats(int op_dtype, const void* A, long long lda, const void* W, int M, int N, int K,
                      const float* bias, const float* row_bias, int rb_div, int ld_rb, const float* residual,
                      long long ld_res, float* out_f32, long long ld_o32, float* col_stats, int rows_per_image,
                      void* stream) {
  B200AD_REQUIRE(col_stats, "gemm_stats: null statistics buffer");
  return gemm_impl(op_dtype, A, lda, W, M, N, K, bias, row_bias, rb_div, ld_rb, residual, ld_res, out_f32, ld_o32,
                   nullptr, 0, 0, col_stats, rows_per_image, stream);
}

// M tile of a conv = 128 output pixels = n_box * h_box * w_box (powers of two), chosen to waste the fewest rows
static void conv_box(int h, int w, int* nb, int* hb, int* wb) {
  auto best_box = [](int extent, int limit) {
    int best = 1;
    double best_eff = -1.0;
    for (int b = limit; b >= 1; b >>= 1) {
      const double eff = (double)extent / (double)(((extent + b - 1) / b) * b);
      if (eff > best_eff + 1e-9) { best_eff = eff; best = b; }
    }
    return best;
  };
  *wb = best_box(w, 128);
  *hb = best_box(h, 128 / *wb);
  *nb = 128 / (*wb * *hb);
}

int b200ad_stats_slots(int conv, int h, int w) {
  if (!conv) {
    const int hw = h * w;
    if (hw % 32 != 0) return -1;
    const int mq = hw % 128 == 0 ? 4 : (hw % 64 == 0 ? 2 : 1);      // 32-row quarters of a tile merged per slot
    return hw / (32 * mq);
  }
  int nb, hb, wb;
  conv_box(h, w, &nb, &hb, &wb);
  if (hb * wb < 32) return -1;
  return ((w + wb - 1) / wb) * ((h + hb - 1) / hb);                // one slot per image and tile
}

static int conv_impl(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout, int mode,
                     const float* bias, const float* row_bias, int rb_div, int ld_rb, const float* residual,
                     long long ld_res, float* out_f32, long long ld_o32, void* out_op, long long ld_oop,
                     float* col_stats, void* stream, const void* sc_src = nullptr, int sc_cin = 0) {
  B200AD_REQUIRE(src && W, "conv3x3: null operand");
  B200AD_REQUIRE(out_f32 || out_op, "conv3x3: no output buffer");
  B200AD_REQUIRE(mode == 0 || mode == 1, "conv3x3: mode %d", mode);
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.A = src; a.W = W;
  a.M = nimg * h * w; a.N = cout; a.K = 9 * cin + sc_cin;
  ConvGeom& g = a.conv;
  g.enabled = 1;
  g.nimg = nimg; g.h = h; g.w = w; g.cin = cin; g.ntaps = 9;
  g.sc_src = sc_src; g.sc_cin = sc_cin;
  B200AD_REQUIRE(sc_cin == 0 || (op_dtype == 0 && mode == 0 && sc_src),
                 "conv3x3: the fused 1x1 shortcut exists on the tcgen05 path, stride 1 only");
  g.src_h = h; g.src_w = w;
  g.src_nimg = mode == 0 ? nimg : 4 * nimg;
  for (int ky = 0; ky < 3; ++ky)
    for (int kx = 0; kx < 3; ++kx) {
      const int t = ky * 3 + kx;
      if (mode == 0) {
        g.tap_dy[t] = ky - 1; g.tap_dx[t] = kx - 1; g.tap_dn[t] = 0;
      } else {
        // input pixel (2y+ky-1, 2x+kx-1): ky=0 -> odd phase row y-1; ky=1 -> even phase row y; ky=2 -> odd phase row y
        const int py = (ky == 1) ? 0 : 1, px = (kx == 1) ? 0 : 1;
        g.tap_dy[t] = (ky == 0) ? -1 : 0;
        g.tap_dx[t] = (kx == 0) ? -1 : 0;
        g.tap_dn[t] = (py * 2 + px) * nimg;
      }
    }
  conv_box(h, w, &g.n_box, &g.h_box, &g.w_box);
  fill_epilogue(a.epi, bias, row_bias, rb_div, ld_rb, residual, ld_res, out_f32, ld_o32, out_op, ld_oop, 0);
  a.epi.col_stats = col_stats;
  B200AD_REQUIRE(!col_stats || op_dtype == 0, "conv3x3: column statistics exist on the tcgen05 path only");
  if (op_dtype == 0) return gemm_bf16_tc(a, (cudaStream_t)stream);
  if (op_dtype == 1) return gemm_f32_simt(a, (cudaStream_t)stream);
  set_error("conv3x3: unknown op_dtype %d", op_dtype);
  return -1;
}

int b200ad_conv3x3(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout, int mode,
                   const float* bias, const float* row_bias, int rb_div, int ld_rb, const float* residual,
                   long long ld_res, float* out_f32, long long ld_o32, void* out_op, long long ld_oop,
                   void* stream) {
  return conv_impl(op_dtype, src, W, nimg, h, w, cin, cout, mode, bias, row_bias, rb_div, ld_rb, residual, ld_res,
                   out_f32, ld_o32, out_op, ld_oop, nullptr, stream);
}

int b200ad_conv3x3_stats(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout,
                         int mode, const float* bias, const float* row_bias, int rb_div, int ld_rb,
                         const float* residual, long long ld_res, float* out_f32, long long ld_o32, float* col_stats,
                         void* stream) {
  B200AD_REQUIRE(col_stats, "conv3x3_stats: null statistics buffer");
  return conv_impl(op_dtype, src, W, nimg, h, w, cin, cout, mode, bias, row_bias, rb_div, ld_rb, residual, ld_res,
                   out_f32, ld_o32, nullptr, 0, col_stats, stream);
}

// conv3x3 (stride 1, pad 1) + a 1x1 convolution of a second image accumulated into the same output tile:
// out = conv3x3(src; W[:, :9cin]) + sc_src W[:, 9cin:]^T + bias (+ row_bias).  ResnetBlock2D.conv2 + conv_shortcut: the
// shortcut's fp32 result is never written and re-read as a residual.
int b200ad_conv3x3_sc(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout,
                      const void* sc_src, int sc_cin, const float* bias, float* out_f32, long long ld_o32, void* out_op,
                      long long ld_oop, float* col_stats, void* stream) {
  B200AD_REQUIRE(sc_src && sc_cin > 0, "conv3x3_sc: null shortcut source");
  B200AD_REQUIRE(!col_stats || (out_f32 && !out_op), "conv3x3_sc: column statistics need an fp32-only output");
  return conv_impl(op_dtype, src, W, nimg, h, w, cin, cout, 0, bias, nullptr, 1, 0, nullptr, 0, out_f32, ld_o32, out_op,
                   ld_oop, col_stats, stream, sc_src, sc_cin);
}

// nearest-x2 upsample followed by conv3x3 (pad 1) as four 2x2 convolutions over the LOW-resolution image, one per output
// phase (py, px): output pixel (2y+py, 2x+px) reads the upsampled pixels (2y+py+dy, 2x+px+dx), dy, dx in {-1,0,1}, i.e.
// the source pixels y + floor((py+dy)/2): rows {y-1, y} for py = 0 and {y, y+1} for py = 1 (columns alike), with the
// 3x3 weights that land on the same source pixel summed beforehand by the caller.  4/9 of the MACs of the direct form
// and the 4x larger upsampled image is never written.  Zero padding agrees: an upsampled pixel outside the image is a
// source pixel outside the image, which the TMA load zero-fills.
int b200ad_upconv3x3(int op_dtype, const void* src, const void* W4, int nimg, int h, int w, int cin, int cout,
                     const float* bias, float* out_f32, long long ld_o32, void* out_op, long long ld_oop,
                     float* col_stats, void* stream) {
  B200AD_REQUIRE(src && W4, "upconv3x3: null operand");
  B200AD_REQUIRE(out_f32 && !out_op, "upconv3x3: writes an fp32 output (out_op must be null)");
  B200AD_REQUIRE(op_dtype == 0, "upconv3x3: tcgen05 (bf16 operand) path only");
  B200AD_REQUIRE(!col_stats || (out_f32 && !out_op), "upconv3x3: column statistics need an fp32-only output");
  int nb, hb, wb;
  conv_box(h, w, &nb, &hb, &wb);
  const int tiles_per_img = ((w + wb - 1) / wb) * ((h + hb - 1) / hb);
  // ONE launch over (phase, pixel tile, column tile): with one launch per phase the 16 -> 32 upsampler ran 3.46 waves
  // rounded up to 4, four times, and paid four prologues and tails (0.69 of the sustained peak against 0.91 for the
  // direct form, tools/kbench.py restructured)
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.A = src;
  a.W = W4;
  a.M = nimg * h * w; a.N = cout; a.K = 4 * cin;
  ConvGeom& g = a.conv;
  g.enabled = 1;
  g.nimg = nimg; g.h = h; g.w = w; g.cin = cin; g.ntaps = 4;
  g.src_h = h; g.src_w = w; g.src_nimg = nimg;
  for (int ty = 0; ty < 2; ++ty)
    for (int tx = 0; tx < 2; ++tx) {
      const int t = ty * 2 + tx;
      g.tap_dy[t] = ty - 1; g.tap_dx[t] = tx - 1; g.tap_dn[t] = 0;      // phase (0, 0); phase (py, px) adds (py, px)
    }
  g.n_box = nb; g.h_box = hb; g.w_box = wb;
  g.out_scale = 2; g.nphase = 4;
  g.stats_slots = 4 * tiles_per_img; g.stats_slot0 = 0;
  fill_epilogue(a.epi, bias, nullptr, 1, 0, nullptr, 0, out_f32, ld_o32, out_op, ld_oop, 0);
  a.epi.col_stats = col_stats;
  return gemm_bf16_tc(a, (cudaStream_t)stream);
}

}  // extern "C"
