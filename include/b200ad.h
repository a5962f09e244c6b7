/* b200ad.h -- C ABI of the B200-native AnimateDiff denoise-step kernels (libb200ad.so).
 *
 * Drop-in boundary.  The reference (tanm-ast/video-diffusion-experiments) has no native code: the hot
 * path is what diffusers/torch execute underneath
 *     experiments/05_grid_search_ablation.py:158-167   pipe(prompt=..., num_frames=16, ...)
 *     experiments/03_trace_forward_pass.py:109-113     unet(sample, timestep, encoder_hidden_states=...)
 * Each entry point below replaces the torch/cuDNN/cuBLAS library kernels one diffusers module reaches
 * (SURVEY.md section 2.4, K1..K12); the diffusers-shaped Python surface that binds them lives in
 * video-diffusion-experiments_b200/ and is loaded with ctypes.
 *
 * Conventions
 *  - Plain pointers and sizes only; every pointer is DEVICE memory owned by the caller unless stated.
 *  - Activations are channels-last: an image batch [N,H,W,C] is the row-major matrix [N*H*W, C];
 *    frame-images are ordered (batch b, frame f): n = b*F + f.
 *  - `op_dtype`: 0 = bf16 operands on tcgen05 tensor cores (fast mode), 1 = fp32 operands on SIMT FFMA
 *    (the "fp32 kernel mode" of the north star).  "operand-typed" buffers hold that type.
 *  - The residual stream, norm statistics, softmax and all epilogue math are fp32 in both modes.
 *  - Kernels never allocate and never synchronise; they run on `stream` (a cudaStream_t) and are
 *    CUDA-graph capturable.  There is no CPU fallback: without a CUDA device every compute call fails.
 *  - Return 0 on success, <0 on error; b200ad_last_error() returns a thread-local message.
 */
#ifndef B200AD_H
#define B200AD_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* b200ad_last_error(void);
int b200ad_version(void);
/* number of kernels launched by this library in this process since load (bench.py "gpu_launches") */
long long b200ad_launch_count(void);

/* diagnostics: device buffer of [grid][8] int64 cycle counters the GEMM kernel fills when non-NULL
 * (producer total / wait-on-empty, UMMA issuer total / wait-on-full / wait-on-accumulator); NULL disables. */
int b200ad_debug_buffer(void* dev_i64);

/* ---- K2: Linear / 1x1 conv (replaces torch.nn.functional.linear / cuBLAS under diffusers
 * Attention.to_q/k/v/to_out, FeedForward, Transformer2DModel.proj_in/out, ResnetBlock2D.conv_shortcut).
 *   C[M,N] = A[M,K(lda)] * W[N,K]^T (+bias[N]) (+row_bias[(row/rb_div)*ld_rb + col]) (+residual[M,ld_res])
 *   writes out_f32[M,ld_o32] and/or out_op[M,ld_oop]; geglu=1: W rows interleaved (h0,g0,h1,g1..) and
 *   out_op[M,N/2] = (h+bh)*gelu_erf(g+bg)   (diffusers GEGLU). */
int b200ad_gemm(int op_dtype, const void* A, long long lda, const void* W, int M, int N, int K,
                const float* bias, const float* row_bias, int rb_div, int ld_rb,
                const float* residual, long long ld_res, float* out_f32, long long ld_o32,
                void* out_op, long long ld_oop, int geglu, void* stream);

/* ---- K1: 3x3 convolution as implicit GEMM over NHWC (replaces cuDNN under ResnetBlock2D.conv1/conv2,
 * Downsample2D.conv, Upsample2D.conv).  src: operand-typed [src_nimg, src_h, src_w, cin]; W: [cout, 9*cin]
 * with k = (ky*3+kx)*cin + c; output rows = nimg*h*w pixels.  mode 0: stride 1 pad 1 (src geometry = output
 * geometry).  mode 1: stride 2 pad 1 on a phase-split source produced by b200ad_cast(mode=1)
 * (src_nimg = 4*nimg, src_h = h, src_w = w). Epilogue arguments as in b200ad_gemm (no geglu). */
int b200ad_conv3x3(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout,
                   int mode, const float* bias, const float* row_bias, int rb_div, int ld_rb,
                   const float* residual, long long ld_res, float* out_f32, long long ld_o32,
                   void* out_op, long long ld_oop, void* stream);

/* ---- K6: GroupNorm (replaces torch.nn.GroupNorm [+SiLU] in ResnetBlock2D.norm1/2, Transformer2DModel.norm,
 * AnimateDiffTransformer3D.norm, conv_norm_out).  Input is the channel-concatenation of x0[rows,c0] and
 * x1[rows,c1] (x1 may be NULL; this is the UNet skip torch.cat), rows = nimg*hw.
 * stats: float2 partial sums [nimg, nblk, groups]; nblk = b200ad_gn_nblk(hw).  The `partial` workspace must hold
 * (nimg*nblk*groups + nimg*groups) float2: gn_apply reduces the partials to (mean, rstd) in its tail.  apply: frames_per_stat=1 for
 * per-frame statistics, =F for the motion module (statistics over all F frames of a batch item).
 * out_op[rows, c0+c1] = maybe_silu(gn(x)); raw_op (optional) = x cast to the operand type (conv_shortcut input). */
int b200ad_gn_nblk(int hw);
int b200ad_gn_stats(const float* x0, int c0, const float* x1, int c1, int nimg, int hw, int groups,
                    float* partial, void* stream);
int b200ad_gn_apply(int op_dtype, const float* x0, int c0, const float* x1, int c1, int nimg, int hw, int groups,
                    const float* partial, int frames_per_stat, float eps, const float* gamma, const float* beta,
                    int silu, void* out_op, void* raw_op, void* stream);

/* diagnostics: how many of the GroupNorm ticket counters are non-zero (must be 0 between calls); synchronises. */
int b200ad_gn_counters_nonzero(void);

/* Statistics + apply in one call (same arguments and workspace as the pair above). */
int b200ad_group_norm(int op_dtype, const float* x0, int c0, const float* x1, int c1, int nimg, int hw, int groups,
                      float* partial, int frames_per_stat, float eps, const float* gamma, const float* beta, int silu,
                      void* out_op, void* raw_op, void* stream);

/* ---- K6 without the statistics pass: the producing GEMM / conv emits per-(32-row slot, column) sums of its fp32
 * output from its epilogue (b200ad_gemm_stats / b200ad_conv3x3_stats: float2 col_stats[(image*spi + k)*N + col], spi =
 * b200ad_stats_slots(conv, h, w) slots per image), and b200ad_gn_apply_stats reduces them (tiny) and applies: the
 * GroupNorm reads its input ONCE.  meanrstd_ws: b200ad_gn_stats_ws_bytes(nimg, groups, frames_per_stat) bytes, 16-byte aligned.  Same module
 * replacements as K6 / K1 / K2; the stats entry points write the fp32 output only. */
int b200ad_stats_slots(int conv, int h, int w);
long long b200ad_gn_stats_ws_bytes(int nimg, int groups, int frames_per_stat);
int b200ad_gemm_stats(int op_dtype, const void* A, long long lda, const void* W, int M, int N, int K,
                      const float* bias, const float* row_bias, int rb_div, int ld_rb, const float* residual,
                      long long ld_res, float* out_f32, long long ld_o32, float* col_stats, int rows_per_image,
                      void* stream);
int b200ad_conv3x3_stats(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout,
                         int mode, const float* bias, const float* row_bias, int rb_div, int ld_rb,
                         const float* residual, long long ld_res, float* out_f32, long long ld_o32, float* col_stats,
                         void* stream);

/* ResnetBlock2D.conv2 + conv_shortcut in one launch (`conv_shortcut(input) + conv2(h)`): the 1x1
 * shortcut is a second K segment of the implicit GEMM -- after the 9*cin tap columns, K continues with the sc_cin
 * channels of sc_src [nimg, h, w, sc_cin] (operand-typed, unshifted).  W: [cout, 9*cin + sc_cin] = (conv2 weight as
 * in b200ad_conv3x3 | shortcut weight [cout, sc_cin]); bias = conv2.bias + conv_shortcut.bias.  cin, sc_cin multiples
 * of 64; bf16 (tcgen05) path, stride 1.  col_stats (optional, fp32-only output) as b200ad_conv3x3_stats. */
int b200ad_conv3x3_sc(int op_dtype, const void* src, const void* W, int nimg, int h, int w, int cin, int cout,
                      const void* sc_src, int sc_cin, const float* bias, float* out_f32, long long ld_o32, void* out_op,
                      long long ld_oop, float* col_stats, void* stream);

/* Upsample2D = F.interpolate(scale_factor=2, mode="nearest") + Conv2d(C, C, 3, padding=1) of the UNet's up blocks and the
 * VAE decoder WITHOUT the upsampled image: four 2x2 convolutions over the low-resolution
 * source, one per output phase (py, px).  src: operand-typed NHWC [nimg, h, w, cin]; outputs are rows of the
 * [nimg, 2h, 2w] image.  W4: [4 phases (py*2+px)][cout][2x2 taps (ty*2+tx)][cin] where tap (ty, tx) of phase (py, px) is
 * the sum of the 3x3 weights w[ky][kx] with floor((py + ky - 1) / 2) == ty - 1 + py (same in x), reading source pixel
 * (y + ty - 1 + py, x + tx - 1 + px).  One launch over (phase, pixel tile, column tile).  bf16 (tcgen05) path only;
 * fp32 output (out_op / ld_oop are reserved: pass NULL / 0).  col_stats (optional): as b200ad_conv3x3_stats,
 * 4 * b200ad_stats_slots(1, h, w) slots per image. */
int b200ad_upconv3x3(int op_dtype, const void* src, const void* W4, int nimg, int h, int w, int cin, int cout,
                     const float* bias, float* out_f32, long long ld_o32, void* out_op, long long ld_oop,
                     float* col_stats, void* stream);
int b200ad_gn_apply_stats(int op_dtype, const float* x0, int c0, const float* st0, int spi0, const float* x1, int c1,
                          const float* st1, int spi1, int nimg, int hw, int groups, int frames_per_stat, float eps,
                          const float* gamma, const float* beta, int silu, float* meanrstd_ws, void* out_op,
                          void* raw_op, void* stream);

/* ---- K7: LayerNorm (+ sinusoidal positional-embedding add of the motion module; replaces nn.LayerNorm and
 * SinusoidalPositionalEmbedding).  x fp32 [rows,C]; pe fp32 [max_seq,C] or NULL; frame index of a row is
 * (row / hw) % frames. */
int b200ad_layernorm(int op_dtype, const float* x, int rows, int C, const float* gamma, const float* beta, float eps,
                     const float* pe, int hw, int frames, void* out_op, void* stream);

/* ---- K9: fp32 -> operand-type copies with re-layout.  mode 0: plain cast [rows*C].  mode 1: space-to-depth
 * phase split of NHWC [nimg,h,w,C] -> [4*nimg, h/2, w/2, C] (input of a stride-2 conv).  mode 2: nearest x2
 * upsample NHWC [nimg,h,w,C] -> [nimg,2h,2w,C] (Upsample2D). */
int b200ad_cast(int op_dtype, const float* x, int nimg, int h, int w, int C, int mode, void* out_op, void* stream);

/* ---- K3/K4: spatial self-attention and text cross-attention (replaces F.scaled_dot_product_attention in
 * AttnProcessor2_0).  q/k/v: operand-typed, head `hd` of batch item n at q[(n*Lq + i)*ldq + hd*dpad + j];
 * k/v batch index = n / kv_div (text K/V shared by the frames of a batch item).  dpad = head stride in the rows: on
 * the tcgen05 path a multiple of 8 with 16*ceil(d/16) <= dpad <= 192, columns d..dpad-1 of q and k zero (the kernels
 * load 64-column slabs, which may run into the next head, but use 16*ceil(d/16) columns of q/k and d (+1) of v only);
 * out_op[(n*Lq+i)*ldo + hd*d + j], softmax scale = d^-1/2. */
int b200ad_attention(int op_dtype, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                     long long ldv, int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div,
                     void* out_op, long long ldo, void* stream);

/* Same with `flags`.  B200AD_ATTN_V_ONES (1): the caller guarantees V[:, hd*dpad + d] == 1 for every key row and head
 * (dpad > d), e.g. written by the bias of the GEMM that produced V; the long self-attention kernel then takes the
 * softmax denominator from the PV product instead of summing the probabilities.  Other kernels ignore the flag (the
 * ones column only feeds an output pad column that is never stored). */
#define B200AD_ATTN_V_ONES 1
int b200ad_attention2(int op_dtype, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                      long long ldv, int nbatch, int heads, int Lq, int Lk, int d, int dpad, int kv_div, int flags,
                      void* out_op, long long ldo, void* stream);

/* ---- K5: motion-module temporal self-attention over the frame axis, read in place from the (b,f,hw)-ordered
 * rows (no permute copy).  qkv: operand-typed [B*F*hw, 3C] = (q|k|v); out_op [B*F*hw, C]. */
int b200ad_temporal_attention(int op_dtype, const void* qkv, int B, int F, int hw, int C, int heads,
                              void* out_op, void* stream);

/* ---- K11: timestep embedding (Timesteps flip_sin_to_cos + small-M Linear with optional SiLU in/out).
 * small_linear: y[M,N] = act_out(W[N,K] * act_in(x[M,K]) + b), M <= 8, fp32. */
int b200ad_timestep_sinusoid(const float* t, int B, int dim, float* out, void* stream);
int b200ad_small_linear(const float* x, int M, int K, const float* W, const float* b, int N, int silu_in,
                        int silu_out, float* y, void* stream);

/* ---- conv_in (4 -> C0, reads the [B,Cin,F,h,w] latent batch directly) and conv_out (C0 -> 4, writes
 * [B,Cout,F,h,w]); W in torch layout [cout,cin,3,3] fp32. */
int b200ad_conv_in(const float* sample, int B, int cin, int F, int h, int w, const float* W, const float* bias,
                   int cout, float* out_nhwc, void* stream);
int b200ad_conv_out(int op_dtype, const void* act_nhwc, int B, int F, int h, int w, int cin, const float* W,
                    const float* bias, int cout, float* out_bcfhw, void* stream);

/* ---- K10 (+K12 tail): classifier-free-guidance combine fused with the scheduler update
 * (replaces noise_pred.chunk(2) / u+g*(c-u) / DDIMScheduler.step or EulerDiscreteScheduler.step and the next
 * iteration's torch.cat([latents]*2) + scale_model_input).
 *   guided = eps_u + g*(eps_c - eps_u);  latents' = cx*latents + ce*guided;
 *   model_in[0..1] = latents' * in_scale (both CFG halves; may be NULL). */
int b200ad_cfg_step(const float* eps_u, const float* eps_c, const float* latents, float g, float cx, float ce,
                    float in_scale, long long n, float* latents_out, float* model_in, int model_in_copies,
                    void* stream);
/* Same update with (g, cx, ce, in_scale) read from 16-byte-aligned DEVICE memory: a CUDA graph captured once
 * is replayed for every denoise step after a 16-byte parameter upload (no per-step re-capture, no host sync). */
int b200ad_cfg_step_dev(const float* eps_u, const float* eps_c, const float* latents, const float* dev_coef4,
                        long long n, float* latents_out, float* model_in, int model_in_copies, void* stream);

/* ---- conv_in with a folded per-pixel affine front end (VAE decoder: latents / scaling_factor -> post_quant_conv
 * (1x1, cin -> cin, + bias) -> decoder.conv_in 3x3).  W: [cout, (cin+1)*9] fp32 in torch layout with an EXTRA input
 * channel cin whose value is 1 inside the image and 0 in the zero padding, so that the 1x1 conv's bias is carried
 * exactly through the padded 3x3 conv:  W[:, c, tap] = sum_j Win[:, j, tap] * Wpq[j, c] / s,  W[:, cin, tap] =
 * sum_j Win[:, j, tap] * bpq[j]  (replaces AutoencoderKL.post_quant_conv + Decoder.conv_in of diffusers). */
int b200ad_conv_in_affine(const float* sample, int B, int cin, int F, int h, int w, const float* W, const float* bias,
                          int cout, float* out_nhwc, void* stream);

/* ---- module-mode glue (the hook-visible eager mode: one diffusers sub-module = one call).
 * mode 0: out[i] = silu(a[i]) (nn.SiLU);  mode 1: out[i] = a[i] + b[(i / div) % mod] (residual add: div = 1, mod = n;
 * timestep-embedding add over NCHW: div = h*w, mod = N*C; positional-embedding add: div = 1, mod = F*C);
 * mode 2: GEGLU gate, a = [rows, 2C] (h | g), n = rows*C, mod = C: out[rows, C] = h * gelu_erf(g);
 * mode 3: out[i] = a[i] * sigmoid(1.702 a[i])  (CLIP's quick_gelu). */
int b200ad_eltwise(int mode, const float* a, const float* b, long long n, long long div, long long mod, float* out,
                   void* stream);

/* ---- f2: CLIP text encoder pieces (replaces transformers' CLIPTextModel under AnimateDiffPipeline.encode_prompt, loaded
 * at 05_grid_search_ablation.py:130-134).  embed_tokens: out[t,:] = token_table[ids[t],:] + position_table[t % seq,:]
 * (fp32).  attention_small: softmax(q k^T / sqrt(d) [+ causal mask]) v for short sequences (77 tokens), head h of batch
 * item n at q[(n*L + i)*ldq + h*d + j]; operand-typed in and out.  The projections / MLP run on b200ad_gemm, the
 * LayerNorms on b200ad_layernorm, quick_gelu on b200ad_eltwise mode 3. */
int b200ad_embed_tokens(const long long* ids, long long ntok, int seq, int vocab, const float* tok_table,
                        const float* pos_table, int dim, float* out, void* stream);
int b200ad_attention_small(int op_dtype, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                           long long ldv, int nbatch, int heads, int L, int d, int causal, void* out_op, long long ldo,
                           void* stream);

/* ---- row softmax of an fp32 score matrix into the operand type: out[r, :] = softmax(scale * s[r, :]), cols <= 8192,
 * cols % 4 == 0 (VAE decoder mid-block attention, 1 head of d = 512: S = Q K^T by b200ad_gemm, this, P V by b200ad_gemm;
 * replaces the softmax inside F.scaled_dot_product_attention / Attention.get_attention_scores of diffusers). */
int b200ad_softmax_rows(int op_dtype, const float* s, long long rows, int cols, long long ld, float scale, void* out_op,
                        long long ldo, void* stream);

/* ---- VideoProcessor.postprocess_video: video [C, F, h, w] fp32 in [-1, 1] -> frames [F, h, w, C] uint8 =
 * round(clamp(x/2 + 0.5, 0, 1) * 255)  (the frames consumed at 05_grid_search_ablation.py:169,178-179). */
int b200ad_video_to_u8(const float* video_cfhw, int C, int F, int h, int w, unsigned char* out_fhwc, void* stream);

/* ---- f4: temporal-consistency metrics of a generated video (replaces the CPU loops of the reference's
 * experiments/06_measure_grid_search.py:209-284,290-352).  frames: uint8 [F, H, W, C] on the device (what
 * b200ad_video_to_u8 writes).  workspace: b200ad_metric_blocks() * 3 * (F-1) doubles.
 *  video_pair_sums: out[2*i] = sum over the frame of (I_i - I_{i+1})^2, out[2*i+1] = sum |I_i - 2 I_{i+1} + I_{i+2}|
 *    (0 for the last pair), i < F-1, in uint8 units, exact integers (-> compute_mse / compute_psnr / compute_flicker_index).
 *  video_warp_error: flow [F-1, H, W, 2] fp32 (dx, dy); out[3*i] = sum (warp(I_i, flow_i) - I_{i+1})^2 in [0,1] units
 *    (warp_frame: bilinear, border padding, align_corners=True), out[3*i+1] = sum |flow_i|, out[3*i+2] = sum |flow_i|^2.
 *  video_gray_u8: the uint8 luminance handed to OpenCV's Farneback flow (06:163-164). */
int b200ad_metric_blocks(void);
int b200ad_video_pair_sums(const unsigned char* frames_fhwc, int F, long long n_per_frame, double* workspace, double* out,
                           void* stream);
int b200ad_video_warp_error(const unsigned char* frames_fhwc, int F, int H, int W, int C, const float* flow,
                            double* workspace, double* out, void* stream);
int b200ad_video_gray_u8(const unsigned char* frames_fhwc, long long npix, int C, unsigned char* gray, void* stream);

/* b200ad_video_pair_sums without a workspace argument and without any allocation.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same inputs, `out` and results as b200ad_video_pair_sums; the per-block
 * partial sums live in a file-scope `static __device__ __align__(256)` array of the library (149 184 bytes on every
 * GPU that loads it, until exit), whose device address is taken once per device with cudaGetSymbolAddress.
 * Limit: F <= b200ad_static_max_frames() = 64; larger F returns -1 with b200ad_last_error() naming the limit.
 * Calls no cudaMalloc / cudaFree / allocating library routine.  Every call on a device shares the array (it keeps the
 * last call's partials; each call overwrites all it reads, so nothing needs clearing): calls that overlap on the GPU
 * (different streams, no ordering between them) race -- serialise them. */
int b200ad_static_max_frames(void);
int b200ad_video_pair_sums_static(const unsigned char* frames_fhwc, int F, long long n_per_frame, double* out,
                                  void* stream);

/* LPIPS (AlexNet) between consecutive frames (replaces lpips.LPIPS(net='alex') at 06:101-131): lpips_input applies the
 * [0,255] -> [-1,1] map and LPIPS' ScalingLayer; conv2d_direct is an fp32 direct convolution over NHWC (weights
 * [cout, kh, kw, cin], any kernel / stride / padding, bias, optional ReLU: AlexNet's 11x11 s4, 5x5 and 3x3 layers);
 * maxpool2d = MaxPool2d(ks, stride); lpips_layer: out[i] = sum over pixels of sum_c lin[c] (f_i/|f_i| - f_{i+1}/|f_{i+1}|)^2
 * for the features [N, hw, C] of N consecutive frames (the caller divides by hw and adds the five layers). */
int b200ad_lpips_input(const unsigned char* frames_nhwc, long long n_elems, float* out, void* stream);
int b200ad_conv2d_direct(const float* x_nhwc, int N, int H, int W, int Cin, const float* w_ohwi, const float* bias,
                         int Cout, int KH, int KW, int stride, int pad, int relu, float* out_nhwc, void* stream);
int b200ad_maxpool2d(const float* x_nhwc, int N, int H, int W, int C, int ks, int stride, float* out_nhwc, void* stream);
int b200ad_lpips_layer(const float* feat_nhwc, int N, int hw, int C, const float* lin, double* workspace, double* out,
                       void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200AD_H */
