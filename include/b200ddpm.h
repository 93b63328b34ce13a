/*
 * b200ddpm.h — C ABI of libb200ddpm.so: the sm_100a kernels behind the DDPM hot path of
 * matwilso/generative_models (gms/diffusion).  The reference is pure PyTorch and has no FFI of its own; each
 * entry point below names the reference expression (file:line under the reference root) it replaces.
 *
 * Conventions:
 *   - every call only ENQUEUES work on `stream` (a cudaStream_t passed as void*); no allocation, no sync,
 *     so every call is CUDA-graph capturable;
 *   - all pointers are device pointers owned by the caller (PyTorch allocator);
 *   - activations are NHWC fp32 ("pixel rows" of C contiguous channels); conv weights are
 *     [Cout][kh][kw][Cin] (== a torch OIHW parameter with channels_last strides);
 *   - return 0 = enqueued, <0 = argument rejected (nothing enqueued), >0 = cudaError_t observed;
 *     b200_last_error() returns a static message for the last non-zero return on this thread.
 */
#ifndef B200DDPM_H_
#define B200DDPM_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 6

int b200_abi_version(void);
/* validates `device` (compute capability 10.x, equal to the caller's current device); never changes the current
 * device; one device per process (a second, different device is refused) */
int b200_init(int device);
const char* b200_last_error(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches). */
unsigned long long b200_launch_count(void);

/* ------------------------------------------------------------------ implicit-GEMM convolution (tcgen05 + TMA im2col)
 * Replaces nn.Conv2d forward / backward-data in gms/diffusion/simple_unet.py:82,85 (Downsample.conv),
 * :117,122 (Upsample.conv), :164,172,177 (ResBlock in_layers.2 / out_layers.3 / skip_connection).
 *   mode 0 (fprop):  out[m, co] = sum_{kh,kw,ci} x[n, p*stride-pad+kh, q*stride-pad+kw, ci] * w[co, kh, kw, ci]
 *                                 (+ bias[co]) (+ emb[n, co]) (+ residual[m, co]),      m = (n, p, q)
 *   mode 1 (dgrad, stride 1): x is dY [N,H,W,Cout]; out[m, ci] for ci in [n_off, n_off+128) =
 *                                 sum_{kh,kw,co} dY[n, h+pad-kh, w+pad-kw, co] * w[co, kh, kw, ci]  (+ residual[m, ci])
 *   mode 2 (dgrad, stride 2, 3x3 pad 1): x is dY [N,H,W,128], w is the transposed tap-flipped shadow [128 ci][9][128 co]
 *                                 written by b200_conv_weights_transpose, out = dX [N,2H,2W,128] (Ho = 2H, Wo = 2W)
 *                                 (+ residual, same layout); one launch, CTA = (128-row tile of dY, output-pixel parity class)
 * x may be a virtual channel concat of two tensors (x: C1 channels, x2: C2 channels), reference torch.cat at
 * simple_unet.py:150.  C1, C2 multiples of 32; Cout (fprop) must be 128; N-tile is 128 columns per launch.
 */
typedef struct {
  const void* x;   /* fp32, or bf16 when op_dtype == 1 */
  const void* x2;  /* NULL when C2 == 0 */
  const void* w;   /* same element type as x */
  void* out;
  const float* bias;     /* [128] or NULL */
  const float* emb;      /* [N][emb_ld] or NULL */
  const float* residual; /* [M][res_ld] or NULL */
  int N, H, W, C1, C2;   /* geometry of the tensor the im2col walks (x) */
  int Ho, Wo;            /* output spatial size */
  int R, S, stride, pad; /* filter of the ORIGINAL convolution */
  int mode;              /* 0 fprop, 1 dgrad (stride 1), 2 dgrad of a stride-2 3x3 conv over the transposed shadow */
  int w_cout, w_cin;     /* weight tensor is [w_cout][R*S*w_cin] */
  int n_off;             /* dgrad only: first input channel produced by this launch */
  int out_ld, emb_ld, res_ld;
  int x_ld, x2_ld;       /* pixel stride (floats) of x / x2; 0 = dense (C1 / C2) */
  int round_out;         /* store the result rounded to TF32 (it feeds another tensor-core operand) */
  int out_dtype;         /* 0 = out is fp32; 1 = out holds bf16 (out_ld in elements; strip-kernel layers only) */
  /* fused skip segment (stride-1 layers): out += Conv1x1(cat(skip_x, skip_x2), skip_w) + bias2, accumulated on the
   * tensor cores as extra K-chunks.  skip_w is [128][skip_c1+skip_c2]; with the 128x128 identity it is the residual
   * add of ResBlock.forward (simple_unet.py:186), with skip_connection.weight it is the 1x1 skip conv (:177). */
  const void* skip_x;
  const void* skip_x2;
  const void* skip_w;
  const float* bias2;
  int skip_c1, skip_c2, skip_x_ld, skip_x2_ld;
  /* tensor-core operand type: 0 = fp32 storage, kind::tf32 MMAs; 1 = x / x2 / w (/ skip_*) hold bf16, kind::f16 MMAs
   * with fp32 accumulation (stride-1 3x3/pad-1 and 1x1/pad-0 layers only; C1, C2 multiples of 64; x_ld in elements).
   * out / bias / emb / residual are fp32 either way. */
  int op_dtype;
  /* optional second output: dense bf16 copy [M][128] of what is stored to out (the operand a following bf16
   * convolution reads); fprop only */
  void* out_bf16;
} b200_conv_args;
int b200_conv_igemm(const b200_conv_args* a, void* ws, size_t ws_bytes, void* stream);
/* which kernel b200_conv_igemm launches for these arguments: 1 = conv_strip_kernel (strip implicit GEMM: every stride-1
 * 3x3 / 1x1 layer with W <= 64), 0 = conv_igemm_tf32_kernel (TMA-im2col: the stride-2 layers).  Used by bench.py's per-kernel roofline. */
int b200_conv_kernel_of(const b200_conv_args* a);

/* weight gradient (nn.Conv2d backward-weight of the same layers):
 *   dw[co, kh, kw, ci] += sum_{n,p,q} dy[(n,p,q), co] * x[n, p*stride-pad+kh, q*stride-pad+kw, ci]
 * Accumulates with fp32 atomics into dw (caller zeroes it once per step). */
typedef struct {
  const void* x;   /* fp32, or bf16 when op_dtype == 1 */
  const void* x2;
  const void* dy;  /* [N*Ho*Wo][128], same element type as x */
  float* dw;       /* [128][R*S*(C1+C2)] */
  int N, H, W, C1, C2;
  int Ho, Wo;
  int R, S, stride, pad;
  int k_slices; /* split of the pixel reduction; 0 = library picks */
  int x_ld, x2_ld; /* pixel stride (floats) of x / x2; 0 = dense */
  int dy_ld;       /* pixel stride of dy; 0 = dense (128) */
  int op_dtype;    /* 0 = fp32 storage / TF32 MMAs, 1 = bf16 operands (fp32 accumulation and dw) */
} b200_wgrad_args;
int b200_conv_wgrad(const b200_wgrad_args* a, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------ GroupNorm(32) + SiLU
 * Replaces nn.GroupNorm(32, C) -> nn.SiLU (simple_unet.py:38-41 out head, :161-163 in_layers, :168-170 out_layers).
 * x may be a virtual concat (x: C1 channels with pixel stride x_ld, x2: C2 channels with stride x2_ld; ld 0 = dense).
 * b200_gn_fwd writes stats[N][32][2] = {mean, rstd} (kept for the backward) and
 * y[N*HW][C1+C2] = silu?((x-mean)*rstd*gamma+beta), optionally rounded to TF32 for the following conv.  One kernel:
 * a CTA holds a whole (sample, channel-slice) in shared memory, so x is read from HBM once (8 B per element). */
typedef struct {
  const float* x;
  const float* x2;
  float* stats;
  const float* gamma;
  const float* beta;
  void* y;         /* fp32, or bf16 when y_dtype == 1 */
  int N, HW, C1, C2;
  int x_ld, x2_ld;
  float eps;
  int silu;
  int round_tf32;
  int y_dtype;     /* 0 = fp32 (optionally TF32-rounded), 1 = bf16 (the operand of a bf16 convolution) */
  int x_dtype;     /* 0 = fp32, 1 = x / x2 hold bf16 (x_ld in elements; shared-memory-resident kernels only) */
  /* nn.Dropout(p) between the SiLU and the convolution of ResBlock.out_layers (simple_unet.py:168-172), fused:
   * y = silu(...) * keep / (1 - p); b200_gn_bwd applies the same keep / (1 - p) to dy.  keep is a hash of
   * (dropout_seed, *dropout_step, position in y), so no mask is stored and a captured graph whose step counter advances
   * draws a fresh mask every replay.  p is quantised to 1/65536; 0 = off (the three fields are then ignored). */
  float dropout_p;
  unsigned dropout_seed;   /* per layer */
  const int* dropout_step; /* optional device counter mixed into the seed; NULL = 0 */
} b200_gn_args;
int b200_gn_fwd(const b200_gn_args* a, void* ws, size_t ws_bytes, void* stream);

/* backward of the same block: dy is the gradient w.r.t. y.  dx[N*HW][C] = GN/SiLU backward (+ add1 + add2 on the
 * first add_c channels: the other gradients that flow into the same tensor).  dgamma/dbeta are accumulated with
 * atomics (caller zeroes the flat gradient buffer once per step).  fwd.stats must hold what b200_gn_fwd wrote.
 * ws: N*64 floats (used only when HW is too large for the shared-memory-resident kernel). */
typedef struct {
  b200_gn_args fwd;
  const float* dy;
  float* dx;
  const float* add1;
  const float* add2;
  int add1_ld, add2_ld, add_c;
  float* dgamma;
  float* dbeta;
  int round_dx; /* store dx rounded to TF32 */
  void* dx_bf16;  /* optional second output: bf16 copy of the first dx_bf16_c channels of dx, dense [N*HW][dx_bf16_c]
                   * (the dY operand of the bf16 dgrad / wgrad of the producing convolution) */
  int dx_bf16_c;
  /* optional fused column sums of dx over its first colsum_c (= 128) channels — what b200_colsum would compute from
   * dx: total1/total2[c] += sum over (n, pixel) (bias gradients of the convolution(s) that produced x),
   * per_sample[n * ps_ld + c] += sum over the pixels of sample n (gradient of the emb broadcast add). */
  float* colsum_per_sample;
  int colsum_ps_ld;
  float* colsum_total1;
  float* colsum_total2;
  int colsum_c;
  int dy_dtype; /* 0 = fp32, 1 = dy holds bf16.  dx may be NULL when only dx_bf16 (and the column sums) are wanted. */
} b200_gn_bwd_args;
int b200_gn_bwd(const b200_gn_bwd_args* a, void* ws, size_t ws_bytes, void* stream);
/* bytes of workspace b200_gn_bwd needs for these arguments (the only entry point that takes one; every other call
 * accepts ws = NULL).  The caller allocates it (PyTorch) and may share one buffer between all launches of a stream. */
size_t b200_gn_bwd_workspace(const b200_gn_bwd_args* a);
/* b200_gn_bwd without a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating library
 * call): for callers under CUDA graph capture, tight serving loops or harnesses that forbid allocation.  No counterpart
 * in the reference.  The streaming path's N*64-float reduction scratch is
 * a static device array of B200_GN_BWD_STATIC_MAX_N * 64 floats (1 MiB, resident on every GPU that loads the library
 * until process exit); fwd.N beyond that limit is rejected (rc != 0, b200_last_error says so).  Same arguments, results
 * and error behaviour as b200_gn_bwd otherwise.  All calls share the one array (cleared on `stream` before use):
 * calls that overlap on the GPU (different streams without a dependency) race on it. */
#define B200_GN_BWD_STATIC_MAX_N 4096
int b200_gn_bwd_static(const b200_gn_bwd_args* a, void* stream);

/* ------------------------------------------------------------------ diffusion math around the net (D = pixels/img)
 * mean_type: 0 = 'eps', 1 = 'x', 2 = 'v' (gaussian_diffusion.py:55-63).  sampler: 0 = 'ddim', 1 = 'noisy'. */

/* q_sample: z = x*sqrt(sigmoid(l)) + sqrt(sigmoid(-l))*eps  (diffusion_utils.py:65-73 + gaussian_diffusion.py:100) */
typedef struct {
  const float* x;
  const float* eps;
  const float* logsnr; /* [B] */
  float* z;
  int B, D;
} b200_q_sample_args;
int b200_q_sample(const b200_q_sample_args* a, void* ws, size_t ws_bytes, void* stream);

/* loss: _run_model + x_mse/eps_mse/max (gaussian_diffusion.py:45-79,156-172) and, when dout != NULL, the gradient of
 * grad_scale * loss_b w.r.t. the raw net output (grad_scale = 1/B for the batch-mean loss of diffusion_model.py:78).
 * loss_mean (optional) receives mean_b(loss_b); mse2 (optional) receives [B][2] = {x_mse, eps_mse}. */
typedef struct {
  const float* out;
  const float* z;
  const float* x;
  const float* eps;
  const float* logsnr;
  float* loss_b;
  float* mse2;
  float* dout;
  float* loss_mean;
  int B, D, mean_type; /* D a multiple of 4 */
  float grad_scale;
  int loss_type; /* 0 = 'snr_trunc': max(x_mse, eps_mse) (gaussian_diffusion.py:166-168); 1 = 'snr': eps_mse (:169-170,
                  * the weighting of distillation step1, :40-41).  x / eps are then the teacher's targets. */
} b200_loss_args;
int b200_loss(const b200_loss_args* a, void* ws, size_t ws_bytes, void* stream);

/* one sampler update: _run_model on the conditional output, optional classifier-free guidance against the
 * unconditional output (gaussian_diffusion.py:174-187), then reverse_dpm_step (:215-243, diffusion_utils.py:34-51
 * with x_logvar='large') or ddim_step (:189-213); z_next = x_pred when the loop index is 0 (:292).
 * sched is the [T][2] table {logsnr_t, logsnr_s} per loop index; the index is `step`, or *step_dev when given. */
typedef struct {
  const float* z;
  const float* out_c;
  const float* out_u;
  const float* cond_w;
  const float* noise;
  const float* sched;
  const int* step_dev;
  float* z_next;
  float* x_hist; /* optional: clipped x prediction of this step (evaluate()'s trajectories) */
  float* e_hist; /* optional: eps prediction of this step */
  int n, D, step, mean_type, sampler; /* D a multiple of 4 (16 B accesses) */
  float* z_dup;  /* optional second copy of z_next: the unconditional rows of a 2n-row guided batch */
  /* whole-trajectory histories (GaussianDiffusion.sample returns three [T, n, ...] stacks, gaussian_diffusion.py:293-296):
   * with hist_T > 0, x_hist / e_hist / z_hist are [hist_T][n*D] and this step writes row hist_T-1-index.  The uint8
   * variants hold DiffusionModel.evaluate's proc(): ((v + 1) * 127.5).clamp(0, 255).to(uint8) (diffusion_model.py:91-92),
   * 1 B/px instead of a 4 B/px fp32 history converted afterwards. */
  float* z_hist;
  int hist_T;
  void* z_u8;
  void* x_u8;
  void* e_u8;
  /* optional per-sample log-SNR pairs [n] (progressive distillation: the teacher's DDIM steps start from each sample's
   * own logsnr_t, gaussian_diffusion.py:113-146).  When given they replace the sched table / loop index. */
  const float* logsnr_t_b;
  const float* logsnr_s_b;
} b200_sampler_step_args;
int b200_sampler_step(const b200_sampler_step_args* a, void* ws, size_t ws_bytes, void* stream);

/* progressive-distillation step2 target (gaussian_diffusion.py:137-154): the x implied by where two teacher DDIM half
 * steps land,  x_target = (z_teacher - f z_t) / (alpha_s - f alpha_t),  f = exp(0.5 (softplus(l_t) - softplus(l_s))),
 * replaced by the teacher's own x prediction where the discrete index i is 0, and eps_target = eps_from_x(z_t, x_target, l_t). */
typedef struct {
  const float* z_teacher;
  const float* z_t;
  const float* x_pred_teacher;
  const float* logsnr_t;   /* [n] */
  const float* logsnr_s;   /* [n] */
  const long long* index;  /* [n] int64: the discrete step index i */
  float* x_target;
  float* eps_target;
  int n, D;
} b200_distill_target_args;
int b200_distill_target(const b200_distill_target_args* a, void* ws, size_t ws_bytes, void* stream);

typedef struct {
  int* step_dev;
} b200_step_advance_args;
int b200_step_advance(const b200_step_advance_args* a, void* ws, size_t ws_bytes, void* stream);

/* also used to bump the device-side Adam step counter */
int b200_int_increment(const b200_step_advance_args* a, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------ self-attention at the lowest UNet level (EXTENSION)
 * The reference's SimpleUnet has no attention (simple_unet.py:16-72); the north star / BASELINE.json configs[4] ask for
 * one on the tensor cores (7x7 = 49 tokens at 28x28, 16x16 = 256 tokens at 64x64; one head of 128 channels).  Shape of
 * the block (guided-diffusion's AttentionBlock, the family SimpleUnet derives from): x + proj(softmax(q k^T / sqrt(C)) v),
 * q / k / v / proj = 1x1 convolutions (b200_conv_igemm) of GroupNorm(x) (b200_gn_fwd, silu = 0).
 *
 * b200_bgemm: C[b] = alpha * op(A[b]) op(B[b])^T, batched, tcgen05 kind::tf32 with fp32 accumulation.
 *   A is [M][K] rows with pitch lda (a_mn = 0) or [K][M] rows (a_mn = 1); B is [N][K] rows with pitch ldb (b_mn = 0) or
 *   [K][N] rows (b_mn = 1); C is [M][N] rows with pitch ldc.  strideX = floats between batch elements.  N <= 256; lda, ldb
 *   and the strides are multiples of 4 floats; A and B hold TF32-rounded values.  M, N, K need not be multiples of
 *   anything (49 tokens): out-of-range operand elements read as zero, out-of-range results are not stored. */
typedef struct {
  const float* A;
  const float* B;
  float* C;
  int M, N, K, batch;
  int lda, ldb, ldc;
  long long strideA, strideB, strideC;
  int a_mn, b_mn;
  float alpha;
  int round_out; /* store C rounded to TF32 (it feeds another tensor-core operand) */
} b200_bgemm_args;
int b200_bgemm(const b200_bgemm_args* a, void* ws, size_t ws_bytes, void* stream);

/* row softmax over T <= 256 scores per row (pitch ld, T <= ld <= 256) and its backward; outputs are rounded to TF32 and
 * the pad columns [T, ld) are written as zeros.  fwd: p = softmax(s).  bwd: ds = p * (dp - sum_j dp_j p_j). */
typedef struct {
  const float* s;
  float* p;        /* fwd: output; bwd: input */
  const float* dp;
  float* ds;
  long long rows;
  int T, ld;
} b200_softmax_args;
int b200_softmax_fwd(const b200_softmax_args* a, void* ws, size_t ws_bytes, void* stream);
int b200_softmax_bwd(const b200_softmax_args* a, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------ edge convolutions (not tensor-core shaped)
 * conv_in: Conv2d(1, 128, 3, padding=1) (simple_unet.py:92-94); x [N,H,W], w [128][3][3], out/dy [N,H,W,128].
 * conv_out: Conv2d(128, 1, 3, padding=1) (simple_unet.py:41); h/dh [N,H,W,128], w [3][3][128], out/dout [N,H,W]. */
typedef struct {
  const float* x;
  const float* w;
  const float* bias;
  float* out;
  const float* dy;
  float* dw; /* accumulated with atomics */
  float* db;
  int N, H, W;
  int round_out; /* fwd: store out rounded to TF32 */
  /* multi-channel images (3x64x64 extension, BASELINE.json configs[4]; the reference is 1-channel): one launch per image
   * channel ci over planar input x[n][ci][H][W] — x points at plane ci of sample 0, x_ld = floats between samples
   * (0 = H*W), the weight is [128][3][3][Cimg] so w / dw point at element ci with w_co_ld = 9*Cimg and w_tap_ld = Cimg
   * (0 = 9 / 1); launches after the first set accumulate (fwd: out += , no bias) and pass db = NULL (bwd). */
  int x_ld, w_co_ld, w_tap_ld, accumulate;
} b200_conv_in_args;
int b200_conv_in_fwd(const b200_conv_in_args* a, void* ws, size_t ws_bytes, void* stream);
int b200_conv_in_bwd(const b200_conv_in_args* a, void* ws, size_t ws_bytes, void* stream);

typedef struct {
  const float* h;
  const float* w;
  const float* bias;
  float* out;
  const float* dout;
  float* dh;
  float* dw;
  float* db;
  int N, H, W;
  int round_dh; /* store dh rounded to TF32 */
  /* multi-channel images: one launch per image channel c with w / dw = weight row c ([3][3][128]), bias / db = element c,
   * out / dout = plane c of sample 0, out_ld = floats between samples (0 = H*W); backward launches after the first set
   * accumulate_dh (dh += ). */
  int out_ld, accumulate_dh;
} b200_conv_out_args;
int b200_conv_out_fwd(const b200_conv_out_args* a, void* ws, size_t ws_bytes, void* stream);
int b200_conv_out_bwd(const b200_conv_out_args* a, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------ glue
 * resample: mode 0 = nearest 2x upsample in [N,H,W,C] -> out [N,2H,2W,C] (F.interpolate, simple_unet.py:121);
 *           mode 1 = its backward: in = dout [N,2H,2W,C] (pixel stride in_ld) -> out = din [N,H,W,C];
 *           mode 2 = zero-stuffing: in = dy [N,ceil(H/2),ceil(W/2),C] (stride in_ld) -> out [N,H,W,C] with dy at
 *                    even pixels (turns the stride-2 Downsample dgrad into a stride-1 one). */
typedef struct {
  const float* in;
  void* out;
  int N, H, W, C;
  int in_ld;
  int mode;
  int round_out; /* mode 1: store the summed gradient rounded to TF32 */
  int out_dtype; /* mode 0: 0 = fp32, 1 = write bf16 (the operand of a bf16 convolution) */
} b200_resample_args;
int b200_upsample2x(const b200_resample_args* a, void* ws, size_t ws_bytes, void* stream);

/* channel sums of a [N,HW,128] gradient (pixel stride ld): per_sample[n*ps_ld + c] += sum_pix, total{1,2}[c] += sum_{n,pix}.
 * These are the conv-bias gradients and the gradient of the emb_out broadcast add (simple_unet.py:183-184). */
typedef struct {
  const float* dy;
  float* per_sample;
  float* total1;
  float* total2;
  int N, HW, ld, ps_ld;
} b200_colsum_args;
int b200_colsum(const b200_colsum_args* a, void* ws, size_t ws_bytes, void* stream);

/* embeddings (simple_unet.py:45-58, 205-224): prep writes sincos[B][64] of t (or of sched[2*step], the sampler's
 * logsnr_t) and onehot[B][10]; combine writes emb = t_emb + masked g_emb and semb = silu(emb); combine_bwd inverts. */
typedef struct {
  const float* t;
  const float* sched;
  const int* step_dev;
  const float* freqs;         /* [32], built by the reference expression on the host */
  const long long* guide;     /* int64 labels, -1 = unconditional; may be NULL */
  float* sincos;
  float* onehot;
  float* t_out;               /* optional [B]: the t actually used */
  const float* t_emb;
  const float* g_emb;
  float* emb;
  float* semb;
  const float* d_semb;
  float* d_emb;
  float* d_g_emb;
  int B, E, step, force_uncond;
  /* optional [B][E]: the cond_w embedding (simple_unet.py:60-64, progressive distillation), added to emb by combine.
   * Its sincos input is b200_embed_prep with t = cond_w and the max_period = 4 frequencies; its gradient is d_emb. */
  const float* w_emb;
} b200_embed_args;
int b200_embed_prep(const b200_embed_args* a, void* ws, size_t ws_bytes, void* stream);
int b200_embed_combine(const b200_embed_args* a, void* ws, size_t ws_bytes, void* stream);
int b200_embed_combine_bwd(const b200_embed_args* a, void* ws, size_t ws_bytes, void* stream);

/* nn.Linear on CUDA cores (fp32): mode 0: Y = act(X) W^T + b; mode 1: dX = (dY W) * act'(X); mode 2: dW += dY^T act(X),
 * db += colsum(dY).  act: 0 = identity, 1 = SiLU on the input (the nn.SiLU preceding the Linear).  W is [N][K]. */
typedef struct {
  const float* X;
  const float* W;
  const float* bias;
  float* Y;
  const float* dY;
  float* dX;
  float* dW;
  float* db;
  int M, N, K, act, mode;
} b200_linear_args;
int b200_linear(const b200_linear_args* a, void* ws, size_t ws_bytes, void* stream);

/* dst[i] = round-to-nearest TF32 of src[i] (n a multiple of 4): the tensor-core shadow copy of the flat parameters.
 * tcgen05 kind::tf32 truncates its operands; pre-rounding them removes the bias (see csrc/dev_common.cuh). */
typedef struct {
  const float* src;
  float* dst;
  size_t n;
} b200_round_args;
int b200_round_tf32(const b200_round_args* a, void* ws, size_t ws_bytes, void* stream);

/* dst[r][c] = bf16(src[r * src_ld + c]) (round to nearest even), dst dense [rows][cols]; cols a multiple of 4.
 * Builds the bf16 weight shadow (rows = 1) and the bf16 copy of a gradient slice that feeds a bf16 dgrad / wgrad.
 * No reference counterpart (torch.autocast's casts, diffusion_model.py:66). */
typedef struct {
  const float* src;
  void* dst;
  size_t rows;
  int cols;
  int src_ld; /* floats between rows of src; 0 = dense (cols) */
} b200_cast_args;
int b200_cast_bf16(const b200_cast_args* a, void* ws, size_t ws_bytes, void* stream);

/* Transposed, tap-flipped shadow of the convolution weights, so that backward-data runs through the forward kernel:
 *   dst[off + (ci*T + t)*128 + co] = round(src[off + (co*T + (T-1-t))*Cin + ci])   for every listed weight
 * (backward-data of a stride-1 'same' convolution == the forward convolution of dY with these weights; no reference
 * counterpart — cuDNN picks its own dgrad algorithm).  table: n_weights rows of {off, T = R*S, Cin, 0} (int32, device
 * memory), all weights have 128 output channels.  dst is fp32 rounded to TF32 (out_dtype 0) or bf16 (out_dtype 1). */
typedef struct {
  const float* src;
  void* dst;
  const int* table;
  int n_weights;
  int max_elems; /* largest 128*T*Cin in the table (sizes the grid) */
  int out_dtype;
} b200_wtranspose_args;
int b200_conv_weights_transpose(const b200_wtranspose_args* a, void* ws, size_t ws_bytes, void* stream);

/* cudaMemsetAsync(ptr, 0, bytes) on the stream (statistics / gradient arenas; replaces optimizer.zero_grad()). */
typedef struct {
  void* ptr;
  size_t bytes;
} b200_zero_args;
int b200_zero(const b200_zero_args* a, void* ws, size_t ws_bytes, void* stream);

/* torch.optim.Adam step (diffusion_model.py:56,71) over flat fp32 buffers; step = 1-based step number (or *step_dev). */
typedef struct {
  float* p;
  const float* g;
  float* m;
  float* v;
  size_t n;
  float lr, beta1, beta2, eps, grad_scale;
  const int* step_dev;
  int step;
  /* optional device-resident {lr, grad_scale} (2 floats): when given they override the two struct fields, so a
   * captured CUDA graph follows a learning-rate schedule / a changed world size without being re-captured */
  const float* hparams_dev;
} b200_adam_args;
int b200_adam(const b200_adam_args* a, void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DDPM_H_ */
