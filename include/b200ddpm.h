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

#define B200_ABI_VERSION 1

int b200_abi_version(void);
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
 * x may be a virtual channel concat of two tensors (x: C1 channels, x2: C2 channels), reference torch.cat at
 * simple_unet.py:150.  C1, C2 multiples of 32; Cout (fprop) must be 128; N-tile is 128 columns per launch.
 */
typedef struct {
  const float* x;
  const float* x2; /* NULL when C2 == 0 */
  const float* w;
  float* out;
  const float* bias;     /* [128] or NULL */
  const float* emb;      /* [N][emb_ld] or NULL */
  const float* residual; /* [M][res_ld] or NULL */
  int N, H, W, C1, C2;   /* geometry of the tensor the im2col walks (x) */
  int Ho, Wo;            /* output spatial size */
  int R, S, stride, pad; /* filter of the ORIGINAL convolution */
  int mode;              /* 0 fprop, 1 dgrad */
  int w_cout, w_cin;     /* weight tensor is [w_cout][R*S*w_cin] */
  int n_off;             /* dgrad only: first input channel produced by this launch */
  int out_ld, emb_ld, res_ld;
} b200_conv_args;
int b200_conv_igemm(const b200_conv_args* a, void* ws, size_t ws_bytes, void* stream);

/* weight gradient (nn.Conv2d backward-weight of the same layers):
 *   dw[co, kh, kw, ci] += sum_{n,p,q} dy[(n,p,q), co] * x[n, p*stride-pad+kh, q*stride-pad+kw, ci]
 * Accumulates with fp32 atomics into dw (caller zeroes it once per step). */
typedef struct {
  const float* x;
  const float* x2;
  const float* dy; /* [N*Ho*Wo][128] */
  float* dw;       /* [128][R*S*(C1+C2)] */
  int N, H, W, C1, C2;
  int Ho, Wo;
  int R, S, stride, pad;
  int k_slices; /* split of the pixel reduction; 0 = library picks */
} b200_wgrad_args;
int b200_conv_wgrad(const b200_wgrad_args* a, void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DDPM_H_ */
