// The small glue kernels of the UNet: nearest-2x upsample (fwd/bwd), zero-stuffing for strided dgrad, per-sample
// channel sums (bias / timestep-embedding gradients), the embedding MLPs (tiny fp32 GEMMs on CUDA cores: 0.02 % of
// the FLOPs, launch-latency territory) and the fused flat Adam step.
// Reference: simple_unet.py:45-58 (embeddings), :112-122 (Upsample), :165-167,183-184 (emb_layers);
// diffusion_model.py:56,70-72 (torch.optim.Adam defaults).
#include "host_common.h"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

namespace b200 {

static int ew_grid2(size_t total) {
  size_t blocks = (total + 255) / 256;
  const size_t cap = 148 * 16;
  return static_cast<int>(blocks < cap ? (blocks ? blocks : 1) : cap);
}
#define GRID_STRIDE(i, total)                                                                  \
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < (total); \
       i += static_cast<size_t>(gridDim.x) * blockDim.x)

// ---------------------------------------------------------------------------------------------- upsample 2x nearest
__global__ void __launch_bounds__(256)
upsample2x_fwd_kernel(const float4* __restrict__ in, float4* __restrict__ out, int N, int H, int W, int C4) {
  const size_t total = static_cast<size_t>(N) * 4 * H * W * C4;
  GRID_STRIDE(i, total) {
    const int c = static_cast<int>(i % C4);
    size_t p = i / C4;
    const int x = static_cast<int>(p % (2 * W));
    p /= (2 * W);
    const int y = static_cast<int>(p % (2 * H));
    const int n = static_cast<int>(p / (2 * H));
    out[i] = in[((static_cast<size_t>(n) * H + (y >> 1)) * W + (x >> 1)) * C4 + c];
  }
}
__global__ void __launch_bounds__(256)
upsample2x_fwd_bf16_kernel(const float4* __restrict__ in, uint2* __restrict__ out, int N, int H, int W, int C4) {
  const size_t total = static_cast<size_t>(N) * 4 * H * W * C4;
  GRID_STRIDE(i, total) {
    const int c = static_cast<int>(i % C4);
    size_t p = i / C4;
    const int x = static_cast<int>(p % (2 * W));
    p /= (2 * W);
    const int y = static_cast<int>(p % (2 * H));
    const int n = static_cast<int>(p / (2 * H));
    out[i] = pack_bf16x4(in[((static_cast<size_t>(n) * H + (y >> 1)) * W + (x >> 1)) * C4 + c]);
  }
}
// din[n,y,x,c] = sum of the 2x2 block of dout; dout pixel stride is ld4 float4s (it may be a slice of a wider buffer)
__global__ void __launch_bounds__(256)
upsample2x_bwd_kernel(const float4* __restrict__ dout, float4* __restrict__ din, int N, int H, int W, int C4, int ld4,
                      int round_out) {
  const size_t total = static_cast<size_t>(N) * H * W * C4;
  GRID_STRIDE(i, total) {
    const int c = static_cast<int>(i % C4);
    size_t p = i / C4;
    const int x = static_cast<int>(p % W);
    p /= W;
    const int y = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    const size_t r0 = ((static_cast<size_t>(n) * 2 * H + 2 * y) * 2 * W + 2 * x) * ld4 + c;
    const size_t r1 = r0 + static_cast<size_t>(2 * W) * ld4;
    const float4 a = dout[r0], b = dout[r0 + ld4], d = dout[r1], e = dout[r1 + ld4];
    float4 o = make_float4((a.x + b.x) + (d.x + e.x), (a.y + b.y) + (d.y + e.y), (a.z + b.z) + (d.z + e.z),
                           (a.w + b.w) + (d.w + e.w));
    if (round_out) o = round_tf32(o);
    din[i] = o;
  }
}
// out[n,y,x,:] = (y,x both even) ? dy[n,y/2,x/2,:] : 0   — turns a stride-2 dgrad into a stride-1 one.
__global__ void __launch_bounds__(256)
zero_stuff_kernel(const float4* __restrict__ dy, float4* __restrict__ out, int N, int H, int W, int Ho, int Wo, int C4,
                  int ld4) {
  const size_t total = static_cast<size_t>(N) * H * W * C4;
  GRID_STRIDE(i, total) {
    const int c = static_cast<int>(i % C4);
    size_t p = i / C4;
    const int x = static_cast<int>(p % W);
    p /= W;
    const int y = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!(x & 1) && !(y & 1) && (y >> 1) < Ho && (x >> 1) < Wo)
      v = dy[((static_cast<size_t>(n) * Ho + (y >> 1)) * Wo + (x >> 1)) * ld4 + c];
    out[i] = v;
  }
}

// ---------------------------------------------------------------------------------------------- channel sums
// per_sample[n*ps_ld + c] = sum_pix dy[n,pix,c] (optional, overwritten via atomics on a zeroed buffer);
// total[c] += the same summed over n (optional, up to two destinations).  grid (chunks, N), block 256 = 8 rows x 32 quads.
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ dy, int ld, int HW, int pix_per_chunk, float* __restrict__ per_sample,
              int ps_ld, float* __restrict__ total1, float* __restrict__ total2) {
  __shared__ float sh[8][128];
  const int n = blockIdx.y;
  const int q = threadIdx.x & 31, r = threadIdx.x >> 5;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const int p0 = blockIdx.x * pix_per_chunk, p1 = min(p0 + pix_per_chunk, HW);
  for (int p = p0 + r; p < p1; p += 8) {
    const float4 v = *reinterpret_cast<const float4*>(dy + (static_cast<size_t>(n) * HW + p) * ld + q * 4);
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  *reinterpret_cast<float4*>(&sh[r][q * 4]) = acc;
  __syncthreads();
  if (threadIdx.x < 128) {
    float s = 0.f;
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) s += sh[rr][threadIdx.x];
    if (per_sample) atomicAdd(&per_sample[static_cast<size_t>(n) * ps_ld + threadIdx.x], s);
    if (total1) atomicAdd(&total1[threadIdx.x], s);
    if (total2) atomicAdd(&total2[threadIdx.x], s);
  }
}

// ---------------------------------------------------------------------------------------------- embeddings
// sincos[b][0:32] = cos(t_b * f_k), [32:64] = sin(t_b * f_k); t_b = t[b] or, when t == NULL, sched[2*step] (the
// sampler's logsnr_t broadcast to the batch).  onehot[b][10] with class 0 for unconditional (-1) rows.
__global__ void embed_prep_kernel(const float* __restrict__ t, const float* __restrict__ sched,
                                  const int* __restrict__ step_dev, int step, const float* __restrict__ freqs,
                                  const long long* __restrict__ guide, int force_uncond, float* __restrict__ sincos,
                                  float* __restrict__ onehot, float* __restrict__ t_out, int B) {
  const int b = blockIdx.x;
  if (b >= B) return;
  float tv;
  if (t) tv = t[b];
  else tv = sched[2 * (step_dev ? *step_dev : step)];
  const int k = threadIdx.x;  // 64 threads
  if (k < 32) sincos[b * 64 + k] = cosf(tv * freqs[k]);
  else if (k < 64) sincos[b * 64 + k] = sinf(tv * freqs[k - 32]);
  if (k < 10 && onehot) {
    long long g = guide ? guide[b] : -1;
    if (force_uncond || g < 0) g = 0;
    onehot[b * 10 + k] = (k == g) ? 1.0f : 0.0f;
  }
  if (k == 0 && t_out) t_out[b] = tv;
}
// emb = t_emb + (uncond ? 0 : g_emb);  semb = silu(emb)
__global__ void __launch_bounds__(256)
embed_combine_kernel(const float* __restrict__ t_emb, const float* __restrict__ g_emb,
                     const long long* __restrict__ guide, int force_uncond, const float* __restrict__ w_emb,
                     float* __restrict__ emb, float* __restrict__ semb, int B, int E) {
  const size_t total = static_cast<size_t>(B) * E;
  GRID_STRIDE(i, total) {
    const int b = static_cast<int>(i / E);
    float v = t_emb[i];
    if (g_emb && !force_uncond && guide[b] >= 0) v += g_emb[i];
    if (w_emb) v += w_emb[i];  // simple_unet.py:60-64: emb += cond_w_embed(...), after the guide embedding
    emb[i] = v;
    semb[i] = v / (1.0f + expf(-v));
  }
}
// d_emb = d_semb * silu'(emb);  d_g_emb = uncond ? 0 : d_emb
__global__ void __launch_bounds__(256)
embed_combine_bwd_kernel(const float* __restrict__ emb, const float* __restrict__ d_semb,
                         const long long* __restrict__ guide, float* __restrict__ d_emb, float* __restrict__ d_g_emb,
                         int B, int E) {
  const size_t total = static_cast<size_t>(B) * E;
  GRID_STRIDE(i, total) {
    const int b = static_cast<int>(i / E);
    const float u = emb[i];
    const float sg = 1.0f / (1.0f + expf(-u));
    const float d = d_semb[i] * sg * (1.0f + u * (1.0f - sg));
    d_emb[i] = d;
    if (d_g_emb) d_g_emb[i] = (guide[b] >= 0) ? d : 0.f;
  }
}

// ---------------------------------------------------------------------------------------------- tiny fp32 GEMMs
// Y[M,N] = act(X)[M,K] * W[N,K]^T + b     (act: 0 none, 1 silu applied to X on load)
__global__ void __launch_bounds__(256)
linear_fwd_kernel(const float* __restrict__ X, const float* __restrict__ Wt, const float* __restrict__ bias,
                  float* __restrict__ Y, int M, int N, int K, int act) {
  __shared__ float sx[16][17], sw[16][17];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m = blockIdx.y * 16 + ty, n = blockIdx.x * 16 + tx;
  float acc = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    const int kx = k0 + tx;
    float xv = (m < M && kx < K) ? X[static_cast<size_t>(m) * K + kx] : 0.f;
    if (act) xv = xv / (1.0f + expf(-xv));
    sx[ty][tx] = xv;
    const int wn = blockIdx.x * 16 + ty;
    sw[ty][tx] = (wn < N && kx < K) ? Wt[static_cast<size_t>(wn) * K + kx] : 0.f;
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) acc += sx[ty][kk] * sw[tx][kk];
    __syncthreads();
  }
  if (m < M && n < N) Y[static_cast<size_t>(m) * N + n] = acc + (bias ? bias[n] : 0.f);
}
// dX[M,K] = dY[M,N] * W[N,K]   (times silu'(X) when act; X is the pre-activation input of the forward)
__global__ void __launch_bounds__(256)
linear_bwd_x_kernel(const float* __restrict__ dY, const float* __restrict__ Wt, const float* __restrict__ X,
                    float* __restrict__ dX, int M, int N, int K, int act) {
  __shared__ float sd[16][17], sw[16][17];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m = blockIdx.y * 16 + ty, k = blockIdx.x * 16 + tx;
  float acc = 0.f;
  for (int n0 = 0; n0 < N; n0 += 16) {
    sd[ty][tx] = (m < M && n0 + tx < N) ? dY[static_cast<size_t>(m) * N + n0 + tx] : 0.f;
    sw[ty][tx] = (n0 + ty < N && k < K) ? Wt[static_cast<size_t>(n0 + ty) * K + k] : 0.f;
    __syncthreads();
#pragma unroll
    for (int nn = 0; nn < 16; ++nn) acc += sd[ty][nn] * sw[nn][tx];
    __syncthreads();
  }
  if (m < M && k < K) {
    if (act) {
      const float u = X[static_cast<size_t>(m) * K + k];
      const float sg = 1.0f / (1.0f + expf(-u));
      acc *= sg * (1.0f + u * (1.0f - sg));
    }
    dX[static_cast<size_t>(m) * K + k] = acc;
  }
}
// dW[N,K] += dY^T[N,M] * act(X)[M,K];  db[N] += sum_m dY[m,n] (by the blockIdx.x == 0 column of CTAs)
__global__ void __launch_bounds__(256)
linear_bwd_w_kernel(const float* __restrict__ dY, const float* __restrict__ X, float* __restrict__ dW,
                    float* __restrict__ db, int M, int N, int K, int act) {
  __shared__ float sd[16][17], sx[16][17];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int n = blockIdx.y * 16 + ty, k = blockIdx.x * 16 + tx;
  float acc = 0.f, bacc = 0.f;
  for (int m0 = 0; m0 < M; m0 += 16) {
    // sd[mm][nn] = dY[m0+mm][n-tile nn];  sx[mm][kk] = act(X[m0+mm][k-tile kk])
    sd[ty][tx] = (m0 + ty < M && blockIdx.y * 16 + tx < N) ? dY[static_cast<size_t>(m0 + ty) * N + blockIdx.y * 16 + tx] : 0.f;
    float xv = (m0 + ty < M && k < K) ? X[static_cast<size_t>(m0 + ty) * K + k] : 0.f;
    if (act) xv = xv / (1.0f + expf(-xv));
    sx[ty][tx] = xv;
    __syncthreads();
#pragma unroll
    for (int mm = 0; mm < 16; ++mm) {
      acc += sd[mm][ty] * sx[mm][tx];
      bacc += sd[mm][ty];
    }
    __syncthreads();
  }
  if (n < N && k < K) atomicAdd(&dW[static_cast<size_t>(n) * K + k], acc);
  if (db && blockIdx.x == 0 && tx == 0 && n < N) atomicAdd(&db[n], bacc);
}

// The batched embedding projection (512 -> 13 x 128 at batch 512) is the one linear big enough to matter: 64x64 output
// tiles, 4x4 outputs per thread, both operand tiles staged reduction-major in shared memory (16 FMAs per two LDS.128).
// C[i][j] = sum_r A(i, r) * B(j, r).  ARC / BRC: the operand's reduction index is its contiguous one (row = output
// index) — otherwise the output index is contiguous (row = reduction index).  I x J x R need not be tile multiples in
// I (the batch); J and R are multiples of 64 / 16 (checked by the host).
template <bool RC>
__device__ __forceinline__ float4 lin_fetch(const float* __restrict__ src, int ld, int o0, int r0, int o_lim, int act) {
  const int t = threadIdx.x;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (RC) {  // src[(o0 + o) * ld + r0 + r]: thread -> (o = t / 4, r = 4 * (t % 4))
    const int o = t >> 2, r4 = (t & 3) * 4;
    if (o0 + o < o_lim) v = __ldg(reinterpret_cast<const float4*>(src + static_cast<size_t>(o0 + o) * ld + r0 + r4));
  } else {  // src[(r0 + r) * ld + o0 + o]: thread -> (r = t / 16, o = 4 * (t % 16)); o_lim bounds the reduction rows
    const int r = t >> 4, o4 = (t & 15) * 4;
    if (r0 + r < o_lim) v = __ldg(reinterpret_cast<const float4*>(src + static_cast<size_t>(r0 + r) * ld + o0 + o4));
  }
  if (act) {
    v.x = v.x / (1.0f + expf(-v.x)); v.y = v.y / (1.0f + expf(-v.y));
    v.z = v.z / (1.0f + expf(-v.z)); v.w = v.w / (1.0f + expf(-v.w));
  }
  return v;
}
template <bool RC>
__device__ __forceinline__ void lin_put(float (*dst)[68], float4 v) {
  const int t = threadIdx.x;
  if (RC) {  // stored transposed: reduction-major
    const int o = t >> 2, r4 = (t & 3) * 4;
    dst[r4][o] = v.x; dst[r4 + 1][o] = v.y; dst[r4 + 2][o] = v.z; dst[r4 + 3][o] = v.w;
  } else {
    const int r = t >> 4, o4 = (t & 15) * 4;
    *reinterpret_cast<float4*>(&dst[r][o4]) = v;
  }
}
// MODE 0: Y = act(X) W^T + b  (A = X rows m, RC; B = W rows n, RC; R = K)
// MODE 1: dX = dY W (* silu'(X))  (A = dY rows m, RC over n; B = W[n][k], output-contiguous; R = N)
// MODE 2: dW += dY^T act(X), db += colsum(dY)  (A = dY[m][n], B = X[m][k], both output-contiguous; R = M)
template <int MODE>
__global__ void __launch_bounds__(256)
linear_tile_kernel(const float* __restrict__ A, const float* __restrict__ B, const float* __restrict__ aux,
                   float* __restrict__ Cout, float* __restrict__ db, int M, int N, int K, int act) {
  __shared__ __align__(16) float sa[16][68], sb[16][68];
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  const int R = MODE == 0 ? K : (MODE == 1 ? N : M);
  float acc[4][4] = {};
  float bacc[4] = {0.f, 0.f, 0.f, 0.f};
  // global -> register fetch of tile r0 + 16 is issued before the FMAs on tile r0 (its latency was exposed once per
  // 16-deep step)
  auto fetch = [&](int r0, float4& fa, float4& fb) {
    if (MODE == 0) {
      fa = lin_fetch<true>(A, K, i0, r0, M, act);
      fb = lin_fetch<true>(B, K, j0, r0, N, 0);
    } else if (MODE == 1) {
      fa = lin_fetch<true>(A, N, i0, r0, M, 0);
      fb = lin_fetch<false>(B, K, j0, r0, N, 0);
    } else {
      fa = lin_fetch<false>(A, N, i0, r0, M, 0);
      fb = lin_fetch<false>(B, K, j0, r0, M, act);
    }
  };
  float4 fa, fb;
  fetch(0, fa, fb);
  for (int r0 = 0; r0 < R; r0 += 16) {
    lin_put<MODE != 2>(sa, fa);
    lin_put<MODE == 0>(sb, fb);
    if (r0 + 16 < R) fetch(r0 + 16, fa, fb);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const float4 a4 = *reinterpret_cast<const float4*>(&sa[r][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&sb[r][tx * 4]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (MODE == 2) bacc[i] += av[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
    }
    __syncthreads();
  }
  const int ldc = MODE == 0 ? N : K;
  const int i_lim = MODE == 2 ? N : M;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gi = i0 + ty * 4 + i;
    if (gi >= i_lim) continue;
    const int gj = j0 + tx * 4;
    float4 o = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    float* dst = Cout + static_cast<size_t>(gi) * ldc + gj;
    if (MODE == 0) {
      if (aux) {
        const float4 b = *reinterpret_cast<const float4*>(aux + gj);
        o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
      }
      *reinterpret_cast<float4*>(dst) = o;
    } else if (MODE == 1) {
      if (act) {
        const float4 u = *reinterpret_cast<const float4*>(aux + static_cast<size_t>(gi) * K + gj);
        const float uv[4] = {u.x, u.y, u.z, u.w};
        float ov[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float sg = 1.0f / (1.0f + expf(-uv[j]));
          ov[j] *= sg * (1.0f + uv[j] * (1.0f - sg));
        }
        o = make_float4(ov[0], ov[1], ov[2], ov[3]);
      }
      *reinterpret_cast<float4*>(dst) = o;
    } else {
      atomicAdd(dst, o.x); atomicAdd(dst + 1, o.y); atomicAdd(dst + 2, o.z); atomicAdd(dst + 3, o.w);
      if (db && blockIdx.x == 0 && tx == 0) atomicAdd(&db[gi], bacc[i]);
    }
  }
}

// ---------------------------------------------------------------------------------------------- Adam
// torch.optim.Adam (defaults: amsgrad=False, weight_decay=0), one pass over the flat buffers.  g is scaled by
// grad_scale first (1/world_size after the allreduce-sum, and/or 1/loss_scale).
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, size_t n,
            float lr, float beta1, float beta2, float eps, float grad_scale, const int* __restrict__ step_dev, int step,
            const float* __restrict__ hparams_dev) {
  const int t = step_dev ? *step_dev : step;
  if (hparams_dev) {
    lr = hparams_dev[0];
    grad_scale = hparams_dev[1];
  }
  const float bc1 = 1.0f - powf(beta1, static_cast<float>(t));
  const float bc2 = 1.0f - powf(beta2, static_cast<float>(t));
  const float step_size = lr / bc1;
  const float bc2_sqrt = sqrtf(bc2);
  GRID_STRIDE(i, n) {
    const float gi = g[i] * grad_scale;
    const float mi = m[i] + (gi - m[i]) * (1.0f - beta1);
    const float vi = v[i] * beta2 + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}
// dst = round_to_nearest_tf32(src): the tensor-core shadow copy of the flat parameter buffer
__global__ void __launch_bounds__(256) round_tf32_kernel(const float4* __restrict__ src, float4* __restrict__ dst, size_t n4) {
  GRID_STRIDE(i, n4) dst[i] = round_tf32(src[i]);
}
__global__ void int_increment_kernel(int* v) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *v += 1;
}

}  // namespace b200

using namespace b200;
#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int b200_upsample2x(const b200_resample_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->in || !a->out || a->C % 4) return fail_arg("upsample2x: bad arguments");
  const int C4 = a->C / 4;
  if (a->mode == 0) {
    const size_t total = static_cast<size_t>(a->N) * 4 * a->H * a->W * C4;
    if (a->out_dtype == 1) {
      upsample2x_fwd_bf16_kernel<<<ew_grid2(total), 256, 0, ST(stream)>>>(reinterpret_cast<const float4*>(a->in),
                                                                           reinterpret_cast<uint2*>(a->out), a->N, a->H, a->W, C4);
      return after_launch("upsample2x_fwd_bf16_kernel");
    }
    upsample2x_fwd_kernel<<<ew_grid2(total), 256, 0, ST(stream)>>>(reinterpret_cast<const float4*>(a->in),
                                                                    reinterpret_cast<float4*>(a->out), a->N, a->H, a->W, C4);
    return after_launch("upsample2x_fwd_kernel");
  } else if (a->mode == 1) {
    const size_t total = static_cast<size_t>(a->N) * a->H * a->W * C4;
    const int ld4 = (a->in_ld ? a->in_ld : a->C) / 4;
    upsample2x_bwd_kernel<<<ew_grid2(total), 256, 0, ST(stream)>>>(reinterpret_cast<const float4*>(a->in),
                                                                    reinterpret_cast<float4*>(a->out), a->N, a->H, a->W, C4, ld4,
                                                                    a->round_out);
    return after_launch("upsample2x_bwd_kernel");
  } else if (a->mode == 2) {
    const size_t total = static_cast<size_t>(a->N) * a->H * a->W * C4;
    const int ld4 = (a->in_ld ? a->in_ld : a->C) / 4;
    zero_stuff_kernel<<<ew_grid2(total), 256, 0, ST(stream)>>>(reinterpret_cast<const float4*>(a->in),
                                                                reinterpret_cast<float4*>(a->out), a->N, a->H, a->W,
                                                                (a->H + 1) / 2, (a->W + 1) / 2, C4, ld4);
    return after_launch("zero_stuff_kernel");
  }
  return fail_arg("upsample2x: mode must be 0 (fwd), 1 (bwd) or 2 (zero-stuff)");
}

extern "C" int b200_colsum(const b200_colsum_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->dy) return fail_arg("colsum: null pointer");
  int chunks = (592 + a->N - 1) / a->N;
  if (chunks > (a->HW + 31) / 32) chunks = (a->HW + 31) / 32;
  if (chunks < 1) chunks = 1;
  const int ppc = (a->HW + chunks - 1) / chunks;
  colsum_kernel<<<dim3(chunks, a->N), 256, 0, ST(stream)>>>(a->dy, a->ld ? a->ld : 128, a->HW, ppc, a->per_sample,
                                                            a->ps_ld, a->total1, a->total2);
  return after_launch("colsum_kernel");
}

extern "C" int b200_embed_prep(const b200_embed_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->freqs || !a->sincos || (!a->t && !a->sched)) return fail_arg("embed_prep: null pointer");
  embed_prep_kernel<<<a->B, 64, 0, ST(stream)>>>(a->t, a->sched, a->step_dev, a->step, a->freqs, a->guide, a->force_uncond,
                                                 a->sincos, a->onehot, a->t_out, a->B);
  return after_launch("embed_prep_kernel");
}

extern "C" int b200_embed_combine(const b200_embed_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->t_emb || !a->emb || !a->semb) return fail_arg("embed_combine: null pointer");
  if (a->g_emb && !a->guide) return fail_arg("embed_combine: g_emb needs guide labels");
  const size_t total = static_cast<size_t>(a->B) * a->E;
  embed_combine_kernel<<<ew_grid2(total), 256, 0, ST(stream)>>>(a->t_emb, a->g_emb, a->guide, a->force_uncond, a->w_emb,
                                                                 a->emb, a->semb, a->B, a->E);
  return after_launch("embed_combine_kernel");
}

extern "C" int b200_embed_combine_bwd(const b200_embed_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->emb || !a->d_semb || !a->d_emb) return fail_arg("embed_combine_bwd: null pointer");
  if (a->d_g_emb && !a->guide) return fail_arg("embed_combine_bwd: d_g_emb needs guide labels");
  const size_t total = static_cast<size_t>(a->B) * a->E;
  embed_combine_bwd_kernel<<<ew_grid2(total), 256, 0, ST(stream)>>>(a->emb, a->d_semb, a->guide, a->d_emb, a->d_g_emb,
                                                                     a->B, a->E);
  return after_launch("embed_combine_bwd_kernel");
}

extern "C" int b200_linear(const b200_linear_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->W) return fail_arg("linear: null pointer");
  const int M = a->M, N = a->N, K = a->K;
  // the 64x64-tile kernel where it gets at least 32 CTAs (batch-512 embedding MLPs; measured 64: 0.343 ms for the 13
  // linears of a step, 32: 0.328, 16: 0.369); the 16x16 one otherwise
  const int out_tiles = a->mode == 0 ? (N / 64) * ((M + 63) / 64) : a->mode == 1 ? (K / 64) * ((M + 63) / 64) : (K / 64) * (N / 64);
  static int min_tiles = -1;  // developer knob B200_LIN_MIN_TILES
  if (min_tiles < 0) {
    const char* e = getenv("B200_LIN_MIN_TILES");
    min_tiles = e ? atoi(e) : 32;
  }
  const bool big = (N % 64 == 0) && (K % 64 == 0) && out_tiles >= min_tiles;
  if (a->mode == 0) {
    if (!a->X || !a->Y) return fail_arg("linear fwd: null pointer");
    if (big) {
      linear_tile_kernel<0><<<dim3(N / 64, (M + 63) / 64), 256, 0, ST(stream)>>>(a->X, a->W, a->bias, a->Y, nullptr, M, N, K, a->act);
      return after_launch("linear_tile_kernel<fwd>");
    }
    linear_fwd_kernel<<<dim3((N + 15) / 16, (M + 15) / 16), 256, 0, ST(stream)>>>(a->X, a->W, a->bias, a->Y, M, N, K, a->act);
    return after_launch("linear_fwd_kernel");
  } else if (a->mode == 1) {
    if (!a->dY || !a->dX || (a->act && !a->X)) return fail_arg("linear bwd_x: null pointer");
    if (big) {
      linear_tile_kernel<1><<<dim3(K / 64, (M + 63) / 64), 256, 0, ST(stream)>>>(a->dY, a->W, a->X, a->dX, nullptr, M, N, K, a->act);
      return after_launch("linear_tile_kernel<bwd_x>");
    }
    linear_bwd_x_kernel<<<dim3((K + 15) / 16, (M + 15) / 16), 256, 0, ST(stream)>>>(a->dY, a->W, a->X, a->dX, M, N, K, a->act);
    return after_launch("linear_bwd_x_kernel");
  } else if (a->mode == 2) {
    if (!a->dY || !a->X || !a->dW) return fail_arg("linear bwd_w: null pointer");
    if (big) {
      linear_tile_kernel<2><<<dim3(K / 64, N / 64), 256, 0, ST(stream)>>>(a->dY, a->X, nullptr, a->dW, a->db, M, N, K, a->act);
      return after_launch("linear_tile_kernel<bwd_w>");
    }
    linear_bwd_w_kernel<<<dim3((K + 15) / 16, (N + 15) / 16), 256, 0, ST(stream)>>>(a->dY, a->X, a->dW, a->db, M, N, K, a->act);
    return after_launch("linear_bwd_w_kernel");
  }
  return fail_arg("linear: mode must be 0 (fwd), 1 (bwd x) or 2 (bwd w)");
}

extern "C" int b200_adam(const b200_adam_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->p || !a->g || !a->m || !a->v) return fail_arg("adam: null pointer");
  adam_kernel<<<ew_grid2(a->n), 256, 0, ST(stream)>>>(a->p, a->g, a->m, a->v, a->n, a->lr, a->beta1, a->beta2, a->eps,
                                                      a->grad_scale, a->step_dev, a->step, a->hparams_dev);
  if (int r = after_launch("adam_kernel")) return r;
  return 0;
}

extern "C" int b200_round_tf32(const b200_round_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->src || !a->dst || a->n % 4) return fail_arg("round_tf32: null pointer or n not a multiple of 4");
  round_tf32_kernel<<<ew_grid2(a->n / 4), 256, 0, ST(stream)>>>(reinterpret_cast<const float4*>(a->src),
                                                               reinterpret_cast<float4*>(a->dst), a->n / 4);
  return after_launch("round_tf32_kernel");
}

// fp32 -> bf16 (RNE) of a strided 2-D slice; 4 elements per thread, 16 B loads / 8 B stores.
__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ src, uint2* __restrict__ dst,
                                                        size_t total4, int cols4, int src_ld) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / cols4;
    const int c = static_cast<int>(i - r * cols4) * 4;
    const float4 v = *reinterpret_cast<const float4*>(src + r * src_ld + c);
    dst[i] = pack_bf16x4(v);
  }
}

extern "C" int b200_cast_bf16(const b200_cast_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->src || !a->dst || a->cols <= 0 || a->cols % 4) return fail_arg("cast_bf16: null pointer or cols not a multiple of 4");
  const int ld = a->src_ld ? a->src_ld : a->cols;
  if (ld % 4) return fail_arg("cast_bf16: src_ld must be a multiple of 4");
  const size_t total4 = a->rows * static_cast<size_t>(a->cols / 4);
  if (total4 == 0) return 0;
  cast_bf16_kernel<<<ew_grid2(total4), 256, 0, ST(stream)>>>(a->src, reinterpret_cast<uint2*>(a->dst), total4, a->cols / 4, ld);
  return after_launch("cast_bf16_kernel");
}

// one thread per output element; consecutive threads = consecutive co (coalesced stores; the strided reads hit L2)
__global__ void __launch_bounds__(256) conv_weights_transpose_kernel(const float* __restrict__ src, void* __restrict__ dst,
                                                                     const int4* __restrict__ table, int out_bf16) {
  const int4 e = table[blockIdx.y];
  const int off = e.x, T = e.y, Cin = e.z;
  const int total = 128 * T * Cin;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int co = i & 127;
    const int r = i >> 7;  // ci * T + t
    const int ci = r / T, t = r - ci * T;
    const float v = src[off + (co * T + (T - 1 - t)) * Cin + ci];
    if (out_bf16) static_cast<uint16_t*>(dst)[off + i] = bf16_bits(v);
    else static_cast<float*>(dst)[off + i] = round_tf32(v);
  }
}

extern "C" int b200_conv_weights_transpose(const b200_wtranspose_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->src || !a->dst || !a->table || a->n_weights <= 0 || a->max_elems <= 0)
    return fail_arg("conv_weights_transpose: bad arguments");
  int bx = (a->max_elems + 255) / 256;
  if (bx > 64) bx = 64;
  conv_weights_transpose_kernel<<<dim3(bx, a->n_weights), 256, 0, ST(stream)>>>(
      a->src, a->dst, reinterpret_cast<const int4*>(a->table), a->out_dtype);
  return after_launch("conv_weights_transpose_kernel");
}

extern "C" int b200_zero(const b200_zero_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->ptr) return fail_arg("zero: null pointer");
  return check_cuda(cudaMemsetAsync(a->ptr, 0, a->bytes, ST(stream)), "cudaMemsetAsync");
}

extern "C" int b200_int_increment(const b200_step_advance_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->step_dev) return fail_arg("int_increment: null pointer");
  int_increment_kernel<<<1, 32, 0, ST(stream)>>>(a->step_dev);
  return after_launch("int_increment_kernel");
}
