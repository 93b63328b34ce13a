// Per-pixel diffusion math around the UNet: q_sample, the clipped x/eps loss (forward + gradient w.r.t. the net
// output) and the fused sampler update (x/eps reconstruction, classifier-free guidance, ancestral or DDIM step).
// One HBM pass each.  Arithmetic mirrors the reference's evaluation order with explicitly rounded mul/add
// (__fmul_rn/__fadd_rn are never contracted into FMAs), because the x<->eps conversions are ill-conditioned at
// |logsnr| ~ 20 and parity has to survive that.
// Reference: gms/diffusion/diffusion_utils.py:34-105, gms/diffusion/gaussian_diffusion.py:45-79,81-100,156-243.
#include "host_common.h"

#include "../../include/b200ddpm.h"

namespace b200 {

__device__ __forceinline__ float mulr(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float addr(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float subr(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float sigmoid_t(float x) { return 1.0f / (1.0f + expf(-x)); }  // torch CUDA sigmoid
__device__ __forceinline__ float clip1(float v) { return fminf(fmaxf(v, -1.0f), 1.0f); }

// Per-sample coefficients of diffusion_utils.py:76-105 for a given logsnr l.
struct Coef {
  float A;      // sqrt(1 + exp(-l))       predict_x_from_eps outer
  float Bc;     // rsqrt(1 + exp(l))       predict_x_from_eps inner
  float Ce;     // sqrt(1 + exp(l))        predict_eps_from_x outer
  float De;     // rsqrt(1 + exp(-l))      predict_eps_from_x inner
  float alpha;  // sqrt(sigmoid(l))
  float sigma;  // sqrt(sigmoid(-l))
};
__device__ __forceinline__ Coef make_coef(float l) {
  Coef c;
  const float en = expf(-l), ep = expf(l);
  c.A = sqrtf(1.0f + en);
  c.Bc = rsqrtf(1.0f + ep);
  c.Ce = sqrtf(1.0f + ep);
  c.De = rsqrtf(1.0f + en);
  c.alpha = sqrtf(sigmoid_t(l));
  c.sigma = sqrtf(sigmoid_t(-l));
  return c;
}
// raw net output -> unclipped x prediction (gaussian_diffusion.py:55-63)
__device__ __forceinline__ float xhat_from_out(const Coef& c, int mean_type, float z, float out) {
  if (mean_type == 0) return mulr(c.A, subr(z, mulr(out, c.Bc)));
  if (mean_type == 1) return out;
  return subr(mulr(c.alpha, z), mulr(c.sigma, out));
}
__device__ __forceinline__ float eps_from_x(const Coef& c, float z, float x) {
  return mulr(c.Ce, subr(z, mulr(x, c.De)));
}

// ------------------------------------------------------------------------------------------------ q_sample
// z = x * sqrt(sigmoid(l)) + sqrt(sigmoid(-l)) * eps      (12 B/px)
// Each thread keeps kEwUnroll independent 16 B loads per tensor in flight (a grid-stride loop with one float4 per
// iteration left the kernel latency-bound at n = 8192: 3.76 TB/s; the per-sample coefficients are recomputed per
// float4, which costs nothing next to the memory time).
constexpr int kEwUnroll = 4;
__global__ void __launch_bounds__(256)
q_sample_kernel(const float* __restrict__ x, const float* __restrict__ eps, const float* __restrict__ logsnr,
                float* __restrict__ z, int B, int D4) {
  const size_t total = static_cast<size_t>(B) * D4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * kEwUnroll;
  for (size_t base = static_cast<size_t>(blockIdx.x) * blockDim.x * kEwUnroll + threadIdx.x; base < total; base += stride) {
    float4 xv[kEwUnroll], ev[kEwUnroll];
    float lv[kEwUnroll];
#pragma unroll
    for (int u = 0; u < kEwUnroll; ++u) {
      const size_t i = base + static_cast<size_t>(u) * blockDim.x;
      if (i < total) {
        xv[u] = __ldcs(reinterpret_cast<const float4*>(x) + i);
        ev[u] = __ldcs(reinterpret_cast<const float4*>(eps) + i);
        lv[u] = __ldg(logsnr + i / D4);
      }
    }
#pragma unroll
    for (int u = 0; u < kEwUnroll; ++u) {
      const size_t i = base + static_cast<size_t>(u) * blockDim.x;
      if (i < total) {
        const float a = sqrtf(sigmoid_t(lv[u])), s = sqrtf(sigmoid_t(-lv[u]));
        float4 o;
        o.x = addr(mulr(xv[u].x, a), mulr(s, ev[u].x));
        o.y = addr(mulr(xv[u].y, a), mulr(s, ev[u].y));
        o.z = addr(mulr(xv[u].z, a), mulr(s, ev[u].z));
        o.w = addr(mulr(xv[u].w, a), mulr(s, ev[u].w));
        reinterpret_cast<float4*>(z)[i] = o;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ loss
__device__ __forceinline__ float block_sum(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float t = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.f;
  if (w == 0) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (l == 0) sh[0] = t;
  }
  __syncthreads();
  return sh[0];
}

// One CTA per sample.  loss_b = max(mean (xc-x)^2, mean (ec-eps)^2); dout = d(grad_scale * loss_b)/d(out).
// The sample (out, z, x, eps: 16 B accesses) is read ONCE and held in registers across the block reduction when it fits
// (D <= 4 * 4 * 256 pixels; 28x28 .. 64x64 do): 20 B/px of HBM traffic, the algorithmic minimum.  Larger images re-read.
// kLossV = float4 chunks per thread and tensor held in registers (1 for 28x28: 40 registers, 6 CTAs per SM; the
// one-size-fits-all kLossV = 4 build ran at 100 registers, 2 CTAs per SM and 1.4 TB/s)
template <int kLossV>
__global__ void __launch_bounds__(256)
loss_kernel(const float* __restrict__ out, const float* __restrict__ z, const float* __restrict__ x,
            const float* __restrict__ eps, const float* __restrict__ logsnr, float* __restrict__ loss_b,
            float* __restrict__ mse2, float* __restrict__ dout, int D, int mean_type, float grad_scale, int loss_type) {
  __shared__ float sh[32];
  const int b = blockIdx.x;
  const Coef c = make_coef(logsnr[b]);
  const size_t base = static_cast<size_t>(b) * D;
  const int D4 = D >> 2;
  const bool cached = D4 <= kLossV * 256;
  const float4* o4 = reinterpret_cast<const float4*>(out + base);
  const float4* z4 = reinterpret_cast<const float4*>(z + base);
  const float4* x4 = reinterpret_cast<const float4*>(x + base);
  const float4* e4 = reinterpret_cast<const float4*>(eps + base);
  float4 ro[kLossV], rz[kLossV], rx[kLossV], re[kLossV];
  float sx = 0.f, se = 0.f;
  if (cached) {
#pragma unroll
    for (int v = 0; v < kLossV; ++v) {
      const int i = threadIdx.x + v * 256;
      if (i < D4) {
        ro[v] = __ldcs(o4 + i);
        rz[v] = __ldcs(z4 + i);
        rx[v] = __ldcs(x4 + i);
        re[v] = __ldcs(e4 + i);
      }
    }
#pragma unroll
    for (int v = 0; v < kLossV; ++v) {
      const int i = threadIdx.x + v * 256;
      if (i < D4) {
        const float ov[4] = {ro[v].x, ro[v].y, ro[v].z, ro[v].w}, zv[4] = {rz[v].x, rz[v].y, rz[v].z, rz[v].w};
        const float xv[4] = {rx[v].x, rx[v].y, rx[v].z, rx[v].w}, ev[4] = {re[v].x, re[v].y, re[v].z, re[v].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float xc = clip1(xhat_from_out(c, mean_type, zv[j], ov[j]));
          const float ec = eps_from_x(c, zv[j], xc);
          const float dx = xc - xv[j], de = ec - ev[j];
          sx += dx * dx;
          se += de * de;
        }
      }
    }
  } else {
    for (int i = threadIdx.x; i < D; i += blockDim.x) {
      const float zv = z[base + i];
      const float xc = clip1(xhat_from_out(c, mean_type, zv, out[base + i]));
      const float ec = eps_from_x(c, zv, xc);
      const float dx = xc - x[base + i], de = ec - eps[base + i];
      sx += dx * dx;
      se += de * de;
    }
  }
  const float x_mse = block_sum(sx, sh) / static_cast<float>(D);
  const float e_mse = block_sum(se, sh) / static_cast<float>(D);
  if (threadIdx.x == 0) {
    // 'snr_trunc' (gaussian_diffusion.py:166-168): max(x_mse, eps_mse); 'snr' (:169-170, distillation step1): eps_mse
    loss_b[b] = loss_type == 1 ? e_mse : fmaxf(x_mse, e_mse);
    if (mse2) {
      mse2[2 * b] = x_mse;
      mse2[2 * b + 1] = e_mse;
    }
  }
  if (!dout) return;
  // torch.maximum backward: the larger branch gets the gradient, exact ties split it in half.
  const float gx = loss_type == 1 ? 0.f : ((x_mse > e_mse) ? 1.f : ((x_mse == e_mse) ? 0.5f : 0.f));
  const float ge = 1.f - gx;
  const float k = 2.0f * grad_scale / static_cast<float>(D);
  const float dxc_de = -(c.Ce * c.De);
  const float dxh_dout = (mean_type == 0) ? -(c.A * c.Bc) : ((mean_type == 1) ? 1.f : -c.sigma);
  auto grad1 = [&](float zv, float ov, float xv, float ev) {
    const float xh = xhat_from_out(c, mean_type, zv, ov);
    const float xc = clip1(xh);
    const float ec = eps_from_x(c, zv, xc);
    float g = k * (gx * (xc - xv) + ge * (ec - ev) * dxc_de);
    if (xh < -1.0f || xh > 1.0f) g = 0.f;  // clip backward is inclusive at +-1
    return g * dxh_dout;
  };
  if (cached) {
#pragma unroll
    for (int v = 0; v < kLossV; ++v) {
      const int i = threadIdx.x + v * 256;
      if (i < D4) {
        float4 g;
        g.x = grad1(rz[v].x, ro[v].x, rx[v].x, re[v].x);
        g.y = grad1(rz[v].y, ro[v].y, rx[v].y, re[v].y);
        g.z = grad1(rz[v].z, ro[v].z, rx[v].z, re[v].z);
        g.w = grad1(rz[v].w, ro[v].w, rx[v].w, re[v].w);
        reinterpret_cast<float4*>(dout + base)[i] = g;
      }
    }
  } else {
    for (int i = threadIdx.x; i < D; i += blockDim.x)
      dout[base + i] = grad1(z[base + i], out[base + i], x[base + i], eps[base + i]);
  }
}

// One WARP per sample for small images (D <= 32 * 4 * kV pixels; 28x28 -> kV = 7, 32x32 -> 8): the whole sample sits in
// the warp's registers, both reductions are shuffle trees, there is no block barrier, and every warp keeps 4 * kV
// independent 16 B loads per lane in flight.  The CTA-per-sample kernel above was bound by its CTA lifetime (load ->
// six barriers -> store, 100 KB in flight per SM: 2.7 TB/s at n = 8192); this one is bound by HBM.
template <int kV>
__global__ void __launch_bounds__(128, 3)
loss_warp_kernel(const float* __restrict__ out, const float* __restrict__ z, const float* __restrict__ x,
                 const float* __restrict__ eps, const float* __restrict__ logsnr, float* __restrict__ loss_b,
                 float* __restrict__ mse2, float* __restrict__ dout, int B, int D, int mean_type, float grad_scale,
                 int loss_type) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (b >= B) return;
  const Coef c = make_coef(logsnr[b]);
  const size_t base = static_cast<size_t>(b) * D;
  const int D4 = D >> 2;
  const float4* o4 = reinterpret_cast<const float4*>(out + base);
  const float4* z4 = reinterpret_cast<const float4*>(z + base);
  const float4* x4 = reinterpret_cast<const float4*>(x + base);
  const float4* e4 = reinterpret_cast<const float4*>(eps + base);
  float4 ro[kV], rz[kV], rx[kV], re[kV];
#pragma unroll
  for (int v = 0; v < kV; ++v) {
    const int i = lane + v * 32;
    if (i < D4) {
      ro[v] = __ldcs(o4 + i);
      rz[v] = __ldcs(z4 + i);
      rx[v] = __ldcs(x4 + i);
      re[v] = __ldcs(e4 + i);
    }
  }
  float sx = 0.f, se = 0.f;
#pragma unroll
  for (int v = 0; v < kV; ++v) {
    const int i = lane + v * 32;
    if (i < D4) {
      const float ov[4] = {ro[v].x, ro[v].y, ro[v].z, ro[v].w}, zv[4] = {rz[v].x, rz[v].y, rz[v].z, rz[v].w};
      const float xv[4] = {rx[v].x, rx[v].y, rx[v].z, rx[v].w}, ev[4] = {re[v].x, re[v].y, re[v].z, re[v].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float xc = clip1(xhat_from_out(c, mean_type, zv[j], ov[j]));
        const float ec = eps_from_x(c, zv[j], xc);
        const float dx = xc - xv[j], de = ec - ev[j];
        sx += dx * dx;
        se += de * de;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sx += __shfl_xor_sync(0xffffffffu, sx, o);
    se += __shfl_xor_sync(0xffffffffu, se, o);
  }
  const float x_mse = sx / static_cast<float>(D), e_mse = se / static_cast<float>(D);
  if (lane == 0) {
    loss_b[b] = loss_type == 1 ? e_mse : fmaxf(x_mse, e_mse);
    if (mse2) {
      mse2[2 * b] = x_mse;
      mse2[2 * b + 1] = e_mse;
    }
  }
  if (!dout) return;
  const float gx = loss_type == 1 ? 0.f : ((x_mse > e_mse) ? 1.f : ((x_mse == e_mse) ? 0.5f : 0.f));
  const float ge = 1.f - gx;
  const float k = 2.0f * grad_scale / static_cast<float>(D);
  const float dxc_de = -(c.Ce * c.De);
  const float dxh_dout = (mean_type == 0) ? -(c.A * c.Bc) : ((mean_type == 1) ? 1.f : -c.sigma);
  auto grad1 = [&](float zv, float ov, float xv, float ev) {
    const float xh = xhat_from_out(c, mean_type, zv, ov);
    const float xc = clip1(xh);
    const float ec = eps_from_x(c, zv, xc);
    float g = k * (gx * (xc - xv) + ge * (ec - ev) * dxc_de);
    if (xh < -1.0f || xh > 1.0f) g = 0.f;  // clip backward is inclusive at +-1
    return g * dxh_dout;
  };
#pragma unroll
  for (int v = 0; v < kV; ++v) {
    const int i = lane + v * 32;
    if (i < D4) {
      float4 g;
      g.x = grad1(rz[v].x, ro[v].x, rx[v].x, re[v].x);
      g.y = grad1(rz[v].y, ro[v].y, rx[v].y, re[v].y);
      g.z = grad1(rz[v].z, ro[v].z, rx[v].z, re[v].z);
      g.w = grad1(rz[v].w, ro[v].w, rx[v].w, re[v].w);
      reinterpret_cast<float4*>(dout + base)[i] = g;
    }
  }
}

// mean over B values, one block, fixed order (deterministic).
__global__ void __launch_bounds__(256) mean_kernel(const float* __restrict__ v, float* __restrict__ out, int B) {
  __shared__ float sh[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) s += v[i];
  const float t = block_sum(s, sh);
  if (threadIdx.x == 0) out[0] = t / static_cast<float>(B);
}

// ------------------------------------------------------------------------------------------------ sampler update
struct SampK {
  const float* z;
  const float* out_c;
  const float* out_u;   // NULL: no guidance
  const float* cond_w;  // [n]
  const float* noise;   // noisy sampler
  const float* sched;   // [T][2] = (logsnr_t, logsnr_s) per loop index i
  const int* step_dev;  // device loop index (graph replay) or NULL
  float* z_next;
  float* z_dup;   // optional second copy of z_next (the unconditional half of a 2n-row guided batch)
  float* x_hist;  // optional [n*D], or [hist_T][n*D] when hist_T > 0 (row hist_T-1-step)
  float* e_hist;
  float* z_hist;
  uint8_t* z_u8;  // optional uint8 histories [hist_T][n*D] of ((v + 1) * 127.5).clamp(0, 255) (evaluate()'s proc)
  uint8_t* x_u8;
  uint8_t* e_u8;
  int hist_T;
  int n, D, step, mean_type, sampler;
  const float* lt_b;  // optional per-sample (logsnr_t, logsnr_s): distillation's teacher steps
  const float* ls_b;
};

// per-step coefficients of diffusion_reverse (diffusion_utils.py:34-51) / ddim (gaussian_diffusion.py:210-212)
struct StepCoef {
  Coef c;
  float k_z, k_x, stdv, alpha_s, sigma_s;
};
__device__ __forceinline__ StepCoef make_step_coef(float lt, float ls) {
  StepCoef s;
  s.c = make_coef(lt);
  const float alpha_st = sqrtf((1.0f + expf(-lt)) / (1.0f + expf(-ls)));
  s.alpha_s = sqrtf(sigmoid_t(ls));
  s.sigma_s = sqrtf(sigmoid_t(-ls));
  const float r = expf(lt - ls);
  const float omr = -expm1f(lt - ls);
  s.k_z = mulr(r, alpha_st);
  s.k_x = mulr(omr, s.alpha_s);
  s.stdv = sqrtf(mulr(omr, sigmoid_t(-lt)));
  return s;
}

// diffusion_model.py:91-92 proc(): ((v + 1) * 127.5).clamp(0, 255).to(torch.uint8) (float -> uint8 truncates)
__device__ __forceinline__ uint32_t proc_u8(float v) {
  const float t = fminf(fmaxf(mulr(addr(v, 1.0f), 127.5f), 0.0f), 255.0f);
  return static_cast<uint32_t>(static_cast<int>(t)) & 0xffu;
}
__device__ __forceinline__ uint32_t proc_u8x4(const float* v) {
  return proc_u8(v[0]) | (proc_u8(v[1]) << 8) | (proc_u8(v[2]) << 16) | (proc_u8(v[3]) << 24);
}

// 16 B per thread and tensor: z, out_c (, out_u), noise in; z_next (, z_dup, x_hist, e_hist) out.  16 B/px algorithmic
// for the plain ancestral step (z, out, noise in; z out).
__global__ void __launch_bounds__(256) sampler_step_kernel(SampK a) {
  const int step = a.lt_b ? 1 : (a.step_dev ? *a.step_dev : a.step);
  StepCoef sc = a.lt_b ? make_step_coef(0.f, 0.f) : make_step_coef(a.sched[2 * step], a.sched[2 * step + 1]);
  const bool last = !a.lt_b && (step == 0);
  const int D4 = a.D >> 2;
  const size_t total = static_cast<size_t>(a.n) * D4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * 2;
  for (size_t base = static_cast<size_t>(blockIdx.x) * blockDim.x * 2 + threadIdx.x; base < total; base += stride) {
    // two independent float4 per tensor in flight per thread
    float4 zq[2], oq[2], uq[2], nq[2];
    float wq[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const size_t i = base + static_cast<size_t>(u) * blockDim.x;
      uq[u] = nq[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      wq[u] = 0.f;
      if (i < total) {
        zq[u] = reinterpret_cast<const float4*>(a.z)[i];
        oq[u] = __ldcs(reinterpret_cast<const float4*>(a.out_c) + i);
        if (a.out_u) {
          uq[u] = __ldcs(reinterpret_cast<const float4*>(a.out_u) + i);
          wq[u] = a.cond_w[i / D4];
        }
        if (a.sampler == 1) nq[u] = __ldcs(reinterpret_cast<const float4*>(a.noise) + i);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const size_t i = base + static_cast<size_t>(u) * blockDim.x;
      if (i >= total) continue;
      const float w = wq[u];
      if (a.lt_b) {  // per-sample log-SNR pair
        const size_t b = i / D4;
        sc = make_step_coef(a.lt_b[b], a.ls_b[b]);
      }
      const Coef& c = sc.c;
      const float zv[4] = {zq[u].x, zq[u].y, zq[u].z, zq[u].w}, ov[4] = {oq[u].x, oq[u].y, oq[u].z, oq[u].w};
      const float uv[4] = {uq[u].x, uq[u].y, uq[u].z, uq[u].w}, nv[4] = {nq[u].x, nq[u].y, nq[u].z, nq[u].w};
      float zo[4], xo[4], eo[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float xc = clip1(xhat_from_out(c, a.mean_type, zv[j], ov[j]));
        float ec = eps_from_x(c, zv[j], xc);
        if (a.out_u) {  // _cf_guidance, gaussian_diffusion.py:174-187
          const float xu = clip1(xhat_from_out(c, a.mean_type, zv[j], uv[j]));
          const float eu = eps_from_x(c, zv[j], xu);
          const float e = addr(mulr(addr(1.0f, w), ec), mulr(-w, eu));
          xc = clip1(mulr(c.A, subr(zv[j], mulr(e, c.Bc))));
          ec = eps_from_x(c, zv[j], xc);
        }
        float zs;
        if (a.sampler == 1) {
          const float mean = addr(mulr(sc.k_z, zv[j]), mulr(sc.k_x, xc));
          zs = addr(mean, mulr(sc.stdv, nv[j]));
        } else {
          zs = addr(mulr(sc.alpha_s, xc), mulr(sc.sigma_s, ec));
        }
        zo[j] = last ? xc : zs;
        xo[j] = xc;
        eo[j] = ec;
      }
      const float4 zr = make_float4(zo[0], zo[1], zo[2], zo[3]);
      reinterpret_cast<float4*>(a.z_next)[i] = zr;
      if (a.z_dup) reinterpret_cast<float4*>(a.z_dup)[i] = zr;
      // histories: row k = T-1-step of a [T][n*D] tensor (gaussian_diffusion.py:293-296 stacks in loop order)
      const size_t h = a.hist_T > 0 ? static_cast<size_t>(a.hist_T - 1 - step) * total + i : i;
      if (a.z_hist) reinterpret_cast<float4*>(a.z_hist)[h] = zr;
      if (a.x_hist) reinterpret_cast<float4*>(a.x_hist)[h] = make_float4(xo[0], xo[1], xo[2], xo[3]);
      if (a.e_hist) reinterpret_cast<float4*>(a.e_hist)[h] = make_float4(eo[0], eo[1], eo[2], eo[3]);
      if (a.z_u8) reinterpret_cast<uint32_t*>(a.z_u8)[h] = proc_u8x4(zo);
      if (a.x_u8) reinterpret_cast<uint32_t*>(a.x_u8)[h] = proc_u8x4(xo);
      if (a.e_u8) reinterpret_cast<uint32_t*>(a.e_u8)[h] = proc_u8x4(eo);
    }
  }
}

// gaussian_diffusion.py:137-154 (step2 of progressive distillation); same evaluation order as the reference.
__global__ void __launch_bounds__(256)
distill_target_kernel(const float* __restrict__ z_teacher, const float* __restrict__ z_t, const float* __restrict__ x_pred,
                      const float* __restrict__ lt_b, const float* __restrict__ ls_b, const long long* __restrict__ index,
                      float* __restrict__ x_target, float* __restrict__ eps_target, int n, int D4) {
  const size_t total = static_cast<size_t>(n) * D4;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / D4;
    const float lt = lt_b[b], ls = ls_b[b];
    const Coef c = make_coef(lt);
    const float alpha_s = sqrtf(sigmoid_t(ls)), alpha_t = sqrtf(sigmoid_t(lt));
    // F.softplus(l) = log1p(exp(l)) below torch's threshold of 20, else l
    const float spt = lt > 20.0f ? lt : log1pf(expf(lt)), sps = ls > 20.0f ? ls : log1pf(expf(ls));
    const float f = expf(mulr(0.5f, subr(spt, sps)));
    const float den = subr(alpha_s, mulr(f, alpha_t));
    const bool first = index[b] == 0;
    const float4 zt4 = reinterpret_cast<const float4*>(z_teacher)[i], z4 = reinterpret_cast<const float4*>(z_t)[i];
    const float4 xp4 = reinterpret_cast<const float4*>(x_pred)[i];
    const float zt[4] = {zt4.x, zt4.y, zt4.z, zt4.w}, zz[4] = {z4.x, z4.y, z4.z, z4.w}, xp[4] = {xp4.x, xp4.y, xp4.z, xp4.w};
    float xo[4], eo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xt = __fdiv_rn(subr(zt[j], mulr(f, zz[j])), den);
      xo[j] = first ? xp[j] : xt;
      eo[j] = eps_from_x(c, zz[j], xo[j]);
    }
    reinterpret_cast<float4*>(x_target)[i] = make_float4(xo[0], xo[1], xo[2], xo[3]);
    reinterpret_cast<float4*>(eps_target)[i] = make_float4(eo[0], eo[1], eo[2], eo[3]);
  }
}

// (logsnr_t broadcast to n) for the embedding kernel + decrement of the device loop index: one tiny launch that
// closes a captured sampler step.
__global__ void step_advance_kernel(int* step_dev) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *step_dev -= 1;
}

static int ew_grid1(size_t total, int per_thread = 1) {
  size_t blocks = (total + 256 * per_thread - 1) / (256 * per_thread);
  const size_t cap = 148 * 8;
  return static_cast<int>(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_q_sample(const b200_q_sample_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->x || !a->eps || !a->logsnr || !a->z) return fail_arg("q_sample: null pointer");
  if (a->D % 4) return fail_arg("q_sample: D must be a multiple of 4");
  const size_t total = static_cast<size_t>(a->B) * (a->D / 4);
  q_sample_kernel<<<ew_grid1(total, kEwUnroll), 256, 0, static_cast<cudaStream_t>(stream)>>>(a->x, a->eps, a->logsnr, a->z, a->B,
                                                                                  a->D / 4);
  return after_launch("q_sample_kernel");
}

extern "C" int b200_loss(const b200_loss_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->out || !a->z || !a->x || !a->eps || !a->logsnr || !a->loss_b) return fail_arg("loss: null pointer");
  if (a->mean_type < 0 || a->mean_type > 2) return fail_arg("loss: mean_type must be 0 (eps), 1 (x) or 2 (v)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a->D % 4) return fail_arg("loss: D must be a multiple of 4");
  if (a->loss_type != 0 && a->loss_type != 1) return fail_arg("loss: loss_type must be 0 (snr_trunc) or 1 (snr)");
  const int D4 = a->D / 4;
  if (D4 <= 224)
    loss_warp_kernel<7><<<(a->B + 3) / 4, 128, 0, st>>>(a->out, a->z, a->x, a->eps, a->logsnr, a->loss_b, a->mse2, a->dout,
                                                        a->B, a->D, a->mean_type, a->grad_scale, a->loss_type);
  else if (D4 <= 256)
    loss_kernel<1><<<a->B, 256, 0, st>>>(a->out, a->z, a->x, a->eps, a->logsnr, a->loss_b, a->mse2, a->dout, a->D,
                                         a->mean_type, a->grad_scale, a->loss_type);
  else if (D4 <= 512)
    loss_kernel<2><<<a->B, 256, 0, st>>>(a->out, a->z, a->x, a->eps, a->logsnr, a->loss_b, a->mse2, a->dout, a->D,
                                         a->mean_type, a->grad_scale, a->loss_type);
  else
    loss_kernel<4><<<a->B, 256, 0, st>>>(a->out, a->z, a->x, a->eps, a->logsnr, a->loss_b, a->mse2, a->dout, a->D,
                                         a->mean_type, a->grad_scale, a->loss_type);
  if (int r = after_launch("loss_kernel")) return r;
  if (a->loss_mean) {
    mean_kernel<<<1, 256, 0, st>>>(a->loss_b, a->loss_mean, a->B);
    return after_launch("mean_kernel");
  }
  return 0;
}

extern "C" int b200_sampler_step(const b200_sampler_step_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->z || !a->out_c || !a->z_next) return fail_arg("sampler_step: null pointer");
  if (!a->sched && !(a->logsnr_t_b && a->logsnr_s_b)) return fail_arg("sampler_step: needs sched or the per-sample log-SNR pair");
  if (a->logsnr_t_b && (!a->logsnr_s_b || a->hist_T > 0)) return fail_arg("sampler_step: per-sample log-SNR needs both vectors and no trajectory history");
  if (a->out_u && !a->cond_w) return fail_arg("sampler_step: guidance needs cond_w");
  if (a->sampler == 1 && !a->noise) return fail_arg("sampler_step: the noisy sampler needs noise");
  if (a->sampler != 0 && a->sampler != 1) return fail_arg("sampler_step: sampler must be 0 (ddim) or 1 (noisy)");
  if (a->mean_type < 0 || a->mean_type > 2) return fail_arg("sampler_step: bad mean_type");
  if (a->D % 4) return fail_arg("sampler_step: D must be a multiple of 4");
  if ((a->z_u8 || a->x_u8 || a->e_u8) && a->hist_T <= 0) return fail_arg("sampler_step: uint8 histories need hist_T");
  if (a->hist_T > 0 && !a->step_dev && (a->step < 0 || a->step >= a->hist_T)) return fail_arg("sampler_step: step outside the history");
  SampK k;
  k.z = a->z; k.out_c = a->out_c; k.out_u = a->out_u; k.cond_w = a->cond_w; k.noise = a->noise; k.sched = a->sched;
  k.step_dev = a->step_dev; k.z_next = a->z_next; k.lt_b = a->logsnr_t_b; k.ls_b = a->logsnr_s_b;
  k.z_dup = a->z_dup; k.z_hist = a->z_hist; k.hist_T = a->hist_T;
  k.z_u8 = static_cast<uint8_t*>(a->z_u8); k.x_u8 = static_cast<uint8_t*>(a->x_u8); k.e_u8 = static_cast<uint8_t*>(a->e_u8);
  k.x_hist = a->x_hist; k.e_hist = a->e_hist;
  k.n = a->n; k.D = a->D; k.step = a->step; k.mean_type = a->mean_type; k.sampler = a->sampler;
  const size_t total = static_cast<size_t>(a->n) * (a->D / 4);
  sampler_step_kernel<<<ew_grid1(total, 2), 256, 0, static_cast<cudaStream_t>(stream)>>>(k);
  return after_launch("sampler_step_kernel");
}

extern "C" int b200_distill_target(const b200_distill_target_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->z_teacher || !a->z_t || !a->x_pred_teacher || !a->logsnr_t || !a->logsnr_s || !a->index || !a->x_target ||
      !a->eps_target)
    return fail_arg("distill_target: null pointer");
  if (a->D % 4) return fail_arg("distill_target: D must be a multiple of 4");
  const size_t total = static_cast<size_t>(a->n) * (a->D / 4);
  distill_target_kernel<<<ew_grid1(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      a->z_teacher, a->z_t, a->x_pred_teacher, a->logsnr_t, a->logsnr_s, a->index, a->x_target, a->eps_target, a->n, a->D / 4);
  return after_launch("distill_target_kernel");
}

extern "C" int b200_step_advance(const b200_step_advance_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->step_dev) return fail_arg("step_advance: null pointer");
  step_advance_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a->step_dev);
  return after_launch("step_advance_kernel");
}
