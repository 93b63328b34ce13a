// GroupNorm(32 groups) + SiLU, forward and backward, over NHWC fp32 rows (optionally a virtual channel concat of two
// tensors).  Bandwidth-bound: every kernel moves float4s with fully coalesced pixel rows; per-(sample, group)
// moments are reduced with warp/smem reductions and a handful of global atomics.
// Reference: nn.GroupNorm(32, C) -> nn.SiLU in gms/diffusion/simple_unet.py:38-41,161-163,168-170.
#include "host_common.h"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

namespace b200 {

struct GnSrc {
  const float* x;
  const float* x2;
  int ld1, ld2;
  int N, HW, C1, C;  // C = C1 + C2
};

__device__ __forceinline__ float4 ld_src(const GnSrc& s, size_t pix, int c) {
  return (c < s.C1) ? *reinterpret_cast<const float4*>(s.x + pix * s.ld1 + c)
                    : *reinterpret_cast<const float4*>(s.x2 + pix * s.ld2 + (c - s.C1));
}
__device__ __forceinline__ float silu_f(float u) { return u / (1.0f + expf(-u)); }

// ---------------------------------------------------------------------------------------------------- forward stats
// grid (chunks, N), block 256.  stats[n][g] = {sum, sumsq}; caller zeroes stats.
__global__ void __launch_bounds__(256) gn_stats_kernel(GnSrc s, float* __restrict__ stats, int pix_per_chunk) {
  __shared__ float sh[64];
  const int n = blockIdx.y;
  const int lanes = s.C >> 2;  // float4 lanes per pixel: 32 or 64
  const int lane4 = threadIdx.x % lanes;
  const int prow = threadIdx.x / lanes;
  const int prows = 256 / lanes;
  const int c = lane4 * 4;
  const int cpg = s.C >> 5;
  const int g = c / cpg;
  if (threadIdx.x < 64) sh[threadIdx.x] = 0.f;
  __syncthreads();
  const int p0 = blockIdx.x * pix_per_chunk;
  const int p1 = min(p0 + pix_per_chunk, s.HW);
  float sum = 0.f, sq = 0.f;
  for (int p = p0 + prow; p < p1; p += prows) {
    const float4 v = ld_src(s, static_cast<size_t>(n) * s.HW + p, c);
    sum += (v.x + v.y) + (v.z + v.w);
    sq += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  atomicAdd(&sh[g * 2], sum);
  atomicAdd(&sh[g * 2 + 1], sq);
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(&stats[n * 64 + threadIdx.x], sh[threadIdx.x]);
}

// ---------------------------------------------------------------------------------------------------- forward apply
__global__ void __launch_bounds__(256)
gn_apply_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                const float* __restrict__ beta, float* __restrict__ y, float eps, int silu, int tf32) {
  const int lanes = s.C >> 2;
  const size_t total = static_cast<size_t>(s.N) * s.HW * lanes;
  const int cpg = s.C >> 5;
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t pix = i / lanes;
    const int c = static_cast<int>(i - pix * lanes) * 4;
    const int n = static_cast<int>(pix / s.HW);
    const int g = c / cpg;
    const float2 st = *reinterpret_cast<const float2*>(stats + (n * 32 + g) * 2);
    const float mean = st.x * inv_cnt;
    const float var = fmaxf(st.y * inv_cnt - mean * mean, 0.f);
    const float rstd = rsqrtf(var + eps);
    const float4 v = ld_src(s, pix, c);
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 be = *reinterpret_cast<const float4*>(beta + c);
    float4 o;
    o.x = (v.x - mean) * rstd * ga.x + be.x;
    o.y = (v.y - mean) * rstd * ga.y + be.y;
    o.z = (v.z - mean) * rstd * ga.z + be.z;
    o.w = (v.w - mean) * rstd * ga.w + be.w;
    if (silu) {
      o.x = silu_f(o.x); o.y = silu_f(o.y); o.z = silu_f(o.z); o.w = silu_f(o.w);
    }
    if (tf32) {
      o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w);
    }
    *reinterpret_cast<float4*>(y + pix * s.C + c) = o;
  }
}

// ---------------------------------------------------------------------------------------------------- backward
// du = dy * silu'(u), dxhat = du * gamma.  Pass 1 reduces, per (n, g): s1 = sum dxhat, s2 = sum dxhat * xhat, and per
// channel dgamma += sum du * xhat, dbeta += sum du.  Pass 2 writes dx = rstd * (dxhat - s1/cnt - xhat * s2/cnt) (+adds).
__device__ __forceinline__ float dsilu_f(float u) {
  const float sg = 1.0f / (1.0f + expf(-u));
  return sg * (1.0f + u * (1.0f - sg));
}

__global__ void __launch_bounds__(256)
gn_bwd_reduce_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                     const float* __restrict__ beta, const float* __restrict__ dy, float* __restrict__ red,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, float eps, int silu, int pix_per_chunk) {
  __shared__ float sh[64];
  __shared__ float shg[256 * 8];
  const int n = blockIdx.y;
  const int lanes = s.C >> 2;
  const int lane4 = threadIdx.x % lanes;
  const int prow = threadIdx.x / lanes;
  const int prows = 256 / lanes;
  const int c = lane4 * 4;
  const int cpg = s.C >> 5;
  const int g = c / cpg;
  if (threadIdx.x < 64) sh[threadIdx.x] = 0.f;
  __syncthreads();
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  const float2 st = *reinterpret_cast<const float2*>(stats + (n * 32 + g) * 2);
  const float mean = st.x * inv_cnt;
  const float rstd = rsqrtf(fmaxf(st.y * inv_cnt - mean * mean, 0.f) + eps);
  const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
  const float4 be = *reinterpret_cast<const float4*>(beta + c);
  const float gav[4] = {ga.x, ga.y, ga.z, ga.w};
  const float bev[4] = {be.x, be.y, be.z, be.w};
  float dg[4] = {0.f, 0.f, 0.f, 0.f}, db[4] = {0.f, 0.f, 0.f, 0.f};
  float s1 = 0.f, s2 = 0.f;
  const int p0 = blockIdx.x * pix_per_chunk;
  const int p1 = min(p0 + pix_per_chunk, s.HW);
  for (int p = p0 + prow; p < p1; p += prows) {
    const size_t pix = static_cast<size_t>(n) * s.HW + p;
    const float4 v = ld_src(s, pix, c);
    const float4 d = *reinterpret_cast<const float4*>(dy + pix * s.C + c);
    const float xv[4] = {v.x, v.y, v.z, v.w};
    const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = (xv[j] - mean) * rstd;
      const float du = silu ? dv[j] * dsilu_f(xh * gav[j] + bev[j]) : dv[j];
      dg[j] += du * xh;
      db[j] += du;
      const float dxh = du * gav[j];
      s1 += dxh;
      s2 += dxh * xh;
    }
  }
  atomicAdd(&sh[g * 2], s1);
  atomicAdd(&sh[g * 2 + 1], s2);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    shg[threadIdx.x * 8 + j] = dg[j];
    shg[threadIdx.x * 8 + 4 + j] = db[j];
  }
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(&red[n * 64 + threadIdx.x], sh[threadIdx.x]);
  if (threadIdx.x < lanes) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < prows; ++r)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += shg[(r * lanes + threadIdx.x) * 8 + j];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      atomicAdd(&dgamma[c + j], acc[j]);
      atomicAdd(&dbeta[c + j], acc[4 + j]);
    }
  }
}

__global__ void __launch_bounds__(256)
gn_bwd_apply_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ dy, const float* __restrict__ red,
                    float* __restrict__ dx, const float* __restrict__ add1, int add1_ld,
                    const float* __restrict__ add2, int add2_ld, int add_c, float eps, int silu, int round_dx) {
  const int lanes = s.C >> 2;
  const size_t total = static_cast<size_t>(s.N) * s.HW * lanes;
  const int cpg = s.C >> 5;
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t pix = i / lanes;
    const int c = static_cast<int>(i - pix * lanes) * 4;
    const int n = static_cast<int>(pix / s.HW);
    const int g = c / cpg;
    const float2 st = *reinterpret_cast<const float2*>(stats + (n * 32 + g) * 2);
    const float2 rd = *reinterpret_cast<const float2*>(red + (n * 32 + g) * 2);
    const float mean = st.x * inv_cnt;
    const float rstd = rsqrtf(fmaxf(st.y * inv_cnt - mean * mean, 0.f) + eps);
    const float m1 = rd.x * inv_cnt, m2 = rd.y * inv_cnt;
    const float4 v = ld_src(s, pix, c);
    const float4 d = *reinterpret_cast<const float4*>(dy + pix * s.C + c);
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 be = *reinterpret_cast<const float4*>(beta + c);
    const float xv[4] = {v.x, v.y, v.z, v.w};
    const float dv[4] = {d.x, d.y, d.z, d.w};
    const float gav[4] = {ga.x, ga.y, ga.z, ga.w};
    const float bev[4] = {be.x, be.y, be.z, be.w};
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = (xv[j] - mean) * rstd;
      const float du = silu ? dv[j] * dsilu_f(xh * gav[j] + bev[j]) : dv[j];
      o[j] = rstd * (du * gav[j] - m1 - xh * m2);
    }
    if (c < add_c) {  // extra gradient flowing into the first add_c channels (residual / skip paths)
      if (add1) {
        const float4 a = *reinterpret_cast<const float4*>(add1 + pix * add1_ld + c);
        o[0] += a.x; o[1] += a.y; o[2] += a.z; o[3] += a.w;
      }
      if (add2) {
        const float4 a = *reinterpret_cast<const float4*>(add2 + pix * add2_ld + c);
        o[0] += a.x; o[1] += a.y; o[2] += a.z; o[3] += a.w;
      }
    }
    float4 ov = make_float4(o[0], o[1], o[2], o[3]);
    if (round_dx) ov = round_tf32(ov);
    *reinterpret_cast<float4*>(dx + pix * s.C + c) = ov;
  }
}

static int gn_chunks(int N, int HW) {
  int chunks = (592 + N - 1) / N;
  const int max_chunks = (HW + 15) / 16;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  return chunks;
}
static int ew_grid(size_t total_threads) {
  size_t blocks = (total_threads + 255) / 256;
  const size_t cap = 148 * 16;
  return static_cast<int>(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace b200

using namespace b200;

static int gn_src(const b200_gn_args* a, GnSrc* s) {
  if (!a || !a->x || !a->stats || !a->gamma || !a->beta) return fail_arg("gn: null pointer");
  const int C = a->C1 + a->C2;
  if (C != 128 && C != 256) return fail_arg("gn: C1+C2 must be 128 or 256");
  if (a->C1 % 4 || a->C2 % 4 || (a->C2 && !a->x2)) return fail_arg("gn: bad channel split");
  s->x = a->x;
  s->x2 = a->x2;
  s->ld1 = a->x_ld ? a->x_ld : a->C1;
  s->ld2 = a->x2_ld ? a->x2_ld : a->C2;
  s->N = a->N;
  s->HW = a->HW;
  s->C1 = a->C1;
  s->C = C;
  return 0;
}

extern "C" int b200_gn_stats(const b200_gn_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  GnSrc s;
  if (int r = gn_src(a, &s)) return r;
  const int chunks = gn_chunks(s.N, s.HW);
  const int ppc = (s.HW + chunks - 1) / chunks;
  gn_stats_kernel<<<dim3(chunks, s.N), 256, 0, static_cast<cudaStream_t>(stream)>>>(s, a->stats, ppc);
  return after_launch("gn_stats_kernel");
}

extern "C" int b200_gn_apply(const b200_gn_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  GnSrc s;
  if (int r = gn_src(a, &s)) return r;
  if (!a->y) return fail_arg("gn_apply: null y");
  const size_t total = static_cast<size_t>(s.N) * s.HW * (s.C / 4);
  gn_apply_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(s, a->stats, a->gamma, a->beta, a->y,
                                                                                 a->eps, a->silu, a->round_tf32);
  return after_launch("gn_apply_kernel");
}

extern "C" int b200_gn_bwd(const b200_gn_bwd_args* b, void* ws, size_t ws_bytes, void* stream) {
  if (!b) return fail_arg("gn_bwd: null args");
  GnSrc s;
  if (int r = gn_src(&b->fwd, &s)) return r;
  if (!b->dy || !b->dx || !b->dgamma || !b->dbeta) return fail_arg("gn_bwd: null pointer");
  const size_t need = static_cast<size_t>(s.N) * 64 * sizeof(float);
  if (!ws || ws_bytes < need) return fail_arg("gn_bwd: workspace of N*64 floats required");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* red = static_cast<float*>(ws);
  if (int r = check_cuda(cudaMemsetAsync(red, 0, need, st), "gn_bwd memset")) return r;
  const int chunks = gn_chunks(s.N, s.HW);
  const int ppc = (s.HW + chunks - 1) / chunks;
  gn_bwd_reduce_kernel<<<dim3(chunks, s.N), 256, 0, st>>>(s, b->fwd.stats, b->fwd.gamma, b->fwd.beta, b->dy, red,
                                                          b->dgamma, b->dbeta, b->fwd.eps, b->fwd.silu, ppc);
  if (int r = after_launch("gn_bwd_reduce_kernel")) return r;
  const size_t total = static_cast<size_t>(s.N) * s.HW * (s.C / 4);
  gn_bwd_apply_kernel<<<ew_grid(total), 256, 0, st>>>(s, b->fwd.stats, b->fwd.gamma, b->fwd.beta, b->dy, red, b->dx,
                                                      b->add1, b->add1_ld, b->add2, b->add2_ld,
                                                      b->add_c ? b->add_c : s.C, b->fwd.eps, b->fwd.silu, b->round_dx);
  return after_launch("gn_bwd_apply_kernel");
}
