// GroupNorm(32 groups) + SiLU, forward and backward, over NHWC fp32 rows (optionally a virtual channel concat of two
// tensors).  Bandwidth-bound, so the design goal is the algorithmic minimum of HBM traffic:
//   forward  = read x once, write y once                  (8 B per element)
//   backward = read x and dy once (+ the residual adds), write dx once   (12 B per element + adds)
// A CTA owns one (sample, 16/32-channel slice): whole groups, all HW pixels.  It pulls the slice into shared memory with
// cp.async (everything in flight at once, no register staging), computes the group moments from smem (two-pass
// variance), then streams the result out — so the second and third read of the classic stats/apply split never leave the
// SM.  Slices are 64/128 B per pixel row: every global access is a whole number of 32 B sectors, and neighbouring
// slices of the same pixel rows are processed by neighbouring CTAs.  Shapes whose slice does not fit in smem (HW > ~1800)
// use the streaming three-kernel path at the end of this file.
// stats is [N][32][2] = {mean, rstd}: written by the forward, read by the backward.
// Reference: nn.GroupNorm(32, C) -> nn.SiLU in gms/diffusion/simple_unet.py:38-41,161-163,168-170.
#include "host_common.h"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

namespace b200 {

// optional fused column sums of dx over channels [0, c): total1/total2[ch] += sum over (n, pixels),
// per_sample[n * ps_ld + ch] += sum over the pixels of sample n
struct GnColsum {
  float* per_sample;
  int ps_ld;
  float* total1;
  float* total2;
  int c;
};

struct GnSrc {
  const float* x;
  const float* x2;
  int ld1, ld2;
  int N, HW, C1, C;  // C = C1 + C2
  int bf;            // x / x2 hold bf16 (ld in elements); fused kernels only
};

// nn.Dropout(p) between the SiLU and the convolution of a ResBlock's out_layers (simple_unet.py:168-172), fused into the
// GroupNorm kernels: y = silu(...) * keep / (1 - p) in the forward, dy * keep / (1 - p) in front of the SiLU derivative
// in the backward.  keep is a pure function of (seed, step, position of the float4 in the dense [N*HW][C] output), so
// forward and backward agree without a stored mask, and a captured graph draws a fresh mask every replay through the
// device step counter.  16 random bits per element: p is quantised to 1/65536 and `scale` is 1 / (1 - p quantised).
struct GnDrop {
  uint32_t thr;     // element kept when its 16 bits >= thr; 0 = dropout off
  float scale;
  uint32_t seed;    // per layer
  const int* step;  // optional device counter mixed into the seed
};
__device__ __forceinline__ uint32_t fmix32(uint32_t h) {  // murmur3 finaliser
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ uint32_t drop_seed(const GnDrop& d) {
  return fmix32(d.seed ^ fmix32(0x9E3779B9u * static_cast<uint32_t>(d.step ? __ldg(d.step) : 0) + 0x7F4A7C15u));
}
// v * keep * scale for the float4 number e4 of the output tensor
__device__ __forceinline__ float4 drop_apply(float4 v, const GnDrop& d, uint32_t seed, uint32_t e4) {
  const uint32_t a = fmix32(e4 * 0x9E3779B1u + seed);
  const uint32_t b = fmix32(a + 0x632BE5ABu);
  v.x = (a & 0xffffu) >= d.thr ? v.x * d.scale : 0.f;
  v.y = (a >> 16) >= d.thr ? v.y * d.scale : 0.f;
  v.z = (b & 0xffffu) >= d.thr ? v.z * d.scale : 0.f;
  v.w = (b >> 16) >= d.thr ? v.w * d.scale : 0.f;
  return v;
}

__device__ __forceinline__ float4 ld_src(const GnSrc& s, size_t pix, int c) {
  return (c < s.C1) ? *reinterpret_cast<const float4*>(s.x + pix * s.ld1 + c)
                    : *reinterpret_cast<const float4*>(s.x2 + pix * s.ld2 + (c - s.C1));
}
// SiLU with the hardware exp2 / reciprocal (2 MUFU ops, ~1e-6 relative): the IEEE expf + division sequence costs about
// as many issue slots as the kernel's whole memory phase.
__device__ __forceinline__ float sigmoid_f(float u) {
  // exactly two MUFU ops (ex2.approx, rcp.approx; ~2^-22 relative): __frcp_rn is the IEEE-rounded reciprocal, a MUFU
  // plus a Newton/fix-up sequence that cost the forward a quarter of its time (68 -> 89 us at 28x28, C=128)
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * u));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return r;
}
__device__ __forceinline__ float silu_f(float u) { return u * sigmoid_f(u); }
// d silu(u) / du
__device__ __forceinline__ float dsilu_f(float u) {
  const float sg = sigmoid_f(u);
  return sg * (1.0f + u * (1.0f - sg));
}
// One-MUFU sigmoid for the bf16 paths: 0.5 tanh(u/2) + 0.5 with tanh.approx (abs error ~2.5e-4, an eighth of a bf16
// ulp of the values it produces); the fp32 paths keep the ex2 + rcp form (1e-6).
__device__ __forceinline__ float sigmoid_fast(float u) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * u));
  return fmaf(0.5f, t, 0.5f);
}
__device__ __forceinline__ float silu_fast(float u) { return u * sigmoid_fast(u); }
__device__ __forceinline__ float dsilu_fast(float u) {
  const float sg = sigmoid_fast(u);
  return sg * (1.0f + u * (1.0f - sg));
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}


// reduce `v` over the threads of a warp that share the same float4 lane (and, for 8-channel groups, its neighbour)
template <int L>
__device__ __forceinline__ float lane_group_sum(float v, bool pair) {
#pragma unroll
  for (int o = 16; o >= L; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (pair) v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}

// Block reduction of per-channel partial sums without shared-memory atomics (atomicAdd on a shared float compiles to a
// compare-and-swap spin loop — ATOMS.CAST.SPIN — and sixteen warps hammering the same 2*CS words made the reduction a
// visible share of the backward's instruction count; ncu: 1536 instructions per warp for 6 loop iterations).
// Every warp writes its lane-reduced values to its own row of `wred`, a barrier, then 2*CS threads add the rows up in
// a fixed order (so the per-CTA sums are also reproducible run to run).
//   in : dg[4], db[4] — this thread's sums of du*xhat and du for its 4 channels (already reduced over the warp's lanes
//        that share the channel quad: lanes 0..L-1 hold the result)
//   out: chan[0..CS) = sum du*xhat per channel, chan[CS..2CS) = sum du per channel; gsum = the same times gamma
// and dgamma / dbeta receive one global atomic per channel.  Ends with a barrier.
template <int CS, int NW>
__device__ __forceinline__ void gn_bwd_block_sums(const float (&dg)[4], const float (&db)[4], int tid, int lane, int l,
                                                  float (*wred)[2 * CS], float* chan, float* gsum,
                                                  const float* __restrict__ gamma, int c0, float* __restrict__ dgamma,
                                                  float* __restrict__ dbeta) {
  constexpr int L = CS / 4;
  if (lane < L) {
    float4* w = reinterpret_cast<float4*>(&wred[tid >> 5][0]);
    w[l] = make_float4(dg[0], dg[1], dg[2], dg[3]);
    w[L + l] = make_float4(db[0], db[1], db[2], db[3]);
  }
  __syncthreads();
  if (tid < 2 * CS) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) v += wred[w][tid];
    chan[tid] = v;
    const int ch = tid < CS ? tid : tid - CS;
    gsum[tid] = v * __ldg(gamma + c0 + ch);
    atomicAdd(tid < CS ? &dgamma[c0 + ch] : &dbeta[c0 + ch], v);
  }
  __syncthreads();
}
// fused column sums of dx: same scheme, CS values per warp
template <int CS, int NW>
__device__ __forceinline__ void gn_bwd_colsum(float (&ca)[4], int tid, int lane, int l, float (*wred)[2 * CS], int n, int c0,
                                              const GnColsum& cs) {
  constexpr int L = CS / 4;
#pragma unroll
  for (int j = 0; j < 4; ++j) ca[j] = lane_group_sum<L>(ca[j], false);
  if (lane < L) reinterpret_cast<float4*>(&wred[tid >> 5][0])[l] = make_float4(ca[0], ca[1], ca[2], ca[3]);
  __syncthreads();
  if (tid < CS) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) v += wred[w][tid];
    if (cs.per_sample) atomicAdd(&cs.per_sample[static_cast<size_t>(n) * cs.ps_ld + c0 + tid], v);
    if (cs.total1) atomicAdd(&cs.total1[c0 + tid], v);
    if (cs.total2) atomicAdd(&cs.total2[c0 + tid], v);
  }
}

// ---------------------------------------------------------------------------------------------------- fused forward
// grid = N * (C / CS); block 256; dynamic smem = HW * CS * 4 bytes.
template <int CS, int kGnThreads, bool DROP>
__global__ void __launch_bounds__(kGnThreads)
gn_fwd_fused_kernel(GnSrc s, float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                    float* __restrict__ y, float eps, int silu, int tf32, GnDrop dp) {
  extern __shared__ float4 gn_smem[];
  __shared__ float fred[2][kGnThreads / 32][8];  // [sum | centred squares][warp][group of the slice]
  constexpr int L = CS / 4;  // float4 lanes per pixel of the slice
  const int slices = s.C / CS;
  const int n = blockIdx.x / slices;
  const int c0 = (blockIdx.x - n * slices) * CS;
  const int cpg = s.C >> 5;
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int l = tid % L;  // fixed per thread: kGnThreads % L == 0
  const int c = c0 + l * 4;
  const int gl = (l * 4) / cpg;  // group within the slice
  const int total = s.HW * L;
  const float* src;
  int ld;
  if (c < s.C1) {
    src = s.x + c;
    ld = s.ld1;
  } else {
    src = s.x2 + (c - s.C1);
    ld = s.ld2;
  }
  src += static_cast<size_t>(n) * s.HW * ld;
  for (int idx = tid; idx < total; idx += kGnThreads) cp_async16(&gn_smem[idx], src + static_cast<size_t>(idx / L) * ld);
  const float4 ga = __ldg(reinterpret_cast<const float4*>(gamma + c));  // in flight with the slice
  const float4 be = __ldg(reinterpret_cast<const float4*>(beta + c));
  cp_async_wait_all();
  __syncthreads();
  const bool pair = (cpg == 8);
  const bool leader = (lane < L) && (!pair || !(lane & 1));
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  float sum = 0.f;
  for (int idx = tid; idx < total; idx += kGnThreads) {
    const float4 v = gn_smem[idx];
    sum += (v.x + v.y) + (v.z + v.w);
  }
  sum = lane_group_sum<L>(sum, pair);
  if (leader) fred[0][tid >> 5][gl] = sum;  // one slot per (warp, group): no shared-memory atomics (CAS spin loops)
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < kGnThreads / 32; ++w) tot += fred[0][w][gl];
  const float mean = tot * inv_cnt;
  float sq = 0.f;
  for (int idx = tid; idx < total; idx += kGnThreads) {
    const float4 v = gn_smem[idx];
    const float a = v.x - mean, b = v.y - mean, cc = v.z - mean, d = v.w - mean;
    sq += (a * a + b * b) + (cc * cc + d * d);
  }
  sq = lane_group_sum<L>(sq, pair);
  if (leader) fred[1][tid >> 5][gl] = sq;
  __syncthreads();
  tot = 0.f;
#pragma unroll
  for (int w = 0; w < kGnThreads / 32; ++w) tot += fred[1][w][gl];
  const float rstd = rsqrtf(tot * inv_cnt + eps);
  if (tid < L && (!pair || !(tid & 1)))
    *reinterpret_cast<float2*>(stats + (static_cast<size_t>(n) * 32 + c / cpg) * 2) = make_float2(mean, rstd);
  const float4 sc = make_float4(rstd * ga.x, rstd * ga.y, rstd * ga.z, rstd * ga.w);
  // tf32: 0 = plain fp32, 1 = fp32 rounded to TF32, 2 = bf16 (y then holds 2-byte elements)
  // running output pointers: a thread's successive elements are kGnThreads / L pixels apart (see the backward kernel)
  const size_t o0 = (static_cast<size_t>(n) * s.HW + tid / L) * s.C + c;
  float* dst = y + o0;
  uint16_t* dstb = reinterpret_cast<uint16_t*>(y) + o0;
  const unsigned ostep = static_cast<unsigned>(kGnThreads / L) * static_cast<unsigned>(s.C);
  const uint32_t dseed = DROP ? drop_seed(dp) : 0u;
  uint32_t e4 = static_cast<uint32_t>(o0 >> 2);
#pragma unroll 2
  for (int idx = tid; idx < total; idx += kGnThreads, dst += ostep, dstb += ostep, e4 += ostep >> 2) {
    const float4 v = gn_smem[idx];
    float4 o;
    o.x = (v.x - mean) * sc.x + be.x;
    o.y = (v.y - mean) * sc.y + be.y;
    o.z = (v.z - mean) * sc.z + be.z;
    o.w = (v.w - mean) * sc.w + be.w;
    if (silu) {
      if (tf32 == 2) {  // bf16 output: the one-MUFU sigmoid is far inside the output rounding
        o.x = silu_fast(o.x); o.y = silu_fast(o.y); o.z = silu_fast(o.z); o.w = silu_fast(o.w);
      } else {
        o.x = silu_f(o.x); o.y = silu_f(o.y); o.z = silu_f(o.z); o.w = silu_f(o.w);
      }
    }
    if (DROP) o = drop_apply(o, dp, dseed, e4);
    if (tf32 == 2) {
      *reinterpret_cast<uint2*>(dstb) = pack_bf16x4(o);
    } else {
      if (tf32) o = round_tf32(o);
      *reinterpret_cast<float4*>(dst) = o;
    }
  }
}

// Phase 2 of the shared-memory-resident backward kernels: dx = elem(idx) (+ add1 + add2) -> bf16 copy / fp32 store, and
// the per-channel sums for the fused column sums.  Which optional tensors exist is a compile-time property of the loop
// (NA add tensors, DX fp32 output, DXB bf16 copy): with run-time pointers every iteration carried four 64-bit null
// tests, predicated pointer moves and the TF32 rounding of a value nobody stores — the ncu source view of the bf16-step
// kernel charged 70 instructions to a float4 of this loop and 58 to one of phase 1, which computes the SiLU derivative.
template <int T, int NA, bool DX, bool DXB, typename Elem>
__device__ __forceinline__ void gn_bwd_phase2_loop(int tid, int total, Elem elem, const float* a1, unsigned a1step,
                                                   const float* a2, unsigned a2step, float* dst, unsigned ostep,
                                                   uint16_t* dstb, unsigned bstep, int round_dx, float cacc[4]) {
#pragma unroll 2
  for (int idx = tid; idx < total; idx += T) {
    float4 o = elem(idx);
    if (NA >= 1) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(a1));
      o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
      a1 += a1step;
    }
    if (NA >= 2) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(a2));
      o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
      a2 += a2step;
    }
    if (DXB) {
      *reinterpret_cast<uint2*>(dstb) = pack_bf16x4(o);
      dstb += bstep;
    }
    if (DX) {
      if (round_dx) o = round_tf32(o);
      *reinterpret_cast<float4*>(dst) = o;
      dst += ostep;
    }
    cacc[0] += o.x; cacc[1] += o.y; cacc[2] += o.z; cacc[3] += o.w;
  }
}
// a1 / a2 / dst / dstb: this thread's running pointers (NULL = absent; a2 without a1 is passed as a1)
template <int T, typename Elem>
__device__ __forceinline__ void gn_bwd_phase2(int tid, int total, Elem elem, const float* a1, unsigned a1step,
                                              const float* a2, unsigned a2step, float* dst, unsigned ostep, uint16_t* dstb,
                                              unsigned bstep, int round_dx, float cacc[4]) {
  if (!a1 && a2) {
    a1 = a2;
    a1step = a2step;
    a2 = nullptr;
  }
  const int na = a1 ? (a2 ? 2 : 1) : 0;
#define GN_P2(NA, DX, DXB) \
  gn_bwd_phase2_loop<T, NA, DX, DXB>(tid, total, elem, a1, a1step, a2, a2step, dst, ostep, dstb, bstep, round_dx, cacc)
  if (dst && dstb) {
    if (na == 0) GN_P2(0, true, true);
    else if (na == 1) GN_P2(1, true, true);
    else GN_P2(2, true, true);
  } else if (dst) {
    if (na == 0) GN_P2(0, true, false);
    else if (na == 1) GN_P2(1, true, false);
    else GN_P2(2, true, false);
  } else if (dstb) {
    if (na == 0) GN_P2(0, false, true);
    else if (na == 1) GN_P2(1, false, true);
    else GN_P2(2, false, true);
  }
#undef GN_P2
}

// ---------------------------------------------------------------------------------------------------- fused backward
// du = dy * silu'(u), dxhat = du * gamma.  Per (n, g): m1 = mean(dxhat), m2 = mean(dxhat * xhat);
// dx = rstd * (dxhat - m1 - xhat * m2) (+ add1 + add2 on the first add_c channels); per channel dgamma += sum du * xhat,
// dbeta += sum du.  grid = N * (C / CS); dynamic smem = 2 * HW * CS * 4 bytes (x slice, then dy slice; overwritten in
// place by xhat and du).  A variant that kept only dy in smem and re-read x through L2 (twice the CTAs per SM) measured
// slower (201 vs 159 us at N512 HW784 C128): the register-staged global loads of phase 1 do not stay in flight the way
// the cp.async bulk does.
template <int CS, int kGnThreads, bool DROP>
__global__ void __launch_bounds__(kGnThreads, 1024 / kGnThreads)  // <= 64 registers: two 512-thread CTAs per SM
gn_bwd_fused_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ dy, float* __restrict__ dx,
                    const float* __restrict__ add1, int add1_ld, const float* __restrict__ add2, int add2_ld, int add_c,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, int silu, int round_dx,
                    uint16_t* __restrict__ dxb, int dxb_c, GnColsum cs, GnDrop dp) {
  extern __shared__ float4 gn_smem[];
  __shared__ __align__(16) float wred[kGnThreads / 32][2 * CS];
  __shared__ float chan[2 * CS];
  __shared__ float gsum[2 * CS];
  constexpr int L = CS / 4;
  const int slices = s.C / CS;
  const int n = blockIdx.x / slices;
  const int c0 = (blockIdx.x - n * slices) * CS;
  const int cpg = s.C >> 5;
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int l = tid % L;
  const int c = c0 + l * 4;
  const int gl = (l * 4) / cpg;
  const int total = s.HW * L;
  float4* xs = gn_smem;
  float4* ds = gn_smem + total;
  const float* src;
  int ld;
  if (c < s.C1) {
    src = s.x + c;
    ld = s.ld1;
  } else {
    src = s.x2 + (c - s.C1);
    ld = s.ld2;
  }
  src += static_cast<size_t>(n) * s.HW * ld;
  const size_t row0 = static_cast<size_t>(n) * s.HW;
  const float* dsrc = dy + row0 * s.C + c;
  // Running pointers instead of per-iteration index arithmetic: the ncu source view of round 1's kernel charged 23
  // instructions to every cp.async and 79 to every float4 of phase 2 (idx / L, 64-bit multiplies, one address per
  // optional tensor); a thread's successive elements are a constant kGnThreads / L pixels apart.
  constexpr int PSTEP = kGnThreads / L;  // pixels between a thread's consecutive elements
  const int pix0 = tid / L;
  {
    const float* sp = src + static_cast<size_t>(pix0) * ld;
    const float* dp = dsrc + static_cast<size_t>(pix0) * s.C;
    const unsigned sstep = PSTEP * ld, dstep = PSTEP * s.C;
    uint32_t xa = static_cast<uint32_t>(__cvta_generic_to_shared(&xs[tid]));
    const uint32_t dsoff = static_cast<uint32_t>(total) * 16u;
#pragma unroll 2
    for (int idx = tid; idx < total; idx += kGnThreads, sp += sstep, dp += dstep, xa += kGnThreads * 16) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(xa), "l"(sp) : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(xa + dsoff), "l"(dp) : "memory");
    }
  }
  // the rows of the add tensors that phase 2 will read are pulled into L2 now (its loads were the top stall)
  if (c0 < add_c) {
    for (int pix = tid; pix < s.HW; pix += kGnThreads) {
      if (add1) asm volatile("prefetch.global.L2 [%0];" ::"l"(add1 + (row0 + pix) * add1_ld + c0));
      if (add2) asm volatile("prefetch.global.L2 [%0];" ::"l"(add2 + (row0 + pix) * add2_ld + c0));
    }
  }
  // statistics and affine parameters are requested while the slice is still in flight (after the wait their
  // latency was fully exposed: 11 % of the stall samples)
  const float2 st = __ldg(reinterpret_cast<const float2*>(stats + (static_cast<size_t>(n) * 32 + c / cpg) * 2));
  const float4 ga4 = __ldg(reinterpret_cast<const float4*>(gamma + c));
  const float4 be4 = __ldg(reinterpret_cast<const float4*>(beta + c));
  cp_async_wait_all();
  __syncthreads();
  const float mean = st.x, rstd = st.y;
  const float ga[4] = {ga4.x, ga4.y, ga4.z, ga4.w};
  const float be[4] = {be4.x, be4.y, be4.z, be4.w};
  float dg[4] = {0.f, 0.f, 0.f, 0.f}, db[4] = {0.f, 0.f, 0.f, 0.f};
  float s1 = 0.f, s2 = 0.f;
  const float nmr = -mean * rstd;
  const uint32_t dseed = DROP ? drop_seed(dp) : 0u;
  uint32_t e4 = static_cast<uint32_t>(((row0 + pix0) * s.C + c) >> 2);  // float4 number of dy[idx] in the dense tensor
  for (int idx = tid; idx < total; idx += kGnThreads, e4 += (PSTEP * s.C) >> 2) {
    const float4 v = xs[idx];
    float4 d = ds[idx];
    if (DROP) d = drop_apply(d, dp, dseed, e4);  // gradient through the dropout that follows the SiLU
    const float xv[4] = {v.x, v.y, v.z, v.w};
    const float dv[4] = {d.x, d.y, d.z, d.w};
    float xh[4], du[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      xh[j] = fmaf(xv[j], rstd, nmr);
      du[j] = silu ? dv[j] * dsilu_f(fmaf(xh[j], ga[j], be[j])) : dv[j];
      dg[j] += du[j] * xh[j];
      db[j] += du[j];
    }
    xs[idx] = make_float4(xh[0], xh[1], xh[2], xh[3]);  // own slot: re-read by this thread only
    ds[idx] = make_float4(du[0], du[1], du[2], du[3]);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    dg[j] = lane_group_sum<L>(dg[j], false);
    db[j] = lane_group_sum<L>(db[j], false);
  }
  gn_bwd_block_sums<CS, kGnThreads / 32>(dg, db, tid, lane, l, wred, chan, gsum, gamma, c0, dgamma, dbeta);
  // sum(du * gamma) and sum(du * gamma * xhat) over the group follow from the per-channel sums (gamma is constant along
  // the pixels): cpg shared-memory words each
  s1 = s2 = 0.f;
  for (int k = 0; k < cpg; ++k) {
    s2 += gsum[gl * cpg + k];
    s1 += gsum[CS + gl * cpg + k];
  }
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  const float m1 = s1 * inv_cnt, m2 = s2 * inv_cnt;
  const bool has_add = c < add_c;
  const float* a1 = (has_add && add1) ? add1 + row0 * add1_ld + c : nullptr;
  const float* a2 = (has_add && add2) ? add2 + row0 * add2_ld + c : nullptr;
  float* dst = dx ? dx + (row0 + pix0) * s.C + c : nullptr;  // dx may be NULL when only the bf16 copy is wanted
  uint16_t* dstb = (dxb && c < dxb_c) ? dxb + (row0 + pix0) * dxb_c + c : nullptr;  // bf16 copy of the first dxb_c channels
  if (a1) a1 += static_cast<size_t>(pix0) * add1_ld;
  if (a2) a2 += static_cast<size_t>(pix0) * add2_ld;
  const unsigned ostep = PSTEP * s.C, bstep = PSTEP * dxb_c;  // elements; < 2^32 (PSTEP <= 128 pixels)
  const unsigned a1step = PSTEP * add1_ld, a2step = PSTEP * add2_ld;
  float ca[4] = {0.f, 0.f, 0.f, 0.f};  // per-channel sum of dx over this sample's pixels (fused colsum)
  const float k1 = -m1 * rstd, k2 = -m2 * rstd;
  const float gr[4] = {ga[0] * rstd, ga[1] * rstd, ga[2] * rstd, ga[3] * rstd};
  auto elem = [&](int idx) {
    const float4 xh = xs[idx];
    const float4 du = ds[idx];
    float4 o;
    o.x = fmaf(xh.x, k2, fmaf(du.x, gr[0], k1));  // rstd * (du * gamma - m1 - xhat * m2)
    o.y = fmaf(xh.y, k2, fmaf(du.y, gr[1], k1));
    o.z = fmaf(xh.z, k2, fmaf(du.z, gr[2], k1));
    o.w = fmaf(xh.w, k2, fmaf(du.w, gr[3], k1));
    return o;
  };
  gn_bwd_phase2<kGnThreads>(tid, total, elem, a1, a1step, a2, a2step, dst, ostep, dstb, bstep, round_dx, ca);
  // fused column sums of dx (bias gradient of the convolution that produced x, and the per-sample gradient of the
  // timestep-embedding broadcast add): this CTA owns all pixels of (n, channel slice)
  if (cs.c > 0 && c0 < cs.c)  // uniform per CTA
    gn_bwd_colsum<CS, kGnThreads / 32>(ca, tid, lane, l, wred, n, c0, cs);
}

// ---------------------------------------------------------------------------------------------------- bf16 sources
// Variants for bf16 x and / or dy (bf16 storage of the tensors only GroupNorm reads).  The slice is staged in shared
// memory in its RAW element type with plain 16 B cp.async chunks (8 bf16 channels each; loader and consumer mappings
// differ, a block barrier separates them) and every pass widens on the fly; nothing is written back to shared memory
// — the backward recomputes xhat and du in its second phase (2 extra MUFU ops per element) instead of parking them —
// so a bf16 slice costs half the footprint: 4 CTAs/SM in the backward when both sources are bf16, 3 when one is.
__device__ __forceinline__ float4 widen_bf16x4(uint2 r) {
  return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                     __uint_as_float(r.y & 0xffff0000u));
}
// stage the [HW][CS] slice of channels [c0, c0+CS) of sample n into smem `dst` (raw, EB bytes per element)
template <int CS, int EB, int T>
__device__ __forceinline__ void gn_stage(uint8_t* dst, const void* base, int ld, int HW, int tid) {
  constexpr int CPR = CS * EB / 16;  // 16 B chunks per pixel row of the slice
  static_assert(T % CPR == 0, "a thread keeps its chunk column");
  // chunk i = tid + k*T: pixel tid / CPR + k * (T / CPR), chunk column tid % CPR — running pointers, no per-chunk division
  const uint8_t* src = static_cast<const uint8_t*>(base) + (static_cast<size_t>(tid / CPR) * ld * EB + (tid % CPR) * 16);
  const unsigned sstep = static_cast<unsigned>(T / CPR) * static_cast<unsigned>(ld) * EB;
  uint32_t da = static_cast<uint32_t>(__cvta_generic_to_shared(dst)) + static_cast<uint32_t>(tid) * 16u;
  const int total = HW * CPR;
#pragma unroll 2
  for (int i = tid; i < total; i += T, src += sstep, da += T * 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da), "l"(src) : "memory");
}
template <int EB>
__device__ __forceinline__ float4 gn_elem(const uint8_t* raw, int idx) {
  if (EB == 4) return reinterpret_cast<const float4*>(raw)[idx];
  return widen_bf16x4(reinterpret_cast<const uint2*>(raw)[idx]);
}

template <int CS, int T, bool DROP>
__global__ void __launch_bounds__(T)
gn_fwd_bfx_kernel(GnSrc s, float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                  float* __restrict__ y, float eps, int silu, int tf32, GnDrop dp) {
  extern __shared__ float4 gn_smem[];
  __shared__ float fred[2][T / 32][8];  // [sum | centred squares][warp][group of the slice]
  const uint8_t* raw = reinterpret_cast<const uint8_t*>(gn_smem);
  constexpr int L = CS / 4;
  const int slices = s.C / CS;
  const int n = blockIdx.x / slices;
  const int c0 = (blockIdx.x - n * slices) * CS;
  const int cpg = s.C >> 5;
  const int tid = threadIdx.x, lane = tid & 31, l = tid % L;
  const int c = c0 + l * 4;
  const int gl = (l * 4) / cpg;
  const int total = s.HW * L;
  const bool first = c0 < s.C1;
  const int ld = first ? s.ld1 : s.ld2;
  const uint16_t* base = reinterpret_cast<const uint16_t*>(first ? s.x : s.x2) + (first ? c0 : c0 - s.C1) +
                         static_cast<size_t>(n) * s.HW * ld;
  gn_stage<CS, 2, T>(reinterpret_cast<uint8_t*>(gn_smem), base, ld, s.HW, tid);
  const float4 ga = __ldg(reinterpret_cast<const float4*>(gamma + c));  // in flight with the slice
  const float4 be = __ldg(reinterpret_cast<const float4*>(beta + c));
  cp_async_wait_all();
  __syncthreads();
  const bool pair = (cpg == 8);
  const bool leader = (lane < L) && (!pair || !(lane & 1));
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  float sum = 0.f;
  for (int idx = tid; idx < total; idx += T) {
    const float4 v = gn_elem<2>(raw, idx);
    sum += (v.x + v.y) + (v.z + v.w);
  }
  sum = lane_group_sum<L>(sum, pair);
  if (leader) fred[0][tid >> 5][gl] = sum;  // one slot per (warp, group): no shared-memory atomics (CAS spin loops)
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < T / 32; ++w) tot += fred[0][w][gl];
  const float mean = tot * inv_cnt;
  float sq = 0.f;
  for (int idx = tid; idx < total; idx += T) {
    const float4 v = gn_elem<2>(raw, idx);
    const float a = v.x - mean, b = v.y - mean, cc = v.z - mean, d = v.w - mean;
    sq += (a * a + b * b) + (cc * cc + d * d);
  }
  sq = lane_group_sum<L>(sq, pair);
  if (leader) fred[1][tid >> 5][gl] = sq;
  __syncthreads();
  tot = 0.f;
#pragma unroll
  for (int w = 0; w < T / 32; ++w) tot += fred[1][w][gl];
  const float rstd = rsqrtf(tot * inv_cnt + eps);
  if (tid < L && (!pair || !(tid & 1)))
    *reinterpret_cast<float2*>(stats + (static_cast<size_t>(n) * 32 + c / cpg) * 2) = make_float2(mean, rstd);
  const float4 sc = make_float4(rstd * ga.x, rstd * ga.y, rstd * ga.z, rstd * ga.w);
  // running output pointers: a thread's successive elements are T / L pixels apart (see the backward kernels)
  const size_t o0 = (static_cast<size_t>(n) * s.HW + tid / L) * s.C + c;
  float* dst = y + o0;
  uint16_t* dstb = reinterpret_cast<uint16_t*>(y) + o0;
  const unsigned ostep = static_cast<unsigned>(T / L) * static_cast<unsigned>(s.C);
  const uint32_t dseed = DROP ? drop_seed(dp) : 0u;
  uint32_t e4 = static_cast<uint32_t>(o0 >> 2);
#pragma unroll 2
  for (int idx = tid; idx < total; idx += T, dst += ostep, dstb += ostep, e4 += ostep >> 2) {
    const float4 v = gn_elem<2>(raw, idx);
    float4 o;
    o.x = (v.x - mean) * sc.x + be.x;
    o.y = (v.y - mean) * sc.y + be.y;
    o.z = (v.z - mean) * sc.z + be.z;
    o.w = (v.w - mean) * sc.w + be.w;
    if (silu) {
      o.x = silu_fast(o.x); o.y = silu_fast(o.y); o.z = silu_fast(o.z); o.w = silu_fast(o.w);
    }
    if (DROP) o = drop_apply(o, dp, dseed, e4);
    if (tf32 == 2) {
      *reinterpret_cast<uint2*>(dstb) = pack_bf16x4(o);
    } else {
      if (tf32) o = round_tf32(o);
      *reinterpret_cast<float4*>(dst) = o;
    }
  }
}

// XB / DB: bytes per element of x / dy (2 or 4; at least one is 2)
// (Parking du over dy in shared memory, in dy's format, so that phase 2 skips the SiLU derivative: another 3-8 % — 88 ->
// 81 us at 28x28, 128 bf16 channels — but du is then rounded to bf16 once more and the fused column sums move by 1e-3;
// not taken.)
template <int CS, int T, int XB, int DB, bool DROP>
__global__ void __launch_bounds__(T, (XB + DB == 4) ? 1024 / T : 768 / T)  // CTAs/SM the slice footprint allows: 4 or 3
gn_bwd_raw_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                  const float* __restrict__ beta, const void* __restrict__ dy, float* __restrict__ dx,
                  const float* __restrict__ add1, int add1_ld, const float* __restrict__ add2, int add2_ld, int add_c,
                  float* __restrict__ dgamma, float* __restrict__ dbeta, int silu, int round_dx,
                  uint16_t* __restrict__ dxb, int dxb_c, GnColsum cs, GnDrop dp) {
  extern __shared__ float4 gn_smem[];
  __shared__ __align__(16) float wred[T / 32][2 * CS];
  __shared__ float chan[2 * CS];
  __shared__ float gsum[2 * CS];
  constexpr int L = CS / 4;
  const int slices = s.C / CS;
  const int n = blockIdx.x / slices;
  const int c0 = (blockIdx.x - n * slices) * CS;
  const int cpg = s.C >> 5;
  const int tid = threadIdx.x, lane = tid & 31, l = tid % L;
  const int c = c0 + l * 4;
  const int gl = (l * 4) / cpg;
  const int total = s.HW * L;
  uint8_t* xraw = reinterpret_cast<uint8_t*>(gn_smem);
  uint8_t* draw = xraw + static_cast<size_t>(s.HW) * CS * XB;
  const bool first = c0 < s.C1;
  const int ld = first ? s.ld1 : s.ld2;
  const size_t row0 = static_cast<size_t>(n) * s.HW;
  const uint8_t* xbase = reinterpret_cast<const uint8_t*>(first ? s.x : s.x2) +
                         (static_cast<size_t>(first ? c0 : c0 - s.C1) + row0 * ld) * XB;
  const uint8_t* dbase = static_cast<const uint8_t*>(dy) + (row0 * s.C + c0) * DB;
  gn_stage<CS, XB, T>(xraw, xbase, ld, s.HW, tid);
  gn_stage<CS, DB, T>(draw, dbase, s.C, s.HW, tid);
  if (c0 < add_c) {  // add-tensor rows of phase 2 into L2 now
    for (int pix = tid; pix < s.HW; pix += T) {
      if (add1) asm volatile("prefetch.global.L2 [%0];" ::"l"(add1 + (row0 + pix) * add1_ld + c0));
      if (add2) asm volatile("prefetch.global.L2 [%0];" ::"l"(add2 + (row0 + pix) * add2_ld + c0));
    }
  }
  // statistics / affine parameters requested while the slice is in flight
  const float2 st = __ldg(reinterpret_cast<const float2*>(stats + (static_cast<size_t>(n) * 32 + c / cpg) * 2));
  const float4 ga4 = __ldg(reinterpret_cast<const float4*>(gamma + c));
  const float4 be4 = __ldg(reinterpret_cast<const float4*>(beta + c));
  cp_async_wait_all();
  __syncthreads();
  const float mean = st.x, rstd = st.y;
  const float ga[4] = {ga4.x, ga4.y, ga4.z, ga4.w};
  const float be[4] = {be4.x, be4.y, be4.z, be4.w};
  const float nmr = -mean * rstd;
  float dg[4] = {0.f, 0.f, 0.f, 0.f}, db[4] = {0.f, 0.f, 0.f, 0.f};
  float s1 = 0.f, s2 = 0.f;
  const uint32_t dseed = DROP ? drop_seed(dp) : 0u;
  const uint32_t e40 = static_cast<uint32_t>(((row0 + tid / L) * s.C + c) >> 2);  // float4 number of dy[idx = tid]
  const uint32_t e4step = static_cast<uint32_t>((T / L) * s.C) >> 2;
  uint32_t e4 = e40;
  for (int idx = tid; idx < total; idx += T, e4 += e4step) {
    const float4 v = gn_elem<XB>(xraw, idx);
    float4 d = gn_elem<DB>(draw, idx);
    if (DROP) d = drop_apply(d, dp, dseed, e4);  // gradient through the dropout that follows the SiLU
    const float xv[4] = {v.x, v.y, v.z, v.w};
    const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float xh = fmaf(xv[j], rstd, nmr);
      const float du = silu ? dv[j] * dsilu_fast(fmaf(xh, ga[j], be[j])) : dv[j];
      dg[j] += du * xh;
      db[j] += du;
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    dg[j] = lane_group_sum<L>(dg[j], false);
    db[j] = lane_group_sum<L>(db[j], false);
  }
  gn_bwd_block_sums<CS, T / 32>(dg, db, tid, lane, l, wred, chan, gsum, gamma, c0, dgamma, dbeta);
  for (int k = 0; k < cpg; ++k) {  // gamma is constant along the pixels: the group sums follow from the channel sums
    s2 += gsum[gl * cpg + k];
    s1 += gsum[CS + gl * cpg + k];
  }
  const float inv_cnt = 1.0f / static_cast<float>(s.HW * cpg);
  const float m1 = s1 * inv_cnt, m2 = s2 * inv_cnt;
  const bool has_add = c < add_c;
  const float* a1 = (has_add && add1) ? add1 + row0 * add1_ld + c : nullptr;
  const float* a2 = (has_add && add2) ? add2 + row0 * add2_ld + c : nullptr;
  constexpr int PSTEP = T / L;  // pixels between a thread's consecutive elements (running pointers, see the fp32 kernel)
  const int pix0 = tid / L;
  float* dst = dx ? dx + (row0 + pix0) * s.C + c : nullptr;
  uint16_t* dstb = (dxb && c < dxb_c) ? dxb + (row0 + pix0) * dxb_c + c : nullptr;
  if (a1) a1 += static_cast<size_t>(pix0) * add1_ld;
  if (a2) a2 += static_cast<size_t>(pix0) * add2_ld;
  const unsigned ostep = PSTEP * s.C, bstep = PSTEP * dxb_c, a1step = PSTEP * add1_ld, a2step = PSTEP * add2_ld;
  const float k1 = -m1 * rstd, k2 = -m2 * rstd;
  const float gr[4] = {ga[0] * rstd, ga[1] * rstd, ga[2] * rstd, ga[3] * rstd};
  float cacc[4] = {0.f, 0.f, 0.f, 0.f};
  auto elem = [&](int idx) {
    const float4 v = gn_elem<XB>(xraw, idx);
    float4 d = gn_elem<DB>(draw, idx);
    if (DROP) d = drop_apply(d, dp, dseed, e40 + static_cast<uint32_t>(idx / T) * e4step);
    const float xv[4] = {v.x, v.y, v.z, v.w};
    const float dv[4] = {d.x, d.y, d.z, d.w};
    float ov[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {  // recomputed, not parked in smem
      const float xh = fmaf(xv[j], rstd, nmr);
      const float du = silu ? dv[j] * dsilu_fast(fmaf(xh, ga[j], be[j])) : dv[j];
      ov[j] = fmaf(xh, k2, fmaf(du, gr[j], k1));  // rstd * (du * gamma - m1 - xhat * m2)
    }
    return make_float4(ov[0], ov[1], ov[2], ov[3]);
  };
  gn_bwd_phase2<T>(tid, total, elem, a1, a1step, a2, a2step, dst, ostep, dstb, bstep, round_dx, cacc);
  if (cs.c > 0 && c0 < cs.c)  // uniform per CTA
    gn_bwd_colsum<CS, T / 32>(cacc, tid, lane, l, wred, n, c0, cs);
}

// ---------------------------------------------------------------------------------------------------- streaming path
// For slices that do not fit in shared memory (e.g. 64x64 images): stats (atomics on {sum, sumsq}) -> finalize
// ({mean, rstd}) -> apply; backward: reduce -> apply.  One more read of x (and dy) than the fused kernels.
// grid (chunks, N), block 256.  acc[n][g] = {sum, sumsq}; zeroed by the caller.
__global__ void __launch_bounds__(256) gn_stats_kernel(GnSrc s, float* __restrict__ acc, int pix_per_chunk) {
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
  // Shifted moments: sums of (x - K) and (x - K)^2 with K = the group's first element of this sample, so that
  // var = E[(x-K)^2] - E[x-K]^2 does not cancel catastrophically when |mean| >> std (E[x^2] - mean^2 did)
  const float K = ld_src(s, static_cast<size_t>(n) * s.HW, g * cpg).x;
  float sum = 0.f, sq = 0.f;
  for (int p = p0 + prow; p < p1; p += prows) {
    float4 v = ld_src(s, static_cast<size_t>(n) * s.HW + p, c);
    v.x -= K; v.y -= K; v.z -= K; v.w -= K;
    sum += (v.x + v.y) + (v.z + v.w);
    sq += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  atomicAdd(&sh[g * 2], sum);
  atomicAdd(&sh[g * 2 + 1], sq);
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(&acc[n * 64 + threadIdx.x], sh[threadIdx.x]);
}

__global__ void gn_finalize_kernel(GnSrc s, float* __restrict__ stats, int n_groups, float inv_cnt, float eps) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_groups) return;
  const int n = i >> 5, g = i & 31;
  const float K = ld_src(s, static_cast<size_t>(n) * s.HW, g * (s.C >> 5)).x;  // the shift gn_stats_kernel used
  const float2 a = *reinterpret_cast<const float2*>(stats + 2 * i);
  const float d = a.x * inv_cnt;  // E[x - K]
  const float var = fmaxf(a.y * inv_cnt - d * d, 0.f);
  *reinterpret_cast<float2*>(stats + 2 * i) = make_float2(K + d, rsqrtf(var + eps));
}

__global__ void __launch_bounds__(256)
gn_apply_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                const float* __restrict__ beta, float* __restrict__ y, int silu, int tf32, GnDrop dp) {
  const uint32_t dseed = dp.thr ? drop_seed(dp) : 0u;
  const int lanes = s.C >> 2;
  const size_t total = static_cast<size_t>(s.N) * s.HW * lanes;
  const int cpg = s.C >> 5;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t pix = i / lanes;
    const int c = static_cast<int>(i - pix * lanes) * 4;
    const int n = static_cast<int>(pix / s.HW);
    const float2 st = *reinterpret_cast<const float2*>(stats + (n * 32 + c / cpg) * 2);
    const float mean = st.x, rstd = st.y;
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
    if (dp.thr) o = drop_apply(o, dp, dseed, static_cast<uint32_t>(i));
    if (tf32 == 2) {
      *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(y) + pix * s.C + c) = pack_bf16x4(o);
    } else {
      if (tf32) o = round_tf32(o);
      *reinterpret_cast<float4*>(y + pix * s.C + c) = o;
    }
  }
}

__global__ void __launch_bounds__(256)
gn_bwd_reduce_kernel(GnSrc s, const float* __restrict__ stats, const float* __restrict__ gamma,
                     const float* __restrict__ beta, const float* __restrict__ dy, float* __restrict__ red,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, int silu, int pix_per_chunk, GnDrop dp) {
  const uint32_t dseed = dp.thr ? drop_seed(dp) : 0u;
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
  const float2 st = *reinterpret_cast<const float2*>(stats + (n * 32 + g) * 2);
  const float mean = st.x, rstd = st.y;
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
    float4 d = *reinterpret_cast<const float4*>(dy + pix * s.C + c);
    if (dp.thr) d = drop_apply(d, dp, dseed, static_cast<uint32_t>(pix * lanes + lane4));
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
                    const float* __restrict__ add2, int add2_ld, int add_c, int silu, int round_dx,
                    uint16_t* __restrict__ dxb, int dxb_c, GnDrop dp) {
  const uint32_t dseed = dp.thr ? drop_seed(dp) : 0u;
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
    const float mean = st.x, rstd = st.y;
    const float m1 = rd.x * inv_cnt, m2 = rd.y * inv_cnt;
    const float4 v = ld_src(s, pix, c);
    float4 d = *reinterpret_cast<const float4*>(dy + pix * s.C + c);
    if (dp.thr) d = drop_apply(d, dp, dseed, static_cast<uint32_t>(i));
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
    if (dxb && c < dxb_c) *reinterpret_cast<uint2*>(dxb + pix * dxb_c + c) = pack_bf16x4(ov);
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
  if (a->N <= 0 || a->HW <= 0) return fail_arg("gn: empty tensor");
  s->x = a->x;
  s->x2 = a->x2;
  s->ld1 = a->x_ld ? a->x_ld : a->C1;
  s->ld2 = a->x2_ld ? a->x2_ld : a->C2;
  s->N = a->N;
  s->HW = a->HW;
  s->C1 = a->C1;
  s->C = C;
  if (a->x_dtype != 0 && a->x_dtype != 1) return fail_arg("gn: x_dtype must be 0 (fp32) or 1 (bf16)");
  s->bf = a->x_dtype;
  return 0;
}

// Shared-memory budget of the fused kernels; 0 = not initialised.  Developer knobs: B200_GN_STREAM=1 forces the
// streaming path, B200_GN_CS=16|32 the slice width, B200_GN_THREADS=256|512 the CTA size.
static int g_gn_max_dyn = 0;
static int g_gn_stream = 0, g_gn_cs = 0, g_gn_threads = 0;
template <typename K>
static cudaError_t gn_optin(K kernel, int bytes) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
#define GN_OPTIN2(KERNEL, ...)                                                   \
  do {                                                                           \
    if (e == cudaSuccess) e = gn_optin(KERNEL<__VA_ARGS__, false>, max_dyn);      \
    if (e == cudaSuccess) e = gn_optin(KERNEL<__VA_ARGS__, true>, max_dyn);       \
  } while (0)
static int gn_init() {
  if (g_gn_max_dyn) return 0;
  int dev = 0, optin = 0;
  cudaGetDevice(&dev);
  cudaError_t e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return check_cuda(e, "cudaDeviceGetAttribute(smem optin)");
  const int max_dyn = optin - 8192;  // room for the kernels' static shared memory (per-warp reduction slots: up to 4.6 KB)
  GN_OPTIN2(gn_fwd_fused_kernel, 32, 256);
  GN_OPTIN2(gn_fwd_fused_kernel, 16, 256);
  GN_OPTIN2(gn_fwd_fused_kernel, 32, 512);
  GN_OPTIN2(gn_fwd_fused_kernel, 16, 512);
  GN_OPTIN2(gn_bwd_fused_kernel, 32, 256);
  GN_OPTIN2(gn_bwd_fused_kernel, 16, 256);
  GN_OPTIN2(gn_bwd_fused_kernel, 32, 512);
  GN_OPTIN2(gn_bwd_fused_kernel, 16, 512);
  GN_OPTIN2(gn_fwd_bfx_kernel, 16, 256);
  GN_OPTIN2(gn_fwd_bfx_kernel, 32, 256);
  GN_OPTIN2(gn_bwd_raw_kernel, 16, 256, 2, 2);
  GN_OPTIN2(gn_bwd_raw_kernel, 32, 256, 2, 2);
  GN_OPTIN2(gn_bwd_raw_kernel, 16, 256, 4, 2);
  GN_OPTIN2(gn_bwd_raw_kernel, 32, 256, 4, 2);
  GN_OPTIN2(gn_bwd_raw_kernel, 16, 256, 2, 4);
  GN_OPTIN2(gn_bwd_raw_kernel, 32, 256, 2, 4);
  GN_OPTIN2(gn_fwd_fused_kernel, 8, 256);
  GN_OPTIN2(gn_bwd_fused_kernel, 8, 256);
  if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gn fused)");
  const char* env = getenv("B200_GN_STREAM");
  g_gn_stream = env ? atoi(env) : 0;
  env = getenv("B200_GN_CS");
  g_gn_cs = env ? atoi(env) : 0;
  env = getenv("B200_GN_THREADS");
  g_gn_threads = env ? atoi(env) : 0;
  g_gn_max_dyn = max_dyn;
  return 0;
}
// channels per slice for `bufs` resident copies of the slice (1 forward, 2 backward); 0 = use the streaming path.
// Several CTAs per SM are wanted (some stream in while others compute / stream out).
static int gn_slice(const GnSrc& s, int bufs) {
  if (g_gn_stream || s.C1 % 32) return 0;
  const size_t b16 = static_cast<size_t>(s.HW) * 16 * 4 * bufs;
  if (b16 > static_cast<size_t>(g_gn_max_dyn)) return 0;
  if (g_gn_cs == 32 && 2 * b16 <= static_cast<size_t>(g_gn_max_dyn)) return 32;
  if (g_gn_cs == 16) return 16;
  if (g_gn_cs == 8) return 8;
  return (2 * b16 <= 50 * 1024) ? 32 : 16;
}
static int gn_threads(size_t smem) {
  if (g_gn_threads == 256 || g_gn_threads == 512) return g_gn_threads;
  return smem > 64 * 1024 ? 512 : 256;
}

// dropout fields of the arguments -> kernel form (thr = 0: off)
static int gn_drop(const b200_gn_args* a, GnDrop* d) {
  d->thr = 0; d->scale = 1.f; d->seed = a->dropout_seed; d->step = a->dropout_step;
  if (a->dropout_p == 0.f) return 0;
  if (!(a->dropout_p > 0.f && a->dropout_p < 1.f)) return fail_arg("gn: dropout_p must be in [0, 1)");
  uint32_t thr = static_cast<uint32_t>(a->dropout_p * 65536.0f + 0.5f);
  if (thr > 65535u) thr = 65535u;
  d->thr = thr;  // (a p below 2^-17 rounds to "off")
  d->scale = 65536.0f / static_cast<float>(65536u - thr);
  return 0;
}

extern "C" int b200_gn_fwd(const b200_gn_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  GnSrc s;
  if (int r = gn_src(a, &s)) return r;
  if (!a->y) return fail_arg("gn_fwd: null y");
  if (a->y_dtype != 0 && a->y_dtype != 1) return fail_arg("gn_fwd: y_dtype must be 0 (fp32) or 1 (bf16)");
  const int ymode = a->y_dtype == 1 ? 2 : (a->round_tf32 ? 1 : 0);
  if (int r = gn_init()) return r;
  GnDrop dp;
  if (int r = gn_drop(a, &dp)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (s.bf) {  // bf16 x: raw-staged kernel, 2 B per element of shared memory
    const size_t b16 = static_cast<size_t>(s.HW) * 16 * 2;
    if (g_gn_stream || s.C1 % 32 || b16 > static_cast<size_t>(g_gn_max_dyn))
      return fail_arg("gn_fwd: a bf16 input needs the shared-memory-resident kernel (HW too large)");
    const int cs = (2 * b16 <= 50 * 1024) ? 32 : 16;
    const int grid = s.N * (s.C / cs);
    const size_t smem = static_cast<size_t>(s.HW) * cs * 2;
#define GN_FWD_BFX(CS, DROP) \
  gn_fwd_bfx_kernel<CS, 256, DROP><<<grid, 256, smem, st>>>(s, a->stats, a->gamma, a->beta, static_cast<float*>(a->y), a->eps, a->silu, ymode, dp)
    if (cs == 32) { if (dp.thr) GN_FWD_BFX(32, true); else GN_FWD_BFX(32, false); }
    else { if (dp.thr) GN_FWD_BFX(16, true); else GN_FWD_BFX(16, false); }
#undef GN_FWD_BFX
    return after_launch("gn_fwd_bfx_kernel");
  }
  const int cs = gn_slice(s, 1);
  if (cs) {
    const int grid = s.N * (s.C / cs);
    const size_t smem = static_cast<size_t>(s.HW) * cs * 4;
    const int th = gn_threads(smem);
#define GN_FWD_D(CS, TH, DROP) \
  gn_fwd_fused_kernel<CS, TH, DROP><<<grid, TH, smem, st>>>(s, a->stats, a->gamma, a->beta, static_cast<float*>(a->y), a->eps, a->silu, ymode, dp)
#define GN_FWD(CS, TH) do { if (dp.thr) GN_FWD_D(CS, TH, true); else GN_FWD_D(CS, TH, false); } while (0)
    if (cs == 8) GN_FWD(8, 256);
    else if (cs == 32 && th == 256) GN_FWD(32, 256);
    else if (cs == 32) GN_FWD(32, 512);
    else if (th == 256) GN_FWD(16, 256);
    else GN_FWD(16, 512);
#undef GN_FWD
#undef GN_FWD_D
    return after_launch("gn_fwd_fused_kernel");
  }
  if (s.bf) return fail_arg("gn_fwd: a bf16 input needs the shared-memory-resident kernel (HW too large)");
  if (int r = check_cuda(cudaMemsetAsync(a->stats, 0, static_cast<size_t>(s.N) * 64 * sizeof(float), st), "gn_fwd memset")) return r;
  const int chunks = gn_chunks(s.N, s.HW);
  const int ppc = (s.HW + chunks - 1) / chunks;
  gn_stats_kernel<<<dim3(chunks, s.N), 256, 0, st>>>(s, a->stats, ppc);
  if (int r = after_launch("gn_stats_kernel")) return r;
  const int ng = s.N * 32;
  gn_finalize_kernel<<<(ng + 255) / 256, 256, 0, st>>>(s, a->stats, ng, 1.0f / static_cast<float>(s.HW * (s.C >> 5)), a->eps);
  if (int r = after_launch("gn_finalize_kernel")) return r;
  const size_t total = static_cast<size_t>(s.N) * s.HW * (s.C / 4);
  gn_apply_kernel<<<ew_grid(total), 256, 0, st>>>(s, a->stats, a->gamma, a->beta, static_cast<float*>(a->y), a->silu, ymode, dp);
  return after_launch("gn_apply_kernel");
}

// Bytes of workspace b200_gn_bwd wants for these arguments (SURVEY.md §8b's `b200_<op>_workspace`): the streaming path
// (slices that do not fit in shared memory) reduces into N*64 floats; the shared-memory-resident kernels need none, but
// which path a shape takes depends on developer knobs, so the upper bound is reported for every shape.
extern "C" size_t b200_gn_bwd_workspace(const b200_gn_bwd_args* b) {
  if (!b || b->fwd.N <= 0) return 0;
  return static_cast<size_t>(b->fwd.N) * 64 * sizeof(float);
}

extern "C" int b200_gn_bwd(const b200_gn_bwd_args* b, void* ws, size_t ws_bytes, void* stream) {
  if (!b) return fail_arg("gn_bwd: null args");
  GnSrc s;
  if (int r = gn_src(&b->fwd, &s)) return r;
  if (!b->dy || !b->dgamma || !b->dbeta) return fail_arg("gn_bwd: null pointer");
  if (!b->dx && !b->dx_bf16) return fail_arg("gn_bwd: neither dx nor dx_bf16 given");
  if (b->dy_dtype != 0 && b->dy_dtype != 1) return fail_arg("gn_bwd: dy_dtype must be 0 (fp32) or 1 (bf16)");
  if (int r = gn_init()) return r;
  GnDrop dp;
  if (int r = gn_drop(&b->fwd, &dp)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int add_c = b->add_c ? b->add_c : s.C;
  uint16_t* dxb = static_cast<uint16_t*>(b->dx_bf16);
  const int dxb_c = dxb ? b->dx_bf16_c : 0;
  if (dxb && (dxb_c <= 0 || dxb_c > s.C || dxb_c % 32)) return fail_arg("gn_bwd: dx_bf16_c must be a multiple of 32 in (0, C]");
  GnColsum csm = {b->colsum_per_sample, b->colsum_ps_ld, b->colsum_total1, b->colsum_total2, 0};
  if (csm.per_sample || csm.total1 || csm.total2) {
    if (b->colsum_c != 128) return fail_arg("gn_bwd: colsum_c must be 128");
    csm.c = 128;
  }
  if (s.bf || b->dy_dtype) {  // raw-staged kernel: 2 or 4 B per element of x and dy in shared memory, recompute in phase 2
    const int xb = s.bf ? 2 : 4, dbb = b->dy_dtype ? 2 : 4;
    const size_t b16 = static_cast<size_t>(s.HW) * 16 * (xb + dbb);
    if (g_gn_stream || s.C1 % 32 || b16 > static_cast<size_t>(g_gn_max_dyn))
      return fail_arg("gn_bwd: bf16 inputs need the shared-memory-resident kernel (HW too large)");
    const int cs = (2 * b16 <= 50 * 1024) ? 32 : 16;
    const int grid = s.N * (s.C / cs);
    const size_t smem = static_cast<size_t>(s.HW) * cs * (xb + dbb);
#define GN_BWD_RAW_D(CS, XB, DB, DROP)                                                                                  \
  gn_bwd_raw_kernel<CS, 256, XB, DB, DROP><<<grid, 256, smem, st>>>(s, b->fwd.stats, b->fwd.gamma, b->fwd.beta, b->dy,   \
                                                                    b->dx, b->add1, b->add1_ld, b->add2, b->add2_ld,     \
                                                                    add_c, b->dgamma, b->dbeta, b->fwd.silu,             \
                                                                    b->round_dx, dxb, dxb_c, csm, dp)
#define GN_BWD_RAW(CS, XB, DB) do { if (dp.thr) GN_BWD_RAW_D(CS, XB, DB, true); else GN_BWD_RAW_D(CS, XB, DB, false); } while (0)
    if (xb == 2 && dbb == 2) { if (cs == 32) GN_BWD_RAW(32, 2, 2); else GN_BWD_RAW(16, 2, 2); }
    else if (xb == 4) { if (cs == 32) GN_BWD_RAW(32, 4, 2); else GN_BWD_RAW(16, 4, 2); }
    else { if (cs == 32) GN_BWD_RAW(32, 2, 4); else GN_BWD_RAW(16, 2, 4); }
#undef GN_BWD_RAW
#undef GN_BWD_RAW_D
    return after_launch("gn_bwd_raw_kernel");
  }
  const int cs = gn_slice(s, 2);
  if (cs) {
    const int grid = s.N * (s.C / cs);
    const size_t smem = static_cast<size_t>(s.HW) * cs * 4 * 2;
    const int th = gn_threads(smem);
#define GN_BWD_D(CS, TH, DROP)                                                                                          \
  gn_bwd_fused_kernel<CS, TH, DROP><<<grid, TH, smem, st>>>(s, b->fwd.stats, b->fwd.gamma, b->fwd.beta, b->dy, b->dx,    \
                                                            b->add1, b->add1_ld, b->add2, b->add2_ld, add_c, b->dgamma,  \
                                                            b->dbeta, b->fwd.silu, b->round_dx, dxb, dxb_c, csm, dp)
#define GN_BWD(CS, TH) do { if (dp.thr) GN_BWD_D(CS, TH, true); else GN_BWD_D(CS, TH, false); } while (0)
    if (cs == 8) GN_BWD(8, 256);
    else if (cs == 32 && th == 256) GN_BWD(32, 256);
    else if (cs == 32) GN_BWD(32, 512);
    else if (th == 256) GN_BWD(16, 256);
    else GN_BWD(16, 512);
#undef GN_BWD
#undef GN_BWD_D
    return after_launch("gn_bwd_fused_kernel");
  }
  if (s.bf || b->dy_dtype || !b->dx)
    return fail_arg("gn_bwd: bf16 inputs / a bf16-only dx need the shared-memory-resident kernel (HW too large)");
  const size_t need = static_cast<size_t>(s.N) * 64 * sizeof(float);
  if (!ws || ws_bytes < need) return fail_arg("gn_bwd: workspace of N*64 floats required for the streaming path");
  float* red = static_cast<float*>(ws);
  if (int r = check_cuda(cudaMemsetAsync(red, 0, need, st), "gn_bwd memset")) return r;
  const int chunks = gn_chunks(s.N, s.HW);
  const int ppc = (s.HW + chunks - 1) / chunks;
  gn_bwd_reduce_kernel<<<dim3(chunks, s.N), 256, 0, st>>>(s, b->fwd.stats, b->fwd.gamma, b->fwd.beta, b->dy, red,
                                                          b->dgamma, b->dbeta, b->fwd.silu, ppc, dp);
  if (int r = after_launch("gn_bwd_reduce_kernel")) return r;
  const size_t total = static_cast<size_t>(s.N) * s.HW * (s.C / 4);
  gn_bwd_apply_kernel<<<ew_grid(total), 256, 0, st>>>(s, b->fwd.stats, b->fwd.gamma, b->fwd.beta, b->dy, red, b->dx,
                                                      b->add1, b->add1_ld, b->add2, b->add2_ld, add_c, b->fwd.silu,
                                                      b->round_dx, dxb, dxb_c, dp);
  if (int r = after_launch("gn_bwd_apply_kernel")) return r;
  if (csm.c) {  // streaming path: the column sums take their own pass over dx
    b200_colsum_args ca = {b->dx, csm.per_sample, csm.total1, csm.total2, s.N, s.HW, s.C, csm.ps_ld};
    return b200_colsum(&ca, nullptr, 0, stream);
  }
  return 0;
}

// ------------------------------------------------------------------------------ allocation-free GroupNorm backward
// b200_gn_bwd for callers that cannot hand over a workspace they allocated (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
// The N*64-float reduction scratch of the streaming path lives in a file-scope device array (1 MiB, held by every GPU
// that loads the library until the process exits).  In HOST code the array's name is only a stand-in for the symbol:
// its device address is looked up once per device with cudaGetSymbolAddress (that device current) and passed on as the
// workspace.  b200_gn_bwd clears the part it uses with cudaMemsetAsync on `stream` before the reduction, so stale
// contents of an earlier call are harmless — but every call shares the one array: calls that overlap on the GPU
// (different streams, no dependency between them) race on it.  Nothing here allocates.
constexpr int kGnStaticMaxN = B200_GN_BWD_STATIC_MAX_N;
static __device__ __align__(256) float g_gn_bwd_static_scratch[static_cast<size_t>(kGnStaticMaxN) * 64];

extern "C" int b200_gn_bwd_static(const b200_gn_bwd_args* b, void* stream) {
  if (!b) return fail_arg("gn_bwd_static: null args");
  if (b->fwd.N > kGnStaticMaxN) {
    snprintf(g_last_error, sizeof(g_last_error),
             "gn_bwd_static: N = %d exceeds the static scratch (B200_GN_BWD_STATIC_MAX_N = %d samples); "
             "use b200_gn_bwd with a caller-allocated workspace", b->fwd.N, kGnStaticMaxN);
    return -1;
  }
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static float* dev_ptr[kMaxDev] = {};
  int dev = 0;
  if (int r = check_cuda(cudaGetDevice(&dev), "gn_bwd_static: cudaGetDevice")) return r;
  if (dev < 0 || dev >= kMaxDev) return fail_arg("gn_bwd_static: device ordinal out of range");
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!dev_ptr[dev]) {
      float* p = nullptr;
      if (int r = check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_bwd_static_scratch),
                             "gn_bwd_static: cudaGetSymbolAddress"))
        return r;
      if (!p) return fail_arg("gn_bwd_static: cudaGetSymbolAddress returned a null pointer");
      dev_ptr[dev] = p;
    }
    scratch = dev_ptr[dev];
  }
  return b200_gn_bwd(b, scratch, sizeof(float) * static_cast<size_t>(kGnStaticMaxN) * 64, stream);
}
