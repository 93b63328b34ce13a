// The two convolutions that are not tensor-core shaped: the 1->128 input conv (K = 9) and the 128->1 output conv
// (N = 1).  Both are bandwidth-bound on the 128-channel side, so they are written as coalesced float4 kernels with
// warp-shuffle / shared-memory reductions.  Forward, weight/bias gradients and (for the output conv) data gradient.
// Reference: simple_unet.py:92-94 (Downsample(1, C, 1) = Conv2d(1, C, 3, padding=1)) and :41 (Conv2d(C, 1, 3, padding=1)).
#include "host_common.h"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

namespace b200 {

// ------------------------------------------------------------------------------------------------ conv_in forward
// out[n,y,x,c] = b[c] + sum_{kh,kw} x[n,y+kh-1,x+kw-1] * w[c][kh][kw].
// One warp = one image row; lane = 4 output channels with its 36 weights in registers.  The 3x3 input window slides
// along the row in registers (3 broadcast loads per pixel), so a pixel costs ~50 instructions for a 512 B store and no
// integer division: the kernel is bound by the HBM write (the per-element version spent its time on index math).
__global__ void __launch_bounds__(256)
conv_in_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                   float* __restrict__ out, int N, int H, int W, int round_out, int x_ld, int w_co_ld, int w_tap_ld,
                   int accumulate) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = lane * 4;
  float wv[4][9];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int t = 0; t < 9; ++t) wv[j][t] = w[(c + j) * w_co_ld + t * w_tap_ld];  // w is [c][tap][image channel]
  // image channels beyond the first (3-channel extension) are further launches that add onto out
  const float4 b4 = accumulate ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<const float4*>(bias + c);
  const int rows = N * H;
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const int n = row / H, y = row - n * H;
    const float* img = x + static_cast<size_t>(n) * x_ld;
    const float* r0 = (y > 0) ? img + (y - 1) * W : nullptr;
    const float* r1 = img + y * W;
    const float* r2 = (y + 1 < H) ? img + (y + 1) * W : nullptr;
    float a[3][3];  // a[kh][kw]: input columns x-1, x, x+1
    a[0][0] = a[1][0] = a[2][0] = 0.f;
    a[0][1] = r0 ? __ldg(r0) : 0.f;
    a[1][1] = __ldg(r1);
    a[2][1] = r2 ? __ldg(r2) : 0.f;
    float* orow = out + static_cast<size_t>(row) * W * 128 + c;
    for (int xx = 0; xx < W; ++xx) {
      const bool in = xx + 1 < W;
      a[0][2] = (in && r0) ? __ldg(r0 + xx + 1) : 0.f;
      a[1][2] = in ? __ldg(r1 + xx + 1) : 0.f;
      a[2][2] = (in && r2) ? __ldg(r2 + xx + 1) : 0.f;
      float acc[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[j] += a[kh][kw] * wv[j][kh * 3 + kw];
      float4 o = make_float4(acc[0], acc[1], acc[2], acc[3]);
      if (accumulate) {
        const float4 p = *reinterpret_cast<const float4*>(orow + static_cast<size_t>(xx) * 128);
        o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
      }
      if (round_out) o = round_tf32(o);
      *reinterpret_cast<float4*>(orow + static_cast<size_t>(xx) * 128) = o;
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        a[kh][0] = a[kh][1];
        a[kh][1] = a[kh][2];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ conv_in backward
// dw[c][tap] += sum_pix dy[pix][c] * x[pix + tap];  db[c] += sum_pix dy[pix][c].  grid (chunks, N), block 256:
// thread = (pixel row r = tid/32, channel quad tid%32); 40 register accumulators, smem reduce over the 8 pixel rows.
// The image is staged zero-padded in shared memory (no bounds tests, no global loads in the tap loop) and the dy rows
// of FOUR pixels are requested before any of them is used: the kernel reads 205 MB and was bound by one exposed load
// latency per pixel (112 us; 31 us is the HBM time).
__global__ void __launch_bounds__(256)
conv_in_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw,
                   float* __restrict__ db, int H, int W, int pix_per_chunk, int x_ld, int w_co_ld, int w_tap_ld) {
  __shared__ float sh[8][32][41];
  extern __shared__ float simg[];  // (H + 2) x (W + 2), zero border
  const int n = blockIdx.y;
  const int q = threadIdx.x & 31, r = threadIdx.x >> 5;
  const int HW = H * W, Wp = W + 2;
  const float* img = x + static_cast<size_t>(n) * x_ld;
  for (int i = threadIdx.x; i < (H + 2) * Wp; i += 256) {
    const int yy = i / Wp - 1, xx = i - (yy + 1) * Wp - 1;
    simg[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? img[yy * W + xx] : 0.f;
  }
  __syncthreads();
  float acc[10][4];
#pragma unroll
  for (int t = 0; t < 10; ++t)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[t][j] = 0.f;
  const int p0 = blockIdx.x * pix_per_chunk, p1 = min(p0 + pix_per_chunk, HW);
  const float* dyn = dy + static_cast<size_t>(n) * HW * 128 + q * 4;
  for (int p = p0 + r; p < p1; p += 32) {
    float4 d[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pp = p + 8 * u;
      d[u] = pp < p1 ? __ldg(reinterpret_cast<const float4*>(dyn + static_cast<size_t>(pp) * 128))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pp = p + 8 * u;
      if (pp >= p1) break;
      const int y = pp / W, xx = pp - y * W;
      const float* sp = simg + y * Wp + xx;  // padded (y, xx) = tap (0, 0)
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const float v = sp[kh * Wp + kw];
          acc[kh * 3 + kw][0] += d[u].x * v; acc[kh * 3 + kw][1] += d[u].y * v;
          acc[kh * 3 + kw][2] += d[u].z * v; acc[kh * 3 + kw][3] += d[u].w * v;
        }
      acc[9][0] += d[u].x; acc[9][1] += d[u].y; acc[9][2] += d[u].z; acc[9][3] += d[u].w;
    }
  }
#pragma unroll
  for (int t = 0; t < 10; ++t)
#pragma unroll
    for (int j = 0; j < 4; ++j) sh[r][q][t * 4 + j] = acc[t][j];
  __syncthreads();
  // 128 channels x 10 values, summed over the 8 pixel rows
  for (int i = threadIdx.x; i < 1280; i += 256) {
    const int c = i / 10, t = i - c * 10;
    float s = 0.f;
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) s += sh[rr][c >> 2][t * 4 + (c & 3)];
    if (t < 9) atomicAdd(&dw[c * w_co_ld + t * w_tap_ld], s);
    else if (db) atomicAdd(&db[c], s);  // db only from the first image channel's launch
  }
}

// ------------------------------------------------------------------------------------------------ conv_out forward
// out[pix] = b + sum_{tap,c} h[pix+tap][c] * w[tap][c].  One CTA per image (HW <= 4096): every input row is read
// once, its 9 per-tap dot products are parked in smem, then each output pixel gathers its 9 partials.
__global__ void __launch_bounds__(256)
conv_out_fwd_kernel(const float* __restrict__ h, const float* __restrict__ w, const float* __restrict__ bias,
                    float* __restrict__ out, int H, int W, int out_ld) {
  extern __shared__ float part[];  // [HW][9]
  const int n = blockIdx.x;
  const int HW = H * W;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4 wv[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wv[t] = *reinterpret_cast<const float4*>(w + t * 128 + lane * 4);
  // Reduction of the 9 per-tap dot products over the 32 lanes by value-splitting butterflies: at every step a lane
  // keeps half of its values and trades the other half with its partner (5 + 3 + 2 + 1 + 1 = 12 shuffles per pixel
  // instead of 9 x 5).  After the steps lane L holds tap (b1 + 2 b2) + 3 b3 + 5 b4 (b_k = bit k of L), pads excluded.
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  const int jj = (b1 ? 1 : 0) + (b2 ? 2 : 0), j5 = jj + (b3 ? 3 : 0), tap = j5 + (b4 ? 5 : 0);
  const bool writer = !(lane & 1) && jj < 3 && j5 < 5 && tap < 9;
  const float* hn = h + static_cast<size_t>(n) * HW * 128 + lane * 4;
  for (int p = warp; p < HW; p += 32) {  // four pixels per iteration: four independent 512 B loads in flight per warp
    float4 vv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      vv[u] = (p + 8 * u < HW) ? __ldg(reinterpret_cast<const float4*>(hn + static_cast<size_t>(p + 8 * u) * 128))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int half = 0; half < 4; ++half) {
      const float4 v = vv[half];
      float d[10];
#pragma unroll
      for (int t = 0; t < 9; ++t) d[t] = v.x * wv[t].x + v.y * wv[t].y + v.z * wv[t].z + v.w * wv[t].w;
      d[9] = 0.f;
      float u[6];
#pragma unroll
      for (int i = 0; i < 5; ++i)
        u[i] = (b4 ? d[5 + i] : d[i]) + __shfl_xor_sync(0xffffffffu, b4 ? d[i] : d[5 + i], 16);
      u[5] = 0.f;
      float q[4];
#pragma unroll
      for (int i = 0; i < 3; ++i)
        q[i] = (b3 ? u[3 + i] : u[i]) + __shfl_xor_sync(0xffffffffu, b3 ? u[i] : u[3 + i], 8);
      q[3] = 0.f;
      float z2[2];
#pragma unroll
      for (int i = 0; i < 2; ++i)
        z2[i] = (b2 ? q[2 + i] : q[i]) + __shfl_xor_sync(0xffffffffu, b2 ? q[i] : q[2 + i], 4);
      float z = (b1 ? z2[1] : z2[0]) + __shfl_xor_sync(0xffffffffu, b1 ? z2[0] : z2[1], 2);
      z += __shfl_xor_sync(0xffffffffu, z, 1);
      const int pp = p + 8 * half;
      if (writer && pp < HW) part[pp * 9 + tap] = z;
    }
  }
  __syncthreads();
  const float b = bias[0];
  for (int p = threadIdx.x; p < HW; p += blockDim.x) {
    const int y = p / W, xx = p - y * W;
    float s = b;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int yy = y + kh - 1, xw = xx + kw - 1;
        if (yy >= 0 && yy < H && xw >= 0 && xw < W) s += part[(yy * W + xw) * 9 + kh * 3 + kw];
      }
    out[static_cast<size_t>(n) * out_ld + p] = s;
  }
}

// ------------------------------------------------------------------------------------------------ conv_out backward
// dh[pix][c] = sum_tap dout[pix - tap + 1] * w[tap][c];  dw[tap][c] += sum_pix h[pix][c] * dout[pix - tap + 1];
// db += sum dout.  One CTA per image; dout of the image lives in smem.
__global__ void __launch_bounds__(256)
conv_out_bwd_kernel(const float* __restrict__ h, const float* __restrict__ w, const float* __restrict__ dout,
                    float* __restrict__ dh, float* __restrict__ dw, float* __restrict__ db, int H, int W, int round_dh,
                    int out_ld, int accumulate) {
  extern __shared__ float sm[];  // [(H+2)(W+2)] zero-padded dout, then [8][9][128] dw partials
  const int n = blockIdx.x;
  const int HW = H * W, Wp = W + 2;
  const int padded = (H + 2) * Wp;
  float* sd = sm;
  float* sdw = sm + ((padded + 3) & ~3);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float bsum = 0.f;
  for (int i = threadIdx.x; i < padded; i += blockDim.x) {
    const int yy = i / Wp - 1, xx = i - (yy + 1) * Wp - 1;
    float v = 0.f;
    if (yy >= 0 && yy < H && xx >= 0 && xx < W) v = dout[static_cast<size_t>(n) * out_ld + yy * W + xx];
    sd[i] = v;
    bsum += v;
  }
  __syncthreads();
  float4 wv[9], aw[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    wv[t] = *reinterpret_cast<const float4*>(w + t * 128 + lane * 4);
    aw[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const size_t base = static_cast<size_t>(n) * HW * 128 + lane * 4;
  for (int p = warp; p < HW; p += 32) {  // four pixels per trip, their 512 B rows of h requested together
    float4 vv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      vv[u] = (p + 8 * u < HW) ? __ldg(reinterpret_cast<const float4*>(h + base + static_cast<size_t>(p + 8 * u) * 128))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pp = p + 8 * u;
      if (pp >= HW) break;
      const float4 v = vv[u];
      const int y = pp / W, xx = pp - y * W;
      const float* sp = sd + (y + 2) * Wp + xx + 2;  // padded position of (y + 1, xx + 1): tap (kh, kw) reads sp[-kh*Wp - kw]
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const float d = sp[-kh * Wp - kw];
          const int t = kh * 3 + kw;
          g.x += d * wv[t].x; g.y += d * wv[t].y; g.z += d * wv[t].z; g.w += d * wv[t].w;
          aw[t].x += d * v.x; aw[t].y += d * v.y; aw[t].z += d * v.z; aw[t].w += d * v.w;
        }
      if (accumulate) {  // image channels beyond the first add onto dh (3-channel extension)
        const float4 p4 = *reinterpret_cast<const float4*>(dh + base + static_cast<size_t>(pp) * 128);
        g.x += p4.x; g.y += p4.y; g.z += p4.z; g.w += p4.w;
      }
      if (round_dh) g = round_tf32(g);
      *reinterpret_cast<float4*>(dh + base + static_cast<size_t>(pp) * 128) = g;
    }
  }
#pragma unroll
  for (int t = 0; t < 9; ++t) *reinterpret_cast<float4*>(sdw + (warp * 9 + t) * 128 + lane * 4) = aw[t];
  __syncthreads();
  for (int i = threadIdx.x; i < 9 * 128; i += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int ww = 0; ww < 8; ++ww) s += sdw[ww * 9 * 128 + i];
    atomicAdd(&dw[i], s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bsum += __shfl_xor_sync(0xffffffffu, bsum, o);
  if (lane == 0) atomicAdd(db, bsum);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_conv_in_fwd(const b200_conv_in_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->x || !a->w || !a->bias || !a->out) return fail_arg("conv_in_fwd: null pointer");
  size_t blocks = (static_cast<size_t>(a->N) * a->H + 7) / 8;  // one warp per image row
  if (blocks > 148 * 8) blocks = 148 * 8;
  conv_in_fwd_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      a->x, a->w, a->bias, a->out, a->N, a->H, a->W, a->round_out, a->x_ld ? a->x_ld : a->H * a->W, a->w_co_ld ? a->w_co_ld : 9,
      a->w_tap_ld ? a->w_tap_ld : 1, a->accumulate);
  return after_launch("conv_in_fwd_kernel");
}

extern "C" int b200_conv_in_bwd(const b200_conv_in_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->x || !a->dy || !a->dw) return fail_arg("conv_in_bwd: null pointer");
  const int HW = a->H * a->W;
  int chunks = (592 + a->N - 1) / a->N;
  if (chunks > (HW + 63) / 64) chunks = (HW + 63) / 64;
  if (chunks < 1) chunks = 1;
  const int ppc = (HW + chunks - 1) / chunks;
  const size_t img_smem = static_cast<size_t>(a->H + 2) * (a->W + 2) * sizeof(float);
  if (img_smem > 64 * 1024) return fail_arg("conv_in_bwd: image too large for the shared-memory copy");
  static bool in_bwd_attr = false;
  if (!in_bwd_attr) {
    cudaError_t e = cudaFuncSetAttribute(conv_in_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_in_bwd)");
    in_bwd_attr = true;
  }
  conv_in_bwd_kernel<<<dim3(chunks, a->N), 256, img_smem, static_cast<cudaStream_t>(stream)>>>(
      a->x, a->dy, a->dw, a->db, a->H, a->W, ppc, a->x_ld ? a->x_ld : HW, a->w_co_ld ? a->w_co_ld : 9, a->w_tap_ld ? a->w_tap_ld : 1);
  return after_launch("conv_in_bwd_kernel");
}

extern "C" int b200_conv_out_fwd(const b200_conv_out_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->h || !a->w || !a->bias || !a->out) return fail_arg("conv_out_fwd: null pointer");
  const int HW = a->H * a->W;
  if (HW > 4096) return fail_arg("conv_out_fwd: H*W must be <= 4096");
  const size_t smem = static_cast<size_t>(HW) * 9 * sizeof(float);
  static bool fwd_attr_set = false;
  if (!fwd_attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_out_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4096 * 9 * 4);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_out_fwd)");
    fwd_attr_set = true;
  }
  conv_out_fwd_kernel<<<a->N, 256, smem, static_cast<cudaStream_t>(stream)>>>(a->h, a->w, a->bias, a->out, a->H, a->W,
                                                                              a->out_ld ? a->out_ld : HW);
  return after_launch("conv_out_fwd_kernel");
}

extern "C" int b200_conv_out_bwd(const b200_conv_out_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->h || !a->w || !a->dout || !a->dh || !a->dw || !a->db) return fail_arg("conv_out_bwd: null pointer");
  const int HW = a->H * a->W;
  if (HW > 4096 || a->H > 64 || a->W > 64) return fail_arg("conv_out_bwd: H and W must be <= 64");
  const size_t smem = (static_cast<size_t>(((a->H + 2) * (a->W + 2) + 3) & ~3) + 8 * 9 * 128) * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_out_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (66 * 66 + 4 + 8 * 9 * 128) * 4);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_out_bwd)");
    attr_set = true;
  }
  conv_out_bwd_kernel<<<a->N, 256, smem, static_cast<cudaStream_t>(stream)>>>(a->h, a->w, a->dout, a->dh, a->dw, a->db,
                                                                              a->H, a->W, a->round_dh, a->out_ld ? a->out_ld : HW,
                                                                              a->accumulate_dh);
  return after_launch("conv_out_bwd_kernel");
}
