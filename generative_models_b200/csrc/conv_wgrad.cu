// Convolution weight gradient on tcgen05, TF32 inputs / fp32 accumulate, split over the pixel reduction.
//
//   dW[co, tap, ci] = sum_pixels dY[pixel, co] * X[pixel shifted by tap, ci]
//   GEMM view per (tap, 128-wide ci block):  D[128 co, 128 ci] = A^T[K = pixels, 128 co] * B[K = pixels, 128 ci]
//   Both operands are "MN-major" (the reduction index — pixels — is the slow dimension of the NHWC rows), which
//   tcgen05 takes directly from 128B-swizzled smem tiles: dY rows come from a 2-D TMA tile, shifted X rows from
//   the same im2col TMA map the forward uses (border pixels arrive as zeros).
//   A CTA owns one filter ROW (kh) of one 128-channel ci block and a slice of the pixels: every pipeline stage loads
//   the dY tile once and the three kw-shifted X tiles, and accumulates three 128x128 tiles in TMEM (384 columns).
//   Sharing dY across the three taps cuts the L2->SM operand feed from 32 KB to 21 KB per 128x128x32 MMA block, which
//   is what bound the one-tap-per-CTA version (ncu: L2->SM fabric saturated, tensor pipe 40 % busy).
//   grid = (pixel slices, kh * ci blocks) sized to one CTA per SM; the epilogue adds the fp32 tiles into dW with
//   red.global.add.v4.f32 (dW is zeroed once per step by the caller) — a quarter of the scalar atomics.
#include "host_common.h"
#include "ptx.cuh"

#include <stdlib.h>

#include "../../include/b200ddpm.h"

namespace b200 {

constexpr int kWgStages = 3;
constexpr int kWgTileBytes = 16384;  // 4 boxes of (32 pixels x 128 B)
constexpr int kWgPix = 32;           // pixels (reduction elements) per pipeline stage, fp32/TF32 operands
constexpr int kWgPixBf = 64;         // ... bf16 operands: same 16 KB tiles, 2 boxes of (64 pixels x 128 B = 64 channels)
constexpr int wg_smem_bytes(int taps) { return kWgStages * (1 + taps) * kWgTileBytes + 1024 + 256; }

struct WgradK {
  float* dw;
  int M_total, HoWo, Wo;
  int stride, pad;
  int S;      // filter width (= height)
  int taps;   // filter taps handled by one CTA: S (one filter row)
  int C1, Cin;
  int steps_total, steps_per_slice;
  int dy3d;   // tmDY is the 3-D channel-block view: the whole dY tile is one TMA instruction
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <bool BF16>
__global__ void __launch_bounds__(192)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                       const __grid_constant__ CUtensorMap tmX2, const WgradK a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stage_bytes = (1 + a.taps) * kWgTileBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kWgStages * stage_bytes);
  uint64_t* empty = full + kWgStages;
  uint64_t* tmem_full = empty + kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int PIX = BF16 ? kWgPixBf : kWgPix;  // pixels per stage
  constexpr int CB = BF16 ? 64 : 32;             // channels per TMA box (128 B)
  constexpr int NB = 128 / CB;                   // boxes per 128-channel tile
  constexpr int BOXB = kWgTileBytes / NB;

  const int step_begin = blockIdx.x * a.steps_per_slice;
  const int step_end = min(step_begin + a.steps_per_slice, a.steps_total);
  const int num_it = step_end - step_begin;
  if (num_it <= 0) return;  // uniform for the whole CTA

  const int ci_blocks = a.Cin / 128;
  const int kh = blockIdx.y / ci_blocks;
  const int cib = blockIdx.y - kh * ci_blocks;
  const uint32_t tmem_cols = a.taps > 2 ? 512u : (a.taps == 2 ? 256u : 128u);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmX2);
    for (int s = 0; s < kWgStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      const int c_glob = cib * 128;
      const CUtensorMap* xm = (c_glob < a.C1) ? &tmX : &tmX2;
      const int c_src = (c_glob < a.C1) ? c_glob : c_glob - a.C1;
      for (int it = 0; it < num_it; ++it) {
        const int s = it % kWgStages;
        const uint32_t ph = (it / kWgStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], stage_bytes);
        const int pix0 = (step_begin + it) * PIX;
        const int n = pix0 / a.HoWo;
        const int rem = pix0 - n * a.HoWo;
        const int p = rem / a.Wo;
        const int q = rem - p * a.Wo;
        uint8_t* sA = smem + s * stage_bytes;
#pragma unroll
        for (int j = 0; j < NB; ++j) {
          if (a.dy3d) {  // all channel blocks of the dY tile in one instruction (tmDY is the 3-D block view)
            tma_load_3d(sA, &tmDY, &full[s], 0, pix0, 0);
            break;
          }
          tma_load_2d(sA + j * BOXB, &tmDY, &full[s], j * CB, pix0);
        }
        for (int kw = 0; kw < a.taps; ++kw) {
          uint8_t* sB = sA + (1 + kw) * kWgTileBytes;
#pragma unroll
          for (int j = 0; j < NB; ++j)
            tma_load_im2col_4d(sB + j * BOXB, xm, &full[s], c_src + j * CB, q * a.stride - a.pad, p * a.stride - a.pad, n,
                               (uint16_t)kw, (uint16_t)kh);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc1 = make_idesc(BF16 ? 1u : 2u, 128u, 128u, 1u, 1u);
      // Two kw taps in ONE instruction: their X tiles are adjacent in the stage buffer and a tile is a run of channel
      // boxes with a constant pitch (the descriptor's LBO), so taps kw and kw+1 form a single N = 256 operand
      // (8 boxes of 32 fp32 / 4 boxes of 64 bf16 channels) whose result lands in TMEM columns [kw*128, kw*128+256) —
      // exactly where the two N = 128 instructions would put it.  One 128x256 instruction reads the dY operand once
      // for both taps: 20 KB of shared-memory operand traffic per 3 taps instead of 24 KB.
      const uint32_t idesc2 = make_idesc(BF16 ? 1u : 2u, 128u, 256u, 1u, 1u);
      for (int it = 0; it < num_it; ++it) {
        const int s = it % kWgStages;
        const uint32_t ph = (it / kWgStages) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * stage_bytes);
        for (int kw = 0; kw < a.taps; kw += 2) {
          const uint32_t b_base = a_base + (1 + kw) * kWgTileBytes;
          const uint32_t idesc = (kw + 1 < a.taps) ? idesc2 : idesc1;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // MN-major operands: TF32 = 32 B-atom layout (4 boxes of 32 channels, K-step = 8 pixel rows);
            // bf16 = SWIZZLE_128B (2 boxes of 64 channels, K-step = 16 pixel rows = two 8-row groups 1024 B apart)
            if (BF16) {
              const uint64_t adesc = make_smem_desc_sw128(a_base + k * 2048, 8192, 1024, 2);
              const uint64_t bdesc = make_smem_desc_sw128(b_base + k * 2048, 8192, 1024, 2);
              mma_f16_ss(tmem_base + kw * 128, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
            } else {
              const uint64_t adesc = make_smem_desc_sw128(a_base + k * 1024, 4096, 512, 1);
              const uint64_t bdesc = make_smem_desc_sw128(b_base + k * 1024, 4096, 512, 1);
              mma_tf32_ss(tmem_base + kw * 128, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
            }
          }
        }
        tc_commit(&empty[s]);
      }
      tc_commit(tmem_full);
    }
  } else {
    const int quarter = warp & 3;
    const int co = quarter * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row_len = a.S * a.S * a.Cin;
    float* drow = a.dw + static_cast<size_t>(co) * row_len + (kh * a.S) * a.Cin + cib * 128;
    for (int kw = 0; kw < a.taps; ++kw) {
#pragma unroll 1
      for (int cc = 0; cc < 4; ++cc) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + kw * 128 + cc * 32, r);
        tmem_ld_wait();
        float* d = drow + kw * a.Cin + cc * 32;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          red_add_v4(d + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                     __uint_as_float(r[j + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_conv_wgrad(const b200_wgrad_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws;
  (void)ws_bytes;
  if (!a || !a->x || !a->dy || !a->dw) return fail_arg("conv_wgrad: null pointer");
  if (a->C1 <= 0 || a->C1 % 128 || a->C2 < 0 || a->C2 % 128) return fail_arg("conv_wgrad: C1/C2 must be multiples of 128");
  if (a->C2 > 0 && !a->x2) return fail_arg("conv_wgrad: C2 > 0 needs x2");
  if (a->R != a->S || (a->R != 1 && a->R != 3)) return fail_arg("conv_wgrad: only 1x1 and 3x3 filters");
  if (a->stride != 1 && a->stride != 2) return fail_arg("conv_wgrad: stride must be 1 or 2");
  const int ho = (a->H + 2 * a->pad - a->R) / a->stride + 1, wo = (a->W + 2 * a->pad - a->S) / a->stride + 1;
  if (ho != a->Ho || wo != a->Wo) return fail_arg("conv_wgrad: Ho/Wo do not match the geometry");
  static int num_sms = 0;
  if (!num_sms) {
    cudaError_t e = cudaFuncSetAttribute(conv_wgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, wg_smem_bytes(3));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(conv_wgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, wg_smem_bytes(3));
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_wgrad)");
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int Cin = a->C1 + a->C2;
  const int M_total = a->N * a->Ho * a->Wo;
  const int lower = -a->pad, upper = a->pad - (a->R - 1);

  if (a->op_dtype != 0 && a->op_dtype != 1) return fail_arg("conv_wgrad: op_dtype must be 0 (fp32/TF32) or 1 (bf16)");
  const bool bf = a->op_dtype == 1;
  const int eb = bf ? 2 : 4, sw32 = bf ? 0 : 1;
  const uint32_t cb = bf ? 64 : 32, pixs = bf ? kWgPixBf : kWgPix;
  CUtensorMap tmDY, tmX, tmX2;
  static int dy3d = -1;  // developer knob B200_WGRAD_DY3D=0: one 2-D box per channel block of the dY tile
  if (dy3d < 0) {
    const char* e = getenv("B200_WGRAD_DY3D");
    dy3d = e ? atoi(e) : 1;
  }
  int r = dy3d ? get_tmap_blocks3d(a->dy, M_total, a->dy_ld ? a->dy_ld : 128, cb, pixs, 128 / cb, eb, sw32, &tmDY)
               : get_tmap_2d(a->dy, M_total, 128, a->dy_ld ? a->dy_ld : 128, cb, pixs, eb, sw32, &tmDY);
  if (r) return r;
  r = get_tmap_im2col(a->x, a->N, a->H, a->W, a->C1, a->x_ld ? a->x_ld : a->C1, lower, upper, a->stride, cb, pixs, eb, sw32, &tmX);
  if (r) return r;
  if (a->C2 > 0) {
    r = get_tmap_im2col(a->x2, a->N, a->H, a->W, a->C2, a->x2_ld ? a->x2_ld : a->C2, lower, upper, a->stride, cb, pixs, eb, sw32, &tmX2);
    if (r) return r;
  } else {
    tmX2 = tmX;
  }

  const int tiles = a->R * (Cin / 128);  // one CTA per (filter row, 128-channel ci block, pixel slice)
  WgradK k;
  k.dw = a->dw;
  k.dy3d = dy3d;
  k.M_total = M_total;
  k.HoWo = a->Ho * a->Wo;
  k.Wo = a->Wo;
  k.stride = a->stride;
  k.pad = a->pad;
  k.S = a->S;
  k.taps = a->S;
  k.C1 = a->C1;
  k.Cin = Cin;
  k.steps_total = (M_total + static_cast<int>(pixs) - 1) / static_cast<int>(pixs);
  int slices = a->k_slices;
  if (slices <= 0) {
    // one CTA per SM for 3-tap CTAs (they own all 512 TMEM columns and 192 KB of smem), two for 1x1 filters
    slices = (num_sms * (k.taps > 1 ? 1 : 2)) / tiles;
    const int max_slices = (k.steps_total + 7) / 8;  // at least 8 pipeline steps per CTA
    if (slices > max_slices) slices = max_slices;
    if (slices < 1) slices = 1;
  }
  k.steps_per_slice = (k.steps_total + slices - 1) / slices;
  slices = (k.steps_total + k.steps_per_slice - 1) / k.steps_per_slice;

  dim3 grid(slices, tiles);
  if (bf) {
    conv_wgrad_kernel<true><<<grid, 192, wg_smem_bytes(k.taps), static_cast<cudaStream_t>(stream)>>>(tmDY, tmX, tmX2, k);
    return after_launch("conv_wgrad_kernel<bf16>");
  }
  conv_wgrad_kernel<false><<<grid, 192, wg_smem_bytes(k.taps), static_cast<cudaStream_t>(stream)>>>(tmDY, tmX, tmX2, k);
  return after_launch("conv_wgrad_kernel<tf32>");
}
