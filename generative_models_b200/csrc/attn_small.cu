// Self-attention over the tokens of the lowest-resolution UNet level (7x7 = 49 tokens at 28x28 input, 16x16 = 256 at
// 64x64), one head of 128 channels — an EXTENSION: the reference's SimpleUnet has no attention (simple_unet.py:16-72);
// the north star and BASELINE.json configs[4] ask for it on the tensor cores.  Built from two pieces:
//
//   b200_bgemm         C[b] = alpha * op(A[b]) * op(B[b])^T on tcgen05 (kind::tf32, fp32 accumulate in TMEM), one CTA per
//                      (128-row tile of M, batch element), TMA-fed, any of the four operand orientations:
//                        A: [M][K] rows ("K-major")  or  [K][M] rows ("MN-major")
//                        B: [N][K] rows ("K-major")  or  [K][N] rows ("MN-major")
//                      so that S = Q K^T, O = P V, dV = P^T dO, dP = dO V^T, dQ = dS K and dK = dS^T Q all run on the
//                      same kernel without a transpose pass.  Ragged sizes (49 tokens) ride on TMA's out-of-bounds zero
//                      fill; rows / columns past M / N are never stored.
//   b200_softmax_fwd / b200_softmax_bwd   row softmax of S and its backward (a warp per row, warp-shuffle reductions),
//                      writing TF32-rounded probabilities / score gradients (they are the next GEMM's operands) and zeros
//                      in the pad columns.
//
// The attention matrices are materialised in HBM ([B][T][Tp] fp32): at these sizes (0.3 % of the UNet's FLOPs) a fused
// flash-style kernel buys nothing, and the backward needs P anyway.
#include "host_common.h"
#include "ptx.cuh"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

namespace b200 {

constexpr int kBgStages = 3;
constexpr int kBgATile = 16384;             // 128 rows x 128 B (K-major) or 4 boxes of 32 K-rows x 128 B (MN-major)
constexpr int kBgBTile = 32768;             // up to 256 rows x 128 B / 8 boxes
constexpr int kBgStage = kBgATile + kBgBTile;
constexpr int kBgSmem = kBgStages * kBgStage + 1024 + 256;

struct BgemmK {
  float* C;
  long long strideC;
  int ldc;
  int M, N, K;
  int a_mn, b_mn;
  int n_tile;  // UMMA N: N rounded up to 32
  float alpha;
  int round_out;
};

__global__ void __launch_bounds__(192)
bgemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const BgemmK a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kBgStages * kBgStage);
  uint64_t* empty = full + kBgStages;
  uint64_t* tmem_full = empty + kBgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * 128;
  const int b = blockIdx.y;
  const int num_it = (a.K + 31) / 32;
  const uint32_t stage_tx = kBgATile + static_cast<uint32_t>(a.n_tile) * 128u;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kBgStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < num_it; ++it) {
        const int s = it % kBgStages;
        mbar_wait(&empty[s], ((it / kBgStages) & 1) ^ 1);
        mbar_expect_tx(&full[s], stage_tx);
        uint8_t* sA = smem + s * kBgStage;
        uint8_t* sB = sA + kBgATile;
        const int k0 = it * 32;
        if (!a.a_mn) {
          tma_load_3d(sA, &tmA, &full[s], k0, m0, b);  // {K, M, batch}: 32 k x 128 rows
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) tma_load_3d(sA + j * 4096, &tmA, &full[s], m0 + j * 32, k0, b);  // {M, K, batch}
        }
        if (!a.b_mn) {
          tma_load_3d(sB, &tmB, &full[s], k0, 0, b);   // {K, N, batch}: 32 k x n_tile rows
        } else {
          for (int j = 0; j < a.n_tile / 32; ++j) tma_load_3d(sB + j * 4096, &tmB, &full[s], j * 32, k0, b);  // {N, K, batch}
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc(2u, 128u, static_cast<uint32_t>(a.n_tile), a.a_mn ? 1u : 0u, a.b_mn ? 1u : 0u);
      for (int it = 0; it < num_it; ++it) {
        const int s = it % kBgStages;
        mbar_wait(&full[s], (it / kBgStages) & 1);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * kBgStage);
        const uint32_t b_base = a_base + kBgATile;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          // K-major: 128B-swizzled rows, a K-step is 32 B along the row; MN-major TF32: 32 B-atom layout, boxes of
          // 32 MN-elements x 32 K-rows 4 KB apart (LBO), a K-step is 8 rows = 1 KB (the layouts of conv_wgrad.cu)
          const uint64_t adesc = a.a_mn ? make_smem_desc_sw128(a_base + k * 1024, 4096, 512, 1)
                                        : make_smem_desc_sw128(a_base + k * 32, 16, 1024);
          const uint64_t bdesc = a.b_mn ? make_smem_desc_sw128(b_base + k * 1024, 4096, 512, 1)
                                        : make_smem_desc_sw128(b_base + k * 32, 16, 1024);
          mma_tf32_ss(tmem_base, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
        }
        tc_commit(&empty[s]);
      }
      tc_commit(tmem_full);
    }
  } else {
    // epilogue: thread = output row (TMEM lane), 32 columns per tcgen05.ld
    const int quarter = warp & 3;
    const int row = m0 + quarter * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    float* crow = a.C + static_cast<long long>(b) * a.strideC + static_cast<long long>(row) * a.ldc;
    const bool vec = (a.N % 4 == 0) && (a.ldc % 4 == 0);
#pragma unroll 1
    for (int cc = 0; cc < a.n_tile / 32; ++cc) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + cc * 32, r);
      tmem_ld_wait();
      if (row >= a.M) continue;
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        v[j] = __uint_as_float(r[j]) * a.alpha;
        if (a.round_out) v[j] = round_tf32(v[j]);
      }
      const int c0 = cc * 32;
      if (vec) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          if (c0 + j < a.N) *reinterpret_cast<float4*>(crow + c0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (c0 + j < a.N) crow[c0 + j] = v[j];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// ---------------------------------------------------------------------------------------------------- softmax
// One warp per row of T <= 256 scores (row pitch ld >= T); p = softmax(s) rounded to TF32, zeros in [T, ld).
__global__ void __launch_bounds__(256)
softmax_fwd_kernel(const float* __restrict__ s, float* __restrict__ p, long long rows, int T, int ld) {
  const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* sr = s + row * ld;
  float* pr = p + row * ld;
  float v[8];
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = lane + 32 * j;
    v[j] = c < T ? sr[c] : -INFINITY;
    mx = fmaxf(mx, v[j]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    v[j] = (lane + 32 * j < T) ? expf(v[j] - mx) : 0.f;
    sum += v[j];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float inv = 1.0f / sum;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = lane + 32 * j;
    if (c < ld) pr[c] = c < T ? round_tf32(v[j] * inv) : 0.f;
  }
}
// ds = p * (dp - sum_j dp_j p_j), rounded to TF32, zeros in [T, ld)
__global__ void __launch_bounds__(256)
softmax_bwd_kernel(const float* __restrict__ p, const float* __restrict__ dp, float* __restrict__ ds, long long rows, int T,
                   int ld) {
  const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* pr = p + row * ld;
  const float* dr = dp + row * ld;
  float* out = ds + row * ld;
  float pv[8], dv[8];
  float dot = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = lane + 32 * j;
    pv[j] = c < T ? pr[c] : 0.f;
    dv[j] = c < T ? dr[c] : 0.f;
    dot += pv[j] * dv[j];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = lane + 32 * j;
    if (c < ld) out[c] = c < T ? round_tf32(pv[j] * (dv[j] - dot)) : 0.f;
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_bgemm(const b200_bgemm_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->A || !a->B || !a->C) return fail_arg("bgemm: null pointer");
  if (a->M <= 0 || a->N <= 0 || a->K <= 0 || a->batch <= 0) return fail_arg("bgemm: empty problem");
  if (a->N > 256) return fail_arg("bgemm: N must be <= 256 (one UMMA tile)");
  if (a->lda % 4 || a->ldb % 4 || a->strideA % 4 || a->strideB % 4) return fail_arg("bgemm: lda / ldb / strides must be multiples of 4 floats (TMA: 16 B)");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(bgemm_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBgSmem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(bgemm)");
    attr_set = true;
  }
  const int n_tile = (a->N + 31) & ~31;
  CUtensorMap tmA, tmB;
  int r;
  // operand tensor maps, batch outermost; extents are the VALID sizes, so everything past them is zero-filled
  if (!a->a_mn) r = get_tmap_3d(a->A, a->K, a->M, a->batch, static_cast<uint64_t>(a->lda) * 4, static_cast<uint64_t>(a->strideA) * 4, 32, 128, 1, 4, 0, &tmA);
  else r = get_tmap_3d(a->A, a->M, a->K, a->batch, static_cast<uint64_t>(a->lda) * 4, static_cast<uint64_t>(a->strideA) * 4, 32, 32, 1, 4, 1, &tmA);
  if (r) return r;
  if (!a->b_mn) r = get_tmap_3d(a->B, a->K, a->N, a->batch, static_cast<uint64_t>(a->ldb) * 4, static_cast<uint64_t>(a->strideB) * 4, 32, n_tile, 1, 4, 0, &tmB);
  else r = get_tmap_3d(a->B, a->N, a->K, a->batch, static_cast<uint64_t>(a->ldb) * 4, static_cast<uint64_t>(a->strideB) * 4, 32, 32, 1, 4, 1, &tmB);
  if (r) return r;
  BgemmK k;
  k.C = a->C; k.strideC = a->strideC; k.ldc = a->ldc;
  k.M = a->M; k.N = a->N; k.K = a->K;
  k.a_mn = a->a_mn; k.b_mn = a->b_mn; k.n_tile = n_tile;
  k.alpha = a->alpha; k.round_out = a->round_out;
  dim3 grid((a->M + 127) / 128, a->batch);
  bgemm_tf32_kernel<<<grid, 192, kBgSmem, static_cast<cudaStream_t>(stream)>>>(tmA, tmB, k);
  return after_launch("bgemm_tf32_kernel");
}

extern "C" int b200_softmax_fwd(const b200_softmax_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->s || !a->p) return fail_arg("softmax_fwd: null pointer");
  if (a->T <= 0 || a->T > 256 || a->ld < a->T || a->ld > 256) return fail_arg("softmax_fwd: need 0 < T <= ld <= 256");
  const long long rows = a->rows;
  softmax_fwd_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(a->s, a->p, rows, a->T, a->ld);
  return after_launch("softmax_fwd_kernel");
}

extern "C" int b200_softmax_bwd(const b200_softmax_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws; (void)ws_bytes;
  if (!a || !a->p || !a->dp || !a->ds) return fail_arg("softmax_bwd: null pointer");
  if (a->T <= 0 || a->T > 256 || a->ld < a->T || a->ld > 256) return fail_arg("softmax_bwd: need 0 < T <= ld <= 256");
  const long long rows = a->rows;
  softmax_bwd_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(a->p, a->dp, a->ds, rows, a->T, a->ld);
  return after_launch("softmax_bwd_kernel");
}
