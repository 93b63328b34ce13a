// Implicit-GEMM convolution (fprop and stride-1 dgrad) on the 5th-gen tensor cores, TF32 inputs / fp32 accumulate.
//
//   GEMM view:  D[M = N*Ho*Wo pixels, 128] = A[M, K = R*S*Cin] * B[K, 128]
//   A is never materialised: one TMA *im2col* load per (filter tap, 32-channel chunk) gathers the 128 shifted
//   pixel rows of the NHWC activation straight into a 128B-swizzled smem tile (zero fill at the borders).
//   B (weights [Cout][kh][kw][Cin]) is a plain 2-D TMA tile: K-major for fprop, MN-major for dgrad, so the same
//   weight memory serves both directions without a transpose pass.
//   One elected thread issues tcgen05.mma (M128 N128 K8, kind::tf32) into a 128-column TMEM accumulator;
//   4 epilogue warps read it back with tcgen05.ld and fuse bias / timestep-embedding / residual adds.
//
// CTA = 192 threads: warp 0 TMA producer, warp 1 TMEM owner + MMA issuer, warps 2-5 epilogue.
// 3-stage smem ring (32 KB per stage) so two CTAs are co-resident per SM (one's epilogue overlaps the other's MMAs).
#include "host_common.h"
#include "ptx.cuh"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

#include <stdlib.h>

namespace b200 {

constexpr int kStages = 3;
constexpr int kTileBytes = 16384;                 // 128 rows x 128 B
constexpr int kStageBytes = 2 * kTileBytes;       // A + B
constexpr int kConvSmem = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

struct ConvK {
  float* out;
  uint16_t* out_bf;  // optional dense bf16 copy [M][128] of out (fprop)
  const float* bias;
  const float* emb;
  const float* residual;
  int M_total, HoWo, Wo;
  int stride, pad_eff;
  int R, S;
  int chunks1, chunks;  // 32-channel chunks of source 1 / of both sources
  int mode;
  int w_rowlen_tap;     // columns per tap in the weight matrix (= Cin of the weight tensor)
  int n_off;
  int out_ld, emb_ld, res_ld;
  int round_out;
  // mode 2 (backward-data of a stride-2 3x3/pad-1 conv): CTA (tile, class) = (blockIdx.x >> 2, blockIdx.x & 3) runs a
  // small R x S forward conv over dY (R = 1 + class/2, S = 1 + class%2) whose result row m = (n, i, j) lands on pixel
  // (2i + class/2, 2j + class%2) of the twice-as-large dX.  The four classes of a tile are neighbours in launch order:
  // one class alone touches 512 B of every KB of dX, i.e. half of the DRAM channels (measured 1.1 TB/s).
  int sub;
};

// output / residual row of GEMM row m
__device__ __forceinline__ size_t out_row(int sub, int sub_a, int sub_b, int Wo, int m) {
  if (!sub) return static_cast<size_t>(m);
  const int g = m / Wo, j = m - g * Wo;  // g = n * Ho + i
  return static_cast<size_t>(2 * g + sub_a) * (2 * Wo) + 2 * j + sub_b;
}

__global__ void __launch_bounds__(192)
conv_igemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                       const __grid_constant__ CUtensorMap tmB, const ConvK a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmA2);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int cls = a.sub ? (blockIdx.x & 3) : 0;
  const int sub_a = cls >> 1, sub_b = cls & 1;
  const int R = a.sub ? 1 + sub_a : a.R, S = a.sub ? 1 + sub_b : a.S;
  const int m0 = (a.sub ? (blockIdx.x >> 2) : blockIdx.x) * 128;
  const int num_it = R * S * a.chunks;

  if (warp == 0) {
    if (lane == 0) {
      // ---------------- TMA producer
      const int n = m0 / a.HoWo;
      const int rem = m0 - n * a.HoWo;
      const int p = rem / a.Wo;
      const int q = rem - p * a.Wo;
      const int w_base = q * a.stride - a.pad_eff;
      const int h_base = p * a.stride - a.pad_eff;
      int tap = 0, ch = 0;
      for (int it = 0; it < num_it; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], kStageBytes);
        const int kh = tap / S, kw = tap - kh * S;
        uint8_t* sA = smem + s * kStageBytes;
        uint8_t* sB = sA + kTileBytes;
        if (ch < a.chunks1)
          tma_load_im2col_4d(sA, &tmA, &full[s], ch * 32, w_base, h_base, n, (uint16_t)kw, (uint16_t)kh);
        else
          tma_load_im2col_4d(sA, &tmA2, &full[s], (ch - a.chunks1) * 32, w_base, h_base, n, (uint16_t)kw,
                             (uint16_t)kh);
        if (a.mode == 0) {
          // mode 2, per axis: dX[2i+1] = dY[i] w[2] + dY[i+1] w[0], dX[2i] = dY[i] w[1]; shadow tap = 8 - original tap
          const int wt = a.sub ? 8 - ((sub_a ? 2 - 2 * kh : 1) * 3 + (sub_b ? 2 - 2 * kw : 1)) : tap;
          tma_load_2d(sB, &tmB, &full[s], wt * a.w_rowlen_tap + ch * 32, 0);
        } else {
          const int wtap = (R - 1 - kh) * S + (S - 1 - kw);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            tma_load_2d(sB + j * 4096, &tmB, &full[s], wtap * a.w_rowlen_tap + a.n_off + j * 32, ch * 32);
        }
        if (++ch == a.chunks) {
          ch = 0;
          ++tap;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ---------------- MMA issuer
      const uint32_t idesc = make_idesc(2u, 128u, 128u, 0u, a.mode == 0 ? 0u : 1u);
      for (int it = 0; it < num_it; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * kStageBytes);
        const uint32_t b_base = a_base + kTileBytes;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t adesc = make_smem_desc_sw128(a_base + k * 32, 16, 1024);
          const uint64_t bdesc = (a.mode == 0) ? make_smem_desc_sw128(b_base + k * 32, 16, 1024)
                                               : make_smem_desc_sw128(b_base + k * 1024, 4096, 512, 1);
          mma_tf32_ss(tmem_base, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
        }
        tc_commit(&empty[s]);
      }
      tc_commit(tmem_full);
    }
  } else {
    // ---------------- epilogue: TMEM -> registers -> smem (transpose) -> (+bias +emb +residual) -> global
    // tcgen05.ld hands a thread one pixel ROW of the tile; writing (and reading the residual) from that layout touches
    // 16 B per 512 B-strided row per lane — half-empty sectors on every access.  The pipeline's stage buffers are idle
    // once the accumulator is complete, so each warp parks its 32 rows there (row pitch 132 floats: conflict-free
    // float4 accesses) and then streams them out one row per instruction: 512 contiguous bytes per warp access for
    // the output, the residual and the embedding row alike.
    const int quarter = warp & 3;  // TMEM lane quarter this warp may read
    if (a.residual) {
      // these warps idle through the main loop: pull the 32 residual rows this warp will add into L2 meanwhile, so the
      // epilogue's loads see L2 latency (the residual add cost 26 us on a 66 us 14x14 layer)
      const int m = m0 + quarter * 32 + lane;
      if (m < a.M_total) {
        const float* rp = a.residual + out_row(a.sub, sub_a, sub_b, a.Wo, m) * a.res_ld;
#pragma unroll
        for (int k = 0; k < 4; ++k) asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + k * 32));
      }
    }
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    constexpr int kPitch = 132;
    // this warp's 32 rows, addressed in the shared window (generic-pointer accesses compile to LD.E / ST.E)
    const uint32_t stg = smem_u32(smem) + quarter * 32 * kPitch * 4;
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + cc * 32, r);
      tmem_ld_wait();
      const uint32_t srow = stg + (lane * kPitch + cc * 32) * 4;
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(srow + j * 4), "r"(r[j]), "r"(r[j + 1]),
                     "r"(r[j + 2]), "r"(r[j + 3])
                     : "memory");
    }
    __syncwarp();
    const int c = lane * 4;
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.bias) b4 = *reinterpret_cast<const float4*>(a.bias + c);
    const int m_first = m0 + quarter * 32;
    const int nrows = min(32, a.M_total - m_first);
#pragma unroll 1
    for (int r0 = 0; r0 < nrows; r0 += 8) {  // batches of 8 rows: 8 (+8) independent 512 B loads in flight per warp
      float4 add[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        add[j] = b4;
        const int m = min(m_first + r0 + j, a.M_total - 1);
        if (a.residual) {
          const float4 e = __ldg(reinterpret_cast<const float4*>(a.residual + out_row(a.sub, sub_a, sub_b, a.Wo, m) * a.res_ld + c));
          add[j].x += e.x; add[j].y += e.y; add[j].z += e.z; add[j].w += e.w;
        }
        if (a.emb) {
          const float4 e = __ldg(reinterpret_cast<const float4*>(a.emb + static_cast<size_t>(m / a.HoWo) * a.emb_ld + c));
          add[j].x += e.x; add[j].y += e.y; add[j].z += e.z; add[j].w += e.w;
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (r0 + j < nrows) {
          float4 v;
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                       : "r"(stg + ((r0 + j) * kPitch + c) * 4)
                       : "memory");
          v.x += add[j].x; v.y += add[j].y; v.z += add[j].z; v.w += add[j].w;
          if (a.round_out) v = round_tf32(v);
          *reinterpret_cast<float4*>(a.out + out_row(a.sub, sub_a, sub_b, a.Wo, m_first + r0 + j) * a.out_ld + c) = v;
          if (a.out_bf) *reinterpret_cast<uint2*>(a.out_bf + static_cast<size_t>(m_first + r0 + j) * 128 + c) = pack_bf16x4(v);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

}  // namespace b200

using namespace b200;

int conv_strip_launch(const b200_conv_args* a, void* stream);  // conv_strip.cu
int conv_strip_applicable(const b200_conv_args* a);

static int strip_enabled() {
  static int use_strip = -1;
  if (use_strip < 0) {
    const char* e = getenv("B200_CONV_STRIP");
    use_strip = e ? atoi(e) : 1;
  }
  return use_strip;
}
static bool strip_shape(const b200_conv_args* a) {
  return a->stride == 1 && ((a->R == 3 && a->pad == 1) || (a->R == 1 && a->pad == 0));
}

extern "C" int b200_conv_kernel_of(const b200_conv_args* a) {
  if (!a) return fail_arg("conv_kernel_of: null args");
  return (strip_enabled() && strip_shape(a) && conv_strip_applicable(a)) ? 1 : 0;
}

// mode 2: backward-data of a stride-2 3x3/pad-1 convolution WITHOUT zero-stuffing dY.  x = dY [N, H, W, 128],
// w = the transposed, tap-flipped weight shadow [128 ci][9][128 co] that b200_conv_weights_transpose writes,
// out = dX [N, 2H, 2W, 128] (+ residual in the same layout).  dX[2i+a, 2j+b] only receives the filter taps whose parity
// matches: (a, b) = (0, 0): tap (1, 1); (0, 1): (1, 0) and (1, 2); (1, 0): (0, 1) and (2, 1); (1, 1): the four corners
// — nine taps in all instead of the 36 tap evaluations of the zero-stuffed stride-1 formulation.  Each parity class is
// a 1x1 / 1x2 / 2x1 / 2x2 forward conv over dY (far-edge OOB = zero) whose rows scatter to the class's pixels of dX;
// one launch, the four classes of a 128-row tile in neighbouring CTAs.
static int conv_dgrad_stride2(const b200_conv_args* a, void* stream) {
  if (a->R != 3 || a->S != 3 || a->pad != 1 || a->stride != 2) return fail_arg("conv_igemm mode 2: 3x3 / pad 1 / stride 2 only");
  if (a->C1 != 128 || a->C2 != 0 || a->w_cout != 128 || a->w_cin != 128)
    return fail_arg("conv_igemm mode 2: dY and dX have 128 channels, w is the [128][9][128] transposed shadow");
  if (a->Ho != 2 * a->H || a->Wo != 2 * a->W) return fail_arg("conv_igemm mode 2: out must be twice the size of dY");
  if (a->bias || a->emb || a->skip_x || a->out_dtype || a->op_dtype || a->out_bf16)
    return fail_arg("conv_igemm mode 2: plain fp32 output (+ residual) only");
  if (a->out_ld % 4 || (a->residual && a->res_ld % 4)) return fail_arg("conv_igemm: leading dimensions must be multiples of 4 floats");
  CUtensorMap tmA, tmB;
  int r = get_tmap_im2col(a->x, a->N, a->H, a->W, 128, a->x_ld ? a->x_ld : 128, 0, 0, 1, 32, 128, 4, 0, &tmA);
  if (r) return r;
  r = get_tmap_2d(a->w, 128, 9 * 128, 9 * 128, 32, 128, 4, 0, &tmB);
  if (r) return r;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_igemm_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kConvSmem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_igemm)");
    attr_set = true;
  }
  ConvK k = {};
  k.out = static_cast<float*>(a->out);
  k.residual = a->residual;
  k.M_total = a->N * a->H * a->W;
  k.HoWo = a->H * a->W;
  k.Wo = a->W;
  k.stride = 1;
  k.pad_eff = 0;
  k.R = k.S = 2;  // per class in the kernel
  k.chunks1 = k.chunks = 4;
  k.mode = 0;
  k.w_rowlen_tap = 128;
  k.out_ld = a->out_ld;
  k.res_ld = a->res_ld;
  k.round_out = a->round_out;
  k.sub = 1;
  const int grid = 4 * ((k.M_total + 127) / 128);
  conv_igemm_tf32_kernel<<<grid, 192, kConvSmem, static_cast<cudaStream_t>(stream)>>>(tmA, tmA, tmB, k);
  if (int rc = after_launch("conv_igemm_tf32_kernel")) return rc;
  return 0;
}

extern "C" int b200_conv_igemm(const b200_conv_args* a, void* ws, size_t ws_bytes, void* stream) {
  (void)ws;
  (void)ws_bytes;
  if (!a || !a->x || !a->w || !a->out) return fail_arg("conv_igemm: null pointer");
  if (a->C1 <= 0 || a->C1 % 32 || a->C2 < 0 || a->C2 % 32) return fail_arg("conv_igemm: C1/C2 must be multiples of 32");
  if (a->C2 > 0 && !a->x2) return fail_arg("conv_igemm: C2 > 0 needs x2");
  if (a->R != a->S || (a->R != 1 && a->R != 3)) return fail_arg("conv_igemm: only 1x1 and 3x3 filters");
  if (a->mode == 2) return conv_dgrad_stride2(a, stream);
  if (a->mode != 0 && a->mode != 1) return fail_arg("conv_igemm: mode must be 0 (fprop), 1 (dgrad) or 2 (stride-2 dgrad)");
  if (a->op_dtype != 0 && a->op_dtype != 1) return fail_arg("conv_igemm: op_dtype must be 0 (fp32/TF32) or 1 (bf16)");
  if (a->out_ld % 4 || (a->residual && a->res_ld % 4) || (a->emb && a->emb_ld % 4))
    return fail_arg("conv_igemm: leading dimensions must be multiples of 4 floats");
  const int Cin = a->C1 + a->C2;
  int lower, upper, pad_eff;
  if (a->mode == 0) {
    if (a->w_cout != 128 || a->w_cin != Cin) return fail_arg("conv_igemm fprop: weight must be [128][R*S*(C1+C2)]");
    if (a->stride != 1 && a->stride != 2) return fail_arg("conv_igemm: stride must be 1 or 2");
    pad_eff = a->pad;
    lower = -a->pad;
    upper = a->pad - (a->R - 1);
    const int ho = (a->H + 2 * a->pad - a->R) / a->stride + 1, wo = (a->W + 2 * a->pad - a->S) / a->stride + 1;
    if (ho != a->Ho || wo != a->Wo) return fail_arg("conv_igemm fprop: Ho/Wo do not match the geometry");
  } else {
    if (a->stride != 1) return fail_arg("conv_igemm dgrad: stride must be 1 (zero-stuff dY for strided layers)");
    if (a->C2 != 0 || a->C1 != a->w_cout) return fail_arg("conv_igemm dgrad: x is dY with w_cout channels");
    if (a->n_off < 0 || a->n_off + 128 > a->w_cin || a->n_off % 32) return fail_arg("conv_igemm dgrad: bad n_off");
    if (a->Ho != a->H || a->Wo != a->W) return fail_arg("conv_igemm dgrad: output size must equal dY size");
    if (a->bias || a->emb) return fail_arg("conv_igemm dgrad: bias/emb not applicable");
    pad_eff = a->R - 1 - a->pad;
    lower = -pad_eff;
    upper = -a->pad;
  }
  // stride-1 3x3/pad-1 and 1x1/pad-0 layers (all but the two Downsample convs) go to the strip kernel
  if (strip_enabled() && strip_shape(a)) {
    const int rs = conv_strip_launch(a, stream);
    if (rs != -100) return rs;
  }
  if (a->skip_x) return fail_arg("conv_igemm: the fused skip segment needs a stride-1 3x3/pad-1 or 1x1/pad-0 layer");
  if (a->out_dtype != 0) return fail_arg("conv_igemm: a bf16 out needs a stride-1 3x3/pad-1 or 1x1/pad-0 layer (strip kernel)");
  if (a->out_bf16 && a->mode != 0) return fail_arg("conv_igemm: out_bf16 is an fprop output");
  if (a->op_dtype != 0) return fail_arg("conv_igemm: bf16 operands need a stride-1 3x3/pad-1 or 1x1/pad-0 layer (strip kernel)");
  CUtensorMap tmA, tmA2, tmB;
  int r = get_tmap_im2col(a->x, a->N, a->H, a->W, a->C1, a->x_ld ? a->x_ld : a->C1, lower, upper, a->stride, 32, 128, 4, 0, &tmA);
  if (r) return r;
  if (a->C2 > 0) {
    r = get_tmap_im2col(a->x2, a->N, a->H, a->W, a->C2, a->x2_ld ? a->x2_ld : a->C2, lower, upper, a->stride, 32, 128, 4, 0, &tmA2);
    if (r) return r;
  } else {
    tmA2 = tmA;
  }
  const uint64_t wcols = static_cast<uint64_t>(a->R) * a->S * a->w_cin;
  r = get_tmap_2d(a->w, a->w_cout, wcols, wcols, 32, a->mode == 0 ? 128 : 32, 4, a->mode == 0 ? 0 : 1, &tmB);
  if (r) return r;

  ConvK k;
  k.out = static_cast<float*>(a->out);
  k.out_bf = static_cast<uint16_t*>(a->out_bf16);
  k.bias = a->bias;
  k.emb = a->emb;
  k.residual = a->residual;
  k.M_total = a->N * a->Ho * a->Wo;
  k.HoWo = a->Ho * a->Wo;
  k.Wo = a->Wo;
  k.stride = a->stride;
  k.pad_eff = pad_eff;
  k.R = a->R;
  k.S = a->S;
  k.chunks1 = a->C1 / 32;
  k.chunks = Cin / 32;
  k.mode = a->mode;
  k.w_rowlen_tap = a->w_cin;
  k.n_off = a->n_off;
  k.out_ld = a->out_ld;
  k.emb_ld = a->emb_ld;
  k.res_ld = a->res_ld;
  k.round_out = a->round_out;
  k.sub = 0;

  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_igemm_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kConvSmem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_igemm)");
    attr_set = true;
  }
  const int grid = (k.M_total + 127) / 128;
  conv_igemm_tf32_kernel<<<grid, 192, kConvSmem, static_cast<cudaStream_t>(stream)>>>(tmA, tmA2, tmB, k);
  return after_launch("conv_igemm_tf32_kernel");
}
