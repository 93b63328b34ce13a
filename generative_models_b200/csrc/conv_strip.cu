// Stride-1 convolutions (3x3 pad 1, 1x1 pad 0; fprop and dgrad) as a "strip" implicit GEMM on tcgen05, TF32.
//
// Why a second kernel: the TMA-im2col kernel (conv_igemm.cu) re-fetches every input pixel once per filter tap and
// every weight tile once per 128 output pixels (32 KB of L2->SM traffic per 128x128x32 MMA block; ncu shows it pinned
// at ~7.1 KB/clk of L2->SM fabric bandwidth with the tensor pipe 40 % busy), and SS-mode TF32 MMAs turned out to be
// bound by shared-memory OPERAND bandwidth (~64 B/clk: a 128x128x8 instruction reads 8 KB).  This kernel attacks both:
//
//   * positions: the batch is viewed as one tall raster with ONE shared zero column per image row and ONE shared zero
//     row per image (pitch Wp = W+1, Hp = H+1).  In that raster a filter tap is a constant row shift
//     (kh-1)*Wp + (kw-1), so a tile of 256 consecutive positions needs one contiguous "strip" of input rows
//     (256 + 2*Wp + 2 positions) per 32-channel chunk, loaded ONCE (TMA, one box per image row, halo = OOB zero fill)
//     and read by all 9 taps as 9 shifted views: the UMMA smem descriptor simply starts 128 B * shift further on
//     (the 128B-swizzle XOR is a function of the absolute smem address, so row-shifted views stay consistent);
//   * orientation: D^T[out channel (128 TMEM lanes)][position (256 columns)] = W * X^T, i.e. the weight tile is the
//     M=128 A operand and the strip view the N=256 B operand: one instruction reads 4 KB + 8 KB for twice the FLOPs
//     of a 128x128x8 one (measured issue rate 480 -> 680 TFLOP/s);
//   * persistent CTAs (one per SM), TMEM accumulators double-buffered (2 x 256 columns) so the epilogue of tile i
//     overlaps the MMAs of tile i+1; epilogue thread = output channel, so a warp stores 128 contiguous bytes per pixel;
//   * two-phase tiling: whole waves of 256-position tiles, then the remaining positions split evenly over the CTAs in
//     one wave of smaller tiles (N of the UMMA = tail_n, a multiple of 16) instead of a last wave at full cost;
//   * optional fused "skip" K-segment (extra 1-tap chunks over a second tensor pair: identity = residual add on the
//     tensor cores, or the ResBlock's 1x1 skip conv).
//   1x1 convolutions use the same kernel in "flat" mode (positions = pixels, one 256-row box per chunk).
//
// CTA = 320 threads: warp 0 TMA producer, warp 1 TMEM owner + MMA issuer, warps 2-9 epilogue.
#include "host_common.h"
#include "ptx.cuh"
#include "dev_common.cuh"

#include "../../include/b200ddpm.h"

#include <stdlib.h>

#include <type_traits>

#include <mutex>

namespace b200 {

constexpr int kSBMax = 10;                // weight-tile ring: as deep as the strips leave room for (6 at W = 64, 8 at 28/14, 9 at 7)
constexpr int kSSMax = 6;                 // strip buffers: 2 for the 3x3 layers; flat (1x1) layers, which are HBM-bound and do one
                                          // tap per 32 KB strip, keep up to 6 in flight (a.nss)
constexpr int kBTile = 16384;             // 128 x 128 B
constexpr int kStripBytes = 58 * 1024;    // >= (256 + 2*Wp + 2 rounded up to whole rows) * 128 B for W <= 64 (7 rows of 65)
constexpr int kTileM = 256;               // positions per CTA tile (the N of one UMMA)
constexpr int kStripTail = 512 + 2 * 512 * 4;  // barriers + the two position tables
constexpr int kStripSmem = 232448;             // the whole 227 KB opt-in maximum; strips first, the ring takes the rest

struct StripK {
  float* out;
  uint16_t* out_bf;  // optional dense bf16 copy [M][128] of out
  int out_is_bf;     // out itself holds bf16
  const float* bias;
  const float* bias2;
  const float* emb;
  const float* residual;
  int N, H, W, Wp, Hp;
  int flat;        // 1: positions are pixels (1x1 conv), strip = 256 rows of a 2-D map
  int P_total;     // number of positions
  int num_tiles;
  int full_tiles;  // tiles [0, full_tiles) hold kTileM positions; the rest ("tail wave") hold tail_n each
  int tail_n;      // positions per tail tile (multiple of 16, <= kTileM)
  int taps;        // 9 or 1
  int chunks1, chunks;
  int skip_chunks1, skip_chunks;  // fused 1x1 "skip" segment over a second pair of tensors (0 = none)
  int mode;        // 0 fprop, 1 dgrad
  int w_rowlen_tap;
  int n_off;
  int out_ld, emb_ld, res_ld;
  int round_out;
  int strip_rows;  // image rows per strip (non-flat)
  int rb;          // image rows per box of the multi-row strip maps (0: single-row boxes only)
  int rb2;         // rows per box of the second, smaller multi-row maps (0: none)
  int wpair;       // ring slots hold TWO filter taps (32 KB), fetched by one 3-D TMA instruction (tmB = {ci, co, tap})
  int sb;          // weight-tile ring depth (slots)
  int strip_alloc; // bytes between the two strip buffers (1 KB multiple)
  int split;       // filter tap before whose weight load the next chunk's strip is requested
  int nss;         // strip buffers in flight (2; more for flat layers)
  int cl;          // launched as 2-CTA clusters: each weight tile is fetched once per pair (half by each CTA, multicast)
  int dbg;         // developer knobs (B200_STRIP_DBG): 1 = skip TMA loads (timing of the MMA issue path only)
};

__device__ __forceinline__ int floor_div(int a, int b) {
  int q = a / b;
  if ((a % b) != 0 && ((a < 0) != (b < 0))) --q;
  return q;
}

// Two-phase tiling: whole waves of kTileM-position tiles, then ONE last wave whose remaining positions are split evenly
// over the CTAs in tail_n-position tiles (a multiple of 16, the UMMA N granularity) — with uniform 256-position tiles
// the last wave is on the critical path at full cost however few CTAs have a tile in it (28x28 at batch 512: 11.36
// waves cost 12; 14x14: 3.04 cost 4).
struct TileGeom {
  int p0, n;
};
__device__ __forceinline__ TileGeom tile_geom(const StripK& a, int t) {
  TileGeom g;
  if (t < a.full_tiles) {
    g.p0 = t * kTileM;
    g.n = kTileM;
  } else {
    g.p0 = a.full_tiles * kTileM + (t - a.full_tiles) * a.tail_n;
    g.n = a.tail_n;
  }
  return g;
}

// BF16 = false: fp32 operands in smem, kind::tf32, 32 channels per 128 B row.
// BF16 = true:  bf16 operands, kind::f16, 64 channels per 128 B row — the byte geometry of every tile, strip, box and
// descriptor step is identical (a K-step is 32 B either way), only the channel count per chunk doubles, so one
// instruction does twice the FLOPs for the same shared-memory operand bytes.
__device__ __forceinline__ int4 lds_int4(uint32_t saddr) {
  int4 q;
  asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(saddr));
  return q;
}
__device__ __forceinline__ int lds_int(uint32_t saddr) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}

// *p = v when m >= 0, as one predicated store
__device__ __forceinline__ void st_if_nonneg(uint64_t p, float v, int m) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ge.s32 q, %2, 0;\n\t@q st.global.f32 [%0], %1;\n\t}" ::"l"(p), "f"(v), "r"(m)
               : "memory");
}
// base + m * ld_bytes as ONE instruction (IMAD.WIDE.U32): the epilogue forms three such addresses per position (residual
// row, fp32 row, bf16 row); written as size_t(max(m, 0)) * ld + co the compiler spent 7 instructions on each load
// address and 14 on each store (ncu source view, bf16 step: 1 106 instructions per 32-position chunk per warp, the two
// epilogue warps of a scheduler issue-bound).  A negative m (halo position) gives a wild address that is only ever
// used by a predicated-off store.
__device__ __forceinline__ uint64_t row_addr(uint64_t base, int m, uint32_t ld_bytes) {
  uint64_t p;
  asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(p) : "r"(m), "r"(ld_bytes), "l"(base));
  return p;
}

template <bool BF16>
__global__ void __launch_bounds__(320, 1)
conv_strip_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                       const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmS,
                       const __grid_constant__ CUtensorMap tmS2, const __grid_constant__ CUtensorMap tmW2,
                       const __grid_constant__ CUtensorMap tmAm, const __grid_constant__ CUtensorMap tmA2m,
                       const __grid_constant__ CUtensorMap tmAs, const __grid_constant__ CUtensorMap tmA2s,
                       const StripK a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* strip0 = smem;
  const uint32_t kSS = static_cast<uint32_t>(a.nss);
  uint8_t* bring = smem + kSS * a.strip_alloc;
  const int slot_bytes = a.wpair ? 2 * kBTile : kBTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(bring + a.sb * slot_bytes);
  uint64_t* s_full = bars;                // [kSSMax]
  uint64_t* s_empty = s_full + kSSMax;    // [kSSMax]
  uint64_t* b_full = s_empty + kSSMax;    // [kSBMax]
  uint64_t* b_empty = b_full + kSBMax;    // [kSBMax]
  uint64_t* acc_full = b_empty + kSBMax;  // [2]
  uint64_t* acc_empty = acc_full + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  int* tab = reinterpret_cast<int*>(bars + 64);  // 512 B past the barriers, 16 B-aligned: [2 acc stages][m[256], n[256]]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmA2);
    tma_prefetch_desc(&tmB);
    if (a.rb) {
      tma_prefetch_desc(&tmAm);
      tma_prefetch_desc(&tmA2m);
    }
    if (a.rb2) {
      tma_prefetch_desc(&tmAs);
      tma_prefetch_desc(&tmA2s);
    }
    if (a.skip_chunks) {
      tma_prefetch_desc(&tmS);
      tma_prefetch_desc(&tmS2);
      tma_prefetch_desc(&tmW2);
    }
    for (int s = 0; s < a.nss; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&s_empty[s], 1);
    }
    for (int s = 0; s < a.sb; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], a.cl ? 2 : 1);  // cluster: a slot is free when BOTH CTAs' MMAs have read it
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 256);  // every epilogue thread arrives
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (a.cl) cluster_sync_all();  // the peer's barriers exist before anything is multicast at them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  constexpr int CK = BF16 ? 64 : 32;  // channels per K-chunk (one 128 B smem row)
  const int CH = a.chunks + a.skip_chunks;  // K-chunks (32 channels each) per tile: main conv, then the skip segment

  if (warp == 0) {
    if (lane == 0) {
      // ---------------------------------------------------------------- TMA producer
      // Order of issue: strip(0); then per chunk: weight taps 0..5, strip(next chunk), weight taps 6..8.  The request
      // has to wait for the MMAs of the chunk before the one in flight to release the buffer, and the weight stream
      // stands still meanwhile: asked for early (before tap 0 / 2) the ring drains and the kernel loses 7-18 %; before
      // tap 6 or 8 the buffer is free by the time the producer (one ring ahead of the MMAs) gets there (28x28:
      // 194.5 us with tap 4, 190.7 with 6 or 8).
      uint32_t bs = 0, bph = 1;  // ring slot and the parity its "empty" barrier is awaited with
      const int tiles_mine = (a.num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
      const int total_chunks = tiles_mine * CH;
      auto load_strip = [&](int q) {  // q = running chunk index of this CTA
        const int t = blockIdx.x + (q / CH) * gridDim.x;
        const int ch = q % CH;
        const uint32_t ss = q % kSS;
        mbar_wait(&s_empty[ss], ((q / kSS) & 1) ^ 1);
        if (a.dbg & 1) {
          mbar_arrive(&s_full[ss]);
          return;
        }
        const TileGeom tg = tile_geom(a, t);
        // image rows this tile's strip spans (a.strip_rows for a full tile)
        const int rows = a.flat ? 0 : 3 + (tg.n + a.Wp) / a.Wp;
        mbar_expect_tx(&s_full[ss], a.flat ? kTileM * 128 : rows * a.Wp * 128);
        uint8_t* sdst = strip0 + ss * a.strip_alloc;
        const CUtensorMap* m;
        const CUtensorMap* mm = nullptr;  // the same tensor in boxes of a.rb image rows
        const CUtensorMap* ms = nullptr;  // ... and of a.rb2 rows
        int c0;
        if (ch < a.chunks) {
          m = (ch < a.chunks1) ? &tmA : &tmA2;
          if (a.rb) mm = (ch < a.chunks1) ? &tmAm : &tmA2m;
          if (a.rb2) ms = (ch < a.chunks1) ? &tmAs : &tmA2s;
          c0 = ((ch < a.chunks1) ? ch : ch - a.chunks1) * CK;
        } else {
          const int sk = ch - a.chunks;
          m = (sk < a.skip_chunks1) ? &tmS : &tmS2;
          c0 = ((sk < a.skip_chunks1) ? sk : sk - a.skip_chunks1) * CK;
        }
        const int p0 = tg.p0;
        if (a.flat) {
          tma_load_2d(sdst, m, &s_full[ss], c0, p0);
        } else {
          const int g0 = floor_div(p0 - a.Wp - 1, a.Wp);  // first padded row of the strip
          // One box per image row, or per a.rb rows where they stay inside one image (its trailing zero row, which is
          // the next image's row -1, may be the last of them): the cost of a strip is per TMA instruction, not per byte
          // (7x7: 32 one-KB boxes per chunk).
          for (int r = 0; r < rows;) {
            const int g = g0 + r;
            const int n = floor_div(g, a.Hp);
            const int hh = g - n * a.Hp - 1;  // -1 = the shared zero row (OOB -> zeros)
            if (mm && r + a.rb <= rows && hh + a.rb - 1 <= a.H) {
              tma_load_4d(sdst + r * a.Wp * 128, mm, &s_full[ss], c0, -1, hh, n);
              r += a.rb;
            } else if (ms && r + a.rb2 <= rows && hh + a.rb2 - 1 <= a.H) {
              tma_load_4d(sdst + r * a.Wp * 128, ms, &s_full[ss], c0, -1, hh, n);
              r += a.rb2;
            } else {
              tma_load_4d(sdst + r * a.Wp * 128, m, &s_full[ss], c0, -1, hh, n);
              r += 1;
            }
          }
        }
      };
      if (a.flat && !a.skip_chunks && !a.cl && !(a.dbg & 1)) {
        // 1x1 layers: one 32 KB strip and one 16 KB weight tile per chunk, four MMAs — the CTA is bound by the bytes it
        // keeps in flight, so strips and weight tiles are two independent streams, each issued as far ahead as its own
        // buffers allow (polling: a blocking wait on one stream would hold back the other)
        int si = 0, wi = 0;
        while (si < total_chunks || wi < total_chunks) {
          bool did = false;
          if (si < total_chunks && mbar_try_wait(&s_empty[si % kSS], ((si / kSS) & 1) ^ 1)) {
            load_strip(si++);
            did = true;
          }
          if (wi < total_chunks && mbar_try_wait(&b_empty[wi % a.sb], ((wi / a.sb) & 1) ^ 1)) {
            const uint32_t slot = wi % a.sb;
            const int ch = wi % CH;
            mbar_expect_tx(&b_full[slot], slot_bytes);
            uint8_t* sB = bring + slot * slot_bytes;
            if (a.mode == 0) {
              tma_load_2d(sB, &tmB, &b_full[slot], ch * CK, 0);
            } else {
#pragma unroll
              for (int jj = 0; jj < 128 / CK; ++jj)
                tma_load_2d(sB + jj * (CK * 128), &tmB, &b_full[slot], a.n_off + jj * CK, ch * CK);
            }
            ++wi;
            did = true;
          }
          if (!did) __nanosleep(20);
        }
      } else {
      if (total_chunks > 0) load_strip(0);
      for (int q = 0; q < total_chunks; ++q) {
        const int ch = q % CH;
        const bool is_skip = ch >= a.chunks;
        const int ntaps = is_skip ? 1 : a.taps;
        const int split = ntaps > 4 ? a.split : 0;
        const int tstep = a.wpair ? 2 : 1;
        for (int tap = 0; tap < ntaps; tap += tstep, bs = (bs + 1 == static_cast<uint32_t>(a.sb)) ? (bph ^= 1, 0u) : bs + 1) {
          if (tap == split && q + 1 < total_chunks) load_strip(q + 1);
          mbar_wait(&b_empty[bs], bph);
          if (a.dbg & 1) {
            mbar_arrive(&b_full[bs]);
            continue;
          }
          mbar_expect_tx(&b_full[bs], slot_bytes);
          uint8_t* sB = bring + bs * slot_bytes;
          if (a.wpair) {
            // taps (tap, tap + 1) of this channel chunk in one instruction; past the last tap the box is zero-filled
            tma_load_3d(sB, &tmB, &b_full[bs], ch * CK, 0, tap);
          } else if (is_skip) {
            const int sk = ch - a.chunks;
            if (a.mode == 0) {
              tma_load_2d(sB, &tmW2, &b_full[bs], sk * CK, 0);
            } else {  // MN-major weight tile: 128 / CK boxes of [CK co rows][CK ci = 128 B]
#pragma unroll
              for (int jj = 0; jj < 128 / CK; ++jj)
                tma_load_2d(sB + jj * (CK * 128), &tmW2, &b_full[bs], jj * CK, sk * CK);
            }
          } else if (a.cl) {
            // this CTA's half (64 output channels) of the tile, delivered to both CTAs of the pair
            const int half = static_cast<int>(cluster_ctarank());
            tma_load_2d_mc(sB + half * (kBTile / 2), &tmB, &b_full[bs], tap * a.w_rowlen_tap + ch * CK, half * 64, 3);
          } else if (a.mode == 0) {
            tma_load_2d(sB, &tmB, &b_full[bs], tap * a.w_rowlen_tap + ch * CK, 0);
          } else {
            const int wtap = a.taps - 1 - tap;  // dgrad walks the filter flipped
#pragma unroll
            for (int jj = 0; jj < 128 / CK; ++jj)
              tma_load_2d(sB + jj * (CK * 128), &tmB, &b_full[bs], wtap * a.w_rowlen_tap + a.n_off + jj * CK, ch * CK);
          }
        }
      }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ---------------------------------------------------------------- MMA issuer
      // D^T[out channel (128 TMEM lanes)][position (256 TMEM columns)] += W[co][k] * X^T[k][position]:
      // A = weight tile (K-major for fprop, MN-major for dgrad), B = the shifted strip view (K-major, N = 256 rows).
      // One N=256 instruction reads 4 KB of A + 8 KB of B for twice the FLOPs of a 128x128x8 one: SS-mode TF32 is
      // bound by smem operand bandwidth (measured: 480 -> 680 TFLOP/s issue rate), hence this orientation.
      const uint32_t idesc_full = make_idesc(BF16 ? 1u : 2u, 128u, 256u, a.mode == 0 ? 0u : 1u, 0u);
      const uint32_t idesc_tail = make_idesc(BF16 ? 1u : 2u, 128u, static_cast<uint32_t>(a.tail_n), a.mode == 0 ? 0u : 1u, 0u);
      uint32_t sc = 0, bs = 0, bph = 0, tc = 0;
      for (int t = blockIdx.x; t < a.num_tiles; t += gridDim.x, ++tc) {
        const uint32_t as = tc & 1;
        mbar_wait(&acc_empty[as], ((tc >> 1) & 1) ^ 1);
        tc_fence_after();
        const int p0 = tile_geom(a, t).p0;
        const uint32_t idesc = t < a.full_tiles ? idesc_full : idesc_tail;
        const int g0 = a.flat ? 0 : floor_div(p0 - a.Wp - 1, a.Wp);
        const int row0 = a.flat ? 0 : p0 - g0 * a.Wp;  // strip row of position p0 (centre tap)
        const uint32_t d_tmem = tmem_base + as * 256;
        for (int ch = 0; ch < CH; ++ch, ++sc) {
          const uint32_t ss = sc % kSS;
          mbar_wait(&s_full[ss], (sc / kSS) & 1);
          tc_fence_after();
          const uint32_t strip_addr = smem_u32(strip0 + ss * a.strip_alloc);
          const bool is_skip = ch >= a.chunks;
          const int ntaps = is_skip ? 1 : a.taps;
          const int tstep = a.wpair ? 2 : 1;
          for (int tap0 = 0; tap0 < ntaps; tap0 += tstep, bs = (bs + 1 == static_cast<uint32_t>(a.sb)) ? (bph ^= 1, 0u) : bs + 1) {
            mbar_wait(&b_full[bs], bph);
            tc_fence_after();
            for (int tt = 0; tt < tstep && tap0 + tt < ntaps; ++tt) {
              const int tap = tap0 + tt;
              const uint32_t w_base = smem_u32(bring + bs * slot_bytes + tt * kBTile);
              int shift = 0;
              if (ntaps == 9) {
                const int kh = tap / 3, kw = tap - kh * 3;
                shift = (kh - 1) * a.Wp + (kw - 1);
              }
              // strip view: start address only 128 B-aligned.  The 128B-swizzle XOR uses absolute smem address bits
              // [7,10) (as TMA's does), so row-shifted views need no descriptor base offset (verified on B200).
              const uint32_t x_base = strip_addr + static_cast<uint32_t>(row0 + shift) * 128u;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                // MN-major A (dgrad): TF32 = 32 B-atom layout, 4 boxes of 32 ci; bf16 = plain SWIZZLE_128B, 2 boxes of
                // 64 ci (LBO = box pitch, SBO = pitch of the 8-row K groups)
                const uint64_t adesc = (a.mode == 0) ? make_smem_desc_sw128(w_base + k * 32, 16, 1024)
                                       : BF16        ? make_smem_desc_sw128(w_base + k * 2048, 8192, 1024, 2)
                                                     : make_smem_desc_sw128(w_base + k * 1024, 4096, 512, 1);
                const uint64_t bdesc = make_smem_desc_sw128(x_base + k * 32, 16, 1024);
                if (BF16) mma_f16_ss(d_tmem, adesc, bdesc, idesc, (ch | tap | k) != 0 ? 1u : 0u);
                else mma_tf32_ss(d_tmem, adesc, bdesc, idesc, (ch | tap | k) != 0 ? 1u : 0u);
              }
            }
            if (a.cl) tc_commit_mc(&b_empty[bs], 3);
            else tc_commit(&b_empty[bs]);
          }
          tc_commit(&s_empty[ss]);
        }
        tc_commit(&acc_full[as]);
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue: TMEM -> regs -> (+bias +emb +res) -> global
    // thread = output channel (TMEM lane), registers = 32 consecutive positions; a warp stores 32 consecutive
    // channels of one pixel = one coalesced 128 B segment per instruction.
    // 8 warps: warp e reads TMEM lane quarter (e & 3) and position columns [128*(e>>2), +128).
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int colhalf = ((ew >> 2) ^ (quarter >> 1)) & 1;  // warps 2..9 -> quarters 2,3,0,1,2,3,0,1; halves 0 then 1
    const int co = quarter * 32 + lane;
    const int et = threadIdx.x - 64;  // 0..255
    const float bias_v = (a.bias ? a.bias[co] : 0.f) + (a.bias2 ? a.bias2[co] : 0.f);
    uint32_t tc = 0;
    for (int t = blockIdx.x; t < a.num_tiles; t += gridDim.x, ++tc) {
      const uint32_t as = tc & 1;
      // position -> (pixel row m, sample n) table of this tile (m = -1: halo position, nothing to store)
      int* tm = tab + as * 512;
      int* tn = tm + 256;
      // the table is read through explicit shared-space loads: from the generic pointer the compiler emits LD.E
      // (generic path, long-scoreboard latency) instead of LDS
      const uint32_t tm_s = smem_u32(tm), tn_s = tm_s + 1024;
      int tile_n;
      {
        const TileGeom tg = tile_geom(a, t);
        tile_n = tg.n;
        const int P = tg.p0 + et;
        int m, n;
        if (et >= tg.n) {  // past the end of a tail tile: the next tile owns the position
          m = -1;
          n = a.N;
        } else if (a.flat) {
          m = P < a.P_total ? P : -1;
          n = P / (a.H * a.W);
        } else {
          const int g = P / a.Wp;
          const int w = P - g * a.Wp - 1;
          n = g / a.Hp;
          const int h = g - n * a.Hp - 1;
          m = ((w >= 0) && (h >= 0) && (n < a.N)) ? (n * a.H + h) * a.W + w : -1;
        }
        tm[et] = m;
        tn[et] = n;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      // Residual rows are fetched one 32-position chunk AHEAD: chunk 0 before the wait for the accumulator (the MMAs of
      // this tile are still running), chunk c+1 before the stores of chunk c — the loads' DRAM latency used to be
      // exposed four times per tile.  The position table is re-read from shared memory instead of living in 32
      // registers, which is what makes room for the second buffer under the 168-register cap.
      const float* res_co = a.residual ? a.residual + co : nullptr;
      const uint64_t res_base = reinterpret_cast<uint64_t>(res_co);
      const uint32_t res_ldb = static_cast<uint32_t>(a.res_ld) * 4u;
      float nxt[32];
      auto fetch_residual = [&](int cb) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int4 q = lds_int4(tm_s + (cb + j) * 4);
          // unconditional (halo positions read row 0): predicates would serialise the loads
          nxt[j] = __ldg(reinterpret_cast<const float*>(row_addr(res_base, max(q.x, 0), res_ldb)));
          nxt[j + 1] = __ldg(reinterpret_cast<const float*>(row_addr(res_base, max(q.y, 0), res_ldb)));
          nxt[j + 2] = __ldg(reinterpret_cast<const float*>(row_addr(res_base, max(q.z, 0), res_ldb)));
          nxt[j + 3] = __ldg(reinterpret_cast<const float*>(row_addr(res_base, max(q.w, 0), res_ldb)));
        }
      };
      const uint64_t out_base = reinterpret_cast<uint64_t>(a.out + co);
      const uint32_t out_ldb = static_cast<uint32_t>(a.out_ld) * 4u;
      if (res_co) fetch_residual(colhalf * 128);
      mbar_wait(&acc_full[as], (tc >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c4 = 0; c4 < 4; ++c4) {
        const int cb = colhalf * 128 + c4 * 32;  // first position column of this chunk
        if (cb >= tile_n) break;                 // tail tile: no positions left (the TMEM columns were not written)
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * 256 + cb, r);
        float add[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) add[j] = bias_v;
        if (res_co) {
#pragma unroll
          for (int j = 0; j < 32; ++j) add[j] += nxt[j];
          if (c4 < 3) fetch_residual(cb + 32);
        }
        if (a.emb) {
          const int n0 = lds_int(tn_s + cb * 4), n1 = lds_int(tn_s + (cb + 31) * 4);
          if (n0 == n1) {
            const float e = (n0 < a.N) ? __ldg(a.emb + static_cast<size_t>(n0) * a.emb_ld + co) : 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) add[j] += e;
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (lds_int(tm_s + (cb + j) * 4) >= 0)
                add[j] += __ldg(a.emb + static_cast<size_t>(lds_int(tn_s + (cb + j) * 4)) * a.emb_ld + co);
          }
        }
        tmem_ld_wait();
        if (!a.out_is_bf) {
          auto store_f32 = [&](auto rnd) {  // rnd: compile-time "round to TF32" (uniform for the launch)
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const int4 q = lds_int4(tm_s + (cb + j) * 4);
              const int mq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                // predicated store (a branch per position costs more than the store)
                const float v = __uint_as_float(r[j + i]) + add[j + i];
                st_if_nonneg(row_addr(out_base, mq[i], out_ldb), decltype(rnd)::value ? round_tf32(v) : v, mq[i]);
              }
            }
          };
          if (a.round_out) store_f32(std::true_type{});
          else store_f32(std::false_type{});
        }
        uint16_t* ob = a.out_is_bf ? reinterpret_cast<uint16_t*>(a.out) : a.out_bf;
        if (ob) {
          // bf16 stores: neighbouring lanes (channels c, c+1) trade one value per position pair so that every lane
          // stores a 4 B word {c, c+1} — even lanes for position j, odd lanes for j+1 — instead of 2 B per lane:
          // half the store instructions, whole 32 B sectors per pixel
          const uint32_t obldb = static_cast<uint32_t>(a.out_is_bf ? a.out_ld : 128) * 2u;
          const uint64_t ob_base = reinterpret_cast<uint64_t>(ob + (co & ~1));
          const int odd = lane & 1;
          // pack this channel's two positions first, trade the packed word, pick the halves with one byte permute
          // (even lane: {own j, neighbour j}; odd lane: {neighbour j+1, own j+1}) — 3 instructions per pair, was 5
          const uint32_t sel = odd ? 0x3276u : 0x5410u;
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const float v0 = __uint_as_float(r[j]) + add[j], v1 = __uint_as_float(r[j + 1]) + add[j + 1];
            uint32_t own;
            asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(own) : "f"(v1), "f"(v0));  // low half = position j
            const uint32_t nbr = __shfl_xor_sync(0xffffffffu, own, 1);
            const uint32_t wrd = __byte_perm(own, nbr, sel);
            const int m = lds_int(tm_s + (cb + j + odd) * 4);
            st_if_nonneg(row_addr(ob_base, m, obldb), __uint_as_float(wrd), m);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[as]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (a.cl) cluster_sync_all();  // no CTA leaves while its peer may still signal its barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace b200

using namespace b200;

namespace {

// Two-phase tiling of P positions over `ctas` persistent CTAs (see tile_geom).
void split_tiles(int P, int ctas, bool tail, int* full_tiles, int* tail_n, int* num_tiles) {
  if (tail) {
    *full_tiles = (P / (kTileM * ctas)) * ctas;
    const int rest = P - *full_tiles * kTileM;
    int tn = (((rest + ctas - 1) / ctas) + 15) & ~15;
    tn = tn < 16 ? 16 : (tn > kTileM ? kTileM : tn);
    *tail_n = tn;
    *num_tiles = *full_tiles + (rest + tn - 1) / tn;
  } else {
    *full_tiles = (P + kTileM - 1) / kTileM;
    *tail_n = kTileM;
    *num_tiles = *full_tiles;
  }
}

bool tail_split_on() {
  static int v = -1;  // developer knob B200_STRIP_TAIL=0: uniform 256-position tiles
  if (v < 0) {
    const char* e = getenv("B200_STRIP_TAIL");
    v = e ? atoi(e) : 1;
  }
  return v != 0;
}

int num_sms_cached() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

// Rows per box of the two multi-row strip maps: a strip costs per TMA instruction (~90 ns each on the producer's
// critical path), so the launch walks its own tiles once per (N, H, W) and keeps the pair (rb, rb2) that covers the
// strips of this geometry with the fewest boxes.  A box may not leave its image (one trailing zero row allowed), which
// is why the best size depends on how H + 1 and the strip height divide.  B200_STRIP_RB=0 turns the boxes off,
// B200_STRIP_RB=r forces a single level of r rows.
void pick_strip_boxes(int N, int H, int W, int* rb_out, int* rb2_out) {
  static int knob = -2;
  if (knob == -2) {
    const char* e = getenv("B200_STRIP_RB");
    knob = e ? atoi(e) : -1;
  }
  const int Wp = W + 1, Hp = H + 1;
  const int cap = 256 / Wp < Hp ? 256 / Wp : Hp;  // TMA box dimensions are <= 256 elements
  *rb_out = *rb2_out = 0;
  if (knob == 0 || cap < 2) return;
  if (knob > 0) {
    *rb_out = knob <= cap ? knob : 0;
    return;
  }
  struct Memo {
    int N, H, W, rb, rb2;
  };
  static Memo memo[64];
  static int memo_n = 0;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < memo_n; ++i)
    if (memo[i].N == N && memo[i].H == H && memo[i].W == W) {
      *rb_out = memo[i].rb;
      *rb2_out = memo[i].rb2;
      return;
    }
  int full_tiles, tail_n, num_tiles;
  split_tiles(N * Hp * Wp, num_sms_cached(), tail_split_on(), &full_tiles, &tail_n, &num_tiles);
  const int step = num_tiles > 1024 ? num_tiles / 1024 : 1;  // a sample of the tiles is enough
  long best = -1;
  for (int r1 = 2; r1 <= cap; ++r1)
    for (int r2 = 0; r2 < r1; r2 = r2 ? r2 + 1 : 2) {
      long boxes = 0;
      for (int t = 0; t < num_tiles; t += step) {
        const int p0 = t < full_tiles ? t * kTileM : full_tiles * kTileM + (t - full_tiles) * tail_n;
        const int n = t < full_tiles ? kTileM : tail_n;
        const int q = p0 - Wp - 1;
        const int g0 = q >= 0 ? q / Wp : -((-q + Wp - 1) / Wp);
        const int rows = 3 + (n + Wp) / Wp;
        for (int r = 0; r < rows; ++boxes) {
          const int g = g0 + r;
          const int hh = ((g % Hp) + Hp) % Hp - 1;
          if (r + r1 <= rows && hh + r1 - 1 <= H) r += r1;
          else if (r2 && r + r2 <= rows && hh + r2 - 1 <= H) r += r2;
          else r += 1;
        }
      }
      if (best < 0 || boxes < best) {
        best = boxes;
        *rb_out = r1;
        *rb2_out = r2;
      }
    }
  if (memo_n < 64) memo[memo_n++] = {N, H, W, *rb_out, *rb2_out};
}

}  // namespace

// Called by b200_conv_igemm for stride-1 3x3/pad-1 and 1x1/pad-0 layers.  Returns -100 when the geometry does not fit
// the strip buffers (caller falls back to the im2col kernel).
// 1 when the strip kernel takes this (stride-1 3x3/pad-1 or 1x1/pad-0) layer, 0 when the im2col kernel does.
int conv_strip_applicable(const b200_conv_args* a) {
  const int flat = (a->R == 1);
  const int Wp = a->W + 1, Hp = a->H + 1;
  const int strip_rows = 3 + (257 + Wp - 1) / Wp;
  if (!flat && strip_rows * Wp * 128 > kStripBytes) return 0;
  if (!flat && Wp > 256) return 0;
  const long long positions = flat ? static_cast<long long>(a->N) * a->H * a->W : static_cast<long long>(a->N) * Hp * Wp;
  const long long num_tiles = (positions + kTileM - 1) / kTileM;
  // Few tiles (small spatial levels, small batches) used to go to the im2col kernel (128-position tiles, 2 CTAs/SM):
  // the measured crossover was ~4 tiles per SM in round 1 and ~2 once the strip rows and filter taps were fetched in
  // multi-row / two-tap TMA boxes.  With the tail wave split evenly over the CTAs it is gone: taking every applicable
  // layer here is faster at every batch measured (train step 4.35 -> 4.01 ms at batch 64, 5.79 -> 5.61 at 128,
  // 10.00 -> 9.84 at 256, 19.23 -> 19.16 at 512; sample(16, y) 1.13 -> 0.94 ms per step).  B200_STRIP_MIN_TILES=n restores
  // a threshold (developer knob; halved for launches with a residual operand).
  static int min_tiles = -1;
  if (min_tiles < 0) {
    const char* e = getenv("B200_STRIP_MIN_TILES");
    min_tiles = e ? atoi(e) : 0;
  }
  const int need = a->residual ? min_tiles / 2 : min_tiles;
  if (!a->skip_x && num_tiles < need && !flat && a->op_dtype == 0) return 0;
  return 1;
}

int conv_strip_launch(const b200_conv_args* a, void* stream) {
  if (!conv_strip_applicable(a)) return -100;
  const int Cin = a->C1 + a->C2;
  const int flat = (a->R == 1);
  const int Wp = a->W + 1, Hp = a->H + 1;
  const int strip_rows = 3 + (257 + Wp - 1) / Wp;
  const bool bf = a->op_dtype == 1;
  const int eb = bf ? 2 : 4;        // operand element bytes
  const uint32_t CK = bf ? 64 : 32;  // channels per 128 B row
  if (a->C1 % CK || a->C2 % CK) return fail_arg("conv_igemm: C1/C2 must be multiples of the K-chunk (32 fp32 / 64 bf16)");

  if (a->out_bf16 && a->mode != 0) return fail_arg("conv_igemm: out_bf16 is an fprop output");

  CUtensorMap tmA, tmA2, tmB, tmAm = {}, tmA2m = {}, tmAs = {}, tmA2s = {};
  int r;
  int rb = 0, rb2 = 0;
  if (!flat) pick_strip_boxes(a->N, a->H, a->W, &rb, &rb2);
  const long long M = static_cast<long long>(a->N) * a->H * a->W;
  if (flat) {
    r = get_tmap_2d(a->x, M, a->C1, a->x_ld ? a->x_ld : a->C1, CK, 256, eb, 0, &tmA);
    if (r) return r;
    if (a->C2 > 0) {
      r = get_tmap_2d(a->x2, M, a->C2, a->x2_ld ? a->x2_ld : a->C2, CK, 256, eb, 0, &tmA2);
      if (r) return r;
    } else {
      tmA2 = tmA;
    }
  } else {
    r = get_tmap_rows4d(a->x, a->N, a->H, a->W, a->C1, a->x_ld ? a->x_ld : a->C1, CK, Wp, &tmA, eb);
    if (r) return r;
    if (a->C2 > 0) {
      r = get_tmap_rows4d(a->x2, a->N, a->H, a->W, a->C2, a->x2_ld ? a->x2_ld : a->C2, CK, Wp, &tmA2, eb);
      if (r) return r;
    } else {
      tmA2 = tmA;
    }
    tmAm = tmAs = tmA;
    tmA2m = tmA2s = tmA2;
    for (int lvl = 0; lvl < 2; ++lvl) {
      const int rows_box = lvl ? rb2 : rb;
      if (!rows_box) continue;
      CUtensorMap* m1 = lvl ? &tmAs : &tmAm;
      CUtensorMap* m2 = lvl ? &tmA2s : &tmA2m;
      r = get_tmap_rows4d(a->x, a->N, a->H, a->W, a->C1, a->x_ld ? a->x_ld : a->C1, CK, Wp, m1, eb, rows_box);
      if (r) return r;
      *m2 = *m1;
      if (a->C2 > 0) {
        r = get_tmap_rows4d(a->x2, a->N, a->H, a->W, a->C2, a->x2_ld ? a->x2_ld : a->C2, CK, Wp, m2, eb, rows_box);
        if (r) return r;
      }
    }
  }
  const uint64_t wcols = static_cast<uint64_t>(a->R) * a->S * a->w_cin;
  // 2-CTA clusters sharing the weight stream (fprop orientation, no skip segment)
  static int cluster_on = -1;  // developer knob B200_STRIP_CLUSTER
  if (cluster_on < 0) {
    const char* e = getenv("B200_STRIP_CLUSTER");
    cluster_on = e ? atoi(e) : 0;
  }
  const int cl = (cluster_on && a->mode == 0 && !a->skip_x) ? 1 : 0;
  // 3x3 fprop orientation: two filter taps per ring slot and per TMA instruction
  static int wpair_on = -1;  // developer knob B200_STRIP_WPAIR=0: one tap per slot
  if (wpair_on < 0) {
    const char* e = getenv("B200_STRIP_WPAIR");
    wpair_on = e ? atoi(e) : 1;
  }
  const int wpair = (wpair_on && a->mode == 0 && !a->skip_x && !cl && a->R * a->S == 9 && a->w_cout == 128) ? 1 : 0;
  if (wpair)
    r = get_tmap_3d(a->w, a->w_cin, a->w_cout, 9, wcols * eb, static_cast<uint64_t>(a->w_cin) * eb, CK, 128, 2, eb, 0, &tmB);
  else
    r = get_tmap_2d(a->w, a->w_cout, wcols, wcols, CK, a->mode == 0 ? (cl ? 64 : 128) : CK, eb, (a->mode == 1 && !bf) ? 1 : 0, &tmB);
  if (r) return r;
  // fused skip segment: a 1x1 convolution over (skip_x | skip_x2) with weights skip_w [128][skip_c1 + skip_c2],
  // accumulated into the same TMEM tile as extra K-chunks (identity weights = a residual add on the tensor cores,
  // exact because every activation is stored pre-rounded to TF32)
  CUtensorMap tmS = tmA, tmS2 = tmA, tmW2 = tmB;
  const int skip_c = a->skip_x ? a->skip_c1 + a->skip_c2 : 0;
  if (skip_c) {
    if (!a->skip_w || a->skip_c1 <= 0 || a->skip_c1 % CK || a->skip_c2 % CK || (a->skip_c2 && !a->skip_x2))
      return fail_arg("conv_igemm: bad skip segment");
    if (a->mode == 1 && skip_c != 128) return fail_arg("conv_igemm dgrad: the skip segment must have 128 channels");
    const int l1 = a->skip_x_ld ? a->skip_x_ld : a->skip_c1, l2 = a->skip_x2_ld ? a->skip_x2_ld : a->skip_c2;
    if (flat) {
      r = get_tmap_2d(a->skip_x, M, a->skip_c1, l1, CK, 256, eb, 0, &tmS);
      if (r) return r;
      if (a->skip_c2) r = get_tmap_2d(a->skip_x2, M, a->skip_c2, l2, CK, 256, eb, 0, &tmS2);
    } else {
      r = get_tmap_rows4d(a->skip_x, a->N, a->H, a->W, a->skip_c1, l1, CK, Wp, &tmS, eb);
      if (r) return r;
      if (a->skip_c2) r = get_tmap_rows4d(a->skip_x2, a->N, a->H, a->W, a->skip_c2, l2, CK, Wp, &tmS2, eb);
    }
    if (r) return r;
    if (!a->skip_c2) tmS2 = tmS;
    r = get_tmap_2d(a->skip_w, 128, skip_c, skip_c, CK, a->mode == 0 ? 128 : CK, eb, (a->mode == 1 && !bf) ? 1 : 0, &tmW2);
    if (r) return r;
  }

  StripK k;
  k.out = static_cast<float*>(a->out);
  k.out_is_bf = a->out_dtype == 1;
  k.out_bf = static_cast<uint16_t*>(a->out_bf16);
  k.bias = a->bias;
  k.bias2 = a->bias2;
  k.emb = a->emb;
  k.residual = a->residual;
  k.N = a->N; k.H = a->H; k.W = a->W; k.Wp = Wp; k.Hp = Hp;
  k.flat = flat;
  k.P_total = flat ? static_cast<int>(M) : a->N * Hp * Wp;
  k.taps = a->R * a->S;
  k.chunks1 = a->C1 / CK;
  k.chunks = Cin / CK;
  k.skip_chunks1 = skip_c ? a->skip_c1 / CK : 0;
  k.skip_chunks = skip_c / CK;
  k.mode = a->mode;
  k.w_rowlen_tap = a->w_cin;
  k.n_off = a->n_off;
  k.out_ld = a->out_ld; k.emb_ld = a->emb_ld; k.res_ld = a->res_ld;
  k.round_out = a->round_out;
  k.strip_rows = strip_rows;
  k.rb = rb;
  {
    static int split = -1;  // developer knob B200_STRIP_SPLIT (even: the two-tap slots step by 2)
    if (split < 0) {
      const char* e = getenv("B200_STRIP_SPLIT");
      split = e ? atoi(e) : 6;
    }
    k.split = split;
  }
  k.rb2 = rb2;
  k.strip_alloc = ((flat ? kTileM * 128 : strip_rows * Wp * 128) + 1023) & ~1023;
  {
    static int sb_cap = -1;  // developer knob B200_STRIP_RING
    if (sb_cap < 0) {
      const char* e = getenv("B200_STRIP_RING");
      sb_cap = e ? atoi(e) : kSBMax;
    }
    k.wpair = wpair;
    // flat (1x1) layers move 32 KB of activations per 16 KB weight tile and 4 MMAs: they are bound by how many bytes the
    // CTA keeps in flight, so they get more strip buffers (measured with 2: 135 us for the 28x28 256->128 skip conv,
    // 4.6 TB/s, against cuDNN's 99 us) and a shorter weight ring
    static int nss_flat = -1;  // developer knob B200_STRIP_NSS_FLAT
    if (nss_flat < 0) {
      const char* e = getenv("B200_STRIP_NSS_FLAT");
      nss_flat = e ? atoi(e) : 5;
      if (nss_flat < 2) nss_flat = 2;
      if (nss_flat > kSSMax) nss_flat = kSSMax;
    }
    k.nss = flat ? nss_flat : 2;
    int sb = (kStripSmem - 1024 - kStripTail - k.nss * k.strip_alloc) / (wpair ? 2 * kBTile : kBTile);
    sb = sb > kSBMax ? kSBMax : sb;
    k.sb = sb > sb_cap ? sb_cap : sb;
    if (k.sb < 2) return fail_arg("conv_strip: no room for the weight ring");
  }

  static int dbg = -1;
  if (dbg < 0) {
    const char* e = getenv("B200_STRIP_DBG");
    dbg = e ? atoi(e) : 0;
  }
  k.dbg = dbg;
  static bool attr_set = false;
  static int num_sms = 148;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_strip_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStripSmem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_strip)");
    e = cudaFuncSetAttribute(conv_strip_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStripSmem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(conv_strip bf16)");
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    attr_set = true;
  }
  // whole waves of 256-position tiles, then the remainder split evenly over the CTAs (see tile_geom)
  split_tiles(k.P_total, num_sms, tail_split_on(), &k.full_tiles, &k.tail_n, &k.num_tiles);
  k.cl = cl && (num_sms % 2 == 0);
  if (k.cl && (k.num_tiles & 1)) ++k.num_tiles;  // both CTAs of a pair walk the same number of weight tiles: the extra
                                                 // tile lies past the last position (zero-filled loads, nothing stored)
  const int grid = k.num_tiles < num_sms ? k.num_tiles : num_sms;
  if (k.cl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(320);
    cfg.dynamicSmemBytes = kStripSmem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t e = bf ? cudaLaunchKernelEx(&cfg, conv_strip_kernel<true>, tmA, tmA2, tmB, tmS, tmS2, tmW2, tmAm, tmA2m, tmAs, tmA2s, k)
                       : cudaLaunchKernelEx(&cfg, conv_strip_kernel<false>, tmA, tmA2, tmB, tmS, tmS2, tmW2, tmAm, tmA2m, tmAs, tmA2s, k);
    if (e != cudaSuccess) return check_cuda(e, "cudaLaunchKernelEx(conv_strip, cluster 2)");
    return after_launch(bf ? "conv_strip_kernel<bf16>" : "conv_strip_kernel<tf32>");
  }
  if (bf) {
    conv_strip_kernel<true><<<grid, 320, kStripSmem, static_cast<cudaStream_t>(stream)>>>(tmA, tmA2, tmB, tmS, tmS2, tmW2, tmAm, tmA2m, tmAs, tmA2s, k);
    return after_launch("conv_strip_kernel<bf16>");
  }
  conv_strip_kernel<false><<<grid, 320, kStripSmem, static_cast<cudaStream_t>(stream)>>>(tmA, tmA2, tmB, tmS, tmS2, tmW2, tmAm, tmA2m, tmAs, tmA2s, k);
  return after_launch("conv_strip_kernel<tf32>");
}
