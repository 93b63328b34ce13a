// Host-side helpers shared by the C-ABI translation units: error reporting, launch counting, TMA descriptor cache.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace b200 {

extern thread_local char g_last_error[512];
extern std::atomic<unsigned long long> g_launches;

inline int fail_arg(const char* msg) {
  snprintf(g_last_error, sizeof(g_last_error), "%s", msg);
  return -1;
}
inline int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what, cudaGetErrorString(e));
  return static_cast<int>(e);
}
// Call after every kernel launch: counts it and reports launch-time errors (no sync).
inline int after_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return check_cuda(cudaGetLastError(), what);
}

// ---- TMA descriptor cache. Buffers in a plan are static, so each (pointer, geometry) is encoded once.
struct TmapKey {
  uint64_t v[8];
  bool operator==(const TmapKey& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < 8; ++i) {
      h ^= k.v[i];
      h *= 1099511628211ull;
    }
    return static_cast<size_t>(h);
  }
};

// swizzle32 != 0 selects CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B (for MN-major TF32 operands), else SWIZZLE_128B.
// fp32 2-D row-major [rows][cols] (cols contiguous), box = box_cols x box_rows, 128B swizzle.
int get_tmap_2d(const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems, uint32_t box_cols,
                uint32_t box_rows, int elem_bytes, int swizzle32, CUtensorMap* out);
// NHWC [N][H][W][C] im2col map: box = box_c channels x box_pix pixels; corners from the conv geometry.
int get_tmap_im2col(const void* base, int N, int H, int W, int C, int ld, int lower, int upper, int stride, uint32_t box_c,
                    uint32_t box_pix, int elem_bytes, int swizzle32, CUtensorMap* out);

// NHWC [N][H][W][C] plain tiled 4-D map {C, W, H, N}: box = box_c channels x box_w pixels of ONE image row
// (out-of-range w / h / n coordinates are zero-filled: the convolution halo).
int get_tmap_3d(const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1, uint64_t s2, uint32_t b0, uint32_t b1,
                uint32_t b2, int elem_bytes, int swizzle32, CUtensorMap* out);
int get_tmap_blocks3d(const void* base, uint64_t rows, uint64_t ld_elems, uint32_t box_cols, uint32_t box_rows,
                      uint32_t nblocks, int elem_bytes, int swizzle32, CUtensorMap* out);
int get_tmap_rows4d(const void* base, int N, int H, int W, int C, int ld, uint32_t box_c, uint32_t box_w, CUtensorMap* out,
                    int elem_bytes = 4, uint32_t box_h = 1);

}  // namespace b200
