// Library-wide state of libb200ddpm.so: ABI version, error string, launch counter, TMA descriptor encoding.
#include "host_common.h"

#include "../../include/b200ddpm.h"

namespace b200 {

thread_local char g_last_error[512] = {0};
std::atomic<unsigned long long> g_launches{0};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn g_encode_tiled = nullptr;
static EncodeIm2colFn g_encode_im2col = nullptr;
static std::mutex g_mu;
static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_cache;

// The driver entry points are fetched through the runtime so the library does not link libcuda directly.
static int load_driver_fns() {
  if (g_encode_tiled && g_encode_im2col) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
    return fail_arg("cudaGetDriverEntryPoint(cuTensorMapEncodeTiled) failed");
  g_encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
    return fail_arg("cudaGetDriverEntryPoint(cuTensorMapEncodeIm2col) failed");
  g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
  return 0;
}

static CUtensorMapDataType dtype_of(int elem_bytes) {
  return elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
}

int get_tmap_2d(const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems, uint32_t box_cols,
                uint32_t box_rows, int elem_bytes, int swizzle32, CUtensorMap* out) {
  TmapKey k = {{reinterpret_cast<uint64_t>(base), rows, cols, ld_elems,
                (static_cast<uint64_t>(box_cols) << 32) | box_rows, static_cast<uint64_t>(elem_bytes), 2, static_cast<uint64_t>(swizzle32)}};
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_cache.find(k);
  if (it != g_cache.end()) {
    *out = it->second;
    return 0;
  }
  if (int r = load_driver_fns()) return r;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * static_cast<uint64_t>(elem_bytes)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = g_encode_tiled(&m, dtype_of(elem_bytes), 2, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_last_error, sizeof(g_last_error), "cuTensorMapEncodeTiled failed (CUresult %d) rows=%llu cols=%llu",
             static_cast<int>(r), (unsigned long long)rows, (unsigned long long)cols);
    return -2;
  }
  g_cache.emplace(k, m);
  *out = m;
  return 0;
}

// Generic 3-D tiled map {d0 (contiguous), d1, d2} with byte strides s1, s2 (multiples of 16; s2 may be smaller than s1:
// the outer box dimension then walks column blocks of a row-major matrix) and box {b0, b1, b2}.
int get_tmap_3d(const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1, uint64_t s2, uint32_t b0, uint32_t b1,
                uint32_t b2, int elem_bytes, int swizzle32, CUtensorMap* out) {
  TmapKey k = {{reinterpret_cast<uint64_t>(base), (d0 << 32) | d1, (d2 << 40) | s1, s2,
                (static_cast<uint64_t>(b0) << 32) | (static_cast<uint64_t>(b1) << 16) | b2, static_cast<uint64_t>(elem_bytes), 3,
                static_cast<uint64_t>(swizzle32)}};
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_cache.find(k);
  if (it != g_cache.end()) {
    *out = it->second;
    return 0;
  }
  if (int r = load_driver_fns()) return r;
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {s1, s2};
  cuuint32_t box[3] = {b0, b1, b2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = g_encode_tiled(&m, dtype_of(elem_bytes), 3, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_last_error, sizeof(g_last_error), "cuTensorMapEncodeTiled(3d) failed (CUresult %d) dims=%llu,%llu,%llu",
             static_cast<int>(r), (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2);
    return -2;
  }
  g_cache.emplace(k, m);
  *out = m;
  return 0;
}

// A [rows][nblocks * box_cols] matrix as {box_cols, rows, nblocks}: ONE box {box_cols, box_rows, nblocks} lands in shared
// memory as nblocks consecutive [box_rows][128 B] sub-tiles — the layout of nblocks separate 2-D boxes, in one instruction.
int get_tmap_blocks3d(const void* base, uint64_t rows, uint64_t ld_elems, uint32_t box_cols, uint32_t box_rows,
                      uint32_t nblocks, int elem_bytes, int swizzle32, CUtensorMap* out) {
  const uint64_t eb = static_cast<uint64_t>(elem_bytes);
  return get_tmap_3d(base, box_cols, rows, nblocks, ld_elems * eb, box_cols * eb, box_cols, box_rows, nblocks, elem_bytes,
                     swizzle32, out);
}

int get_tmap_im2col(const void* base, int N, int H, int W, int C, int ld, int lower, int upper, int stride, uint32_t box_c,
                    uint32_t box_pix, int elem_bytes, int swizzle32, CUtensorMap* out) {
  TmapKey k = {{reinterpret_cast<uint64_t>(base),
                (static_cast<uint64_t>(N) << 32) | static_cast<uint32_t>(C),
                (static_cast<uint64_t>(H) << 32) | static_cast<uint32_t>(W),
                (static_cast<uint64_t>(static_cast<uint32_t>(lower)) << 32) | static_cast<uint32_t>(upper),
                static_cast<uint64_t>(stride) | (static_cast<uint64_t>(ld) << 32), (static_cast<uint64_t>(box_c) << 32) | box_pix,
                static_cast<uint64_t>(elem_bytes) | (static_cast<uint64_t>(swizzle32) << 8), 4}};
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_cache.find(k);
  if (it != g_cache.end()) {
    *out = it->second;
    return 0;
  }
  if (int r = load_driver_fns()) return r;
  const uint64_t eb = static_cast<uint64_t>(elem_bytes);
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {ld * eb, (uint64_t)W * ld * eb, (uint64_t)H * W * ld * eb};
  int lo[2] = {lower, lower};
  int hi[2] = {upper, upper};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUtensorMap m;
  CUresult r = g_encode_im2col(&m, dtype_of(elem_bytes), 4, const_cast<void*>(base), dims, strides, lo, hi, box_c,
                               box_pix, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_last_error, sizeof(g_last_error),
             "cuTensorMapEncodeIm2col failed (CUresult %d) N=%d H=%d W=%d C=%d lo=%d hi=%d stride=%d", static_cast<int>(r),
             N, H, W, C, lower, upper, stride);
    return -2;
  }
  g_cache.emplace(k, m);
  *out = m;
  return 0;
}

int get_tmap_rows4d(const void* base, int N, int H, int W, int C, int ld, uint32_t box_c, uint32_t box_w, CUtensorMap* out,
                    int elem_bytes, uint32_t box_h) {
  TmapKey k = {{reinterpret_cast<uint64_t>(base), (static_cast<uint64_t>(N) << 32) | static_cast<uint32_t>(C),
                (static_cast<uint64_t>(H) << 32) | static_cast<uint32_t>(W), static_cast<uint64_t>(ld),
                (static_cast<uint64_t>(box_c) << 32) | box_w, static_cast<uint64_t>(elem_bytes), box_h, 44}};
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_cache.find(k);
  if (it != g_cache.end()) {
    *out = it->second;
    return 0;
  }
  if (int r = load_driver_fns()) return r;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  const uint64_t eb4 = static_cast<uint64_t>(elem_bytes);
  cuuint64_t strides[3] = {(uint64_t)ld * eb4, (uint64_t)W * ld * eb4, (uint64_t)H * W * ld * eb4};
  cuuint32_t box[4] = {box_c, box_w, box_h, 1};  // box_h > 1: several image rows of the strip raster per instruction
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMap m;
  CUresult r = g_encode_tiled(&m, dtype_of(elem_bytes), 4, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_last_error, sizeof(g_last_error), "cuTensorMapEncodeTiled(4d rows) failed (CUresult %d) N=%d H=%d W=%d C=%d",
             static_cast<int>(r), N, H, W, C);
    return -2;
  }
  g_cache.emplace(k, m);
  *out = m;
  return 0;
}

}  // namespace b200

extern "C" {

int b200_abi_version(void) { return B200_ABI_VERSION; }

// Validates the device; does NOT change the caller's current device (the process model is one process per GPU: the
// per-process caches of function attributes and the SM count assume every launch of this process targets one device,
// so a second, different device is refused instead of silently mis-launching).
int b200_init(int device) {
  static int bound_device = -1;
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return b200::check_cuda(e, "cudaGetDeviceProperties");
  if (prop.major != 10) return b200::fail_arg("libb200ddpm.so requires a compute-capability 10.x (Blackwell) device");
  if (bound_device >= 0 && bound_device != device)
    return b200::fail_arg("libb200ddpm.so is bound to one device per process (one process per GPU); b200_init was "
                          "called with a second device");
  int cur = -1;
  e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return b200::check_cuda(e, "cudaGetDevice");
  if (cur != device)
    return b200::fail_arg("b200_init: the calling thread's current CUDA device differs from the requested one "
                          "(call torch.cuda.set_device first)");
  bound_device = device;
  return 0;
}

const char* b200_last_error(void) { return b200::g_last_error; }

unsigned long long b200_launch_count(void) { return b200::g_launches.load(); }

}  // extern "C"
