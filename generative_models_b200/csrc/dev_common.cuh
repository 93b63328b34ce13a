// Small device helpers shared by the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// Round-to-nearest (ties away) fp32 -> TF32, kept in an fp32 word (low 13 mantissa bits zero).  tcgen05 kind::tf32
// reads only the top 19 bits of each operand word, i.e. it TRUNCATES: feeding it unrounded data biases every product
// towards zero (~5e-4 relative), which accumulates multiplicatively along the backward chain.  Every tensor that
// feeds a tensor-core operand is therefore stored pre-rounded by its producer.
__device__ __forceinline__ float round_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}
__device__ __forceinline__ float4 round_tf32(float4 v) {
  return make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
}

__device__ __forceinline__ uint16_t bf16_bits(float v) {
  uint16_t r;
  asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(r) : "f"(v));
  return r;
}
// 4 floats -> 4 bf16 (round to nearest even) packed in 8 bytes, element 0 in the low half of .x
__device__ __forceinline__ uint2 pack_bf16x4(float4 v) {
  uint2 r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r.x) : "f"(v.y), "f"(v.x));
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r.y) : "f"(v.w), "f"(v.z));
  return r;
}

}  // namespace b200
