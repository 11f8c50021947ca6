// Shared device helpers for the B200 paged-decode kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/b200_paged_attn.h"
#include "../../include/b200_paged_attn_testing.h"

#ifndef B200PA_TESTING
#define B200PA_TESTING 0
#endif

namespace b200pa {

constexpr int kWarp = 32;
constexpr int kPartitionSize = 512;  // _PARTITION_SIZE, paged_attn.py:21 of the reference

// ---- scalar_t traits -------------------------------------------------------
template <int ST> struct Scalar;
template <> struct Scalar<B200PA_F32> {
  using type = float;
  static __device__ __forceinline__ float to_float(float v) { return v; }
  static __device__ __forceinline__ float from_float(float v) { return v; }
};
template <> struct Scalar<B200PA_F16> {
  using type = __half;
  static __device__ __forceinline__ float to_float(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_float(float v) { return __float2half_rn(v); }
};
template <> struct Scalar<B200PA_BF16> {
  using type = __nv_bfloat16;
  static __device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_float(float v) { return __float2bfloat16(v); }
};

// ---- fp8 primitives --------------------------------------------------------
// Quantise: uint8 = cvt_satfinite_rn(float(x) / scale) with a true IEEE divide,
// as the reference's scaled_vec_conversion (quant_utils.cuh:469-502).
template <int KV>
__device__ __forceinline__ uint8_t fp8_quant(float x, float scale) {
  const float q = __fdiv_rn(x, scale);
  uint16_t r;
  if constexpr (KV == B200PA_KV_FP8_E5M2) {
    asm("{cvt.rn.satfinite.e5m2x2.f32 %0, %2, %1;}" : "=h"(r) : "f"(q), "f"(0.0f));
  } else {
    asm("{cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;}" : "=h"(r) : "f"(q), "f"(0.0f));
  }
  return (uint8_t)(r & 0xff);
}

// Two quantised values packed (lo = a, hi = b).
template <int KV>
__device__ __forceinline__ uint16_t fp8_quant2(float a, float b, float scale) {
  const float qa = __fdiv_rn(a, scale), qb = __fdiv_rn(b, scale);
  uint16_t r;
  if constexpr (KV == B200PA_KV_FP8_E5M2) {
    asm("{cvt.rn.satfinite.e5m2x2.f32 %0, %2, %1;}" : "=h"(r) : "f"(qa), "f"(qb));
  } else {
    asm("{cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;}" : "=h"(r) : "f"(qa), "f"(qb));
  }
  return r;
}

// fp8x2 -> f16x2 (exact for e4m3 and e5m2).
template <int KV>
__device__ __forceinline__ uint32_t fp8x2_to_half2(uint16_t v) {
  uint32_t r;
  if constexpr (KV == B200PA_KV_FP8_E5M2) {
    asm("{cvt.rn.f16x2.e5m2x2 %0, %1;}" : "=r"(r) : "h"(v));
  } else {
    asm("{cvt.rn.f16x2.e4m3x2 %0, %1;}" : "=r"(r) : "h"(v));
  }
  return r;
}

// Exact value of one fp8 code as float (fp8 -> half -> float).
template <int KV>
__device__ __forceinline__ float fp8_to_float(uint8_t b) {
  const uint32_t h2 = fp8x2_to_half2<KV>((uint16_t)b);
  return __half2float(__ushort_as_half((unsigned short)(h2 & 0xffff)));
}

// Bit-faithful dequant of the reference: scalar_t(float(half(fp8)) * scale)
// (quant_utils.cuh:306-313, 361-372, 418-428), returned as float.
template <int KV, int ST>
__device__ __forceinline__ float fp8_dequant_faithful(uint8_t b, float scale) {
  const float v = fp8_to_float<KV>(b) * scale;
  return Scalar<ST>::to_float(Scalar<ST>::from_float(v));
}

// ---- misc ------------------------------------------------------------------
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, m));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

// Block-sparse predicate of the reference (attention_kernels.cuh:229-235).
struct BlockSparse {
  int enabled;         // blocksparse_vert_stride > 1
  int local_blocks;
  int vert_stride;
  int block_size;
  int head_sliding_step;
  int tp_rank;
};

// The new token of every sequence for the fused "write the cache + attend" launch
// (b200pa_decode_step): key/value [num_seqs, num_kv_heads, head_size] scalar_t, token b being
// position seq_lens[b] - 1 of sequence b, stored at slot_mapping[b] (< 0: not stored).
struct FusedWrite {
  const void* key;
  const void* value;
  int64_t key_stride, value_stride;  // elements per token
  const int64_t* slot_mapping;
};

}  // namespace b200pa

// ---- host-side error plumbing (defined in c_abi.cu) -------------------------
namespace b200pa {
int set_error(int code, const char* fmt, ...);
int check_launch(const char* what);
void count_launch(int n = 1);
void* testing_trace_buffer();   // nullptr unless b200pa_testing_set_trace armed one
}  // namespace b200pa
